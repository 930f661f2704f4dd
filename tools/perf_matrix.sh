#!/bin/bash
# diagnostic sweep of the tcgen05 scoring kernel (probe perf mode)
P=tools/umma_probe
for NQ in 32 128 256 1024 2048; do
  timeout 60 $P perf 1000000 $NQ 768 0 0 0 | tail -1
done
for NQ in 32 128 256 1024; do
  timeout 60 $P perf 1000000 $NQ 768 1 0 0 | tail -1
done
timeout 60 $P perf 1000000 1024 768 1 7 0 | tail -1
