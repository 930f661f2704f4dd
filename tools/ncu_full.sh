#!/bin/bash
# One `ncu --set full` capture of the main tcgen05 scoring launch (the largest-grid launch of the step).
# usage: tools/ncu_full.sh <batch> <mode> <tag>
set -u
B=$1; MODE=$2; TAG=$3
CMD="python bench.py --steps 1 --warmup 3 --batch $B --mode $MODE --sweep= --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_umma -s 7 -c 1 -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncufull.log 2>&1
ls -la gpurun_out/${TAG}_full.ncu-rep
