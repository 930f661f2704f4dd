#!/bin/bash
# `ncu --set full` capture of one kernel by name regex.  usage: tools/ncu_kernel.sh <batch> <mode> <regex> <skip> <tag>
set -u
B=$1; MODE=$2; RX=$3; SKIP=$4; TAG=$5
CMD="python bench.py --steps 1 --warmup 3 --batch $B --mode $MODE --sweep= --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RX -s $SKIP -c 1 -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
ls -la gpurun_out/${TAG}.ncu-rep
