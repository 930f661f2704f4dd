#!/bin/bash
# One `ncu --set full` capture of a main tcgen05 scoring launch, after the same command has exited 0 without ncu.
# A search launches score_umma_kernel per 2048-query chunk: sampling, main, and the range launches that return at once
# (two in exact mode, one in bf16 mode), chunk-major per kind.  <skip> picks a main launch: exact 4*s+1 (batch <=
# 2048) or 8*s+2 (batch 4096); bf16 3*s+1 or 6*s+2.
# usage: tools/ncu_full.sh <tag> <skip> <bench.py args...>
set -u
TAG=$1; SKIP=$2; shift 2
CMD="python bench.py --steps 1 --warmup 3 --sweep= --extras= --scale-configs= --no-cpu --no-peaks $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_umma -s $SKIP -c 1 -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncufull.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full.csv 2>/dev/null
python - <<PY
import csv
r=list(csv.reader(open("gpurun_out/${TAG}_full.csv"))); h,u,v=r[0],r[1],r[2]
want=["Kernel Name","Grid Size","gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","dram__bytes_read.sum.per_second",
      "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active","sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active",
      "smsp__cycles_active.avg","lts__t_sector_hit_rate.pct","sm__cycles_elapsed.avg.per_second"]
print("== ${TAG}")
for k in want:
    if k in h: print("  %-70s %s %s" % (k, v[h.index(k)][:60], u[h.index(k)]))
PY

