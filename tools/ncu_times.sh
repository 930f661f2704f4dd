#!/bin/bash
# Per-launch device times of our kernels for one bench configuration (cold-cache, serialised: compare SHARES).
# usage: tools/ncu_times.sh <batch> <mode> <tag>
set -u
B=$1; MODE=$2; TAG=$3
CMD="python bench.py --steps 2 --warmup 3 --batch $B --mode $MODE --sweep= --extras= --scale-configs= --no-cpu --no-peaks"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'score_umma|score_fma|score_exact|finalize|seed_thr|prep_queries|merge_packed' -c 60 \
    --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>10 and r[0].isdigit()]
for r in rows[-20:]:
    print(r[4][:60], r[-1], r[-2])
PY
