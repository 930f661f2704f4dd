#!/bin/bash
# Per-launch device times of our kernels for one bench configuration (cold-cache, serialised: compare SHARES).
# usage: tools/ncu_times.sh <tag> <bench.py args...>
set -u
TAG=$1; shift
CMD="python bench.py --steps 2 --warmup 3 --sweep= --extras= --scale-configs= --no-cpu --no-peaks $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'score_umma|score_fma|score_exact|finalize|seed_thr|tighten|prep_queries|merge_packed' -c 80 \
    --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>10 and r[0].isdigit()]
n=len(rows)//5 if len(rows)>=5 else len(rows)
print("== ${TAG}: last search of", len(rows), "launches")
for r in rows[-n:]:
    print("  %-58s %9s ns" % (r[4][:58], r[-1]))
PY
