#!/bin/bash
# like ncu_times.sh with free bench arguments.  usage: tools/ncu_times2.sh <tag> <bench args...>
TAG=$1; shift
CMD="python bench.py --steps 2 --warmup 3 --sweep= --no-cpu $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'score_umma|score_fma|finalize|seed_select|prep_queries|merge_packed' -c 60 \
    --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>10 and r[0].isdigit()]
for r in rows[-7:]:
    print(r[4][:60], r[-1], r[-2])
PY
