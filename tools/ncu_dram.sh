#!/bin/bash
# DRAM bytes and duration of every scoring launch of one bench configuration (quick ncu pass, two metrics).
# usage: tools/ncu_dram.sh <tag> <bench.py args...>     (environment such as B2R_PAIR is passed through)
set -u
TAG=$1; shift
CMD="python bench.py --steps 1 --warmup 3 --sweep= --extras= --scale-configs= --no-cpu --no-peaks $*"
ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:score_umma -c 24 --csv --log-file gpurun_out/${TAG}_dram.csv $CMD > gpurun_out/${TAG}_dram.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_dram.csv")) if len(r)>10 and r[0].isdigit()]
by={}
for r in rows: by.setdefault(r[0],{"k":r[4][:44]})[r[-3]]=(r[-1],r[-2])
print("== ${TAG}")
for i,d in list(by.items())[-8:]:
    print("  ", d["k"], {k:v for k,v in d.items() if k!="k"})
PY
