#!/bin/bash
# GPU tests, then the batch sweep with programmatic dependent launch off and on (B2R_PDL=0/1).
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_tests.log 2>&1; echo tests rc=$?; tail -4 gpurun_out/r2h_tests.log
for PDL in 0 1; do
  B2R_PDL=$PDL timeout 300 python bench.py --no-cpu --no-peaks --extras= --scale-configs= --sweep 1,8,32,128,1024 --steps 50 > gpurun_out/r2h_bench_pdl$PDL.log 2>gpurun_out/r2h_bench.err; echo bench rc=$?
  python - <<PY
import json
d=json.loads([x for x in open("gpurun_out/r2h_bench_pdl$PDL.log") if x.startswith("{")][-1])
print("PDL=$PDL", round(d["value"]), d["ms_per_step"], round(d["e2e"]["value"]), d["roofline"]["kernel_ms"])
for s in d["sweep"]: print(s["batch"], round(s["qps"]), round(s["score_ms"],4))
PY
done
