#!/bin/bash
# Parity tests, DRAM traffic of the main launch (ncu) and the large-batch sweep after a change to the pair kernel.
timeout 500 python -m pytest tests -m gpu -x -q -k "config or parity or golden or random or multi_round or unproven" > gpurun_out/r2r_tests.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/r2r_tests.log
bash tools/ncu_dram.sh r2r_ex1024 --batch 1024 | tail -5
bash tools/ncu_dram.sh r2r_ex2048 --batch 2048 | tail -5
timeout 200 python bench.py --no-cpu --extras= --scale-configs= --sweep 256,512,1024,2048,4096 > gpurun_out/r2r_bench.log 2>gpurun_out/r2r_bench.err
python - <<PY
import json
d=json.loads([x for x in open("gpurun_out/r2r_bench.log") if x.startswith("{")][-1])
print(round(d["value"]), d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], d["roofline"]["tf32_peak_measured"]["tf32_tflops"])
for s in d["sweep"]: print(s["batch"], round(s["qps"]), round(s["score_ms"],4))
PY
