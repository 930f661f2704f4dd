#!/bin/bash
# GPU tests, then finalize-kernel timings with the warp-per-query finalize off / on (B2R_WARP_FIN=0/1).
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r2m_tests.log
for WF in 0 1; do
  echo "WARP_FIN=$WF"
  B2R_WARP_FIN=$WF bash tools/ncu_times.sh r2m_n8shard_wf$WF --rows 125000 --batch 1024 | grep -E "finalize_" | tail -1
  B2R_WARP_FIN=$WF bash tools/ncu_times.sh r2m_b1024_wf$WF --batch 1024 | grep -E "finalize_" | tail -1
  B2R_WARP_FIN=$WF bash tools/ncu_times.sh r2m_b32_wf$WF --batch 32 | grep -E "finalize_" | tail -1
  B2R_WARP_FIN=$WF bash tools/ncu_times.sh r2m_bf16_wf$WF --batch 1024 --mode bf16 | grep -E "finalize_" | tail -1
done
