#!/bin/bash
# BASELINE.json configs 3-5 on N GPUs of one box.  usage: tools/run_configs.sh <N> <tag>
N=$1; TAG=$2
run() {  # name, extra args
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --sweep= --extras= --scale-configs= "$@" > gpurun_out/${TAG}_${name}_n$N.log 2>&1
  tail -1 gpurun_out/${TAG}_${name}_n$N.log | python -c "
import json,sys
try:
    l=json.loads(sys.stdin.read())
    print('$name N=$N', round(l['value']), 'q/s  e2e', round(l['e2e']['value']), ' ms/step', round(l['ms_per_step'],3), ' main kernel ms', round(l['roofline']['kernel_ms'],3), l['roofline']['bound'], round(l['roofline']['frac'],3), l['engine'], 'kprime', l['kprime'])
except Exception as e:
    print('$name N=$N FAILED', e)
"
}
case "$3" in
  c5) run c5 --rows 100000000 --dim 768 --batch 4096 --k 10 --mode bf16 ;;
  c3) run c3 --rows 10000000 --dim 1024 --batch 256 --k 100 --mode exact ;;
  c4) run c4 --rows 5000000 --dim 4096 --batch 256 --k 4 --mode exact ;;
  c4b) run c4b --rows 5000000 --dim 4096 --batch 256 --k 4 --mode bf16 ;;
esac
