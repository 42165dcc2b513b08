#!/bin/bash
# Quick scaling check on one box: ours at N=1,2,4,8 (device-timed only), the step timeline at the largest N, and the
# fused multi-rank test at that N.
STEPS=${STEPS:-500}; mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -gt $NG ] && break
  if [ $n -eq 1 ]; then timeout 200 python bench.py --steps $STEPS --warmup 20 --skip-e2e $EXTRA > gpurun_out/sq_$n.log 2>&1
  else timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps $STEPS --warmup 20 --skip-e2e $EXTRA > gpurun_out/sq_$n.log 2>&1; fi
  echo "N=$n rc=$? $(grep '^{' gpurun_out/sq_$n.log | python -c 'import sys,json; d=json.loads(sys.stdin.read() or "{}"); print(d.get("value"), d.get("ms_per_step"))' 2>/dev/null)"
done
if [ $NG -ge 2 ]; then
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29711 tools/step_trace.py --out gpurun_out/trace_n$NG > gpurun_out/trace_n$NG.log 2>&1; echo trace rc=$?
fi
