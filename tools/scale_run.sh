#!/bin/bash
# Scaling run on one box: ours and the DDP+NCCL stand-in at N = 1,2,4,8 (whatever the box has), logs into gpurun_out/.
STEPS=${STEPS:-2000}; WARM=${WARM:-50}; EXTRA=${EXTRA:-}
NG=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
for n in 1 2 4 8; do
  [ $n -gt $NG ] && break
  for impl in ours standin; do
    port=$((29600 + n + (RANDOM % 200)))
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --gpus 1 --steps $STEPS --warmup $WARM --impl $impl $EXTRA > gpurun_out/scale_${impl}_$n.log 2>&1
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
        bench.py --gpus $n --steps $STEPS --warmup $WARM --impl $impl $EXTRA > gpurun_out/scale_${impl}_$n.log 2>&1
    fi
    echo "$impl N=$n rc=$? $(grep '^{' gpurun_out/scale_${impl}_$n.log | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read() or "{}"); print(d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"))' 2>/dev/null)"
  done
done
