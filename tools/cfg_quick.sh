#!/bin/bash
# BASELINE.json configs on whatever the box has: ours vs the DDP+NCCL stand-in (eager and whole-step CUDA graph).
# usage: cfg_quick.sh "<N list>" "<tags>" ; logs gpurun_out/cfg_<impl>_<tag>_<N>.log
NS=${1:-"1 2"}; TAGS=${2:-"b128 b1024 w6"}; STEPS=${STEPS:-300}; mkdir -p gpurun_out
run() { # impl tag n extra...
  impl=$1; tag=$2; n=$3; shift 3
  if [ $n -eq 1 ]; then timeout 300 python bench.py --steps $STEPS --warmup 20 --impl $impl "$@" > gpurun_out/cfg_${impl}_${tag}_$n.log 2>&1
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + RANDOM % 100)) bench.py --gpus $n --steps $STEPS --warmup 20 --impl $impl "$@" > gpurun_out/cfg_${impl}_${tag}_$n.log 2>&1; fi
  echo "$impl $tag N=$n rc=$? $(grep '^{' gpurun_out/cfg_${impl}_${tag}_$n.log | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read() or "{}"); print(d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"))' 2>/dev/null)"
}
for n in $NS; do
  for t in $TAGS; do
    case $t in
      b128) X="";; b1024) X="--batch 1024";; w6) X="--width 6";; w16) X="--width 16";; b100) X="--batch 100";; k128) X="--arch keras128";;
    esac
    run ours $t $n $X
    run standin ${t}graph $n $X --graph
    [ $t = b128 ] && run standin ${t}eager $n $X
  done
done
