#!/bin/bash
# Run on the GPU box (under gpurun): each GPU test file in its own process (a trapping kernel poisons only
# that process), logs into gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
status=0
for f in "$@"; do
  name=$(basename "$f" .py)
  timeout 600 python -m pytest "$f" -m gpu -q --timeout 300 > "gpurun_out/$name.log" 2>&1
  rc=$?
  echo "$name rc=$rc: $(tail -1 gpurun_out/$name.log)"
  [ $rc -ne 0 ] && status=1
done
exit $status
