#!/bin/bash
# Same-box A/B of libb200dp.so builds: put the variants at gpurun_ab/lib<NAME>.so (built from different commits with
# tools/build.py in a scratch checkout), then `gpurun -- 'bash tools/ab_libs.sh [bench.py args]'`.  Boxes differ by a few
# percent from call to call, so a step-level comparison is only meaningful inside one call.
cp b200dp/lib/libb200dp.so /tmp/lib_orig.so
for round in 1 2; do
  for f in gpurun_ab/lib*.so; do
    v=$(basename "$f" .so)
    cp "$f" b200dp/lib/libb200dp.so
    echo -n "$v round$round "
    timeout 100 python bench.py --steps 1500 --warmup 30 --skip-e2e "$@" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
  done
done
cp /tmp/lib_orig.so b200dp/lib/libb200dp.so
