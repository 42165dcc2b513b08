#!/bin/bash
# On the GPU box: isolated kernel times of the step, then one `ncu --set full` capture of the kernel named by $1 (regex)
# selected with kernel_bench's --only $2.  The ncu command is the plain command that just exited 0.
mkdir -p gpurun_out
KREGEX=${1:-gemm_tc_kernel}; ONLY=${2:-dense_wgrad+adam}; TAG=${3:-wgrad_adam}
timeout 200 python tools/kernel_bench.py --out gpurun_out/kb.json > gpurun_out/kb.log 2>&1; echo kb rc=$?; tail -22 gpurun_out/kb.log
timeout 200 python tools/kernel_bench.py --only "$ONLY" --reps 2 --out gpurun_out/kb_only.json > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 6 -c 2 -f -o gpurun_out/prof_$TAG \
  python tools/kernel_bench.py --only "$ONLY" --reps 2 --out gpurun_out/kb_only.json > gpurun_out/ncu_$TAG.log 2>&1; echo ncu rc=$?; tail -3 gpurun_out/ncu_$TAG.log
