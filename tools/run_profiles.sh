#!/bin/bash
# One 1-GPU call: isolated kernel times, the launch list and a full ncu capture of one eager training step, and a full ncu
# capture of the collective kernels with eight ranks emulated on this GPU.  Each ncu line repeats a command that just exited 0.
mkdir -p gpurun_out
timeout 200 python tools/kernel_bench.py --out gpurun_out/kernel_bench.json > gpurun_out/kb.log 2>&1; echo kb rc=$?
STEP="python bench.py --steps 2 --warmup 3 --no-graph --skip-e2e"
timeout 200 $STEP > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:_kernel -s 39 -c 26 --csv --log-file gpurun_out/launches.csv $STEP > gpurun_out/ncu_launch.log 2>&1; echo launches rc=$?
timeout 200 $STEP > gpurun_out/plain_step.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:_kernel -s 39 -c 13 -f -o gpurun_out/prof_step $STEP > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
timeout 200 python tools/run_comm_emulated_once.py > gpurun_out/plain_comm.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"allgather_acts|fused_allreduce|gemm_tc" -s 10 -c 10 -f -o gpurun_out/prof_comm python tools/run_comm_emulated_once.py > gpurun_out/ncu_comm.log 2>&1; echo comm rc=$?
tail -2 gpurun_out/plain_comm.log
