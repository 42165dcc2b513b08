#!/usr/bin/env python
"""Headline benchmark: MNIST CNN training throughput (images/s) on N B200s, device-timed, max over ranks.

    python bench.py --gpus 1 --steps 2000 --warmup 50
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
        bench.py --gpus 8 --steps 2000 --warmup 50

Model/config: the reference's CNN as coded (horovod/tensorflow_mnist.py:38-73), bf16 compute with fp32 master
weights + Adam, 128 images per rank (BASELINE.json config 2), LR scaled by world size, synthetic 28x28 inputs,
random-init weights.  Weak scaling: the per-GPU batch is fixed.

Arms:  --impl ours (default) | reference (the unmodified upstream from baseline/_ref, if installable) |
       standin (PyTorch DDP + NCCL + cuDNN/cuBLAS, bf16 autocast: the stand-in documented in BASELINE.md)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "standin"])
    ap.add_argument("--batch", type=int, default=128, help="images per rank")
    ap.add_argument("--width", type=float, default=1.0)
    ap.add_argument("--arch", default="mnist_cnn", choices=["mnist_cnn", "keras128"],
                    help="mnist_cnn: the network as coded in the reference scripts (headline); keras128: the Keras-example shape")
    ap.add_argument("--algo", default="auto", choices=["auto", "oneshot", "twoshot", "nvls"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-tc", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--adasum", action="store_true")
    ap.add_argument("--graph", action="store_true", help="--impl standin: capture the whole DDP step (NCCL all-reduce included) in a CUDA graph")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the benchmark runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.marks = {}
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self, name):
        self.marks[name] = time.time()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        t0, t1 = self.marks.get("timed_start", 0), self.marks.get("timed_end", float("inf"))
        sel = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.06]
        scope = "timed"
        if not sel:
            w0 = self.marks.get("warm_start", 0)
            sel = [r for (t, r) in self.rows if w0 <= t <= t1 + 0.2]
            scope = "warmup+timed"
        sm, smax, reasons = [], None, set()
        for r in sel:
            f = [c.strip() for c in r.split(",")]
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except (ValueError, IndexError):
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm), "scope": scope}


def reference_arm(args):
    """The unmodified upstream needs TensorFlow + Horovod + Open MPI; none is in this image and
    /root/reference carries no installable package (no setup.py / pyproject.toml)."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    why = None
    try:
        sys.path.insert(0, ref)
        import tensorflow  # noqa: F401
        import horovod.tensorflow  # noqa: F401
    except Exception as e:  # noqa: BLE001
        why = f"reference needs tensorflow+horovod ({type(e).__name__}: {e}); pip install of /root/reference fails: no setup.py/pyproject.toml"
    if why is None:
        why = "reference scripts import but download MNIST at start; no network on this box"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why[:300]}))
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)
    import numpy as np
    import torch
    import torch.distributed as dist

    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world_size != args.gpus and not (world_size == 1 and args.gpus == 1):
        if rank == 0:
            print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world_size}; launch with torch.distributed.run", file=sys.stderr)
        return 2
    if args.impl == "standin":
        sys.path.insert(0, os.path.join(ROOT, "bench"))
        import ddp_standin
        return ddp_standin.run(args)

    import b200dp as hvd
    from b200dp.ops import kernels as K
    os.environ.setdefault("B200DP_LOG_LEVEL", "WARNING")
    hvd.init(device="cuda")
    dev = hvd.device()
    B = args.batch
    spec = hvd.get_spec(args.arch, args.width)
    opt = hvd.DistributedOptimizer(hvd.Adam.keras(lr=0.001 * hvd.lr_scaler(args.adasum)),
                                   op=hvd.Adasum if args.adasum else hvd.Average, algo=args.algo)
    model = hvd.Model(spec, engine="b200", batch_size=B, use_graph=not args.no_graph, use_tc=not args.no_tc)
    model.compile(optimizer=opt, loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    eng = model.engine
    eng.broadcast_state(0)

    # synthetic input pool larger than L2 (126 MB): every timed step reads a batch that is not L2-resident
    pool_n = max(8, int(160e6 // (B * 784 * 4)) + 1)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    # learnable synthetic digits (class-conditional blobs + noise + shifts, b200dp.data.mnist.synthetic_mnist): the loss has
    # to FALL during the run, so final_loss says something about the step being timed; a different draw per rank
    from b200dp.data import mnist as _mn
    hp = 64
    (sx, sy), _ = _mn.synthetic_mnist((pool_n + hp) * B, 16, seed=1234 + rank)
    sx = torch.from_numpy(_mn.preprocess(sx)).view(pool_n + hp, B, 784)
    sy = torch.from_numpy(sy.astype(np.int32)).view(pool_n + hp, B)
    pool_x, pool_y = sx[:pool_n].to(dev), sy[:pool_n].to(dev)

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    pxs, pys = list(pool_x.unbind(0)), list(pool_y.unbind(0))

    def step_dev(i):
        j = i % pool_n
        eng.stage_device_batch(pxs[j], pys[j])       # pool -> input buffer on the copy stream, underneath the previous step
        eng.train_step_device(B)

    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    if sampler:
        sampler.mark("warm_start")
    for i in range(max(args.warmup, 3)):
        step_dev(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0 = K.LAUNCHES[0]
    if sampler:
        sampler.mark("timed_start")
    e0.record()
    for i in range(args.steps):
        step_dev(args.warmup + i)
    e1.record()
    barrier()
    if sampler:
        sampler.mark("timed_end")
    ms = e0.elapsed_time(e1)
    launches = eng.launches_per_step * args.steps        # kernels of ours inside the timed region (graph replays included)
    t = torch.tensor([ms], dtype=torch.float64)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = args.steps * B * world_size / (ms / 1e3)

    # ---- end to end through the public API: pinned host batch -> H2D -> step -> loss D2H, every step ----
    e2e = None
    if not args.skip_e2e:
        hx = sx[pool_n:].contiguous().pin_memory()
        hy = sy[pool_n:].contiguous().pin_memory()
        host_loss = torch.zeros(args.steps + 8, dtype=torch.float32).pin_memory()
        for i in range(max(args.warmup, 3)):
            model.train_on_batch(hx[i % hp], hy[i % hp])
        barrier()
        t0 = time.perf_counter()
        e0.record()
        mirror = bool(getattr(eng, "_mirror", False))    # the step's last kernel stores loss + correct count (8 B) to pinned host memory
        for i in range(args.steps):
            r = model.train_on_batch(hx[i % hp], hy[i % hp])
            if not mirror:
                host_loss[i:i + 1].copy_(r.loss.t.view(1), non_blocking=True)
        e1.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ems = max(e0.elapsed_time(e1), 0.0)
        t = torch.tensor([ems, wall], dtype=torch.float64)
        if world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ems = float(t[0].item())
        if mirror:
            host_loss[args.steps - 1] = float(r.loss)                # read from the host ring (waits for that step only)
            assert bool(torch.isfinite(eng.host_loss).all())
        assert bool(torch.isfinite(host_loss[:args.steps]).all())
        e2e = {"value": round(args.steps * B * world_size / (ems / 1e3), 1), "unit": "images/s",
               "h2d_bytes_per_step": B * 784 * 4 + B * 4, "d2h_bytes_per_step": 8 if mirror else 4,
               "d2h": "loss + correct count stored to pinned host memory by the step's last kernel" if mirror else "cudaMemcpyAsync of the loss",
               "ms_per_step": round(ems / args.steps, 5),
               "wall_ms_per_step": round(float(t[1].item()) / args.steps, 5), "final_loss": round(float(host_loss[args.steps - 1]), 4)}
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        out = {
            "metric": "MNIST CNN training throughput (images/sec), device-timed, max over ranks", "value": round(value, 1),
            "unit": "images/s", "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic 28x28 inputs (learnable class-conditional blobs + noise), random-init weights", "impl": "ours",
            "config": {"model": (f"mnist_cnn(conv5x5x{eng.C1}-pool-conv5x5x{eng.C2}-pool-dense{eng.D}-dropout0.5-dense10, "
                                 if args.arch == "mnist_cnn" else
                                 f"keras128(conv3x3x{eng.C1}-conv3x3x{eng.C2}-pool-dropout0.25-dense{eng.D}-dropout0.5-dense10, ")
                                + f"{spec.num_params} params)", "global_batch": B * world_size, "per_gpu_batch": B, "seq_len": None,
                       "parallelism": f"dp{world_size}",
                       "optimizer": ("Adam(lr=0.001*size) in the dense wgrad GEMM's epilogue and in the closing collective" if eng.dense_fused
                                     else "Adam(lr=0.001*size) fused into the NVLink allreduce"),
                       "allreduce": ({"dense.kernel": "none: all-gather of dz + activation slices (bf16) -> wgrad GEMM over the global batch "
                                                      "-> Adam in the GEMM epilogue -> bf16 weights stored to all ranks",
                                      "rest": ("one push-form one-shot kernel at the end of the step (inbox stores in front of the barrier flag)"
                                               if not eng.comm_rest else "oneshot (conv2/out/bias underneath the conv1 wgrad; conv1 at the end)"),
                                      "wire": "bf16 activations / fp32 small grads"}
                                     if eng.dense_fused else
                                     {"dense.kernel": ["oneshot", "twoshot", "nvls", "gemm+reduce-scatter fused, then two-shot tail"][eng.comm_dense[0][2]],
                                      "rest": ["oneshot", "twoshot", "nvls"][eng.comm_rest[1][2]], "wire": "fp32"}),
                       "cuda_graph": bool(eng.use_graph), "tcgen05": {"dense": bool(eng.tc_dense), "conv2": bool(eng.tc_conv)},
                       "l2": f"inputs rotate through a {pool_n * B * 784 * 4 / 1e6:.0f} MB device pool (> 126 MB L2); "
                             f"e2e inputs come from pinned host memory every step"},
            "gpu_launches": int(launches), "launches_per_step": int(eng.launches_per_step), "clocks": clocks, "e2e": e2e,
        }
        print(json.dumps(out))
    hvd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
