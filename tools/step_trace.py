#!/usr/bin/env python
"""Per-kernel, per-stream DEVICE timeline of graph-replayed training steps, on 1..8 GPUs.

The step runs exactly as in bench.py (CUDA-graph replay, inputs from the device pool); kernel start/end
timestamps come from the CUPTI activity records that torch.profiler collects (kernels inside graph replays
included).  A timeline is not a benchmark number: it answers "what overlaps what, and where does the step wait".

    python tools/step_trace.py --out gpurun_out/trace_n1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/step_trace.py --out gpurun_out/trace_n2

Every rank writes <out>_r<rank>.json; rank 0 also writes <out>.md (its own timeline + the step period of every rank).
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def collect(trace_path, anchor):
    ev = json.load(open(trace_path))["traceEvents"]
    ks = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
    ks.sort(key=lambda e: e["ts"])
    starts = [i for i, e in enumerate(ks) if anchor in e["name"]]
    steps = []
    for a, b in zip(starts[:-1], starts[1:]):
        t0 = ks[a]["ts"]
        steps.append({"period_us": ks[b]["ts"] - t0,
                      "kernels": [{"name": e["name"], "stream": e.get("args", {}).get("stream"), "start_us": e["ts"] - t0,
                                   "dur_us": e["dur"], "grid": e.get("args", {}).get("grid"),
                                   "regs": e.get("args", {}).get("registers per thread"),
                                   "smem": e.get("args", {}).get("shared memory")} for e in ks[a:b]]})
    return steps


def short(name):
    n = name.replace("void ", "").replace("(anonymous namespace)::", "")
    n = n.split("(")[0]
    return n[:48]


def summarize(steps):
    """Mean start/duration per kernel slot (slot = position in the step; the sequence is identical in every replay)."""
    by_len = {}
    for s in steps:
        by_len.setdefault(len(s["kernels"]), []).append(s)
    common = max(by_len.values(), key=len)
    n = len(common)
    rows = []
    for i in range(len(common[0]["kernels"])):
        ks = [s["kernels"][i] for s in common]
        rows.append({"name": short(ks[0]["name"]), "stream": ks[0]["stream"], "grid": ks[0]["grid"], "regs": ks[0]["regs"],
                     "smem": ks[0]["smem"],
                     "start_us": sum(k["start_us"] for k in ks) / n, "dur_us": sum(k["dur_us"] for k in ks) / n})
    per = sorted(s["period_us"] for s in common)
    return {"steps_used": n, "period_us_mean": sum(per) / n, "period_us_median": per[n // 2], "period_us_min": per[0],
            "period_us_max": per[-1], "rows": rows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--width", type=float, default=1.0)
    ap.add_argument("--arch", default="mnist_cnn")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--out", default="gpurun_out/trace")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from torch.profiler import ProfilerActivity, profile
    import b200dp as hvd

    os.environ.setdefault("B200DP_LOG_LEVEL", "WARNING")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    hvd.init(device="cuda")
    dev = hvd.device()
    B = a.batch
    spec = hvd.get_spec(a.arch, a.width)
    opt = hvd.DistributedOptimizer(hvd.Adam.keras(lr=0.001 * hvd.size()), op=hvd.Average, algo=a.algo)
    model = hvd.Model(spec, engine="b200", batch_size=B)
    model.compile(optimizer=opt, loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    eng = model.engine
    eng.broadcast_state(0)
    pool_n = max(8, int(160e6 // (B * 784 * 4)) + 1)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    px = torch.rand(pool_n, B, 784, generator=g).to(dev)
    py = torch.randint(0, 10, (pool_n, B), generator=g, dtype=torch.int32).to(dev)

    pxs, pys = list(px.unbind(0)), list(py.unbind(0))

    def step(i):
        eng.stage_device_batch(pxs[i % pool_n], pys[i % pool_n])
        eng.train_step_device(B)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(a.warmup):
        step(i)
    barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(a.steps):
            step(a.warmup + i)
        barrier()
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    tmp = tempfile.mktemp(suffix=".json")
    prof.export_chrome_trace(tmp)
    anchor = "conv1_valid3_fwd" if a.arch == "keras128" else "conv1_pool_fwd"
    steps = collect(tmp, anchor)
    os.unlink(tmp)
    summ = summarize(steps[2:]) if len(steps) > 4 else {"steps_used": 0, "rows": []}
    summ.update(rank=rank, world=world, batch=B, arch=a.arch, width=a.width)
    json.dump(summ, open(f"{a.out}_r{rank}.json", "w"), indent=1)
    barrier()
    if rank == 0:
        alls = [json.load(open(f"{a.out}_r{r}.json")) for r in range(world)]
        with open(a.out + ".md", "w") as f:
            f.write(f"# Step timeline, {a.arch} batch {B}/rank width {a.width}, N={world} (CUPTI kernel records of graph replays; "
                    f"{summ['steps_used']} steps averaged)\n\n")
            f.write("Step period per rank (us, start of conv1 fwd to the next one): " +
                    ", ".join(f"r{s['rank']}: median {s.get('period_us_median', 0):.1f} (min {s.get('period_us_min', 0):.1f}, "
                              f"max {s.get('period_us_max', 0):.1f})" for s in alls) + "\n\n")
            f.write("| # | kernel | stream | grid | regs | smem | start us | dur us | end us |\n|---|---|---|---|---|---|---|---|---|\n")
            for i, r in enumerate(summ["rows"]):
                f.write(f"| {i} | {r['name']} | {r['stream']} | {r['grid']} | {r['regs']} | {r['smem']} | {r['start_us']:.1f} | "
                        f"{r['dur_us']:.1f} | {r['start_us'] + r['dur_us']:.1f} |\n")
            f.write(f"\nSum of kernel durations: {sum(r['dur_us'] for r in summ['rows']):.1f} us\n")
        print(open(a.out + ".md").read())
    hvd.shutdown()


if __name__ == "__main__":
    main()
