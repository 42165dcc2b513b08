#!/usr/bin/env python
"""Per-kernel device times of one training step on the eager path (warm caches), from CUDA events.
    python tools/step_timeline.py [--batch 128] [--steps 20] [--out gpurun_out/timeline]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200dp as hvd
from b200dp.utils.timeline import Timeline

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--width", type=float, default=1.0)
ap.add_argument("--out", default="gpurun_out/timeline")
a = ap.parse_args()
hvd.init(device="cuda")
eng = hvd.make_engine(hvd.get_spec("mnist_cnn", a.width), hvd.DistributedOptimizer(hvd.Adam(1e-3)), engine="b200", batch_size=a.batch, use_graph=False)
x = torch.rand(a.batch, 784, device="cuda"); y = torch.randint(0, 10, (a.batch,), device="cuda", dtype=torch.int32)
for _ in range(5):
    eng.train_step(x, y)
torch.cuda.synchronize()
with Timeline() as tl:
    for _ in range(a.steps):
        eng.train_step(x, y)
s = tl.summary()
tot = 0.0
print(f"{'kernel':34s} {'calls/step':>10s} {'mean_us':>9s} {'min_us':>9s} {'us/step':>9s}")
for k, v in sorted(s.items(), key=lambda kv: -kv[1]['total_us']):
    per = v['total_us'] / a.steps
    tot += per
    print(f"{k:34s} {v['calls'] / a.steps:10.1f} {v['mean_us']:9.2f} {v['min_us']:9.2f} {per:9.2f}")
print(f"{'sum of kernel times per step':34s} {'':10s} {'':9s} {'':9s} {tot:9.2f}")
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
tl.save_chrome_trace(a.out + ".trace.json")
json.dump({"batch": a.batch, "steps": a.steps, "summary": s, "sum_us_per_step": tot}, open(a.out + ".json", "w"), indent=1)
