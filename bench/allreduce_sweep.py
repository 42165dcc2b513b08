#!/usr/bin/env python
"""Fusion-buffer allreduce + broadcast sweep, 1 KB – 1 GB, ours (peer-memory kernels) vs NCCL (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        bench/allreduce_sweep.py --out gpurun_out/sweep_N.json [--max-bytes 1073741824]

Device-timed (CUDA events around `iters` back-to-back launches), max over ranks.  Bus bandwidth uses the NCCL
convention: allreduce S*2(N-1)/N/t, broadcast S/t.  Roofline term: bytes that must cross NVLink per GPU per
direction divided by the measured 770 GB/s peer bandwidth (900 nominal).  Also writes the one-shot/two-shot
crossover per world size into profiles/allreduce_table.json (read by b200dp/parallel/autoselect.py).
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import b200dp as hvd
from b200dp.ops import kernels as K

PEER_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=1024)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--write-table", action="store_true")
    ap.add_argument("--only-bcast", action="store_true", help="measure the broadcast arms only")
    a = ap.parse_args()
    os.environ.setdefault("B200DP_LOG_LEVEL", "WARNING")
    hvd.init(device="cuda")
    n, r, dev = hvd.size(), hvd.rank(), hvd.device()
    arena = hvd.world().arena
    nccl = None
    if not a.no_nccl and n > 1:
        nccl = dist.new_group(backend="nccl")
    sizes = []
    s = a.min_bytes
    while s <= a.max_bytes:
        sizes.append(s)
        s *= 4
    if sizes[-1] != a.max_bytes:
        sizes.append(a.max_bytes)
    maxel = a.max_bytes // 4
    seg = arena.alloc(maxel * 4 * 2 + 4096)
    src = arena.view(seg, 0, maxel, torch.float32)
    dst_off = maxel * 4
    src.copy_(torch.full((maxel,), float(r + 1), device=dev))
    tn = torch.empty(maxel, device=dev) if nccl is not None else None
    seg_mc = None
    if arena.nvls_supported():
        try:
            seg_mc = arena.alloc_multicast(maxel * 4 * 2 + 4096)
            arena.view(seg_mc, 0, maxel, torch.float32).copy_(torch.full((maxel,), float(r + 1), device=dev))
        except Exception as e:  # noqa: BLE001
            if r == 0:
                print(json.dumps({"nvls": "unavailable", "why": str(e)[:200]}), flush=True)
            seg_mc = None
    stream = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def shape(numel):
        nb = max(1, min(296, numel // 4096))
        return nb, 512

    def timed(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) * 1e3      # us

    rows = []
    for S in sizes:
        numel = max(4, S // 4 // 4 * 4)
        iters = 200 if S <= (1 << 20) else (50 if S <= (1 << 26) else 10)
        nb, nt = shape(numel)
        row = {"bytes": numel * 4}
        for name, algo in (() if a.only_bcast else (("oneshot", 0), ("twoshot", 1))):
            if algo == 0 and numel * 4 * n > (3 << 30):
                continue                                   # one-shot reads N*S bytes: pointless at this size
            f = lambda algo=algo: K.check(K.lib().b200_allreduce(arena.h, 5, nb, nt, algo, seg, 0, seg, dst_off, numel, 1.0 / n, stream()))
            f()
            torch.cuda.synchronize()
            got = arena.view(seg, dst_off, numel, torch.float32)
            want = (n + 1) / 2.0
            assert abs(float(got[0]) - want) < 1e-5 and abs(float(got[numel - 1]) - want) < 1e-5, (name, S, float(got[0]), want)
            us = timed(f, iters)
            row[name + "_us"] = us
            row[name + "_busbw_gbs"] = numel * 4 * 2 * (n - 1) / n / us / 1e3
        if seg_mc is not None and numel * 4 >= 4096 and not a.only_bcast:
            fn = lambda: K.check(K.lib().b200_allreduce_nvls(arena.h, 7, nb, nt, seg_mc, 0, seg_mc, dst_off, numel, 1.0 / n, stream()))
            fn()
            torch.cuda.synchronize()
            got = arena.view(seg_mc, dst_off, numel, torch.float32)
            want = (n + 1) / 2.0
            assert abs(float(got[0]) - want) < 1e-5 and abs(float(got[numel - 1]) - want) < 1e-5, ("nvls", S, float(got[0]), want)
            us = timed(fn, iters)
            row["nvls_us"] = us
            row["nvls_busbw_gbs"] = numel * 4 * 2 * (n - 1) / n / us / 1e3
        fb = lambda: K.check(K.lib().b200_broadcast(arena.h, 6, nb, nt, 0, seg, 0, seg, dst_off, numel, stream()))
        us = timed(fb, iters)
        row["bcast_us"] = us
        row["bcast_gbs"] = numel * 4 / us / 1e3
        if seg_mc is not None and numel * 4 >= 4096:
            fbn = lambda: K.check(K.lib().b200_broadcast_nvls(arena.h, 8, nb, nt, 0, seg_mc, 0, seg_mc, dst_off, numel, stream()))
            fbn()
            torch.cuda.synchronize()
            got = arena.view(seg_mc, dst_off, numel, torch.float32)
            assert float(got[0]) == 1.0 and float(got[numel - 1]) == 1.0, ("nvls bcast", S, float(got[0]))
            us = timed(fbn, iters)
            row["nvls_bcast_us"] = us
            row["nvls_bcast_gbs"] = numel * 4 / us / 1e3
        if nccl is not None:
            t = tn[:numel]
            if not a.only_bcast:
                us = timed(lambda: dist.all_reduce(t, group=nccl), iters)
                row["nccl_us"] = us
                row["nccl_busbw_gbs"] = numel * 4 * 2 * (n - 1) / n / us / 1e3
            us = timed(lambda: dist.broadcast(t, src=0, group=nccl), iters)
            row["nccl_bcast_us"] = us
            row["nccl_bcast_gbs"] = numel * 4 / us / 1e3
        # roofline: two-shot moves 2(N-1)/N*S per direction per GPU over NVLink
        cands = [v for k, v in row.items() if k in ("oneshot_us", "twoshot_us", "nvls_us")]
        if cands:
            best = min(cands)
            row["best_us"] = best
            row["nvlink_roofline_us"] = numel * 4 * 2 * (n - 1) / n / (PEER_GBS * 1e3)
            row["frac_of_roofline"] = row["nvlink_roofline_us"] / best
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
    if r == 0:
        cross = 0
        for row in rows:
            if "oneshot_us" in row and row["oneshot_us"] <= row.get("twoshot_us", float("inf")):
                cross = row["bytes"]
        nvls_min = None
        for row in reversed(rows):
            if "nvls_us" in row and row["nvls_us"] < row.get("twoshot_us", float("inf")) and row["nvls_us"] < row.get("oneshot_us", float("inf")):
                nvls_min = row["bytes"]
            else:
                break
        out = {"n_gpus": n, "rows": rows, "oneshot_max_bytes": cross, "nvls_min_bytes": nvls_min, "peer_gbs_denominator": PEER_GBS}
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)
        if a.write_table:
            tp = os.path.join(os.path.dirname(a.out) or ".", "allreduce_table.json")
            tab = json.load(open(tp)) if os.path.exists(tp) else {"oneshot_max_bytes": {}}
            tab["oneshot_max_bytes"][str(n)] = cross
            if nvls_min is not None:
                tab.setdefault("nvls_min_bytes", {})[str(n)] = nvls_min
            json.dump(tab, open(tp, "w"), indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
