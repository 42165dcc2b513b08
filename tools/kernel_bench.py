#!/usr/bin/env python
"""Device time of every kernel of the training step in isolation: each kernel is captured R times back to back
in a CUDA graph (no CPU launch cost in the measurement), replayed, and timed with CUDA events.
    python tools/kernel_bench.py [--batch 128] [--width 1.0] [--reps 20]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200dp as hvd
from b200dp.ops import kernels as K

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--width", type=float, default=1.0)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--arch", default="mnist_cnn", choices=["mnist_cnn", "keras128"])
ap.add_argument("--dense-splits", default="", help="comma list: also time the dense forward GEMM at these split-K factors")
ap.add_argument("--only", default="", help="substring: time only the kernels whose name contains it (short runs under ncu)")
ap.add_argument("--out", default="gpurun_out/kernel_bench.json")
a = ap.parse_args()
hvd.init(device="cuda")
e = hvd.make_engine(hvd.get_spec(a.arch, a.width), hvd.DistributedOptimizer(hvd.Adam(1e-3)), engine="b200", batch_size=a.batch, use_graph=False)
B = a.batch
e.x.copy_(torch.rand(B, 784, device="cuda")); e.y.copy_(torch.randint(0, 10, (B,), device="cuda", dtype=torch.int32))
for _ in range(3):
    e.train_step_device(B)
torch.cuda.synchronize()
w, g = e.w16, e.grads[0]
a2f = e.a2[:B].view(B, e.F); dpool = e.dpool[:B].view(B, e.F)
split = max(1, min(e.F // 64, (148 * 2) // max(1, (e.D // 64) * ((B + 127) // 128)), 32))
T4 = e.T * 4
if e.keras:
    conv1 = {"conv1_valid_fwd": lambda: K.conv1_valid_fwd(e.x[:B], e.pv("conv1.kernel", w), e.pv("conv1.bias", w), e.a1[:B]),
             "feature_dropout": lambda: K.dropout_bf16(e.a2[:B], e.a2[:B], e.feat_drop, 1, e.step_dev, 0)}
    conv1w = {"conv1_valid_wgrad": lambda: K.conv1_valid_wgrad(e.x[:B], e.da1[:B], e.a1[:B], e.pv("conv1.kernel", g), e.pv("conv1.bias", g))}
else:
    conv1 = {"conv1_pool_fwd": lambda: K.conv1_pool_fwd(e.x[:B], e.pv("conv1.kernel", w), e.pv("conv1.bias", w), e.a1[:B], e.am1[:B])}
    conv1w = {"conv1_wgrad": lambda: K.conv1_wgrad(e.x[:B], e.da1[:B], e.a1[:B], e.am1[:B], e.pv("conv1.kernel", g), e.pv("conv1.bias", g))}
ops = {
    **conv1,
    "conv2_fwd(tc)": lambda: K.conv_tc_fwd(e.a1[:B], e.pv("conv2.kernel", w), e.pv("conv2.bias", w), e.a2[:B], e.am2[:B], relu=True, pool=True, pad=e.PAD2),
    "dense_fwd(tc,splitK)": lambda: K.gemm_tc(a2f, e.pv("dense.kernel", w), e.z[:B], B, e.D, e.F, b_mn_major=True, out_mode=2, split_k=split, block_n=64),
    "head_fwd_bwd": lambda: K.head_fwd_bwd(e.z[:B], e.pv("dense.bias", w), e.pv("out.kernel", w), e.pv("out.bias", w), e.y[:B], e.h[:B], e.dz[:B],
                                           e.dlogits[:B], None, e.loss_ring, e.correct_ring, drop_rate=0.5, training=True, seed=1, step_ptr=e.step_dev, ring=1024),
    "head_wgrad": lambda: K.head_wgrad(e.h[:B], e.dz[:B], e.dlogits[:B], e.pv("out.kernel", g), e.pv("out.bias", g), e.pv("dense.bias", g)),
    "dense_wgrad(tc)": lambda: K.gemm_tc(a2f, e.dz[:B], e.pv("dense.kernel", g), e.F, e.D, B, a_mn_major=True, b_mn_major=True, out_mode=0, block_n=128),
    "dense_dgrad(tc)": lambda: K.gemm_tc(e.dz[:B], e.pv("dense.kernel", w), dpool, B, e.F, e.D, out_mode=1, block_n=32),
    "unpool_relu_bwd": lambda: K.unpool_relu_bwd(e.dpool[:B], e.a2[:B], e.am2[:B], e.dy2[:B], e.pv("conv2.bias", g)),
    "conv2_wgrad(tc)": lambda: K.conv_tc_wgrad(e.a1[:B], e.dy2[:B], e.pv("conv2.kernel", g), None, pad=e.PAD2),
    "conv2_dgrad(tc)": lambda: K.conv_tc_dgrad(e.dy2[:B], e.pv("conv2.kernel", w), e.da1[:B], pad=e.PAD2),
    **conv1w,
}
if e.dense_fused:
    ops["dense_wgrad+adam(tc, fused epilogue)"] = lambda: e._dense_wgrad_adam(B)
    ops["fused_adam(conv2,out,dense.bias)"] = lambda: e._comm(1, e.comm_rest, 0, nblocks=64, nthreads=256)
    ops["fused_adam(conv1: step tail)"] = lambda: e._comm(2, e.comm_tail, 0, nblocks=1, nthreads=256)
    del ops["dense_wgrad(tc)"]
    # the same kernel in the shape one of 8 ranks runs: K = 8*B gathered rows, M = F/8 owned rows (no peers to store to here)
    Fs8 = e.F // 8
    A8 = torch.randn(8 * B, Fs8, device="cuda").to(torch.bfloat16)
    B8 = (torch.randn(8 * B, e.D, device="cuda") * 0.01).to(torch.bfloat16)
    dk = e.spec.param("dense.kernel")
    for bn in (64, 128):
        if e.D % bn:
            continue
        ops[f"dense_wgrad+adam as 1 of 8 owners (K=8B, bn={bn})"] = (lambda bn=bn: K.gemm_tc_wgrad_adam(
            e.arena.h, A8, B8, Fs8, e.D, 8 * B, Fs8, e.D, state_seg=e.seg_state, master_off=0, m_off=T4, v_off=2 * T4,
            w16_seg=e.seg_w16, w16_off=0, ctrl_seg=e.seg_ctrl, step_off=0, hyper_off=16, param_start=dk.offset, row0=0,
            gscale=0.125, block_n=bn))
else:
    ops["fused_adam(dense bucket)"] = lambda: e._comm(0, e.comm_dense, 0)
    ops["fused_adam(rest)"] = lambda: e._comm(1, e.comm_rest, 0)
for sp in [int(v) for v in a.dense_splits.split(",") if v]:
    ops[f"dense_fwd split={sp}"] = (lambda sp=sp: K.gemm_tc(a2f, e.pv("dense.kernel", w), e.z[:B], B, e.D, e.F, b_mn_major=True, out_mode=2, split_k=sp, block_n=64))
    if e.D % 128 == 0:
        ops[f"dense_fwd split={sp} bn128"] = (lambda sp=sp: K.gemm_tc(a2f, e.pv("dense.kernel", w), e.z[:B], B, e.D, e.F, b_mn_major=True, out_mode=2, split_k=sp, block_n=128))
if a.only:
    ops = {k: v for k, v in ops.items() if a.only in k}
res = {}
print(f"{'kernel':28s} {'us/launch':>10s}")
for name, fn in ops.items():
    fn(); torch.cuda.synchronize()
    gph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gph):
        for _ in range(a.reps):
            fn()
    gph.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        gph.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (5 * a.reps)
    res[name] = us
    print(f"{name:28s} {us:10.2f}")
print(f"{'sum':28s} {sum(v for k, v in res.items() if 'split=' not in k):10.2f}")
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
json.dump({"arch": a.arch, "batch": B, "width": a.width, "us_per_launch": res}, open(a.out, "w"), indent=1)
