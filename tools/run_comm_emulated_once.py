#!/usr/bin/env python
"""Eight EMULATED ranks on one GPU (blockIdx.y = rank, cooperative launches) running the collective kernels of one
training step: the dz/activation gather, the dense wgrad GEMM with the Adam epilogue as each owner, and the push-form
end-of-step collective.  A target for `ncu -k regex:...` on a single GPU (a multi-rank job must never run under ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.ops import kernels as K
from b200dp.parallel.arena import SymmetricArena

bf16 = torch.bfloat16
n, B, F, D = 8, 128, 3136, 1024
small = [832, 51264, 11276]
T = sum(small) + F * D
a = SymmetricArena.emulated(n)
a.set_timeout_ms(5000)
ctrl, state, w16 = a.alloc(256), a.alloc(T * 4 * 3), a.alloc(T * 2)
grad = a.alloc(sum(small) * 4 * 2)
A2, DZ = 0, B * F * 2
A2G, DZG = DZ + B * D * 2, DZ + B * D * 2 + B * F * 2
act = a.alloc(DZG + n * B * D * 2)
bpp = (n * sum(small) * 4 + 255) // 256 * 256
inbox = a.alloc(2 * bpp)
K.check(K.lib().b200_comm_set_inbox(a.h, 2, inbox, bpp), "set_inbox")
for r in range(n):
    a.view(ctrl, 16, 5, torch.float64, r).copy_(torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.0], dtype=torch.float64))
    a.view(act, A2, B * F, bf16, r).copy_(torch.randn(B * F, device="cuda").relu().to(bf16))
    a.view(act, DZ, B * D, bf16, r).copy_((torch.randn(B * D, device="cuda") * 0.1).to(bf16))
    a.view(grad, 0, sum(small) * 2, torch.float32, r).normal_()
Fs, NS = F // n, sum(small)
BN = int(os.environ.get("BN", "64"))
STEPS = int(os.environ.get("STEPS", "3"))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
tg = 0.0
for step in range(1, STEPS + 1):
    for r in range(n):
        a.view(ctrl, 0, 1, torch.int32, r).fill_(step)
    K.allgather_acts(a.h, 3, 64, act, A2, act, DZ, act, A2G, DZG, B, F, D)
    ev[0].record()
    for r in range(n):
        Ag = a.view(act, A2G, n * B * Fs, bf16, r).view(n * B, Fs)
        Bg = a.view(act, DZG, n * B * D, bf16, r).view(n * B, D)
        K.gemm_tc_wgrad_adam(a.h, Ag, Bg, Fs, D, n * B, Fs, D, state_seg=state, master_off=0, m_off=T * 4, v_off=2 * T * 4,
                             w16_seg=w16, w16_off=0, ctrl_seg=ctrl, step_off=0, hyper_off=16, param_start=NS, row0=r * Fs,
                             as_rank=r, gscale=1.0 / n, block_n=BN)
    ev[1].record()
    torch.cuda.synchronize()
    if step > 2:
        tg += ev[0].elapsed_time(ev[1]) * 1e3 / n
    par = step & 1
    ranges, off = [], 0
    for sz in small:
        ranges.append((off, sz, 0, 1))
        off += sz
    a.fused_allreduce_adam(channel=2, nblocks=32, nthreads=256, reduce=True, grad_seg=grad, grad_cur_off=par * NS * 4,
                           grad_other_off=(1 - par) * NS * 4, state_seg=state, master_off=0, m_off=T * 4, v_off=2 * T * 4,
                           w16_seg=w16, w16_off=0, ctrl_seg=ctrl, step_off=0, hyper_off=16, scale=1.0 / n, ranges=ranges)
torch.cuda.synchronize()
w0 = a.view(w16, 0, T, bf16, 0)
assert all(torch.equal(a.view(w16, 0, T, bf16, r), w0) for r in range(n)), "replicas differ"
print("ok", float(w0.float().abs().mean()), f"wgrad+adam+push: {tg / max(1, STEPS - 2):.2f} us per rank-launch (eager launches back to back, bn={BN})")
a.close()
