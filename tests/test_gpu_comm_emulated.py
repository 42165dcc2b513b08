"""Peer-memory collectives with all ranks emulated on ONE GPU (blockIdx.y = rank, cooperative launch).
Covers the flag/epoch protocol across repeated steps, one-shot and two-shot reductions, the fused
Adam epilogue (bitwise equal on every rank), broadcast, all-gather and Adasum (SURVEY §4.3 'Collective')."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from b200dp.models.torch_ref import adam_step_
from b200dp.ops import kernels as K
from b200dp.parallel.arena import SymmetricArena
from b200dp.parallel.collectives import adasum_tree

bf16 = torch.bfloat16


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class Rig:
    def __init__(self, n, numel):
        self.n, self.numel = n, numel
        self.a = SymmetricArena.emulated(n)
        self.a.set_timeout_ms(2000)
        self.ctrl = self.a.alloc(256)
        self.state = self.a.alloc(numel * 4 * 3)
        self.grad = self.a.alloc(numel * 4 * 2)
        self.w16 = self.a.alloc(numel * 2)
        for r in range(n):
            self.a.view(self.ctrl, 16, 5, torch.float64, r).copy_(torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.0], dtype=torch.float64))

    def t(self, seg, off_elems, dtype, r, es=4):
        return self.a.view(seg, off_elems * es, self.numel, dtype, r)

    def set_step(self, s):
        for r in range(self.n):
            self.a.view(self.ctrl, 0, 1, torch.int32, r).fill_(s)

    def fused(self, parity, ranges, reduce=True, scale=None, nblocks=8):
        N = self.numel
        self.a.fused_allreduce_adam(channel=1, nblocks=nblocks, nthreads=256, reduce=reduce, grad_seg=self.grad,
                                    grad_cur_off=parity * N * 4, grad_other_off=(1 - parity) * N * 4, state_seg=self.state,
                                    master_off=0, m_off=N * 4, v_off=2 * N * 4, w16_seg=self.w16, w16_off=0, ctrl_seg=self.ctrl,
                                    step_off=0, hyper_off=16, scale=(1.0 / self.n if scale is None else scale), ranges=ranges)

    def close(self):
        self.a.close()


@pytest.mark.parametrize("n", [1, 2, 4, 8])
@pytest.mark.parametrize("algo", [0, 1])
def test_fused_allreduce_adam_steps(n, algo):
    torch.manual_seed(n * 10 + algo)
    N = 4096 * 3 + 64
    rig = Rig(n, N)
    try:
        master0 = torch.randn(N, device="cuda")
        for r in range(n):
            rig.t(rig.state, 0, torch.float32, r).copy_(master0)
        pr, mr, vr = master0.clone(), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        ranges = [(0, 4096, algo, 1), (4096, N - 4096, 1 - algo if n > 1 else 0, 1)]
        for step in range(1, 6):
            par = step & 1
            rig.set_step(step)
            grads = [torch.randn(N, device="cuda") for _ in range(n)]
            for r in range(n):
                rig.t(rig.grad, par * N, torch.float32, r).copy_(grads[r])
                rig.t(rig.grad, (1 - par) * N, torch.float32, r).fill_(7.0)      # stale: must come back zeroed
            rig.fused(par, ranges)
            torch.cuda.synchronize()
            g = grads[0].clone()
            for r in range(1, n):
                g += grads[r]
            g *= 1.0 / n
            adam_step_(pr, g, mr, vr, step, lr=1e-3)
            w0 = rig.t(rig.w16, 0, bf16, 0, es=2)
            for r in range(n):
                # bf16 working weights: complete and bitwise identical on every rank
                assert torch.equal(rig.t(rig.w16, 0, bf16, r, es=2), w0), f"step {step}: rank {r} weights differ from rank 0"
                assert torch.count_nonzero(rig.t(rig.grad, (1 - par) * N, torch.float32, r)).item() == 0
            assert torch.allclose(w0.float(), pr.to(bf16).float(), rtol=1e-2, atol=1e-3)
            # fp32 master: one-shot ranges are replicated; two-shot ranges live on their owner
            for (st, ln, al, _) in ranges:
                for r in range(n):
                    mine = rig.t(rig.state, 0, torch.float32, r)[st:st + ln]
                    if al == 0 or n == 1:
                        assert torch.allclose(mine, pr[st:st + ln], rtol=1e-5, atol=1e-6)
                    else:
                        chunk = (math.ceil(ln / n) + 3) // 4 * 4
                        lo, hi = min(r * chunk, ln), min((r + 1) * chunk, ln)
                        assert torch.allclose(mine[lo:hi], pr[st + lo:st + hi], rtol=1e-5, atol=1e-6)
        # gather the sharded master back (checkpoint path) and compare everything
        if n > 1:
            for (st, ln, al, _) in ranges:
                if al == 1:
                    rig.a.allgather_slices_(rig.state, st * 4, ln)
            torch.cuda.synchronize()
            for r in range(n):
                assert torch.allclose(rig.t(rig.state, 0, torch.float32, r), pr, rtol=1e-5, atol=1e-6)
    finally:
        rig.close()


@pytest.mark.parametrize("n", [4, 8])
def test_fused_steps_race_stress_with_rank_skew(n):
    """Hundreds of back-to-back steps with every barrier arrival delayed by a pseudo-random time (up to 20 us):
    a fast rank must never overwrite a gradient buffer or a weight slice a slow peer is still reading
    (double-buffered gradients + epoch flags), and the replicas must stay bitwise identical throughout."""
    torch.manual_seed(n)
    N = 4096 * 2 + 320                       # odd-sized second range: unaligned slice tails
    rig = Rig(n, N)
    try:
        rig.a.set_debug_skew_ns(20000)
        master0 = torch.randn(N, device="cuda")
        for r in range(n):
            rig.t(rig.state, 0, torch.float32, r).copy_(master0)
        pr, mr, vr = master0.clone(), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        ranges = [(0, 4096, 0, 1), (4096, N - 4096, 1, 1)]
        base = [torch.randn(N, device="cuda") for _ in range(n)]
        for step in range(1, 241):
            par = step & 1
            rig.set_step(step)
            for r in range(n):                 # gradient of rank r at this step = base[r] * step (checksummable)
                rig.t(rig.grad, par * N, torch.float32, r).copy_(base[r] * float(step % 7 + 1))
            rig.fused(par, ranges, nblocks=6)
            g = sum(base) * float(step % 7 + 1) / n
            adam_step_(pr, g, mr, vr, step, lr=1e-3)
            if step % 60 == 0:
                torch.cuda.synchronize()
                w0 = rig.t(rig.w16, 0, bf16, 0, es=2)
                for r in range(1, n):
                    assert torch.equal(rig.t(rig.w16, 0, bf16, r, es=2), w0), f"step {step}: rank {r} diverged"
                assert torch.allclose(w0.float(), pr.to(bf16).float(), rtol=2e-2, atol=2e-3), f"step {step}"
    finally:
        rig.close()


@pytest.mark.parametrize("n", [2, 4, 8])
def test_gemm_reduce_scatter_then_fused_adam(n):
    """GEMM -> reduce-scatter fusion: every rank's wgrad GEMM adds its tiles into the owner ranks' buffers; the fused
    kernel (range algo 3) then updates each owned slice without reading peers."""
    torch.manual_seed(n + 7)
    M, Nn, Kk = 128 * n, 128, 96
    N = M * Nn
    rig = Rig(n, N)
    try:
        master0 = torch.randn(N, device="cuda")
        prods = []
        for r in range(n):
            rig.t(rig.state, 0, torch.float32, r).copy_(master0)
            rig.t(rig.grad, 0, torch.float32, r).zero_()
        for r in range(n):                       # "rank r" runs its GEMM; the epilogue scatters rows to their owners
            A = torch.randn(Kk, M, device="cuda").to(bf16)        # stored [K, M]  (activations^T, MN-major)
            Bm = torch.randn(Kk, Nn, device="cuda").to(bf16)      # stored [K, N]
            prods.append(A.float().t() @ Bm.float())
            K.gemm_tc_reduce_scatter(rig.a.h, A, Bm, rig.grad, 0, n, M, Nn, Kk, a_mn_major=True, b_mn_major=True, block_n=128)
        torch.cuda.synchronize()
        total = sum(prods)
        rows = M // n
        for o in range(n):
            got = rig.t(rig.grad, 0, torch.float32, o).view(M, Nn)
            assert torch.allclose(got[o * rows:(o + 1) * rows], total[o * rows:(o + 1) * rows], rtol=1e-3, atol=1e-2), f"owner {o}"
            rest = torch.cat([got[:o * rows], got[(o + 1) * rows:]])
            assert torch.count_nonzero(rest).item() == 0                  # nothing lands outside the owned rows
        rig.set_step(1)
        rig.fused(0, [(0, N, 3, 1)])
        torch.cuda.synchronize()
        pr, mr, vr = master0.clone(), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        adam_step_(pr, total.reshape(-1) / n, mr, vr, 1, lr=1e-3)
        w0 = rig.t(rig.w16, 0, bf16, 0, es=2)
        for r in range(n):
            assert torch.equal(rig.t(rig.w16, 0, bf16, r, es=2), w0)
        assert torch.allclose(w0.float(), pr.to(bf16).float(), rtol=1e-2, atol=1e-3)
    finally:
        rig.close()


def test_fused_local_only_path_after_adasum():
    rig = Rig(4, 1024)
    try:
        for r in range(4):
            rig.t(rig.state, 0, torch.float32, r).fill_(1.0)
            rig.t(rig.grad, 0, torch.float32, r).fill_(float(r + 1))
        rig.set_step(1)
        rig.fused(0, [(0, 1024, 1, 0)], reduce=False, scale=1.0)
        torch.cuda.synchronize()
        vals = [rig.t(rig.state, 0, torch.float32, r)[0].item() for r in range(4)]
        assert all(abs(v - (1.0 - 1e-3)) < 1e-6 for v in vals)       # first Adam step moves by lr regardless of |g|
    finally:
        rig.close()


@pytest.mark.parametrize("n", [2, 8])
@pytest.mark.parametrize("algo", [0, 1])
def test_plain_allreduce_broadcast(n, algo):
    torch.manual_seed(n)
    N = 10_000
    a = SymmetricArena.emulated(n)
    a.set_timeout_ms(2000)
    try:
        seg = a.alloc(N * 4 * 2)
        xs = [torch.randn(N, device="cuda") for _ in range(n)]
        for it in range(3):
            for r in range(n):
                a.view(seg, 0, N, torch.float32, r).copy_(xs[r] * (it + 1))
            K.check(K.lib().b200_allreduce(a.h, 2, 4, 256, algo, seg, 0, seg, N * 4, N, 1.0 / n, _stream()))
            torch.cuda.synchronize()
            want = sum(xs) * (it + 1) / n
            first = a.view(seg, N * 4, N, torch.float32, 0)
            for r in range(n):
                got = a.view(seg, N * 4, N, torch.float32, r)
                assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)
                assert torch.equal(got, first)
        root = n - 1
        K.check(K.lib().b200_broadcast(a.h, 3, 4, 256, root, seg, 0, seg, N * 4, N, _stream()))
        torch.cuda.synchronize()
        for r in range(n):
            assert torch.equal(a.view(seg, N * 4, N, torch.float32, r), xs[root] * 3)
    finally:
        a.close()


@pytest.mark.parametrize("n,root", [(4, 1), (8, 0), (8, 5)])
def test_large_broadcast_scatter_allgather(n, root):
    """>= 1 MB and more than two ranks takes the two-phase path (scatter from the root, all-gather between ranks)."""
    N = 300_000 + 4 * n
    a = SymmetricArena.emulated(n)
    a.set_timeout_ms(2000)
    try:
        seg = a.alloc(N * 4 * 2)
        for r in range(n):
            a.view(seg, 0, N, torch.float32, r).copy_(torch.arange(N, device="cuda", dtype=torch.float32) * (r + 1))
            a.view(seg, N * 4, N, torch.float32, r).fill_(-1.0)
        for _ in range(2):
            K.check(K.lib().b200_broadcast(a.h, 3, 8, 256, root, seg, 0, seg, N * 4, N, _stream()))
        torch.cuda.synchronize()
        want = torch.arange(N, device="cuda", dtype=torch.float32) * (root + 1)
        for r in range(n):
            assert torch.equal(a.view(seg, N * 4, N, torch.float32, r), want), f"rank {r}"
    finally:
        a.close()


@pytest.mark.parametrize("n", [2, 4, 8])
def test_adasum_levels_match_tree(n):
    torch.manual_seed(n + 100)
    sizes = [800, 32, 5000, 64, 10]
    N = sum(sizes)
    N4 = (N + 3) // 4 * 4
    a = SymmetricArena.emulated(n)
    a.set_timeout_ms(2000)
    try:
        E = N4 * 4
        seg = a.alloc(E * 2 + 65536)
        xs = [torch.randn(N, device="cuda") * (r + 1) for r in range(n)]
        for r in range(n):
            a.view(seg, 0, N, torch.float32, r).copy_(xs[r])
        starts = [sum(sizes[:i]) for i in range(len(sizes))]
        res = a.adasum_levels(seg, 0, E, 2 * E, starts, sizes, nblocks=8)
        torch.cuda.synchronize()
        want = torch.cat([adasum_tree([x[s:s + l].cpu() for x in xs]) for s, l in zip(starts, sizes)])
        first = a.view(seg, res, N, torch.float32, 0)
        for r in range(n):
            got = a.view(seg, res, N, torch.float32, r)
            assert torch.equal(got, first), f"rank {r} differs bitwise from rank 0"
            assert torch.allclose(got.cpu(), want, rtol=1e-4, atol=1e-5)
    finally:
        a.close()


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("nblocks", [1, 7, 32])
def test_oneshot_bulk_prefetch_path(n, nblocks):
    """All ranges one-shot + blocks of <= 256 threads: the peers' slices come in by TMA bulk copies (shared-memory
    staging, block-contiguous partition).  Ranges of awkward sizes, several steps (both gradient-buffer parities)."""
    torch.manual_seed(n * 100 + nblocks)
    sizes = [832, 51264, 11276]
    N = sum(sizes)
    rig = Rig(n, N)
    try:
        master0 = torch.randn(N, device="cuda")
        for r in range(n):
            rig.t(rig.state, 0, torch.float32, r).copy_(master0)
        pr, mr, vr = master0.clone(), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        ranges, off = [], 0
        for sz in sizes:
            ranges.append((off, sz, 0, 1))
            off += sz
        for step in range(1, 5):
            par = step & 1
            rig.set_step(step)
            grads = [torch.randn(N, device="cuda") for _ in range(n)]
            for r in range(n):
                rig.t(rig.grad, par * N, torch.float32, r).copy_(grads[r])
                rig.t(rig.grad, (1 - par) * N, torch.float32, r).fill_(7.0)
            rig.fused(par, ranges, nblocks=nblocks)
            torch.cuda.synchronize()
            g = grads[0].clone()
            for r in range(1, n):
                g += grads[r]
            g *= 1.0 / n
            adam_step_(pr, g, mr, vr, step, lr=1e-3)
            w0 = rig.t(rig.w16, 0, bf16, 0, es=2)
            for r in range(n):
                assert torch.equal(rig.t(rig.w16, 0, bf16, r, es=2), w0), f"step {step}: rank {r} weights differ from rank 0"
                assert torch.count_nonzero(rig.t(rig.grad, (1 - par) * N, torch.float32, r)).item() == 0
                assert torch.allclose(rig.t(rig.state, 0, torch.float32, r), pr, rtol=1e-5, atol=1e-6)
    finally:
        rig.close()


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("nblocks", [1, 32])
def test_oneshot_push_inbox_path(n, nblocks):
    """Push form of the small one-shot buckets: every rank stores its slice into the peers' inboxes in front of its
    barrier flag; the sum (own gradient + inbox rows, rank order) must equal the pulled one bit for bit on every rank."""
    import ctypes
    torch.manual_seed(n * 7 + nblocks)
    sizes = [832, 51264, 11276]
    N = sum(sizes)
    rig = Rig(n, N)
    try:
        bpp = (n * N * 4 + 255) // 256 * 256
        inbox = rig.a.alloc(2 * bpp)
        for r in range(n):
            rig.a.view(inbox, 0, 2 * bpp // 4, torch.float32, r).fill_(float("nan"))     # never read before written
        K.check(K.lib().b200_comm_set_inbox(rig.a.h, 1, inbox, bpp), "set_inbox")
        master0 = torch.randn(N, device="cuda")
        for r in range(n):
            rig.t(rig.state, 0, torch.float32, r).copy_(master0)
        pr, mr, vr = master0.clone(), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        ranges, off = [], 0
        for sz in sizes:
            ranges.append((off, sz, 0, 1))
            off += sz
        for step in range(1, 6):
            par = step & 1
            rig.set_step(step)
            grads = [torch.randn(N, device="cuda") for _ in range(n)]
            for r in range(n):
                rig.t(rig.grad, par * N, torch.float32, r).copy_(grads[r])
            rig.fused(par, ranges, nblocks=nblocks)
            torch.cuda.synchronize()
            g = grads[0].clone()
            for r in range(1, n):
                g += grads[r]
            g *= 1.0 / n
            adam_step_(pr, g, mr, vr, step, lr=1e-3)
            w0 = rig.t(rig.w16, 0, bf16, 0, es=2)
            for r in range(n):
                assert torch.equal(rig.t(rig.w16, 0, bf16, r, es=2), w0), f"step {step}: rank {r} weights differ from rank 0"
                assert torch.allclose(rig.t(rig.state, 0, torch.float32, r), pr, rtol=1e-5, atol=1e-6)
    finally:
        K.lib().b200_comm_set_inbox(rig.a.h, -1, 0, 0)
        rig.close()
