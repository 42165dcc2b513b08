"""The fused dense-layer update: all-gather of dz / activation slices -> wgrad GEMM over the global batch with Adam in the
epilogue -> bf16 weights stored into every rank (csrc/gemm_tc.cu out_mode 4, csrc/comm.cu allgather_acts_kernel).

Replaces, for dense.kernel, the gradient allreduce + ApplyAdam pair of the reference
(horovod/tensorflow_mnist.py:130-136).  Ranks are emulated on one GPU; the oracle is fp32 PyTorch on the same bf16 inputs:
grad = (1/n) * sum_r a2_r^T dz_r, then the reference Adam step."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200dp.models.torch_ref import adam_step_
from b200dp.ops import kernels as K
from b200dp.parallel.arena import SymmetricArena

bf16 = torch.bfloat16


class Rig:
    def __init__(self, n, B, F, D, pre=64):
        self.n, self.B, self.F, self.D, self.pre = n, B, F, D, pre
        self.T = pre + F * D                      # a few parameters in front, so param_start != 0
        self.a = SymmetricArena.emulated(n) if n > 1 else self._solo()
        if n > 1:
            self.a.set_timeout_ms(2000)
        T = self.T
        self.ctrl = self.a.alloc(256)
        self.state = self.a.alloc(T * 4 * 3)
        self.w16 = self.a.alloc(T * 2)
        self.A2, self.DZ = 0, B * F * 2
        self.A2G = self.DZ + B * D * 2
        self.DZG = self.A2G + B * F * 2
        self.act = self.a.alloc(self.DZG + n * B * D * 2)
        for r in range(n):
            self.a.view(self.ctrl, 16, 5, torch.float64, r).copy_(torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.0], dtype=torch.float64))

    @staticmethod
    def _solo():
        import ctypes
        h = ctypes.c_void_p()
        K.check(K.lib().b200_comm_create(0, 1, 0, ctypes.byref(h)), "comm_create")
        return SymmetricArena(h, 0, 1, False, torch.device("cuda", torch.cuda.current_device()))

    def v(self, seg, off, numel, dt, r):
        return self.a.view(seg, off, numel, dt, r if self.n > 1 else None)

    def step(self, t, a2s, dzs, scale):
        n, B, F, D, T = self.n, self.B, self.F, self.D, self.T
        Fs = F // n
        for r in range(n):
            self.v(self.ctrl, 0, 1, torch.int32, r).fill_(t)
            self.v(self.act, self.A2, B * F, bf16, r).copy_(a2s[r].reshape(-1))
            self.v(self.act, self.DZ, B * D, bf16, r).copy_(dzs[r].reshape(-1))
        if n > 1:
            K.allgather_acts(self.a.h, 3, 4, self.act, self.A2, self.act, self.DZ, self.act, self.A2G, self.DZG, B, F, D)
        torch.cuda.synchronize()
        for r in range(n):
            if n > 1:
                A = self.v(self.act, self.A2G, n * B * Fs, bf16, r).view(n * B, Fs)
                Bm = self.v(self.act, self.DZG, n * B * D, bf16, r).view(n * B, D)
            else:
                A, Bm = self.v(self.act, self.A2, B * F, bf16, r).view(B, F), self.v(self.act, self.DZ, B * D, bf16, r).view(B, D)
            K.gemm_tc_wgrad_adam(self.a.h, A, Bm, Fs, D, n * B, Fs, D, state_seg=self.state, master_off=0, m_off=T * 4,
                                 v_off=2 * T * 4, w16_seg=self.w16, w16_off=0, ctrl_seg=self.ctrl, step_off=0, hyper_off=16,
                                 param_start=self.pre, row0=r * Fs, as_rank=r if n > 1 else -1, gscale=scale,
                                 block_n=128 if D % 128 == 0 else 64)
        torch.cuda.synchronize()

    def close(self):
        self.a.close()


@pytest.mark.parametrize("n,B,F,D", [(1, 128, 3136, 1024), (1, 100, 512, 192), (2, 128, 1024, 256), (4, 37, 1024, 128),
                                     (8, 128, 3136, 1024), (8, 16, 9216, 128)])
def test_gather_wgrad_adam_matches_fp32_reference(n, B, F, D):
    torch.manual_seed(n * 1000 + B)
    rig = Rig(n, B, F, D)
    try:
        T, pre = rig.T, rig.pre
        master0 = torch.randn(T, device="cuda") * 0.05
        for r in range(n):
            rig.v(rig.state, 0, T, torch.float32, r).copy_(master0)
            rig.v(rig.w16, 0, T, bf16, r).copy_(master0.to(bf16))
        pr = master0[pre:].clone()
        mr, vr = torch.zeros_like(pr), torch.zeros_like(pr)
        Fs = F // n
        for t in range(1, 4):
            a2s = [torch.randn(B, F, device="cuda").relu().to(bf16) for _ in range(n)]
            dzs = [(torch.randn(B, D, device="cuda") * 0.1).to(bf16) for _ in range(n)]
            rig.step(t, a2s, dzs, 1.0 / n)
            g = sum(a.float().t() @ d.float() for a, d in zip(a2s, dzs)) / n
            adam_step_(pr, g.reshape(-1), mr, vr, t, lr=1e-3)
            w0 = rig.v(rig.w16, 0, T, bf16, 0)
            for r in range(n):
                wr = rig.v(rig.w16, 0, T, bf16, r)
                assert torch.equal(wr, w0), f"step {t}: bf16 weights of rank {r} differ from rank 0"
                # untouched parameters in front of the range stay as they were
                assert torch.equal(wr[:pre], master0[:pre].to(bf16))
                lo, hi = r * Fs * D, (r + 1) * Fs * D
                mine = rig.v(rig.state, 0, T, torch.float32, r)[pre:]
                err = (mine[lo:hi] - pr[lo:hi]).abs().max().item()
                assert err < 2e-4, f"step {t}: rank {r} master slice off by {err}"
                mm = rig.v(rig.state, T * 4, T, torch.float32, r)[pre:]
                assert torch.allclose(mm[lo:hi], mr[lo:hi], rtol=1e-3, atol=1e-5)
            assert (w0[pre:].float() - pr).abs().max().item() < 2e-3
    finally:
        rig.close()
