"""Property tests over shapes (SURVEY §4.3 'Kernel (GPU)': hypothesis is installed): the tcgen05 GEMM and the
implicit-GEMM convolution against fp32 PyTorch for randomly drawn, supported shapes."""
import pytest
import torch
import torch.nn.functional as F
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu

from b200dp.ops import kernels as K

bf16 = torch.bfloat16
COMMON = dict(deadline=None, max_examples=25, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture],
              derandomize=True)


@settings(**COMMON)
@given(m=st.integers(1, 40).map(lambda v: v * 8), n=st.integers(1, 24).map(lambda v: v * 16), k=st.integers(1, 40).map(lambda v: v * 8),
       a_mn=st.booleans(), b_mn=st.booleans(), bn=st.sampled_from([64, 128]), out_bf16=st.booleans())
def test_gemm_tc_random_shapes(m, n, k, a_mn, b_mn, bn, out_bf16):
    if b_mn and n % 8:
        n = (n + 7) // 8 * 8
    torch.manual_seed(m * 1000003 + n * 1009 + k)
    A = torch.randn((k, m) if a_mn else (m, k), device="cuda").to(bf16)
    Bm = torch.randn((k, n) if b_mn else (n, k), device="cuda").to(bf16)
    ref = (A.float().t() if a_mn else A.float()) @ (Bm.float() if b_mn else Bm.float().t())
    C = torch.full((m, n), float("nan"), device="cuda", dtype=bf16 if out_bf16 else torch.float32)
    K.gemm_tc(A, Bm, C, m, n, k, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn)
    err = (C.float() - ref).abs().max().item()
    tol = (2e-2 * ref.abs().max().item() + 1e-2) if out_bf16 else 2e-3 * (k ** 0.5)
    assert err <= tol, (m, n, k, a_mn, b_mn, bn, err, tol)


@settings(**COMMON)
@given(B=st.integers(1, 9), H=st.sampled_from([8, 10, 12, 14]), k=st.sampled_from([3, 5]), halo=st.booleans())
def test_conv_tc_random_shapes(B, H, k, halo, monkeypatch):
    monkeypatch.setenv("B200DP_CONV_HALO", "1" if halo else "0")
    Ci, Co = 32, 64
    torch.manual_seed(B * 131 + H * 7 + k)
    x = torch.randn(B, H, H, Ci, device="cuda").to(bf16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(bf16)
    b = (torch.randn(Co, device="cuda") * 0.1).to(bf16)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.float().permute(3, 2, 0, 1).requires_grad_(True)
    pre = F.conv2d(xr, wr, b.float(), padding=k // 2)
    out = torch.empty(B, H // 2, H // 2, Co, device="cuda", dtype=bf16)
    am = torch.empty(B, H // 2, H // 2, Co, device="cuda", dtype=torch.uint8)
    K.conv_tc_fwd(x, w, b, out, am, relu=True, pool=True)
    ref = F.max_pool2d(F.relu(pre), 2).permute(0, 2, 3, 1)
    assert (out.float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item() + 2e-2
    dy = torch.randn(B, H, H, Co, device="cuda").to(bf16)
    pre.backward(dy.float().permute(0, 3, 1, 2))
    din = torch.empty(B, H, H, Ci, device="cuda", dtype=bf16)
    K.conv_tc_dgrad(dy, w, din)
    gx = xr.grad.permute(0, 2, 3, 1)
    assert (din.float() - gx).abs().max().item() <= 2e-2 * gx.abs().max().item() + 2e-2
    dw = torch.zeros(k, k, Ci, Co, device="cuda")
    K.conv_tc_wgrad(x, dy, dw, None)
    gw = wr.grad.permute(2, 3, 1, 0)
    assert (dw - gw).abs().max().item() <= 5e-3 * gw.abs().max().item() + 2e-2
