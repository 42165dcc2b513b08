"""tcgen05 implicit-GEMM convolution (fwd / dgrad / wgrad) vs fp32 PyTorch on bf16-rounded operands."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from b200dp.ops import kernels as K

bf16 = torch.bfloat16


@pytest.fixture(params=["halo", "resident", "tma_shift"], autouse=True)
def conv_path(request, monkeypatch):
    """Run every case through the three tensor-core paths: the smem-resident persistent kernel for small SAME images
    (conv_halo.cu), the general smem-resident kernel (conv_res.cu) and the shifted-TMA-box kernel (conv_tc.cu).  Each
    specialised kernel hands shapes it does not cover down the chain."""
    monkeypatch.setenv("B200DP_CONV_HALO", "1" if request.param == "halo" else "0")
    monkeypatch.setenv("B200DP_CONV_RES", "0" if request.param == "tma_shift" else "1")
    return request.param


def _err(a, b):
    return (a.float() - b.float()).abs().max().item(), b.float().abs().max().item()


def _setup(B, H, Ci, Co, k, seed):
    torch.manual_seed(seed)
    x = torch.randn(B, H, H, Ci, device="cuda").to(bf16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(bf16)
    b = (torch.randn(Co, device="cuda") * 0.1).to(bf16)
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    wr = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    return x, w, b, xr, wr


@pytest.mark.parametrize("B,H,Ci,Co,k", [(2, 14, 32, 64, 5), (128, 14, 32, 64, 5), (3, 14, 64, 128, 5), (2, 14, 192, 384, 5), (4, 12, 32, 64, 3), (100, 14, 32, 64, 5), (700, 14, 32, 64, 5)])
def test_conv_tc_fwd_pool(B, H, Ci, Co, k):
    x, w, b, xr, wr = _setup(B, H, Ci, Co, k, 1)
    pre = F.relu(F.conv2d(xr, wr, b.float(), padding=k // 2))
    out = torch.full((B, H, H, Co), float("nan"), device="cuda", dtype=bf16)
    K.conv_tc_fwd(x, w, b, out, None, relu=True, pool=False)
    e, s = _err(out, pre.permute(0, 2, 3, 1))
    assert e <= 2e-2 * s + 2e-2, f"unpooled: err {e} scale {s}"
    pooled = torch.full((B, H // 2, H // 2, Co), float("nan"), device="cuda", dtype=bf16)
    am = torch.full((B, H // 2, H // 2, Co), 255, device="cuda", dtype=torch.uint8)
    K.conv_tc_fwd(x, w, b, pooled, am, relu=True, pool=True)
    ref = F.max_pool2d(pre, 2).permute(0, 2, 3, 1)
    e, s = _err(pooled, ref)
    assert e <= 2e-2 * s + 2e-2, f"pooled: err {e} scale {s}"
    assert am.max().item() <= 3
    # the recorded argmax must point at the window element that equals the pooled value
    win = out.float().view(B, H // 2, 2, H // 2, 2, Co).permute(0, 1, 3, 5, 2, 4).reshape(B, H // 2, H // 2, Co, 4)
    picked = torch.gather(win, 4, am.long().unsqueeze(-1)).squeeze(-1)
    assert torch.equal(picked, pooled.float())


@pytest.mark.parametrize("B,H,Ci,Co,k", [(2, 14, 32, 64, 5), (128, 14, 32, 64, 5), (3, 14, 64, 128, 5), (2, 14, 192, 384, 5), (4, 12, 32, 64, 3), (700, 14, 32, 64, 5)])
def test_conv_tc_dgrad_wgrad(B, H, Ci, Co, k):
    x, w, b, xr, wr = _setup(B, H, Ci, Co, k, 2)
    dy = torch.randn(B, H, H, Co, device="cuda").to(bf16)
    F.conv2d(xr, wr, None, padding=k // 2).backward(dy.float().permute(0, 3, 1, 2))
    din = torch.full((B, H, H, Ci), float("nan"), device="cuda", dtype=bf16)
    K.conv_tc_dgrad(dy, w, din)
    e, s = _err(din, xr.grad.permute(0, 2, 3, 1))
    assert e <= 2e-2 * s + 2e-2, f"dgrad: err {e} scale {s}"
    dw = torch.zeros(k, k, Ci, Co, device="cuda")
    db = torch.zeros(Co, device="cuda")
    K.conv_tc_wgrad(x, dy, dw, db)
    e, s = _err(dw, wr.grad.permute(2, 3, 1, 0))
    assert e <= 5e-3 * s + 1e-2, f"wgrad: err {e} scale {s}"
    e, s = _err(db, dy.float().sum((0, 1, 2)))
    assert e <= 1e-3 * s + 1e-2


# ---- explicit padding (VALID) and images wider than one 16-pixel tile: the keras128 second layer and friends ----
@pytest.mark.parametrize("B,H,Ci,Co,k,pad", [(2, 26, 32, 64, 3, 0), (128, 26, 32, 64, 3, 0), (3, 20, 64, 128, 5, 0),
                                             (2, 28, 32, 64, 3, 1), (5, 18, 32, 64, 3, 2)])
def test_conv_tc_padded_fwd(B, H, Ci, Co, k, pad):
    x, w, b, xr, wr = _setup(B, H, Ci, Co, k, 5)
    Ho = H + 2 * pad - k + 1
    pre = F.relu(F.conv2d(xr, wr, b.float(), padding=pad))
    out = torch.full((B, Ho, Ho, Co), float("nan"), device="cuda", dtype=bf16)
    K.conv_tc_fwd(x, w, b, out, None, relu=True, pool=False, pad=pad)
    e, s = _err(out, pre.permute(0, 2, 3, 1))
    assert e <= 2e-2 * s + 2e-2, f"unpooled: err {e} scale {s}"
    if Ho % 2 == 0:
        pooled = torch.full((B, Ho // 2, Ho // 2, Co), float("nan"), device="cuda", dtype=bf16)
        am = torch.full((B, Ho // 2, Ho // 2, Co), 255, device="cuda", dtype=torch.uint8)
        K.conv_tc_fwd(x, w, b, pooled, am, relu=True, pool=True, pad=pad)
        e, s = _err(pooled, F.max_pool2d(pre, 2).permute(0, 2, 3, 1))
        assert e <= 2e-2 * s + 2e-2, f"pooled: err {e} scale {s}"
        win = out.float().view(B, Ho // 2, 2, Ho // 2, 2, Co).permute(0, 1, 3, 5, 2, 4).reshape(B, Ho // 2, Ho // 2, Co, 4)
        picked = torch.gather(win, 4, am.long().unsqueeze(-1)).squeeze(-1)
        assert torch.equal(picked, pooled.float())


@pytest.mark.parametrize("B,H,Ci,Co,k,pad", [(2, 26, 32, 64, 3, 0), (128, 26, 32, 64, 3, 0), (3, 20, 64, 128, 5, 0),
                                             (2, 28, 32, 64, 3, 1), (5, 18, 32, 64, 3, 2)])
def test_conv_tc_padded_dgrad_wgrad(B, H, Ci, Co, k, pad):
    x, w, b, xr, wr = _setup(B, H, Ci, Co, k, 6)
    Ho = H + 2 * pad - k + 1
    dy = torch.randn(B, Ho, Ho, Co, device="cuda").to(bf16)
    F.conv2d(xr, wr, None, padding=pad).backward(dy.float().permute(0, 3, 1, 2))
    din = torch.full((B, H, H, Ci), float("nan"), device="cuda", dtype=bf16)
    K.conv_tc_dgrad(dy, w, din, pad=pad)
    e, s = _err(din, xr.grad.permute(0, 2, 3, 1))
    assert e <= 2e-2 * s + 2e-2, f"dgrad: err {e} scale {s}"
    dw = torch.zeros(k, k, Ci, Co, device="cuda")
    db = torch.zeros(Co, device="cuda")
    K.conv_tc_wgrad(x, dy, dw, db, pad=pad)
    ref = wr.grad.permute(2, 3, 1, 0)
    e, s = _err(dw, ref)
    assert e <= 1e-2 * s + 1e-2, f"wgrad: err {e} scale {s}"
    e, s = _err(db, dy.float().sum((0, 1, 2)))
    assert e <= 1e-2 * s + 1e-2, f"dbias: err {e} scale {s}"
