"""CUDA-core kernels vs plain PyTorch fp32 references (SURVEY §4.3 'Kernel (GPU)')."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from b200dp.ops import kernels as K
from b200dp.utils.philox import dropout_keep_mask

bf16 = torch.bfloat16


def q(t):
    return t.to(bf16).float()


def close(a, b, rtol=2e-2, atol=2e-2):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item()
    ref = b.abs().max().item()
    assert err <= atol + rtol * ref, f"max err {err} vs ref scale {ref}"


@pytest.mark.parametrize("B,C1", [(1, 32), (100, 32), (128, 32), (5, 64), (3, 192)])
def test_conv1_pool_fwd_and_wgrad(B, C1):
    torch.manual_seed(0)
    x = torch.rand(B, 784, device="cuda")
    w = (torch.randn(5, 5, 1, C1, device="cuda") * 0.2).to(bf16)
    b = (torch.randn(C1, device="cuda") * 0.1).to(bf16)
    out = torch.empty(B, 14, 14, C1, device="cuda", dtype=bf16)
    am = torch.empty(B, 14, 14, C1, device="cuda", dtype=torch.uint8)
    step = torch.zeros(1, device="cuda", dtype=torch.int32)
    K.conv1_pool_fwd(x, w, b, out, am, step, bump_step=True)
    assert step.item() == 1
    xr = x.view(B, 1, 28, 28).clone().requires_grad_(False)
    wr = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    br = b.float().clone().requires_grad_(True)
    conv = F.relu(F.conv2d(xr, wr, br, padding=2))
    ref = F.max_pool2d(conv, 2).permute(0, 2, 3, 1)
    close(out, ref, 1e-2, 1e-2)
    # backward: weight gradient from a random pooled-output gradient
    da1 = (torch.randn(B, 14, 14, C1, device="cuda") * 0.1).to(bf16)
    # reference uses the *kernel's* bf16 activations for the ReLU mask, like the engine does
    (F.max_pool2d(conv, 2) * da1.float().permute(0, 3, 1, 2)).sum().backward()
    dw = torch.zeros(5, 5, 1, C1, device="cuda")
    db = torch.zeros(C1, device="cuda")
    K.conv1_wgrad(x, da1, out, am, dw, db)
    close(dw, wr.grad.permute(2, 3, 1, 0), 2e-2, 1e-3 * B)
    close(db, br.grad, 2e-2, 1e-3 * B)


@pytest.mark.parametrize("B,D,training,rate", [(128, 1024, True, 0.5), (1024, 1024, True, 0.5), (100, 1024, True, 0.5), (7, 256, True, 0.0), (64, 1024, False, 0.5), (4, 6144, True, 0.25)])
def test_head_fwd_bwd_and_wgrad(B, D, training, rate):
    torch.manual_seed(1)
    NC = 10
    z = torch.randn(B, D, device="cuda")
    z0 = z.clone()
    b1 = (torch.randn(D, device="cuda") * 0.1).to(bf16)
    w2 = (torch.randn(D, NC, device="cuda") * 0.05).to(bf16)
    b2 = (torch.randn(NC, device="cuda") * 0.1).to(bf16)
    y = torch.randint(0, NC, (B,), device="cuda", dtype=torch.int32)
    h = torch.empty(B, D, device="cuda", dtype=bf16)
    dz = torch.empty(B, D, device="cuda", dtype=bf16)
    dl = torch.empty(B, NC, device="cuda")
    lg = torch.empty(B, NC, device="cuda")
    loss = torch.zeros(1, device="cuda")
    corr = torch.zeros(1, device="cuda", dtype=torch.int32)
    step = torch.full((1,), 3, device="cuda", dtype=torch.int32)
    seed = 0x1234_5678_9ABC
    K.head_fwd_bwd(z, b1, w2, b2, y, h, dz, dl, lg, loss, corr, drop_rate=rate, training=training, seed=seed, step_ptr=step)
    assert torch.count_nonzero(z).item() == 0                      # split-K accumulator cleared
    zr = z0.clone().requires_grad_(True)
    hr = F.relu(zr + b1.float())
    if training and rate > 0:
        keep = dropout_keep_mask(B * D, rate, seed, 3, 0).view(B, D).cuda()
        hr = hr * keep / (1 - rate)
    hq = hr + (q(hr) - hr).detach()                                # bf16 rounding, straight-through
    w2r = w2.float().clone().requires_grad_(True)
    logits = hq @ w2r + b2.float()
    ref_loss = F.cross_entropy(logits, y.long(), reduction="sum")
    close(h, hr, 1e-2, 1e-2)
    close(lg, logits, 1e-2, 2e-2)
    assert abs(loss.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item()) + 1e-2
    assert corr.item() == (logits.argmax(1) == y.long()).sum().item()
    if not training:
        return
    (ref_loss / B).backward()
    close(dz, zr.grad, 2e-2, 1e-4)
    dw2 = torch.empty(D, NC, device="cuda"); db2 = torch.empty(NC, device="cuda"); db1 = torch.empty(D, device="cuda")
    K.head_wgrad(h, dz, dl, dw2, db2, db1)
    close(dw2, w2r.grad, 2e-2, 1e-4)
    close(db1, dz.float().sum(0), 1e-3, 1e-5)
    close(db2, dl.sum(0), 1e-4, 1e-6)


def test_pool_and_unpool():
    torch.manual_seed(2)
    B, H, W, Cc = 9, 14, 14, 64
    x = F.relu(torch.randn(B, H, W, Cc, device="cuda")).to(bf16)
    out = torch.empty(B, 7, 7, Cc, device="cuda", dtype=bf16)
    am = torch.empty(B, 7, 7, Cc, device="cuda", dtype=torch.uint8)
    K.maxpool_fwd(x, out, am)
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    ref = F.max_pool2d(xr, 2)
    assert torch.equal(out.float(), ref.permute(0, 2, 3, 1))
    dp = torch.randn(B, 7, 7, Cc, device="cuda").to(bf16)
    dy = torch.empty(B, H, W, Cc, device="cuda", dtype=bf16)
    db = torch.zeros(Cc, device="cuda")
    K.unpool_relu_bwd(dp, out, am, dy, db)
    (ref * dp.float().permute(0, 3, 1, 2)).sum().backward()
    mask = (x.float() > 0)                                         # relu' on the pre-pool activation
    want = (xr.grad.permute(0, 2, 3, 1)) * mask
    # ties (exact zeros) carry no gradient in either implementation because of the relu mask
    assert torch.allclose(dy.float(), want.to(bf16).float())
    assert torch.allclose(db, dy.float().sum((0, 1, 2)), rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("k,pad,Ci,Co,H", [(5, 2, 32, 64, 14), (3, 0, 8, 16, 12), (3, 1, 16, 8, 9)])
def test_direct_convs(k, pad, Ci, Co, H):
    torch.manual_seed(3)
    B = 4
    x = torch.randn(B, H, H, Ci, device="cuda").to(bf16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.1).to(bf16)
    b = (torch.randn(Co, device="cuda") * 0.1).to(bf16)
    Ho = H + 2 * pad - k + 1
    out = torch.empty(B, Ho, Ho, Co, device="cuda", dtype=bf16)
    K.conv_direct_fwd(x, w, b, out, pad, relu=True)
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    wr = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    pre = F.conv2d(xr, wr, b.float(), padding=pad)
    close(out, F.relu(pre).permute(0, 2, 3, 1), 1e-2, 1e-2)
    dout = torch.randn(B, Ho, Ho, Co, device="cuda").to(bf16)
    pre.backward(dout.float().permute(0, 3, 1, 2))
    din = torch.empty_like(x)
    K.conv_direct_dgrad(dout, w, din, pad)
    close(din, xr.grad.permute(0, 2, 3, 1), 1e-2, 1e-2)
    dw = torch.zeros(k, k, Ci, Co, device="cuda"); db = torch.zeros(Co, device="cuda")
    K.conv_direct_wgrad(x, dout, dw, db, pad)
    close(dw, wr.grad.permute(2, 3, 1, 0), 1e-3, 1e-3)
    close(db, dout.float().sum((0, 1, 2)), 1e-3, 1e-3)


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm_simple(ta, tb):
    torch.manual_seed(4)
    M, N, Kd = 100, 70, 130
    A = torch.randn((Kd, M) if ta else (M, Kd), device="cuda").to(bf16)
    Bm = torch.randn((N, Kd) if tb else (Kd, N), device="cuda").to(bf16)
    ref = (A.float().t() if ta else A.float()) @ (Bm.float().t() if tb else Bm.float())
    C = torch.empty(M, N, device="cuda")
    K.gemm_simple(A, Bm, C, M, N, Kd, transA=ta, transB=tb)
    close(C, ref, 1e-4, 1e-3)
    K.gemm_simple(A, Bm, C, M, N, Kd, transA=ta, transB=tb, accumulate=True)
    close(C, 2 * ref, 1e-4, 1e-3)
    Cb = torch.empty(M, N, device="cuda", dtype=bf16)
    K.gemm_simple(A, Bm, Cb, M, N, Kd, transA=ta, transB=tb)
    close(Cb, ref, 1e-2, 1e-1)


@pytest.mark.parametrize("mode", ["tf", "torch"])
def test_adam_local_matches_oracle(mode):
    from b200dp.models.torch_ref import adam_step_
    torch.manual_seed(5)
    n = 10_000
    p = torch.randn(n, device="cuda"); g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    w16 = torch.empty(n, device="cuda", dtype=bf16)
    pr, mr, vr = p.clone(), m.clone(), v.clone()
    for step in (1, 2, 3):
        K.adam_local(p, m, v, g, w16, step, 1e-3, 0.9, 0.999, 1e-8, mode, grad_scale=0.5)
        adam_step_(pr, g * 0.5, mr, vr, step, lr=1e-3, eps_mode=mode)
    assert torch.allclose(p, pr, rtol=1e-5, atol=1e-6) and torch.allclose(m, mr, rtol=1e-5, atol=1e-7)
    assert torch.equal(w16, p.to(bf16))


def test_topology_reports_devices():
    t = K.topology()
    assert len(t["gpus"]) == torch.cuda.device_count() and t["gpus"][0]["sms"] > 0
    assert t["peer_access"][0][0] == 1


# ---- keras128 family: VALID first layer, feature dropout, dropout-aware unpool --------------------------------
@pytest.mark.parametrize("B,C1,k", [(1, 32, 3), (128, 32, 3), (100, 32, 3), (5, 64, 3), (3, 192, 3), (4, 32, 5)])
def test_conv1_valid_fwd_and_wgrad(B, C1, k):
    torch.manual_seed(3)
    Ho = 28 - k + 1
    x = torch.rand(B, 784, device="cuda")
    w = (torch.randn(k, k, 1, C1, device="cuda") * 0.2).to(bf16)
    b = (torch.randn(C1, device="cuda") * 0.1).to(bf16)
    out = torch.full((B, Ho, Ho, C1), float("nan"), device="cuda", dtype=bf16)
    step = torch.zeros(1, device="cuda", dtype=torch.int32)
    ring_l = torch.ones(8, device="cuda")
    ring_c = torch.ones(8, device="cuda", dtype=torch.int32)
    K.conv1_valid_fwd(x, w, b, out, step, bump_step=True, loss_ring=ring_l, correct_ring=ring_c)
    assert step.item() == 1 and ring_l[1].item() == 0 and ring_c[1].item() == 0 and ring_l[2].item() == 1
    wr = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    br = b.float().clone().requires_grad_(True)
    conv = F.relu(F.conv2d(x.view(B, 1, 28, 28), wr, br))
    close(out, conv.permute(0, 2, 3, 1), 1e-2, 1e-2)
    da1 = (torch.randn(B, Ho, Ho, C1, device="cuda") * 0.1).to(bf16)
    # gradient through the ReLU mask of the kernel's own bf16 activations
    pre = F.conv2d(x.view(B, 1, 28, 28), wr, br)
    mask = (out.float() > 0).permute(0, 3, 1, 2)
    (pre * mask * da1.float().permute(0, 3, 1, 2)).sum().backward()
    dw = torch.zeros(k, k, 1, C1, device="cuda")
    db = torch.zeros(C1, device="cuda")
    K.conv1_valid_wgrad(x, da1, out, dw, db)
    close(dw, wr.grad.permute(2, 3, 1, 0), 2e-2, 1e-3 * B)
    close(db, br.grad, 2e-2, 1e-3 * B)


@pytest.mark.parametrize("n,rate,layer", [(128 * 9216, 0.25, 0), (4 * 12 * 12 * 64, 0.25, 0), (1000, 0.5, 3), (64, 0.0, 1)])
def test_dropout_bf16_matches_host_philox(n, rate, layer):
    torch.manual_seed(4)
    x = torch.randn(n, device="cuda").to(bf16)
    step = torch.tensor([7], device="cuda", dtype=torch.int32)
    out = torch.empty_like(x)
    K.dropout_bf16(x, out, rate, 1234, step, layer)
    keep = dropout_keep_mask(n, rate, 1234, 7, layer).cuda()
    import numpy as np
    sc = float(np.float32(1.0) / (np.float32(1.0) - np.float32(rate)))         # the kernel's fp32 keep-scale
    ref = torch.where(keep, x.float() * sc, torch.zeros((), device="cuda")).to(bf16)
    assert torch.equal(out, ref)
    K.dropout_bf16(x, x, rate, 1234, step, layer)          # in place
    assert torch.equal(x, ref)


def test_unpool_relu_bwd_scaled():
    torch.manual_seed(5)
    B, HP, C = 6, 12, 64
    pooled = torch.relu(torch.randn(B, HP, HP, C, device="cuda")).to(bf16)
    pooled[torch.rand_like(pooled.float()) < 0.25] = 0          # dropped entries
    am = torch.randint(0, 4, (B, HP, HP, C), device="cuda", dtype=torch.uint8)
    dpool = torch.randn(B, HP, HP, C, device="cuda").to(bf16)
    dy = torch.full((B, 2 * HP, 2 * HP, C), float("nan"), device="cuda", dtype=bf16)
    db = torch.zeros(C, device="cuda")
    sc = 1.0 / 0.75
    K.unpool_relu_bwd(dpool, pooled, am, dy, db, gscale=sc)
    g = torch.where(pooled.float() > 0, dpool.float() * sc, torch.zeros((), device="cuda")).to(bf16).float()
    ref = torch.zeros(B, HP, 2, HP, 2, C, device="cuda")
    for qd in range(4):
        ref[:, :, qd >> 1, :, qd & 1, :] = torch.where(am == qd, g, torch.zeros((), device="cuda"))
    assert torch.equal(dy.float(), ref.view(B, 2 * HP, 2 * HP, C))
    close(db, torch.where(pooled.float() > 0, dpool.float() * sc, torch.zeros((), device="cuda")).sum((0, 1, 2)), 1e-2, 1e-2)
