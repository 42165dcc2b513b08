"""tcgen05 GEMM vs fp32 PyTorch on the same bf16-rounded operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200dp.ops import kernels as K

bf16 = torch.bfloat16


def _ref(A, Bm, a_mn, b_mn):
    a = A.float().t() if a_mn else A.float()
    b = Bm.float() if b_mn else Bm.float().t()
    return a @ b


def _mk(M, N, Kd, a_mn, b_mn, scale=1.0):
    A = (torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda") * scale).to(bf16)
    Bm = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda") * scale).to(bf16)
    return A, Bm


def _check(C, ref, Kd):
    err = (C.float() - ref).abs().max().item()
    tol = 2e-2 * ref.abs().max().item() + 1e-3 if C.dtype == bf16 else 1e-3 * (Kd ** 0.5)
    assert err <= tol, f"max err {err} (tol {tol})"


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,Kd,bn", [(128, 128, 64, 128), (128, 1024, 3136, 64), (3136, 1024, 128, 128), (104, 192, 200, 64), (256, 256, 512, 256)])
def test_gemm_tc_all_majors(a_mn, b_mn, M, N, Kd, bn):
    torch.manual_seed(M + N + Kd)
    A, Bm = _mk(M, N, Kd, a_mn, b_mn)
    ref = _ref(A, Bm, a_mn, b_mn)
    C = torch.full((M, N), float("nan"), device="cuda")
    K.gemm_tc(A, Bm, C, M, N, Kd, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn)
    torch.cuda.synchronize()
    _check(C, ref, Kd)


@pytest.mark.parametrize("bn", [16, 32, 48, 96])
def test_gemm_tc_small_block_n_k_major_b(bn):
    torch.manual_seed(bn)
    M, N, Kd = 128, 3136, 1024            # dense dgrad shape: dx = dz . W^T
    A, Bm = _mk(M, N, Kd, False, False, 0.5)
    C = torch.empty(M, N, device="cuda", dtype=bf16)
    K.gemm_tc(A, Bm, C, M, N, Kd, block_n=bn)
    _check(C, _ref(A, Bm, False, False), Kd)


@pytest.mark.parametrize("split", [2, 7, 49])
def test_gemm_tc_split_k_atomic(split):
    torch.manual_seed(split)
    M, N, Kd = 128, 1024, 3136            # dense forward shape at batch 128
    A, Bm = _mk(M, N, Kd, False, True, 0.5)
    C = torch.zeros(M, N, device="cuda")
    K.gemm_tc(A, Bm, C, M, N, Kd, b_mn_major=True, out_mode=2, split_k=split, block_n=64)
    _check(C, _ref(A, Bm, False, True), Kd)
    K.gemm_tc(A, Bm, C, M, N, Kd, b_mn_major=True, out_mode=2, split_k=split, block_n=64)
    _check(C, 2 * _ref(A, Bm, False, True), Kd)


def test_gemm_tc_argument_errors():
    A = torch.zeros(128, 64, device="cuda", dtype=bf16)
    C = torch.zeros(128, 128, device="cuda")
    with pytest.raises(K.B200Error):
        K.gemm_tc(A, A, C, 128, 128, 64, split_k=2, out_mode=0)
    with pytest.raises(K.B200Error):
        K.gemm_tc(A, A, C, 128, 128, 64, block_n=20)
    At = torch.zeros(64, 100, device="cuda", dtype=bf16)         # MN-major A with a pitch that is not 16-byte aligned
    with pytest.raises(K.B200Error):
        K.gemm_tc(At, A, C, 100, 128, 64, a_mn_major=True)
