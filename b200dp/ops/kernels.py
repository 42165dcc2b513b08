"""Thin typed wrappers: torch tensors in, raw pointers + current CUDA stream into libb200dp.

No PyTorch fallbacks here — if the library is missing on a GPU box these raise, loudly.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib

bf16 = torch.bfloat16


class B200Error(RuntimeError):
    pass


def lib():
    return _lib.load()


LAUNCHES = [0]        # entry-point calls that launched kernels (each launches >= 1 kernel of ours)


def check(rc: int, what: str = "") -> None:
    LAUNCHES[0] += 1
    if rc != 0:
        msg = lib().b200_last_error().decode(errors="replace")
        raise B200Error(f"{what or 'libb200dp'} failed (code {rc}): {msg}")


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _chk(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise B200Error(f"{name}: need contiguous CUDA {dtype}, got {t.dtype} cuda={t.is_cuda} contiguous={t.is_contiguous()}")
    return t


# ------------------------------------------------------------------------------- CUDA-core family
def conv1_pool_fwd(x, w, bias, out, argmax, step_ptr: Optional[torch.Tensor] = None, bump_step: bool = False,
                   loss_ring: Optional[torch.Tensor] = None, correct_ring: Optional[torch.Tensor] = None):
    B, C1 = x.shape[0], w.shape[-1]
    _chk(x, torch.float32, "x"); _chk(w, bf16, "w"); _chk(bias, bf16, "bias"); _chk(out, bf16, "out"); _chk(argmax, torch.uint8, "argmax")
    check(lib().b200_conv1_pool_fwd(_p(x), _p(w), _p(bias), _p(out), _p(argmax), B, C1, _p(step_ptr), int(bump_step), _p(loss_ring), _p(correct_ring),
                                    0 if loss_ring is None else loss_ring.numel(), _stream()), "conv1_pool_fwd")


def conv1_wgrad(x, da1, a1, argmax, dw, dbias):
    B, C1 = x.shape[0], a1.shape[-1]
    check(lib().b200_conv1_wgrad(_p(_chk(x, torch.float32, "x")), _p(_chk(da1, bf16, "da1")), _p(_chk(a1, bf16, "a1")),
                                 _p(argmax), _p(_chk(dw, torch.float32, "dw")), _p(_chk(dbias, torch.float32, "dbias")), B, C1, _stream()),
          "conv1_wgrad")


def head_fwd_bwd(z, b1, w2, b2, labels, h, dz, dlogits, logits, loss_sum, correct, *, drop_rate: float, training: bool,
                 seed: int, step_ptr: Optional[torch.Tensor], layer_id: int = 0, zero_z: bool = True, ring: int = 0):
    B, D = z.shape
    NC = w2.shape[1]
    _chk(z, torch.float32, "z"); _chk(labels, torch.int32, "labels")
    check(lib().b200_head_fwd_bwd(_p(z), _p(_chk(b1, bf16, "b1")), _p(_chk(w2, bf16, "w2")), _p(_chk(b2, bf16, "b2")), _p(labels),
                                  _p(h), _p(dz), _p(dlogits), _p(logits), _p(loss_sum), _p(correct), B, D, NC,
                                  float(drop_rate), int(training), ctypes.c_uint64(seed & (2 ** 64 - 1)), _p(step_ptr), layer_id,
                                  int(zero_z), int(ring), _stream()), "head_fwd_bwd")


def head_wgrad(h, dz, dlogits, dw2, db2, db1):
    B, D = h.shape
    NC = dlogits.shape[1]
    check(lib().b200_head_wgrad(_p(_chk(h, bf16, "h")), _p(_chk(dz, bf16, "dz")), _p(_chk(dlogits, torch.float32, "dlogits")),
                                _p(dw2), _p(db2), _p(db1), B, D, NC, _stream()), "head_wgrad")


def unpool_relu_bwd(dpool, pooled, argmax, dy, dbias=None, gscale: float = 1.0):
    B, HP, WP, Cc = pooled.shape
    check(lib().b200_unpool_relu_bwd_scaled(_p(_chk(dpool, bf16, "dpool")), _p(_chk(pooled, bf16, "pooled")), _p(argmax),
                                            _p(_chk(dy, bf16, "dy")), _p(dbias), B, HP, WP, Cc, float(gscale), _stream()), "unpool_relu_bwd")


def conv1_valid_fwd(x, w, bias, out, step_ptr: Optional[torch.Tensor] = None, bump_step: bool = False,
                    loss_ring: Optional[torch.Tensor] = None, correct_ring: Optional[torch.Tensor] = None):
    """keras128 first layer: k x k VALID conv of the 28x28 image + bias + ReLU -> [B, 28-k+1, 28-k+1, C1] bf16."""
    B, k, C1 = x.shape[0], w.shape[0], w.shape[-1]
    _chk(x, torch.float32, "x"); _chk(w, bf16, "w"); _chk(bias, bf16, "bias"); _chk(out, bf16, "out")
    check(lib().b200_conv1_valid_fwd(_p(x), _p(w), _p(bias), _p(out), B, C1, k, _p(step_ptr), int(bump_step), _p(loss_ring),
                                     _p(correct_ring), 0 if loss_ring is None else loss_ring.numel(), _stream()), "conv1_valid_fwd")


def conv1_valid_wgrad(x, da1, a1, dw, dbias):
    B, k, C1 = x.shape[0], dw.shape[0], a1.shape[-1]
    check(lib().b200_conv1_valid_wgrad(_p(_chk(x, torch.float32, "x")), _p(_chk(da1, bf16, "da1")), _p(_chk(a1, bf16, "a1")),
                                       _p(_chk(dw, torch.float32, "dw")), _p(_chk(dbias, torch.float32, "dbias")), B, C1, k, _stream()),
          "conv1_valid_wgrad")


def dropout_bf16(inp, out, rate: float, seed: int, step_ptr: Optional[torch.Tensor], layer_id: int):
    check(lib().b200_dropout_bf16(_p(_chk(inp, bf16, "in")), _p(_chk(out, bf16, "out")), inp.numel(), float(rate),
                                  ctypes.c_uint64(seed & (2 ** 64 - 1)), _p(step_ptr), layer_id, _stream()), "dropout_bf16")


def maxpool_fwd(inp, out, argmax):
    B, H, W, Cc = inp.shape
    check(lib().b200_maxpool_fwd(_p(_chk(inp, bf16, "in")), _p(_chk(out, bf16, "out")), _p(argmax), B, H, W, Cc, _stream()), "maxpool_fwd")


def conv_direct_fwd(inp, w, bias, out, pad: int, relu: bool):
    B, H, W, Ci = inp.shape
    k, Co = w.shape[0], w.shape[3]
    check(lib().b200_conv_direct_fwd(_p(_chk(inp, bf16, "in")), _p(_chk(w, bf16, "w")), _p(bias), _p(_chk(out, bf16, "out")),
                                     B, H, W, Ci, Co, k, pad, int(relu), _stream()), "conv_direct_fwd")


def conv_direct_dgrad(dout, w, din, pad: int):
    B, H, W, Ci = din.shape
    k, Co = w.shape[0], w.shape[3]
    check(lib().b200_conv_direct_dgrad(_p(_chk(dout, bf16, "dout")), _p(_chk(w, bf16, "w")), _p(_chk(din, bf16, "din")),
                                       B, H, W, Ci, Co, k, pad, _stream()), "conv_direct_dgrad")


def conv_direct_wgrad(inp, dout, dw, dbias, pad: int):
    B, H, W, Ci = inp.shape
    k, Co = dw.shape[0], dw.shape[3]
    check(lib().b200_conv_direct_wgrad(_p(_chk(inp, bf16, "in")), _p(_chk(dout, bf16, "dout")), _p(_chk(dw, torch.float32, "dw")),
                                       _p(dbias), B, H, W, Ci, Co, k, pad, _stream()), "conv_direct_wgrad")


def gemm_simple(A, Bm, Cm, M, N, K, *, transA=False, transB=False, accumulate=False):
    check(lib().b200_gemm_simple(_p(_chk(A, bf16, "A")), _p(_chk(Bm, bf16, "B")), _p(Cm), M, N, K, int(transA), int(transB),
                                 int(accumulate), int(Cm.dtype == bf16), _stream()), "gemm_simple")


def cast_f32_to_bf16(src, dst):
    check(lib().b200_cast_f32_to_bf16(_p(_chk(src, torch.float32, "src")), _p(_chk(dst, bf16, "dst")), src.numel(), _stream()), "cast")


def bias_relu_bf16(z, bias, out, relu=True):
    rows, cols = z.shape
    check(lib().b200_bias_relu_bf16(_p(_chk(z, torch.float32, "z")), _p(bias), _p(_chk(out, bf16, "out")), rows, cols, int(relu), _stream()),
          "bias_relu")


def adam_local(master, m, v, grad, w16, step, lr, b1, b2, eps, eps_mode="tf", grad_scale=1.0):
    check(lib().b200_adam_local(_p(master), _p(m), _p(v), _p(grad), _p(w16), master.numel(), int(step), lr, b1, b2, eps,
                                0 if eps_mode == "tf" else 1, float(grad_scale), _stream()), "adam_local")


def l2_flush(scratch: torch.Tensor):
    check(lib().b200_l2_flush(_p(scratch), scratch.numel() * scratch.element_size(), _stream()), "l2_flush")


# ------------------------------------------------------------------------------- tcgen05 family
def gemm_tc(A, Bm, Cm, M, N, K, *, a_mn_major=False, b_mn_major=False, out_mode: Optional[int] = None, split_k=1,
            block_n=128, lda=None, ldb=None, ldc=None):
    """D[M,N] = A.B on tensor cores.  ``A`` is stored [M,K] (or [K,M] if a_mn_major), ``Bm`` is stored [N,K]
    (or [K,N] if b_mn_major).  out_mode: 0 fp32 store, 1 bf16 store, 2 fp32 atomic-add."""
    if out_mode is None:
        out_mode = 1 if Cm.dtype == bf16 else 0
    lda = lda if lda is not None else A.stride(0)
    ldb = ldb if ldb is not None else Bm.stride(0)
    ldc = ldc if ldc is not None else Cm.stride(0)
    check(lib().b200_gemm_tc(_p(_chk(A, bf16, "A")), _p(_chk(Bm, bf16, "B")), _p(Cm), M, N, K, lda, ldb, ldc, int(a_mn_major),
                             int(b_mn_major), out_mode, split_k, block_n, _stream()), "gemm_tc")


def gemm_tc_reduce_scatter(comm_handle, A, Bm, c_seg: int, c_off: int, nranks: int, M, N, K, *, a_mn_major=False,
                           b_mn_major=False, block_n=128, lda=None, ldb=None, ldc=None):
    """D = A.B red.add'ed tile by tile into the owner ranks' fp32 buffers at (c_seg, c_off); rank r owns rows r*M/n..."""
    lda = lda if lda is not None else A.stride(0)
    ldb = ldb if ldb is not None else Bm.stride(0)
    ldc = ldc if ldc is not None else N
    check(lib().b200_gemm_tc_reduce_scatter(comm_handle, _p(_chk(A, bf16, "A")), _p(_chk(Bm, bf16, "B")), c_seg, c_off, nranks,
                                            M, N, K, lda, ldb, ldc, int(a_mn_major), int(b_mn_major), block_n, _stream()),
          "gemm_tc_reduce_scatter")


def allgather_acts(comm_handle, channel: int, nblocks: int, act_seg: int, act_off: int, dz_seg: int, dz_off: int, out_seg: int,
                   act_out_off: int, dz_out_off: int, B: int, F: int, D: int):
    """Every rank's dz [B,D] and this rank's column slice of every rank's a2 [B,F] -> local [n*B,D] / [n*B,F/n]."""
    check(lib().b200_allgather_acts(comm_handle, channel, nblocks, act_seg, act_off, dz_seg, dz_off, out_seg, act_out_off,
                                    dz_out_off, B, F, D, _stream()), "allgather_acts")


def gemm_tc_wgrad_adam(comm_handle, A, Bm, M: int, N: int, K: int, lda: int, ldb: int, *, state_seg: int, master_off: int, m_off: int,
                       v_off: int, w16_seg: int, w16_off: int, ctrl_seg: int, step_off: int, hyper_off: int, param_start: int,
                       row0: int, as_rank: int = -1, gscale: float = 1.0, block_n: int = 128):
    """Dense wgrad over the global batch with Adam in the epilogue; new bf16 weights go to every rank (see b200dp.h)."""
    check(lib().b200_gemm_tc_wgrad_adam(comm_handle, _p(_chk(A, bf16, "A")), _p(_chk(Bm, bf16, "B")), M, N, K, lda, ldb, state_seg,
                                        master_off, m_off, v_off, w16_seg, w16_off, ctrl_seg, step_off, hyper_off, param_start,
                                        row0, as_rank, float(gscale), block_n, _stream()), "gemm_tc_wgrad_adam")


def conv_tc_fwd(inp, w, bias, out, argmax=None, *, relu=True, pool=False, pad: Optional[int] = None):
    """``pad=None``: SAME (k//2); ``pad=0``: VALID.  ``out`` is [B, H+2p-k+1, W+2p-k+1, Co] (halved when pooling)."""
    B, H, W, Ci = inp.shape
    k, Co = w.shape[0], w.shape[3]
    check(lib().b200_conv_tc_fwd_pad(_p(_chk(inp, bf16, "in")), _p(_chk(w, bf16, "w")), _p(bias), _p(_chk(out, bf16, "out")), _p(argmax),
                                     B, H, W, Ci, Co, k, k // 2 if pad is None else pad, int(relu), int(pool), _stream()), "conv_tc_fwd")


def conv_tc_dgrad(dout, w, din, *, pad: Optional[int] = None):
    B, H, W, Ci = din.shape
    k, Co = w.shape[0], w.shape[3]
    check(lib().b200_conv_tc_dgrad_pad(_p(_chk(dout, bf16, "dout")), _p(_chk(w, bf16, "w")), _p(_chk(din, bf16, "din")),
                                       B, H, W, Ci, Co, k, k // 2 if pad is None else pad, _stream()), "conv_tc_dgrad")


def conv_tc_wgrad(inp, dout, dw, dbias, *, pad: Optional[int] = None):
    B, H, W, Ci = inp.shape
    k, Co = dw.shape[0], dw.shape[3]
    check(lib().b200_conv_tc_wgrad_pad(_p(_chk(inp, bf16, "in")), _p(_chk(dout, bf16, "dout")), _p(_chk(dw, torch.float32, "dw")), _p(dbias),
                                       B, H, W, Ci, Co, k, k // 2 if pad is None else pad, _stream()), "conv_tc_wgrad")


def topology() -> dict:
    import json
    buf = ctypes.create_string_buffer(1 << 16)
    n = lib().b200_topology_json(buf, len(buf))
    if n < 0:
        raise B200Error("topology buffer too small")
    return json.loads(buf.value.decode())
