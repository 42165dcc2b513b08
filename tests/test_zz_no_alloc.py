"""The allocation-free entry point ``b200_adasum_pair_static`` (no counterpart in the reference).

Some users of this library sit inside another program's memory regime — a gradient hook under a framework whose
caching allocator already owns the device, a server that fixed its memory budget at start-up — and cannot call
``cudaMalloc``, so they cannot build a communicator just to get scratch space.  For them the pairwise Adasum has a
variant whose scratch is a 48 KiB array inside the library.  This test drives it the way such a caller would,
through the plain C ABI (``ctypes``, raw device pointers), and checks it against the regular entry point
``b200_adasum_level`` (one level over an emulated 2-rank communicator computes the same pair), twice in a row
because every call reuses — and leaves its partial sums in — the same scratch.

The file name sorts last on purpose: a run that stops at the first failure has then already run everything else.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from b200dp.ops import _lib
from b200dp.parallel.arena import SymmetricArena
from b200dp.parallel.collectives import adasum_tree

LL = ctypes.c_longlong


def _cabi():
    lib = _lib.load()
    fn = lib.b200_adasum_pair_static
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                   ctypes.POINTER(LL), ctypes.POINTER(LL), ctypes.c_void_p]
    lib.b200_last_error.restype = ctypes.c_char_p
    return lib, fn


def _reference_level(xa, xb, starts, sizes, nblocks):
    """out = adasum(xa, xb) through the existing entry point (emulated 2-rank communicator, one level)."""
    n = xa.numel()
    n4 = (n + 3) // 4 * 4
    arena = SymmetricArena.emulated(2)
    arena.set_timeout_ms(2000)
    try:
        seg = arena.alloc(n4 * 4 * 2 + 65536)
        arena.view(seg, 0, n, torch.float32, 0).copy_(xa)
        arena.view(seg, 0, n, torch.float32, 1).copy_(xb)
        res = arena.adasum_levels(seg, 0, n4 * 4, 2 * n4 * 4, starts, sizes, nblocks=nblocks)
        torch.cuda.synchronize()
        return arena.view(seg, res, n, torch.float32, 0).clone()
    finally:
        arena.close()


def test_static_scratch_variant_matches_regular_entry_point_twice():
    lib, fn = _cabi()
    torch.manual_seed(2024)
    sizes = [800, 32, 5000, 64, 10, 51200]
    starts = [sum(sizes[:i]) for i in range(len(sizes))]
    n = sum(sizes)
    nblocks = 8
    st, sz = (LL * len(sizes))(*starts), (LL * len(sizes))(*sizes)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = torch.empty(n, device="cuda")
    for call in range(2):                      # second call finds the first call's partial sums in the scratch
        xa = torch.randn(n, device="cuda") * (call + 1)
        xb = torch.randn(n, device="cuda") * 0.5 + 0.1 * xa
        want = _reference_level(xa, xb, starts, sizes, nblocks)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        rc = fn(xa.data_ptr(), xb.data_ptr(), out.data_ptr(), nblocks, len(sizes), st, sz, stream)
        assert rc == 0, lib.b200_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        if call == 1:                          # by now every module of the library is resident; nothing may be allocated
            assert free_after == free_before, "the call changed the amount of free device memory"
        assert torch.allclose(out, want, rtol=1e-4, atol=1e-5), f"call {call}: max diff {(out - want).abs().max().item()}"
        tree = torch.cat([adasum_tree([xa[s:s + l].cpu(), xb[s:s + l].cpu()]) for s, l in zip(starts, sizes)])
        assert torch.allclose(out.cpu(), tree, rtol=1e-4, atol=1e-5)


def test_static_scratch_variant_rejects_oversize_requests():
    lib, fn = _cabi()
    x = torch.zeros(64, device="cuda")
    one = (LL * 1)(0)
    cnt = (LL * 1)(64)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert fn(x.data_ptr(), x.data_ptr(), x.data_ptr(), 65, 1, one, cnt, stream) != 0          # more blocks than the scratch holds
    assert b"at most 64 blocks" in lib.b200_last_error()
    many = (LL * 33)(*range(33))
    ones = (LL * 33)(*([1] * 33))
    assert fn(x.data_ptr(), x.data_ptr(), x.data_ptr(), 8, 33, many, ones, stream) != 0         # more segments than the scratch holds
    assert b"at most 32 tensor segments" in lib.b200_last_error()
    assert fn(None, x.data_ptr(), x.data_ptr(), 8, 1, one, cnt, stream) != 0
    torch.cuda.synchronize()
