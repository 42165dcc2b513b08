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


def _pair_on(fn, lib, dev, seed, sizes, nblocks=8):
    """One call on `dev` (made current, as the entry point requires); returns (out, xa, xb, starts)."""
    starts = [sum(sizes[:i]) for i in range(len(sizes))]
    n = sum(sizes)
    st, sz = (LL * len(sizes))(*starts), (LL * len(sizes))(*sizes)
    with torch.cuda.device(dev):
        g = torch.Generator(device="cpu").manual_seed(seed)
        xa = torch.randn(n, generator=g).to(f"cuda:{dev}")
        xb = (torch.randn(n, generator=g) * 0.5).to(f"cuda:{dev}") + 0.1 * xa
        out = torch.empty(n, device=f"cuda:{dev}")
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = fn(xa.data_ptr(), xb.data_ptr(), out.data_ptr(), nblocks, len(sizes), st, sz, stream)
        assert rc == 0, lib.b200_last_error().decode()
        torch.cuda.synchronize()
    return out, xa, xb, starts


@pytest.mark.multigpu(2)
def test_static_scratch_is_resolved_per_device():
    """Every GPU that loads the library carries its own copy of the scratch array, and the entry point looks its address
    up separately for each device: a call made with device 1 current must use device 1's array (device 0's address
    would fault there or land in the wrong memory).  Calls alternate between the two devices, twice each."""
    lib, fn = _cabi()
    sizes = [4096, 32, 1000]
    for rnd in range(2):
        for dev in (1, 0):
            out, xa, xb, starts = _pair_on(fn, lib, dev, 100 + 10 * rnd + dev, sizes)
            tree = torch.cat([adasum_tree([xa[s:s + l].cpu(), xb[s:s + l].cpu()]) for s, l in zip(starts, sizes)])
            assert torch.allclose(out.cpu(), tree, rtol=1e-4, atol=1e-5), (rnd, dev)


def test_static_scratch_calls_from_two_host_threads_on_one_stream():
    """Two host threads of one process may both call the entry point as long as their calls do not overlap on the GPU:
    here both launch into the SAME stream, which serialises the kernels, while the host-side set-up (the per-device
    address cache) is hit concurrently."""
    import threading
    lib, fn = _cabi()
    sizes = [2048, 64, 7000]
    starts = [sum(sizes[:i]) for i in range(len(sizes))]
    n = sum(sizes)
    st, sz = (LL * len(sizes))(*starts), (LL * len(sizes))(*sizes)
    dev = torch.cuda.current_device()
    s = torch.cuda.Stream()
    per_thread, errs = 6, []
    ins = [[(torch.randn(n, device="cuda"), torch.randn(n, device="cuda")) for _ in range(per_thread)] for _ in range(2)]
    outs = [[torch.empty(n, device="cuda") for _ in range(per_thread)] for _ in range(2)]
    torch.cuda.synchronize()

    def work(t):
        try:
            torch.cuda.set_device(dev)
            for k in range(per_thread):
                xa, xb = ins[t][k]
                rc = fn(xa.data_ptr(), xb.data_ptr(), outs[t][k].data_ptr(), 8, len(sizes), st, sz, ctypes.c_void_p(s.cuda_stream))
                if rc != 0:
                    errs.append(lib.b200_last_error().decode())
        except Exception as e:                      # noqa: BLE001
            errs.append(repr(e))

    ths = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    s.synchronize()
    assert not errs, errs
    for t in range(2):
        for k in range(per_thread):
            xa, xb = ins[t][k]
            tree = torch.cat([adasum_tree([xa[a:a + l].cpu(), xb[a:a + l].cpu()]) for a, l in zip(starts, sizes)])
            assert torch.allclose(outs[t][k].cpu(), tree, rtol=1e-4, atol=1e-5), (t, k)
