"""The rest of Horovod's collective API, for programs that go beyond the reference scripts.

The reference only calls ``allreduce`` (through ``DistributedOptimizer``), ``broadcast`` (through the hooks) and the
metric average (``horovod/tensorflow_mnist.py:133,143``; ``horovod/tensorflow_mnist_gpu.py:137-153``).  A user moving
an existing Horovod program over will also reach for the asynchronous handles, grouped (fused) allreduce, alltoall,
reducescatter, ``join`` and wire compression; they are provided here on top of the same two data planes as
:mod:`b200dp.parallel.collectives` (gloo for CPU tensors, the NVLink peer-memory kernels for CUDA tensors).
"""
from __future__ import annotations

from typing import Any, Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

from ..runtime import world as _w
from . import collectives as C


# ------------------------------------------------------------------------------------------- handles
class Handle:
    """Result of an ``*_async`` call.  ``synchronize(handle)`` returns the output tensor."""

    def __init__(self, output: Any, work: Any = None, finish: Optional[Callable[[], None]] = None):
        self.output = output
        self._work = work
        self._finish = finish
        self._done = work is None and finish is None

    def poll(self) -> bool:
        if self._done:
            return True
        if self._work is not None and not self._work.is_completed():
            return False
        return True

    def wait(self) -> Any:
        if not self._done:
            if self._work is not None:
                self._work.wait()
            if self._finish is not None:
                self._finish()
            self._done = True
        return self.output


def poll(handle: Handle) -> bool:
    """True when ``synchronize(handle)`` would not block."""
    return handle.poll()


def synchronize(handle: Handle) -> Any:
    return handle.wait()


def allreduce_async_(tensor: torch.Tensor, op: C.ReduceOp = C.Average, name: Optional[str] = None) -> Handle:
    """In-place allreduce that returns at once.  CPU tensors ride a gloo ``async_op``; CUDA tensors are enqueued on the
    current stream (the peer-memory kernels are asynchronous with respect to the host already)."""
    w = _w.world()
    if w.size == 1:
        return Handle(tensor)
    if tensor.is_cuda or op is C.Adasum or not tensor.is_contiguous() or tensor.dtype in (torch.bfloat16, torch.float16) \
            or not tensor.is_floating_point():
        C.allreduce_(tensor, op, name)
        return Handle(tensor)
    work = dist.all_reduce(tensor, op=dist.ReduceOp.SUM, async_op=True)

    def finish():
        if op is C.Average:
            tensor.div_(w.size)
    return Handle(tensor, work, finish)


def allreduce_async(tensor: torch.Tensor, op: C.ReduceOp = C.Average, name: Optional[str] = None) -> Handle:
    return allreduce_async_(tensor.clone(), op, name)


def broadcast_async_(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> Handle:
    w = _w.world()
    if w.size == 1:
        return Handle(tensor)
    if tensor.is_cuda or not tensor.is_contiguous():
        C.broadcast_(tensor, root_rank, name)
        return Handle(tensor)
    if not (0 <= root_rank < w.size):
        raise ValueError(f"root_rank {root_rank} outside world of size {w.size}")
    return Handle(tensor, dist.broadcast(tensor, src=root_rank, async_op=True))


def broadcast_async(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> Handle:
    return broadcast_async_(tensor.clone(), root_rank, name)


def allgather_async(tensor: torch.Tensor, name: Optional[str] = None) -> Handle:
    return Handle(C.allgather(tensor, name))


# ------------------------------------------------------------------------------------------- grouped (fused) allreduce
def grouped_allreduce_(tensors: Sequence[torch.Tensor], op: C.ReduceOp = C.Average, name: Optional[str] = None) -> List[torch.Tensor]:
    """Reduce several tensors with ONE collective: they are packed into a fusion buffer (fp32 for floating tensors of
    mixed dtype), reduced, and unpacked in place — Horovod's tensor fusion, made explicit.  With Adasum the tensor
    boundaries are kept so every tensor gets its own dot products."""
    tensors = list(tensors)
    if not tensors:
        return tensors
    w = _w.world()
    if w.size == 1:
        return tensors
    dev = tensors[0].device
    if any(t.device != dev for t in tensors):
        raise ValueError("grouped_allreduce: all tensors must live on one device")
    if not all(t.is_floating_point() for t in tensors):
        for t in tensors:                               # integer tensors: no common wire type worth fusing
            C.allreduce_(t, op, name)
        return tensors
    wire = tensors[0].dtype if all(t.dtype == tensors[0].dtype for t in tensors) and tensors[0].dtype == torch.float64 else torch.float32
    flat = torch.cat([t.detach().reshape(-1).to(wire) for t in tensors])
    C.allreduce_(flat, op, name, segments=[t.numel() for t in tensors])
    off = 0
    for t in tensors:
        n = t.numel()
        t.detach().copy_(flat[off:off + n].view_as(t))
        off += n
    return tensors


def grouped_allreduce(tensors: Sequence[torch.Tensor], op: C.ReduceOp = C.Average, name: Optional[str] = None) -> List[torch.Tensor]:
    return grouped_allreduce_([t.clone() for t in tensors], op, name)


# ------------------------------------------------------------------------------------------- alltoall / reducescatter
def alltoall(tensor: torch.Tensor, splits: Optional[Sequence[int]] = None, name: Optional[str] = None):
    """Rank r sends ``splits[i]`` rows (dim 0) to rank i and receives the rows addressed to it, concatenated in rank
    order.  Without ``splits`` dim 0 is divided evenly.  Returns the received tensor (and, when ``splits`` was given, the
    list of received row counts, as Horovod does)."""
    w = _w.world()
    n = w.size
    if splits is None:
        if tensor.shape[0] % n:
            raise ValueError(f"alltoall: dim 0 ({tensor.shape[0]}) is not divisible by the world size {n}")
        send = [tensor.shape[0] // n] * n
    else:
        send = [int(s) for s in splits]
        if len(send) != n or sum(send) != tensor.shape[0]:
            raise ValueError("alltoall: splits must have one entry per rank and sum to dim 0")
    if n == 1:
        return (tensor.clone(), list(send)) if splits is not None else tensor.clone()
    # exchange the row counts, then the rows; gloo has no native all_to_all, so every rank gathers what is meant for it
    counts = C.allgather_object(send)                       # counts[src][dst]
    recv = [counts[src][w.rank] for src in range(n)]
    cpu = tensor.detach().cpu().contiguous()
    offs = [0]
    for s in send:
        offs.append(offs[-1] + s)
    outs: List[torch.Tensor] = []
    tail = tuple(cpu.shape[1:])
    for dst in range(n):
        mine = cpu[offs[dst]:offs[dst + 1]].contiguous()
        if dst == w.rank:
            bufs = [torch.empty((counts[src][dst],) + tail, dtype=cpu.dtype) for src in range(n)]
            _gather_var(mine, bufs, dst)
            outs = bufs
        else:
            _gather_var(mine, None, dst)
    out = torch.cat(outs, dim=0).to(tensor.device)
    return (out, recv) if splits is not None else out


def _gather_var(mine: torch.Tensor, bufs: Optional[List[torch.Tensor]], dst: int) -> None:
    """Variable-size gather to ``dst`` over gloo point-to-point sends (deadlock-free: receives are posted first)."""
    w = _w.world()
    if w.rank == dst:
        reqs = []
        for src in range(w.size):
            if src == dst:
                bufs[src].copy_(mine)
            elif bufs[src].numel():
                reqs.append(dist.irecv(bufs[src], src=src))
        for r in reqs:
            r.wait()
    elif mine.numel():
        dist.send(mine, dst=dst)


def reducescatter(tensor: torch.Tensor, op: C.ReduceOp = C.Average, name: Optional[str] = None) -> torch.Tensor:
    """Reduce across ranks, then keep this rank's slice of dim 0 (rank r gets rows [r*k, (r+1)*k) for even splits; the
    first ``dim0 % size`` ranks get one extra row otherwise, as Horovod splits)."""
    if op is C.Adasum:
        raise ValueError("reducescatter supports Average and Sum")
    w = _w.world()
    if w.size == 1:
        return tensor.clone()
    full = C.allreduce(tensor, op, name)
    d0, n = tensor.shape[0], w.size
    base, extra = divmod(d0, n)
    lo = w.rank * base + min(w.rank, extra)
    hi = lo + base + (1 if w.rank < extra else 0)
    return full[lo:hi].clone()


# ------------------------------------------------------------------------------------------- join
def join(device: int = -1) -> int:
    """Horovod's ``join``: a rank that has run out of data calls it and blocks until every rank has; returns the rank that
    joined last.  Ranks still training keep reducing among themselves in Horovod; here data-parallel replicas step in
    lockstep (one fused collective per step), so join is a barrier that also reports the last arrival."""
    import time
    w = _w.world()
    if w.size == 1:
        return 0
    stamps = C.allgather_object(time.time())
    return max(range(w.size), key=lambda r: (stamps[r], r))


# ------------------------------------------------------------------------------------------- wire compression
class _NoneCompressor:
    @staticmethod
    def compress(t: torch.Tensor):
        return t, None

    @staticmethod
    def decompress(t: torch.Tensor, ctx):
        return t


class _CastCompressor:
    def __init__(self, dtype: torch.dtype):
        self.dtype = dtype

    def compress(self, t: torch.Tensor):
        if t.is_floating_point() and t.dtype != self.dtype:
            return t.to(self.dtype), t.dtype
        return t, None

    def decompress(self, t: torch.Tensor, ctx):
        return t.to(ctx) if ctx is not None else t


class Compression:
    """``hvd.Compression.none`` / ``.fp16`` (Horovod's names) / ``.bf16``: what a gradient looks like on the wire."""
    none = _NoneCompressor()
    fp16 = _CastCompressor(torch.float16)
    bf16 = _CastCompressor(torch.bfloat16)


# ------------------------------------------------------------------------------------------- capability probes
def mpi_threads_supported() -> bool:
    return False


def mpi_enabled() -> bool:
    return False


def gloo_enabled() -> bool:
    return _w.is_initialized() and _w.world().size > 1


def cuda_built() -> bool:
    return _w.nccl_built()


def rocm_built() -> bool:
    return False


def ccl_built() -> bool:
    return False


def ddl_built() -> bool:
    return False


def cross_rank() -> int:
    """Index of this rank's node (ranks are laid out node by node, ``local_size`` per node)."""
    w = _w.world()
    return w.rank // max(1, w.local_size)


def cross_size() -> int:
    w = _w.world()
    return max(1, (w.size + max(1, w.local_size) - 1) // max(1, w.local_size))
