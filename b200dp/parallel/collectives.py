"""Public collectives: allreduce (Average / Sum / Adasum), broadcast, allgather.

These are the primitives under ``hvd.DistributedOptimizer`` and the broadcast
hooks (reference: ``horovod/tensorflow_mnist.py:133,143``;
``horovod/tensorflow_mnist_gpu.py:137-138,150,153``).

* CPU tensors go over the gloo control plane (plumbing / tests / multi-host).
* CUDA tensors go through the symmetric peer-memory arena: the tensor is staged
  into the arena's fusion region and reduced by the hand-written NVLink kernels
  in ``csrc/comm.cu`` (one-shot or two-shot picked by size, see
  :mod:`b200dp.parallel.autoselect`).  No NCCL call on this path.
"""
from __future__ import annotations

import enum
from typing import Any, Iterable, List, Optional

import torch
import torch.distributed as dist

from ..runtime import world as _w


class ReduceOp(enum.Enum):
    Average = "average"
    Sum = "sum"
    Adasum = "adasum"
    Min = "min"
    Max = "max"
    Product = "product"


Average = ReduceOp.Average
Sum = ReduceOp.Sum
Adasum = ReduceOp.Adasum
Min = ReduceOp.Min
Max = ReduceOp.Max
Product = ReduceOp.Product
_ORDER_OPS = {ReduceOp.Min: dist.ReduceOp.MIN, ReduceOp.Max: dist.ReduceOp.MAX, ReduceOp.Product: dist.ReduceOp.PRODUCT}


def _is_pow2(n: int) -> bool:
    return n > 0 and (n & (n - 1)) == 0


# --------------------------------------------------------------------------
# Adasum reference math (CPU / oracle).  Pairwise rule, applied over a binary
# tree of ranks (distance doubling), per tensor:
#     a' = (1 - a.b / (2 |a|^2)) a + (1 - a.b / (2 |b|^2)) b
# --------------------------------------------------------------------------
def adasum_pair(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    af = a.double().flatten()
    bf = b.double().flatten()
    dot = torch.dot(af, bf)
    na = torch.dot(af, af)
    nb = torch.dot(bf, bf)
    ca = 1.0 - (dot / (2.0 * na) if na > 0 else 0.0)
    cb = 1.0 - (dot / (2.0 * nb) if nb > 0 else 0.0)
    return (ca * a.double() + cb * b.double()).to(a.dtype)


def adasum_tree(tensors: List[torch.Tensor]) -> torch.Tensor:
    """Reduce a list (one entry per rank, rank order) with the Adasum tree."""
    n = len(tensors)
    if not _is_pow2(n):
        raise ValueError(f"Adasum needs a power-of-two world size, got {n}")
    cur = list(tensors)
    while len(cur) > 1:
        cur = [adasum_pair(cur[i], cur[i + 1]) for i in range(0, len(cur), 2)]
    return cur[0]


# --------------------------------------------------------------------------
def _cpu_allreduce_(t: torch.Tensor, op: ReduceOp) -> torch.Tensor:
    w = _w.world()
    if op is ReduceOp.Adasum:
        parts = [torch.empty_like(t) for _ in range(w.size)]
        dist.all_gather(parts, t.contiguous())
        t.copy_(adasum_tree(parts))
        return t
    work = t if t.is_contiguous() else t.contiguous()
    if op in _ORDER_OPS:
        acc = work.float() if work.dtype in (torch.bfloat16, torch.float16) else work
        dist.all_reduce(acc, op=_ORDER_OPS[op])
        t.copy_(acc.to(t.dtype))
        return t
    if work.dtype in (torch.bfloat16, torch.float16):
        acc = work.float()
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
        if op is ReduceOp.Average:
            acc /= w.size
        t.copy_(acc.to(t.dtype))
        return t
    dist.all_reduce(work, op=dist.ReduceOp.SUM)
    if op is ReduceOp.Average:
        if work.is_floating_point():
            work /= w.size
        else:
            work //= w.size
    if work is not t:
        t.copy_(work)
    return t


def allreduce_(tensor: torch.Tensor, op: ReduceOp = Average, name: Optional[str] = None,
               segments: Optional[Iterable[int]] = None) -> torch.Tensor:
    """In-place allreduce.  ``segments`` (element counts) marks tensor boundaries
    inside a fused buffer so Adasum computes its dot products per tensor."""
    w = _w.world()
    if w.size == 1:
        return tensor
    if tensor.is_cuda and op in _ORDER_OPS:
        # Min / Max / Product are metric-sized in practice: they ride the control plane (the NVLink kernels implement the
        # gradient reductions: sum, average, Adasum)
        host = tensor.detach().cpu()
        _cpu_allreduce_(host, op)
        tensor.copy_(host.to(tensor.device))
        return tensor
    if tensor.is_cuda:
        if w.arena is None:
            raise RuntimeError("CUDA tensor given but the peer-memory arena is not up (init(device='cuda'))")
        return w.arena.allreduce_(tensor, op, segments=segments)
    if op is ReduceOp.Adasum and segments is not None:
        flat = tensor.view(-1)
        off = 0
        for n in segments:
            _cpu_allreduce_(flat[off:off + n], op)
            off += n
        return tensor
    return _cpu_allreduce_(tensor, op)


def allreduce(tensor: torch.Tensor, op: ReduceOp = Average, name: Optional[str] = None) -> torch.Tensor:
    out = tensor.clone()
    return allreduce_(out, op, name)


def broadcast_(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    w = _w.world()
    if w.size == 1:
        return tensor
    if not (0 <= root_rank < w.size):
        raise ValueError(f"root_rank {root_rank} outside world of size {w.size}")
    if tensor.is_cuda:
        if w.arena is None:
            raise RuntimeError("CUDA tensor given but the peer-memory arena is not up")
        return w.arena.broadcast_(tensor, root_rank)
    work = tensor if tensor.is_contiguous() else tensor.contiguous()
    dist.broadcast(work, src=root_rank)
    if work is not tensor:
        tensor.copy_(work)
    return tensor


def broadcast(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    return broadcast_(tensor.clone(), root_rank, name)


def allgather(tensor: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
    """Concatenate along dim 0 across ranks.  Dim 0 may differ per rank (Horovod's allgather allows it); the other
    dimensions must agree."""
    w = _w.world()
    if w.size == 1:
        return tensor.clone()
    cpu = tensor.detach().cpu().contiguous()
    if cpu.dim() == 0:
        cpu = cpu.reshape(1)
    shapes = allgather_object(tuple(cpu.shape))
    if any(sh[1:] != shapes[0][1:] for sh in shapes):
        raise ValueError(f"allgather: trailing dimensions differ across ranks: {shapes}")
    rows = [sh[0] for sh in shapes]
    if len(set(rows)) == 1:
        parts = [torch.empty_like(cpu) for _ in range(w.size)]
        dist.all_gather(parts, cpu)
        return torch.cat(parts, dim=0).to(tensor.device)
    # ragged: pad every contribution to the longest, gather, trim
    top = max(rows)
    padded = torch.zeros((top,) + tuple(cpu.shape[1:]), dtype=cpu.dtype)
    padded[:cpu.shape[0]] = cpu
    parts = [torch.empty_like(padded) for _ in range(w.size)]
    dist.all_gather(parts, padded)
    return torch.cat([p[:r] for p, r in zip(parts, rows)], dim=0).to(tensor.device)


def broadcast_object(obj: Any, root_rank: int = 0) -> Any:
    w = _w.world()
    if w.size == 1:
        return obj
    box = [obj if w.rank == root_rank else None]
    dist.broadcast_object_list(box, src=root_rank)
    return box[0]


def allgather_object(obj: Any) -> List[Any]:
    w = _w.world()
    if w.size == 1:
        return [obj]
    out: List[Any] = [None] * w.size
    dist.all_gather_object(out, obj)
    return out


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """Broadcast a ``state_dict``-like mapping or ``named_parameters()`` iterable."""
    if hasattr(params, "items"):
        items = list(params.items())
    else:
        items = list(params)
    for _, p in items:
        if isinstance(p, torch.Tensor):
            t = p.data if isinstance(p, torch.nn.Parameter) else p
            broadcast_(t, root_rank)


def broadcast_optimizer_state(optimizer: torch.optim.Optimizer, root_rank: int = 0) -> None:
    """Make every rank's optimizer state equal to the root's (tensors in place,
    python scalars via object broadcast)."""
    w = _w.world()
    if w.size == 1:
        return
    sd = optimizer.state_dict()
    meta = broadcast_object({"param_groups": sd["param_groups"],
                             "keys": {k: {kk: (None if isinstance(vv, torch.Tensor) else vv)
                                          for kk, vv in v.items()} for k, v in sd["state"].items()},
                             "shapes": {k: {kk: (tuple(vv.shape), str(vv.dtype)) for kk, vv in v.items()
                                            if isinstance(vv, torch.Tensor)} for k, v in sd["state"].items()}},
                            root_rank)
    if w.rank != root_rank:
        state = {}
        for k, shapes in meta["shapes"].items():
            ent = dict(meta["keys"][k])
            old = sd["state"].get(k, {})
            for kk, (shape, dtype) in shapes.items():
                cur = old.get(kk)
                if isinstance(cur, torch.Tensor) and tuple(cur.shape) == shape:
                    ent[kk] = cur
                else:
                    ent[kk] = torch.zeros(shape, dtype=getattr(torch, dtype.split(".")[-1]))
            state[k] = ent
        sd = {"state": state, "param_groups": meta["param_groups"]}
    for k in sorted(sd["state"].keys(), key=str):
        for kk in sorted(sd["state"][k].keys()):
            v = sd["state"][k][kk]
            if isinstance(v, torch.Tensor):
                broadcast_(v, root_rank)
    if w.rank != root_rank:
        optimizer.load_state_dict(sd)
