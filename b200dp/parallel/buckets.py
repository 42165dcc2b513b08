"""Static bucket planning.

Horovod negotiates, every cycle, which tensors are ready on all ranks and packs
them into its fusion buffer (pulled in at ``horovod/Dockerfile:64-65``).
Data-parallel replicas run the *same* model, so the order gradients become
ready is known when the optimizer is wrapped: the plan is computed once, is
identical on every rank, and no negotiation traffic exists at run time.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple


def plan_buckets(sizes_bytes: Sequence[int], cap_bytes: int) -> List[List[int]]:
    """Greedy packing of consecutive tensors (given in ready order) into buckets
    of at most ``cap_bytes``; a tensor larger than the cap gets its own bucket."""
    buckets: List[List[int]] = []
    cur: List[int] = []
    cur_b = 0
    for i, s in enumerate(sizes_bytes):
        if cur and cur_b + s > cap_bytes:
            buckets.append(cur)
            cur, cur_b = [], 0
        cur.append(i)
        cur_b += s
    if cur:
        buckets.append(cur)
    return buckets


@dataclass(frozen=True)
class ArenaBucket:
    """A contiguous element range [start, end) of the flat gradient arena."""
    name: str
    start: int
    end: int
    segments: Tuple[Tuple[int, int], ...]      # (offset, numel) of the tensors inside, absolute

    @property
    def numel(self) -> int:
        return self.end - self.start


def plan_arena_buckets(spec, ready_order: Sequence[Sequence[str]]) -> List[ArenaBucket]:
    """Buckets over the flat layout of ``spec``.  ``ready_order`` lists groups of
    parameter-name prefixes in the order backward finishes them; each group must
    be contiguous in the layout (true for layer-ordered layouts)."""
    out: List[ArenaBucket] = []
    params = list(spec.params)
    for gi, group in enumerate(ready_order):
        idx = [i for i, p in enumerate(params) if any(p.name.startswith(g) for g in group)]
        if not idx:
            continue
        lo, hi = min(idx), max(idx)
        if idx != list(range(lo, hi + 1)):
            raise ValueError(f"bucket group {group} is not contiguous in the parameter layout")
        start = params[lo].offset
        end = params[hi + 1].offset if hi + 1 < len(params) else spec.total
        segs = tuple((params[i].offset, params[i].numel) for i in idx)
        out.append(ArenaBucket("+".join(group), start, end, segs))
    covered = sorted((b.start, b.end) for b in out)
    pos = 0
    for s, e in covered:
        if s != pos:
            raise ValueError("buckets do not tile the arena")
        pos = e
    if pos != spec.total:
        raise ValueError("buckets do not cover the arena")
    return out
