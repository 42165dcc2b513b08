"""Allreduce algorithm choice per message size, from a measured table.

Horovod autotunes its fusion threshold / cycle time at run time ([upstream], pulled in at
``horovod/Dockerfile:64-65``).  Here the only knob is one-shot vs two-shot (vs NVLS): one-shot costs
one barrier but every rank reads all N buffers ((N-1)·S bytes in), two-shot costs two barriers and
moves 2·(N-1)/N·S.  ``bench/allreduce_sweep.py`` measures both over 1 KB–1 GB and writes the
crossover per world size to ``profiles/allreduce_table.json``; without a table the defaults below
(derived from the 770 GB/s measured peer bandwidth and ~4 µs barrier) are used.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

ONESHOT, TWOSHOT, NVLS = 0, 1, 2
_NAMES = {"oneshot": ONESHOT, "twoshot": TWOSHOT, "nvls": NVLS}

# bytes at or below which one-shot wins, per world size (fallback model)
_DEFAULT_CROSSOVER = {2: 1 << 62, 4: 512 * 1024, 8: 256 * 1024}
_table: Optional[Dict[int, int]] = None


def _load() -> Dict[int, int]:
    global _table
    if _table is not None:
        return _table
    _table = dict(_DEFAULT_CROSSOVER)
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    for p in (os.environ.get("B200DP_ALLREDUCE_TABLE"), os.path.join(here, "profiles", "allreduce_table.json")):
        if p and os.path.exists(p):
            try:
                data = json.load(open(p))
                for k, v in data.get("oneshot_max_bytes", {}).items():
                    _table[int(k)] = int(v)
                break
            except (OSError, ValueError):
                pass
    return _table


def crossover_bytes(nranks: int) -> int:
    t = _load()
    if nranks in t:
        return t[nranks]
    keys = sorted(t)
    for k in keys:
        if nranks <= k:
            return t[k]
    return t[keys[-1]]


_nvls_min: Optional[Dict[int, int]] = None


def nvls_min_bytes(nranks: int) -> Optional[int]:
    """Message size from which the in-switch (NVLS) reduction beats two-shot, if a sweep recorded one."""
    global _nvls_min
    if _nvls_min is None:
        _nvls_min = {}
        here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        for p in (os.environ.get("B200DP_ALLREDUCE_TABLE"), os.path.join(here, "profiles", "allreduce_table.json")):
            if p and os.path.exists(p):
                try:
                    for k, v in json.load(open(p)).get("nvls_min_bytes", {}).items():
                        _nvls_min[int(k)] = int(v)
                    break
                except (OSError, ValueError):
                    pass
    return _nvls_min.get(nranks)


def choose(nbytes: int, nranks: int, forced: Optional[str] = None, nvls_ok: bool = False) -> int:
    forced = forced or os.environ.get("B200DP_ALLREDUCE_ALGO")
    if nranks <= 1:
        return ONESHOT
    if forced and forced != "auto":
        a = _NAMES[forced]
        if a == NVLS:
            # tiny ranges stay one-shot (a multicast round trip buys nothing below the latency floor)
            return NVLS if (nvls_ok and nbytes >= (1 << 16)) else (TWOSHOT if nbytes > crossover_bytes(nranks) else ONESHOT)
        return a
    if nbytes <= crossover_bytes(nranks):
        return ONESHOT
    mn = nvls_min_bytes(nranks)
    if nvls_ok and mn is not None and nbytes >= mn:
        return NVLS
    return TWOSHOT


def reset() -> None:
    global _table, _nvls_min
    _table = None
    _nvls_min = None
