"""Philox4x32-10 on the host, bit-identical to the device generator in
``csrc/common.cuh``.  Lets the CPU oracle draw exactly the dropout mask the
sm_100a head kernel draws (reference op: ``layers.dropout(..., rate=0.5)``,
``horovod/tensorflow_mnist.py:65-67``).

Element ``i`` of a layer's activation uses output lane ``i % 4`` of the block with
counter ``(i // 4, 0, step, layer_id)`` and key ``(seed_lo, seed_hi)``.
"""
from __future__ import annotations

import numpy as np
import torch

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint32(0x9E3779B9)
_W1 = np.uint32(0xBB67AE85)
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1):
    c0 = c0.astype(np.uint32); c1 = c1.astype(np.uint32)
    c2 = c2.astype(np.uint32); c3 = c3.astype(np.uint32)
    k0 = np.uint32(k0); k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = c0.astype(np.uint64) * _M0
            p1 = c2.astype(np.uint64) * _M1
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32); lo0 = (p0 & _MASK).astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32); lo1 = (p1 & _MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32(k0 + _W0); k1 = np.uint32(k1 + _W1)
    return c0, c1, c2, c3


def dropout_keep_mask(numel: int, rate: float, seed: int, step: int, layer_id: int = 0) -> torch.Tensor:
    """bool[numel]: True where the activation is kept."""
    nblk = (numel + 3) // 4
    q = np.arange(nblk, dtype=np.uint64)
    z = np.zeros(nblk, dtype=np.uint32)
    r = philox4x32(q & _MASK, z, z + np.uint32(step & 0xFFFFFFFF), z + np.uint32(layer_id),
                   seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    u = np.stack(r, axis=1).reshape(-1)[:numel]
    thresh = np.uint64(min(int(rate * 4294967296.0), 0xFFFFFFFF))
    return torch.from_numpy(u.astype(np.uint64) >= thresh)
