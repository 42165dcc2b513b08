"""Minibatch loader over an in-memory dataset, native when the library is built.

Semantics are the reference generator's (``horovod/tensorflow_mnist.py:76-85``): an endless stream, a fresh permutation
every pass, consecutive full batches, tail dropped.  The native path (``csrc/loader.cpp``) assembles batches on a
background thread into a small ring of host buffers — pinned when CUDA is available, so the engine's H2D copy is a straight
DMA — and stays ``nslots - hold`` batches ahead of the training loop.  The pure-Python path draws the same permutations
(splitmix64 Fisher-Yates, bit for bit), which is how the tests pin the native one down.

    loader = BatchLoader(x_train, y_train, 128, seed=hvd.rank())
    for xb, yb in loader:            # torch tensors, valid until `hold` more batches have been drawn
        engine.train_step(xb, yb)
"""
from __future__ import annotations

import ctypes
from typing import Iterator, List, Optional, Tuple

import numpy as np
import torch

_MASK = (1 << 64) - 1


def _splitmix64(state: int) -> Tuple[int, int]:
    state = (state + 0x9E3779B97F4A7C15) & _MASK
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
    return state, z ^ (z >> 31)


def epoch_permutation(n: int, seed: int, epoch: int, shuffle: bool = True) -> np.ndarray:
    """The permutation ``csrc/loader.cpp`` uses for pass ``epoch`` (host twin of ``Loader::make_perm``)."""
    perm = np.arange(n, dtype=np.int64)
    if not shuffle:
        return perm
    s = (seed ^ ((0xD1B54A32D192ED03 * (epoch + 1)) & _MASK)) & _MASK
    p = perm.tolist()
    for i in range(n - 1, 0, -1):
        s, r = _splitmix64(s)
        j = r % (i + 1)
        p[i], p[j] = p[j], p[i]
    return np.asarray(p, dtype=np.int64)


class BatchLoader:
    def __init__(self, x, y, batch_size: int, *, shuffle: bool = True, seed: int = 0, nslots: int = 4, hold: int = 1,
                 native: Optional[bool] = None, pin: Optional[bool] = None):
        self.x = np.ascontiguousarray(np.asarray(x, dtype=np.float32).reshape(len(x), -1))
        self.y = np.ascontiguousarray(np.asarray(y).astype(np.int32).reshape(-1))
        if len(self.x) != len(self.y):
            raise ValueError("x and y differ in length")
        self.n, self.row = self.x.shape
        self.batch = int(batch_size)
        if not (1 <= self.batch <= self.n):
            raise ValueError(f"batch_size {batch_size} outside [1, {self.n}]")
        if not (1 <= hold < nslots):
            raise ValueError("need 1 <= hold < nslots")
        self.shuffle, self.seed, self.nslots, self.hold = bool(shuffle), int(seed) & _MASK, int(nslots), int(hold)
        self.batches_per_epoch = self.n // self.batch
        if pin is None:
            pin = torch.cuda.is_available()
        self._sx = [torch.empty(self.batch, self.row, dtype=torch.float32) for _ in range(self.nslots)]
        self._sy = [torch.empty(self.batch, dtype=torch.int32) for _ in range(self.nslots)]
        if pin:
            self._sx = [t.pin_memory() for t in self._sx]
            self._sy = [t.pin_memory() for t in self._sy]
        self._h = None
        self._lib = None
        if native is None or native:
            try:
                from ..ops import _lib
                self._lib = _lib.load()
            except (RuntimeError, OSError):
                if native:
                    raise
                self._lib = None
        self.native = self._lib is not None
        self._held: List[int] = []
        self._b = 0                          # next batch number (Python path)
        self._perm_epoch, self._perm = -1, None
        if self.native:
            arr_x = (ctypes.c_void_p * self.nslots)(*[t.data_ptr() for t in self._sx])
            arr_y = (ctypes.c_void_p * self.nslots)(*[t.data_ptr() for t in self._sy])
            h = ctypes.c_void_p()
            rc = self._lib.b200_loader_create(ctypes.c_void_p(self.x.ctypes.data), ctypes.c_void_p(self.y.ctypes.data), self.n,
                                              self.row * 4, self.batch, self.nslots, int(self.shuffle), ctypes.c_uint64(self.seed),
                                              arr_x, arr_y, ctypes.byref(h))
            if rc != 0:
                raise RuntimeError(self._lib.b200_last_error().decode())
            self._h = h

    # -------------------------------------------------------------------------------------------------
    def _release_old(self) -> None:
        while len(self._held) > self.hold - 1:
            s = self._held.pop(0)
            if self.native:
                self._lib.b200_loader_release(self._h, s)

    def next(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """The next batch.  The returned tensors alias a ring slot and stay valid until ``hold`` more batches were drawn."""
        self._release_old()
        if self.native:
            slot, idx = ctypes.c_int(), ctypes.c_longlong()
            if self._lib.b200_loader_next(self._h, ctypes.byref(slot), ctypes.byref(idx)) != 0:
                raise RuntimeError(self._lib.b200_last_error().decode())
            s = slot.value
        else:
            b = self._b
            self._b += 1
            epoch, k = divmod(b, self.batches_per_epoch)
            if epoch != self._perm_epoch:
                self._perm, self._perm_epoch = epoch_permutation(self.n, self.seed, epoch, self.shuffle), epoch
            rows = self._perm[k * self.batch:(k + 1) * self.batch]
            s = b % self.nslots
            self._sx[s].copy_(torch.from_numpy(self.x[rows]))
            self._sy[s].copy_(torch.from_numpy(self.y[rows]))
        self._held.append(s)
        return self._sx[s], self._sy[s]

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        while True:
            yield self.next()

    def close(self) -> None:
        if self._h is not None:
            self._lib.b200_loader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
