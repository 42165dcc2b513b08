"""Optimizers and the distributed wrapper.

``DistributedOptimizer(opt, op=Average|Adasum)`` is the reference's only
parallelism strategy: replicate the model, reduce every gradient across ranks
before the update (``horovod/tensorflow_mnist.py:132-136``;
``horovod/tensorflow_mnist_gpu.py:136-138``).

Two forms:

* wrapping a ``torch.optim.Optimizer`` — generic: gradients are packed into
  flat buckets (reverse parameter order, like backward produces them), reduced
  with :func:`b200dp.allreduce_`, unpacked, then the inner optimizer steps;
* wrapping :class:`Adam` (a plain config) — the engines fuse the reduction, the
  1/size scale and the Adam update into one peer-memory kernel per bucket.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, replace
from typing import List, Optional

import torch

from ..runtime import world as _w
from . import collectives as C
from .buckets import plan_buckets


@dataclass(frozen=True)
class Adam:
    """Adam hyper-parameters.  Defaults follow ``tf.train.AdamOptimizer``
    (epsilon 1e-8, TF epsilon placement); ``Adam.keras()`` gives Keras' 1e-7."""
    lr: float = 0.001
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8
    eps_mode: str = "tf"

    @staticmethod
    def keras(lr: float = 0.001) -> "Adam":
        return Adam(lr=lr, epsilon=1e-7)

    def scaled(self, factor: float) -> "Adam":
        return replace(self, lr=self.lr * factor)


def lr_scaler(use_adasum: bool = False) -> int:
    """The reference's rule (``horovod/tensorflow_mnist.py:123-127``): scale the LR
    by the world size; with Adasum by ``local_size`` when the GPU collective path
    is built, else by 1."""
    if use_adasum:
        return _w.local_size() if _w.nccl_built() else 1
    return _w.size()


@dataclass(frozen=True)
class DistributedAdam:
    """What ``DistributedOptimizer(Adam(...))`` returns: config consumed by engines."""
    base: Adam
    op: C.ReduceOp = C.Average
    algo: str = "auto"                # auto | oneshot | twoshot | nvls

    @property
    def lr(self) -> float:
        return self.base.lr


def fusion_threshold_bytes() -> int:
    v = os.environ.get("B200DP_FUSION_THRESHOLD", os.environ.get("HOROVOD_FUSION_THRESHOLD"))
    return int(v) if v else 64 * 1024 * 1024


class _DistributedTorchOptimizer:
    """Generic wrapper around any ``torch.optim.Optimizer``.

    Gradients are packed into flat buckets in the order backward produces them (reverse parameter order).  A
    post-accumulate-grad hook on every parameter counts readiness; the moment a bucket is complete its allreduce is
    launched — asynchronously (gloo ``async_op``) for CPU tensors, on the peer-memory kernels for CUDA tensors — so
    the reduction of late layers runs underneath the backward of early ones.  ``step()`` waits, unpacks, steps.
    ``backward_passes_per_step`` > 1 accumulates locally and reduces on the last pass (Horovod's option of that name).
    """

    def __init__(self, optimizer: torch.optim.Optimizer, named_parameters=None, op: C.ReduceOp = C.Average,
                 bucket_bytes: Optional[int] = None, backward_passes_per_step: int = 1, compression=None):
        self._opt = optimizer
        self.op = op
        self._compression = compression
        params: List[torch.nn.Parameter] = []
        for g in optimizer.param_groups:
            params.extend(p for p in g["params"] if p.requires_grad)
        self._params = params
        sizes = [p.numel() * 4 for p in reversed(params)]
        self._bucket_ix = plan_buckets(sizes, bucket_bytes or fusion_threshold_bytes())
        self._rev = list(reversed(params))
        self._synced = False
        self._passes = max(1, int(backward_passes_per_step))
        self._pass = 0
        self._bucket_of = {}
        for bi, ix in enumerate(self._bucket_ix):
            for i in ix:
                self._bucket_of[id(self._rev[i])] = bi
        self._ready = [0] * len(self._bucket_ix)
        self._launched = [False] * len(self._bucket_ix)
        self._pending = {}
        self.overlapped_launches = 0
        self._hooks = []
        if hasattr(torch.Tensor, "register_post_accumulate_grad_hook"):
            for p in params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))

    # torch optimizer surface -------------------------------------------------
    @property
    def param_groups(self):
        return self._opt.param_groups

    @property
    def state(self):
        return self._opt.state

    def state_dict(self):
        return self._opt.state_dict()

    def load_state_dict(self, sd):
        return self._opt.load_state_dict(sd)

    def zero_grad(self, set_to_none: bool = True):
        self._synced = False
        return self._opt.zero_grad(set_to_none=set_to_none)

    def __getattr__(self, item):
        return getattr(self._opt, item)

    # -------------------------------------------------------------------------
    def _on_grad(self, p) -> None:
        if not _w.is_initialized() or _w.size() == 1:
            return
        bi = self._bucket_of.get(id(p))
        if bi is None:
            return
        self._ready[bi] += 1
        if self._ready[bi] == len(self._bucket_ix[bi]):
            self._ready[bi] = 0
            if bi == len(self._bucket_ix) - 1:
                self._pass += 1                      # the last bucket closes a backward pass
            last_pass = (self._pass % self._passes == 0) if bi == len(self._bucket_ix) - 1 else ((self._pass + 1) % self._passes == 0)
            if last_pass and not self._launched[bi]:
                self._launch(bi)
                self.overlapped_launches += 1

    def _launch(self, bi: int) -> None:
        ps = [self._rev[i] for i in self._bucket_ix[bi]]
        ps = [p for p in ps if p.grad is not None]
        self._launched[bi] = True
        if not ps:
            return
        dev = ps[0].grad.device
        flat = torch.cat([p.grad.detach().reshape(-1).float() for p in ps]).to(dev)
        if self._passes > 1:
            flat /= self._passes
        work = None
        ctx = None
        if self._compression is not None:
            # Horovod's `compression=`: the bucket is rounded to the wire type before the reduction and widened after it
            wire, ctx = self._compression.compress(flat)
            if ctx is not None:
                C.allreduce_(wire, self.op, segments=[p.numel() for p in ps])
                flat = self._compression.decompress(wire, ctx)
                self._pending[bi] = (None, flat, ps)
                return
        if dev.type == "cpu" and self.op in (C.Average, C.Sum):
            import torch.distributed as dist
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True)
        else:
            C.allreduce_(flat, self.op, segments=[p.numel() for p in ps])
        self._pending[bi] = (work, flat, ps)

    def synchronize(self) -> None:
        """Finish the reduction of every bucket (launching the ones whose hooks never fired) and unpack."""
        if not _w.is_initialized() or _w.size() == 1:
            self._synced = True
            return
        for bi in range(len(self._bucket_ix)):
            if not self._launched[bi]:
                self._launch(bi)
        for bi in sorted(self._pending):
            work, flat, ps = self._pending[bi]
            if work is not None:
                work.wait()
                if self.op is C.Average:
                    flat /= _w.size()
            off = 0
            for p in ps:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p.grad))
                off += n
        self._pending.clear()
        self._launched = [False] * len(self._bucket_ix)
        self._ready = [0] * len(self._bucket_ix)
        self._synced = True

    def step(self, closure=None):
        if not self._synced:
            self.synchronize()
        self._synced = False
        return self._opt.step(closure) if closure is not None else self._opt.step()


def DistributedOptimizer(optimizer, named_parameters=None, op: C.ReduceOp = C.Average, *,
                         algo: str = "auto", bucket_bytes: Optional[int] = None,
                         backward_passes_per_step: int = 1, compression=None):
    """Horovod-named factory (see module docstring)."""
    if isinstance(optimizer, DistributedAdam):
        return optimizer
    if isinstance(optimizer, Adam):
        algo = os.environ.get("B200DP_ALLREDUCE_ALGO", algo)
        return DistributedAdam(base=optimizer, op=op, algo=algo)
    if isinstance(optimizer, torch.optim.Optimizer):
        return _DistributedTorchOptimizer(optimizer, named_parameters, op, bucket_bytes, backward_passes_per_step, compression)
    raise TypeError(f"cannot wrap optimizer of type {type(optimizer).__name__}")
