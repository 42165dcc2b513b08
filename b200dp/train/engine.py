"""Training engines: the object that owns parameters + optimizer state and runs
one synchronous data-parallel step.

``TorchEngine`` is the CPU/gloo engine and the oracle (plain PyTorch fp32 ops).
``B200Engine`` (:mod:`b200dp.models.b200_engine`) is the sm_100a engine: hand-written
kernels, flat gradient arena, allreduce+Adam fused per bucket over NVLink, whole
step replayed from a CUDA graph.  Both share the flat layout of
:mod:`b200dp.models.spec`, so checkpoints move freely between them.

One ``train_step`` = the reference's ``mon_sess.run(train_op, feed_dict=…)``
(``horovod/tensorflow_mnist.py:168-171``): forward, backward, gradient reduction
across ranks, Adam apply, ``global_step += 1``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, Optional, Union

import numpy as np
import torch

from ..models import torch_ref
from ..models.spec import ModelSpec, get_spec, init_flat
from ..parallel import collectives as C
from ..parallel.optimizer import Adam, DistributedAdam, DistributedOptimizer
from ..runtime import world as _w

ArrayLike = Union[np.ndarray, torch.Tensor]


@dataclass
class StepResult:
    loss: Any            # 0-dim tensor or float
    correct: Any         # 0-dim tensor or int
    n: int

    def loss_value(self) -> float:
        return float(self.loss)

    def accuracy(self) -> float:
        return float(self.correct) / max(self.n, 1)


def _as_dist_adam(opt) -> Optional[DistributedAdam]:
    if opt is None:
        return None
    if isinstance(opt, DistributedAdam):
        return opt
    if isinstance(opt, Adam):
        # unwrapped optimizer: local update only (no gradient reduction)
        return None
    raise TypeError(f"engine optimizers are b200dp.Adam configs (optionally wrapped), got {type(opt).__name__}")


class Engine:
    spec: ModelSpec
    global_step: int

    def train_step(self, x: ArrayLike, y: ArrayLike) -> StepResult: ...
    def eval_batch(self, x: ArrayLike, y: ArrayLike) -> StepResult: ...
    def predict(self, x: ArrayLike) -> torch.Tensor: ...
    def state(self) -> Dict[str, Any]: ...
    def load_state(self, tensors: Dict[str, torch.Tensor], meta: Dict[str, Any]) -> None: ...
    def broadcast_state(self, root_rank: int = 0) -> None: ...

    def finalize(self) -> None:
        """Collective hook at the end of training (session exit / end of fit); engines with rank-sharded state gather it."""

    # shared helpers ---------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        """Change the learning rate without rebuilding anything (LR schedules / warm-up callbacks)."""
        from dataclasses import replace
        if getattr(self, "adam", None) is None:
            raise RuntimeError("no optimizer set")
        self.adam = replace(self.adam, lr=float(lr))
        if getattr(self, "dist", None) is not None:
            self.dist = replace(self.dist, base=self.adam)

    def weights(self) -> Dict[str, torch.Tensor]:
        flat = self.state()["tensors"]["params"]
        return {k: v.clone() for k, v in self.spec.views(flat).items()}

    def meta(self) -> Dict[str, Any]:
        return {"arch": self.spec.arch, "width": self.spec.width, "num_params": self.spec.num_params,
                "world_size": _w.size() if _w.is_initialized() else 1}


class _NoSpan:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def _no_span(*_a, **_k):
    return _NoSpan()


class TorchEngine(Engine):
    """fp32 PyTorch engine (CPU by default).  Gradients come out of autograd
    already flat (the parameters are views of one leaf), get reduced in the
    planned buckets, and Adam runs on the flat buffers."""

    def __init__(self, spec: Optional[ModelSpec] = None, optimizer=None, *, seed: int = 0,
                 device: Union[str, torch.device, None] = None, dropout: bool = True,
                 compute_dtype: Optional[torch.dtype] = None, bucket_bytes: Optional[int] = None):
        self.spec = spec or get_spec()
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.seed = seed
        self.dropout = dropout
        self.compute_dtype = compute_dtype
        self.params = init_flat(self.spec, seed, self.device).requires_grad_(True)
        self.m = torch.zeros_like(self.params.data)
        self.v = torch.zeros_like(self.params.data)
        self.global_step = 0
        self.set_optimizer(optimizer)
        self._bucket_bytes = bucket_bytes
        self.host_timeline = None            # a utils.timeline.HostTimeline when B200DP_TIMELINE / HOROVOD_TIMELINE is set

    def set_optimizer(self, optimizer) -> None:
        self.dist = _as_dist_adam(optimizer)
        self.adam: Optional[Adam] = self.dist.base if self.dist else (optimizer if isinstance(optimizer, Adam) else None)

    # -------------------------------------------------------------------------
    def _to_dev(self, x: ArrayLike, y: Optional[ArrayLike] = None):
        xt = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x)
        xt = xt.to(self.device, dtype=torch.float32).reshape(xt.shape[0], -1)
        if y is None:
            return xt, None
        yt = torch.as_tensor(np.asarray(y) if not isinstance(y, torch.Tensor) else y).to(self.device).long()
        return xt, yt

    def _forward(self, x, training: bool):
        views = self.spec.views(self.params)
        # dropout counter = 1-based index of the step being run (same convention as the CUDA engine)
        return torch_ref.forward(self.spec, views, x, training=training, seed=self.seed,
                                 step=self.global_step + 1, dropout=self.dropout,
                                 compute_dtype=self.compute_dtype)

    def train_step(self, x: ArrayLike, y: ArrayLike) -> StepResult:
        if self.adam is None:
            raise RuntimeError("no optimizer set; pass optimizer=Adam(...) or compile() the model")
        xt, yt = self._to_dev(x, y)
        self.params.grad = None
        span = self.host_timeline.span if self.host_timeline is not None else _no_span
        with span("forward", "compute"):
            logits = self._forward(xt, training=True)
            loss, correct = torch_ref.loss_and_metrics(logits, yt)
        with span("backward", "compute"):
            loss.backward()
        g = self.params.grad
        if self.dist is not None and _w.is_initialized() and _w.size() > 1:
            if self.dist.op is C.Adasum:
                # per-tensor Adasum over the fused buffer
                for p in self.spec.params:
                    with span(f"adasum[{p.name}]", "collective"):
                        C.allreduce_(g[p.offset:p.offset + p.numel], C.Adasum)
            else:
                with span(f"allreduce[fused {g.numel() * 4} B]", "collective"):
                    C.allreduce_(g, self.dist.op)
        with torch.no_grad(), span("adam", "compute"):
            a = self.adam
            torch_ref.adam_step_(self.params.data, g, self.m, self.v, self.global_step + 1,
                                 lr=a.lr, beta1=a.beta1, beta2=a.beta2, eps=a.epsilon, eps_mode=a.eps_mode)
        self.global_step += 1
        return StepResult(loss.detach(), correct.detach(), xt.shape[0])

    @torch.no_grad()
    def eval_batch(self, x: ArrayLike, y: ArrayLike) -> StepResult:
        xt, yt = self._to_dev(x, y)
        logits = self._forward(xt, training=False)
        loss, correct = torch_ref.loss_and_metrics(logits, yt)
        return StepResult(loss, correct, xt.shape[0])

    @torch.no_grad()
    def predict(self, x: ArrayLike) -> torch.Tensor:
        xt, _ = self._to_dev(x)
        return self._forward(xt, training=False)

    # -------------------------------------------------------------------------
    def state(self) -> Dict[str, Any]:
        meta = self.meta()
        meta.update(global_step=self.global_step, seed=self.seed, engine="torch")
        if self.adam is not None:
            meta["optimizer"] = dict(lr=self.adam.lr, beta1=self.adam.beta1, beta2=self.adam.beta2,
                                     epsilon=self.adam.epsilon, eps_mode=self.adam.eps_mode)
        return {"tensors": {"params": self.params.data.detach().cpu().clone(),
                            "adam_m": self.m.cpu().clone(), "adam_v": self.v.cpu().clone()},
                "meta": meta}

    def load_state(self, tensors: Dict[str, torch.Tensor], meta: Dict[str, Any]) -> None:
        if meta.get("arch", self.spec.arch) != self.spec.arch or tensors["params"].numel() != self.spec.total:
            raise ValueError("checkpoint does not match this model (arch/width differ)")
        with torch.no_grad():
            self.params.data.copy_(tensors["params"].to(self.device))
            if "adam_m" in tensors:
                self.m.copy_(tensors["adam_m"].to(self.device))
                self.v.copy_(tensors["adam_v"].to(self.device))
        self.global_step = int(meta.get("global_step", meta.get("step", 0)))
        if "seed" in meta:
            self.seed = int(meta["seed"])

    def broadcast_state(self, root_rank: int = 0) -> None:
        """All 'global variables' from the root: weights, Adam slots, step, seed
        (``hvd.BroadcastGlobalVariablesHook(0)``, ``horovod/tensorflow_mnist.py:139-143``)."""
        if not _w.is_initialized() or _w.size() == 1:
            return
        with torch.no_grad():
            for t in (self.params.data, self.m, self.v):
                C.broadcast_(t, root_rank)
        self.global_step, self.seed = C.broadcast_object((self.global_step, self.seed), root_rank)


def make_engine(spec: Optional[ModelSpec] = None, optimizer=None, *, engine: str = "auto", seed: int = 0,
                **kw) -> Engine:
    """``engine``: "auto" (B200 engine on CUDA, else torch), "b200", "torch"."""
    use_cuda = torch.cuda.is_available() and (not _w.is_initialized() or _w.world().is_cuda)
    if engine == "auto":
        engine = "b200" if use_cuda else "torch"
    if engine == "b200" and spec is not None and spec.arch not in ("mnist_cnn", "keras128"):
        # architectures the sm_100a engine has no kernels for train through the PyTorch engine on the same device,
        # with the same collectives underneath
        from ..utils.logging import get_logger
        get_logger().warning("engine_fallback", arch=spec.arch, engine="torch", device="cuda" if use_cuda else "cpu")
        kw.pop("batch_size", None); kw.pop("use_graph", None); kw.pop("use_tc", None)
        return TorchEngine(spec, optimizer, seed=seed, device="cuda" if use_cuda else "cpu", **kw)
    if engine == "b200":
        from ..models.b200_engine import B200Engine
        return B200Engine(spec, optimizer, seed=seed, **kw)
    if engine == "torch":
        kw.pop("batch_size", None)
        kw.pop("use_graph", None)
        kw.pop("use_tc", None)
        return TorchEngine(spec, optimizer, seed=seed, **kw)
    raise ValueError(f"unknown engine {engine!r}")
