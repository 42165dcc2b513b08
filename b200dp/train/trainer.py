"""Keras-style front end: ``Model.compile / fit / evaluate / save`` and callbacks.

Covers the surface ``tensorflow_mnist_gpu.py`` uses
(``horovod/tensorflow_mnist_gpu.py:126-191``): ``compile(optimizer, loss, metrics)``,
``fit(x, y, batch_size, epochs, steps_per_epoch, validation_data, validation_steps,
callbacks, verbose)``, ``evaluate`` -> ``[loss, accuracy]``, ``save(dir)``; callbacks
``BroadcastGlobalVariablesCallback`` (broadcast after the first batch),
``MetricAverageCallback`` (allreduce-average epoch logs; must precede callbacks that
read metrics), ``TensorBoard`` and ``ModelCheckpoint(save_best_only=True)`` for rank 0.

Like Keras given arrays + ``steps_per_epoch``, every rank iterates the FULL arrays
it was handed (the reference does not shard) with an epoch-wise
reshuffle; per-epoch validation runs on every rank.
"""
from __future__ import annotations

import math
import os
import sys
import time
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..ckpt.manager import export_model, save_tensors_atomic
from ..models.spec import ModelSpec, get_spec
from ..parallel import collectives as C
from ..parallel.optimizer import Adam, DistributedAdam
from ..runtime import world as _w
from ..utils.eventlog import EventWriter
from ..utils.logging import get_logger
from .engine import Engine, StepResult, make_engine


# ----------------------------------------------------------------------------- callbacks
class Callback:
    model: "Model" = None  # type: ignore

    def set_model(self, model: "Model") -> None:
        self.model = model

    def on_train_begin(self, logs=None): ...
    def on_train_end(self, logs=None): ...
    def on_epoch_begin(self, epoch: int, logs=None): ...
    def on_epoch_end(self, epoch: int, logs=None): ...
    def on_train_batch_end(self, batch: int, logs=None): ...


class BroadcastGlobalVariablesCallback(Callback):
    """Broadcast model + optimizer variables from ``root_rank`` once, at the end of
    the first batch (when every variable, optimizer slots included, exists) —
    ``hvd.callbacks.BroadcastGlobalVariablesCallback(0)``,
    ``horovod/tensorflow_mnist_gpu.py:149-150``."""

    def __init__(self, root_rank: int = 0):
        self.root_rank = root_rank
        self.done = False

    def on_train_batch_end(self, batch, logs=None):
        if not self.done:
            self.model.engine.broadcast_state(self.root_rank)
            self.done = True


class MetricAverageCallback(Callback):
    """Average every numeric entry of the epoch ``logs`` across ranks, in place
    (``horovod/tensorflow_mnist_gpu.py:152-153``).  One packed allreduce for all
    metrics."""

    def on_epoch_end(self, epoch, logs=None):
        if not logs or not _w.is_initialized() or _w.size() == 1:
            return
        keys = sorted(k for k, v in logs.items() if isinstance(v, (int, float)))
        dev = _w.device()
        buf = torch.tensor([float(logs[k]) for k in keys], dtype=torch.float32, device=dev)
        C.allreduce_(buf, C.Average)
        for k, v in zip(keys, buf.tolist()):
            logs[k] = v


class TensorBoard(Callback):
    def __init__(self, log_dir: str = "./logs"):
        self.log_dir = log_dir
        self.writer: Optional[EventWriter] = None

    def on_train_begin(self, logs=None):
        self.writer = EventWriter(self.log_dir)

    def on_epoch_end(self, epoch, logs=None):
        if self.writer and logs:
            self.writer.add_scalars({f"epoch_{k}": v for k, v in logs.items() if isinstance(v, (int, float))},
                                    step=epoch + 1)

    def on_train_end(self, logs=None):
        if self.writer:
            self.writer.close()
            self.writer = None


class ModelCheckpoint(Callback):
    """``filepath`` may contain ``{epoch}`` (1-based, like Keras) and metric names.
    ``save_best_only`` keeps a file only when ``monitor`` (default ``val_loss``) improves."""

    def __init__(self, filepath: str, monitor: str = "val_loss", save_best_only: bool = False,
                 mode: str = "auto", save_weights_only: bool = False):
        self.filepath = filepath
        self.monitor = monitor
        self.save_best_only = save_best_only
        if mode == "auto":
            mode = "max" if ("acc" in monitor or monitor.startswith("fmeasure")) else "min"
        self.mode = mode
        self.best = -math.inf if mode == "max" else math.inf
        self.save_weights_only = save_weights_only
        self.saved: List[str] = []

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self.save_best_only:
            cur = logs.get(self.monitor)
            if cur is None:
                get_logger().warning("checkpoint_monitor_missing", monitor=self.monitor)
                return
            better = cur > self.best if self.mode == "max" else cur < self.best
            if not better:
                return
            self.best = cur
        path = self.filepath.format(epoch=epoch + 1, **{k: v for k, v in logs.items() if isinstance(v, (int, float))})
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        st = self.model.engine.state()
        tensors = st["tensors"] if not self.save_weights_only else {"params": st["tensors"]["params"]}
        meta = dict(st["meta"], epoch=epoch + 1, **{k: v for k, v in logs.items() if isinstance(v, (int, float))})
        save_tensors_atomic(path, tensors, meta)
        self.saved.append(path)
        get_logger().info("epoch_checkpoint_saved", path=path, epoch=epoch + 1, monitor=self.monitor, value=logs.get(self.monitor))


class History(Callback):
    def __init__(self):
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


class LearningRateScheduleCallback(Callback):
    """``hvd.callbacks.LearningRateScheduleCallback``: lr = initial_lr * multiplier for epochs in [start_epoch, end_epoch).
    ``multiplier`` is a number or ``f(epoch)``; with ``staircase=False`` the epoch passed to ``f`` is fractional and the
    rate is updated every batch."""

    def __init__(self, initial_lr: float, multiplier, start_epoch: int = 0, end_epoch: Optional[int] = None,
                 staircase: bool = True, steps_per_epoch: Optional[int] = None):
        self.initial_lr = float(initial_lr)
        self.multiplier = multiplier if callable(multiplier) else (lambda epoch: multiplier)
        self.start_epoch, self.end_epoch, self.staircase = start_epoch, end_epoch, staircase
        self.steps_per_epoch = steps_per_epoch
        self._epoch = 0
        self.history: List[float] = []

    def _active(self, epoch: float) -> bool:
        return epoch >= self.start_epoch and (self.end_epoch is None or epoch < self.end_epoch)

    def _apply(self, epoch: float) -> None:
        if self._active(epoch):
            lr = self.initial_lr * float(self.multiplier(epoch))
            self.model.engine.set_lr(lr)
            self.history.append(lr)

    def on_epoch_begin(self, epoch, logs=None):
        self._epoch = epoch
        self._apply(epoch)

    def on_train_batch_end(self, batch, logs=None):
        if not self.staircase:
            spe = self.steps_per_epoch or getattr(self.model, "_steps_per_epoch", None)
            if spe:
                self._apply(self._epoch + (batch + 1) / spe)

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = self.model.engine.adam.lr


class LearningRateWarmupCallback(LearningRateScheduleCallback):
    """``hvd.callbacks.LearningRateWarmupCallback``: ramp from ``initial_lr / size`` to ``initial_lr`` over the first
    ``warmup_epochs`` epochs, batch by batch (Goyal et al.) — the companion of the lr x size rule the reference applies
    (``horovod/tensorflow_mnist.py:123-130``)."""

    def __init__(self, initial_lr: float, warmup_epochs: int = 5, steps_per_epoch: Optional[int] = None, verbose: int = 0):
        size = _w.size() if _w.is_initialized() else 1

        def mult(epoch):
            return 1.0 / size * (epoch * (size - 1) / warmup_epochs + 1)
        super().__init__(initial_lr, mult, start_epoch=0, end_epoch=warmup_epochs, staircase=False, steps_per_epoch=steps_per_epoch)
        self.verbose = verbose
        self.warmup_epochs = warmup_epochs

    def on_epoch_end(self, epoch, logs=None):
        super().on_epoch_end(epoch, logs)
        if epoch == self.warmup_epochs - 1:
            self.model.engine.set_lr(self.initial_lr)
            if self.verbose:
                get_logger().info("lr_warmup_done", epoch=epoch + 1, lr=self.initial_lr)


class MetricsCallback(Callback):
    """Feeds the Prometheus exporter (:mod:`b200dp.utils.metrics`) every ``every_n_batches`` batches and at epoch end."""

    def __init__(self, exporter, every_n_batches: int = 50):
        self.exporter = exporter
        self.every = max(1, every_n_batches)
        self._n = 0

    def on_train_batch_end(self, batch, logs=None):
        self._n += 1
        if self._n % self.every == 0:
            eng = self.model.engine
            self.exporter.observe_step(eng.global_step, getattr(self.model, "_last_batch", 0) or 0)

    def on_epoch_end(self, epoch, logs=None):
        eng = self.model.engine
        self.exporter.observe_step(eng.global_step, getattr(self.model, "_last_batch", 0) or 0,
                                   loss=(logs or {}).get("loss"), accuracy=(logs or {}).get("accuracy"))


class FaultInjectionCallback(Callback):
    def on_train_batch_end(self, batch, logs=None):
        spec = os.environ.get("B200DP_FAULT", "")
        if not spec or os.environ.get("B200DP_RESTART_COUNT", "0") != "0":
            return
        r, s = (int(v) for v in spec.split(":"))
        me = _w.rank() if _w.is_initialized() else 0
        if me == r and self.model.engine.global_step == s:
            get_logger().error("fault_injected", step=s)
            os._exit(17)


# ----------------------------------------------------------------------------- losses
class SparseCategoricalCrossentropy:
    def __init__(self, from_logits: bool = True):
        if not from_logits:
            raise NotImplementedError("only from_logits=True (as the reference uses) is implemented")
        self.from_logits = from_logits


# ----------------------------------------------------------------------------- model
class Model:
    """A compiled, trainable network.  ``MNISTModel()`` below is the reference's."""

    def __init__(self, spec: Optional[ModelSpec] = None, *, engine: str = "auto", seed: int = 0,
                 dtype: str = "bf16", **engine_kw):
        self.spec = spec or get_spec()
        self._engine_kind = engine
        self._seed = seed
        self._dtype = dtype
        self._engine_kw = engine_kw
        self.engine: Optional[Engine] = None
        self.optimizer = None
        self.metrics_names = ["loss"]
        self.stop_training = False

    def compile(self, optimizer, loss: Any = None, metrics: Optional[Sequence[str]] = None) -> None:
        if loss is not None and not isinstance(loss, (SparseCategoricalCrossentropy, str)):
            raise TypeError("loss must be SparseCategoricalCrossentropy(from_logits=True)")
        self.optimizer = optimizer
        self.metrics_names = ["loss"] + [("accuracy" if m in ("accuracy", "acc") else m) for m in (metrics or [])]
        if self.engine is None:
            self.engine = make_engine(self.spec, optimizer, engine=self._engine_kind, seed=self._seed, **self._engine_kw)
        else:
            self.engine.set_optimizer(optimizer)

    def _ensure_engine(self):
        if self.engine is None:
            self.engine = make_engine(self.spec, None, engine=self._engine_kind, seed=self._seed, **self._engine_kw)

    # -------------------------------------------------------------------------
    def fit(self, x, y, batch_size: int = 32, epochs: int = 1, steps_per_epoch: Optional[int] = None,
            validation_data: Optional[Tuple[Any, Any]] = None, validation_steps: Optional[int] = None,
            callbacks: Optional[List[Callback]] = None, verbose: int = 1, shuffle: bool = True,
            initial_epoch: int = 0) -> History:
        if self.engine is None or self.optimizer is None:
            raise RuntimeError("compile() the model before fit()")
        x = np.asarray(x)
        y = np.asarray(y)
        n = len(x)
        if steps_per_epoch is None:
            steps_per_epoch = max(1, n // batch_size)
        hist = History()
        cbs: List[Callback] = list(callbacks or []) + [hist, FaultInjectionCallback()]
        from ..utils.metrics import exporter_from_env
        exp = exporter_from_env(_w.rank(), _w.size(), _w.local_rank()) if _w.is_initialized() else exporter_from_env(0, 1, 0)
        if exp is not None and not any(isinstance(c, MetricsCallback) for c in cbs):
            cbs.append(MetricsCallback(exp))
        self._last_batch = batch_size
        self._steps_per_epoch = steps_per_epoch
        for cb in cbs:
            cb.set_model(self)
        for cb in cbs:
            cb.on_train_begin()
        rng = np.random.RandomState()       # unseeded: per-rank order, as the reference
        log = get_logger()
        order = rng.permutation(n) if shuffle else np.arange(n)
        cursor = 0
        # batches are assembled by the native loader thread (csrc/loader.cpp) while the previous step is being issued;
        # B200DP_NATIVE_LOADER=0 keeps the gather in this loop
        loader = None
        if os.environ.get("B200DP_NATIVE_LOADER", "1") != "0" and batch_size <= n and x.dtype == np.float32:
            try:
                from ..data.loader import BatchLoader
                loader = BatchLoader(x, y, batch_size, shuffle=shuffle, seed=int(rng.randint(0, 2 ** 31 - 1)))
            except (RuntimeError, OSError, ValueError):
                loader = None
        self.stop_training = False
        for epoch in range(initial_epoch, epochs):
            for cb in cbs:
                cb.on_epoch_begin(epoch)
            t0 = time.time()
            loss_sum = 0.0
            correct = 0
            seen = 0
            pending: List[StepResult] = []
            for b in range(steps_per_epoch):
                if loader is not None:
                    xb, yb = loader.next()
                    # numpy views: the engine copies them into its own staging before returning, so the slot is free again
                    res = self.engine.train_step(xb.numpy(), yb.numpy())
                else:
                    if cursor + batch_size > n:
                        order = rng.permutation(n) if shuffle else np.arange(n)
                        cursor = 0
                    idx = order[cursor:cursor + batch_size]
                    cursor += batch_size
                    res = self.engine.train_step(x[idx], y[idx])
                pending.append(res)
                if len(pending) >= 64 or b == steps_per_epoch - 1:
                    for r in pending:       # one sync for many steps
                        loss_sum += r.loss_value() * r.n
                        correct += int(r.correct)
                        seen += r.n
                    pending.clear()
                batch_logs = None
                for cb in cbs:
                    cb.on_train_batch_end(b, batch_logs)
                if verbose and (b + 1) % max(1, steps_per_epoch // 10) == 0 and seen:
                    sys.stderr.write(f"\rEpoch {epoch + 1}/{epochs} {b + 1}/{steps_per_epoch} "
                                     f"- loss: {loss_sum / seen:.4f} - accuracy: {correct / seen:.4f}")
                    sys.stderr.flush()
            logs: Dict[str, float] = {"loss": loss_sum / max(seen, 1)}
            if "accuracy" in self.metrics_names:
                logs["accuracy"] = correct / max(seen, 1)
            if validation_data is not None:
                vl = self.evaluate(validation_data[0], validation_data[1], batch_size=batch_size,
                                   steps=validation_steps, verbose=0)
                logs["val_loss"] = vl[0]
                if "accuracy" in self.metrics_names:
                    logs["val_accuracy"] = vl[1]
            dt = time.time() - t0
            # collective point (every rank runs the same epochs): bring rank-sharded optimizer state home everywhere, so
            # rank-0-only callbacks (ModelCheckpoint) and a later save()/evaluate() never read a peer's memory
            self.engine.finalize()
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
            if verbose:
                sys.stderr.write("\n")
                ws = _w.size() if _w.is_initialized() else 1
                log.info("epoch_end", epoch=epoch + 1, epochs=epochs, step=self.engine.global_step,
                         epoch_sec=round(dt, 3),
                         images_per_sec=round(steps_per_epoch * batch_size * ws / max(dt, 1e-9), 1), **logs)
            if self.stop_training:
                break
        if loader is not None:
            loader.close()
        for cb in cbs:
            cb.on_train_end()
        return hist

    def evaluate(self, x, y, batch_size: int = 32, steps: Optional[int] = None, verbose: int = 0) -> List[float]:
        self._ensure_engine()
        x = np.asarray(x)
        y = np.asarray(y)
        n = len(x)
        nb = math.ceil(n / batch_size)
        if steps is not None:
            nb = min(nb, steps)
        loss_sum = 0.0
        correct = 0
        seen = 0
        results = []
        for b in range(nb):
            sl = slice(b * batch_size, min(n, (b + 1) * batch_size))
            results.append(self.engine.eval_batch(x[sl], y[sl]))
        for r in results:
            loss_sum += r.loss_value() * r.n
            correct += int(r.correct)
            seen += r.n
        out = [loss_sum / max(seen, 1)]
        if "accuracy" in self.metrics_names or len(self.metrics_names) == 1:
            out.append(correct / max(seen, 1))
        return out

    def predict(self, x, batch_size: int = 256) -> np.ndarray:
        self._ensure_engine()
        x = np.asarray(x)
        outs = [self.engine.predict(x[i:i + batch_size]).float().cpu() for i in range(0, len(x), batch_size)]
        return torch.cat(outs).numpy()

    def train_on_batch(self, x, y) -> StepResult:
        return self.engine.train_step(x, y)

    def save(self, directory: str) -> str:
        """Final model export (weights only + description), rank-0 use."""
        self._ensure_engine()
        st = self.engine.state()
        return export_model(directory, {"params": st["tensors"]["params"]},
                            dict(st["meta"], layout=[{"name": p.name, "shape": list(p.shape), "offset": p.offset}
                                                     for p in self.spec.params]))

    def load_weights(self, path: str) -> None:
        from ..ckpt.manager import load_tensors
        self._ensure_engine()
        if os.path.isdir(path):
            path = os.path.join(path, "model.safetensors")
        tensors, meta = load_tensors(path)
        self.engine.load_state(tensors, meta)


def MNISTModel(width: float = 1.0, arch: str = "mnist_cnn", **kw) -> Model:
    """The reference's ``MNISTModel`` (``horovod/tensorflow_mnist_gpu.py:40-88``)."""
    return Model(get_spec(arch, width), **kw)
