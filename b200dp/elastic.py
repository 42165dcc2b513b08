"""Elastic-style state management (``hvd.elastic`` in Horovod).

The reference's sub-README points at an *external* elastic MPIJob manifest (``horovod/README.md:20-22``); its own
scripts recover through rank-0 checkpoints (``horovod/tensorflow_mnist.py:157-167``).  This module gives the Horovod
elastic programming model — a state object with ``commit`` / ``restore`` / ``sync`` and a ``run`` decorator — on top of
what this framework actually does when a rank dies: the launcher tears the job down at once and, with
``b200run --max-restarts N``, starts it again, and the new incarnation picks up from the last commit.  A resize works the
same way (``b200run --resize-file PATH``: the file holds the new rank count; the next commit exits with code 75 and the
launcher relaunches with that many ranks).  World membership therefore changes *between* incarnations, never inside a running process:
peer-mapped NVLink segments cannot lose a peer and keep going.

    state = hvd.elastic.EngineState(engine, epoch=0, batch=0)

    @hvd.elastic.run
    def train(state):
        for state.epoch in range(state.epoch, epochs):
            ...
            state.commit()          # in-memory snapshot; rank 0 also persists it when a directory is configured

    train(state)

Persistence: ``B200DP_ELASTIC_DIR`` (or ``State(..., directory=...)``).  Without it commits live in memory only, which
still gives ``restore()`` after a caught :class:`HorovodInternalError` but not across a relaunch.
"""
from __future__ import annotations

import copy
import functools
import os
import pickle
from typing import Any, Callable, Dict, List, Optional

import torch

from .parallel import collectives as C
from .runtime import world as _w
from .utils.logging import get_logger


class HorovodInternalError(RuntimeError):
    """A collective failed (peer lost, device-side wait timed out).  ``elastic.run`` restores the last commit and
    re-raises so the launcher's supervisor can relaunch the job."""


class HostsUpdatedInterrupt(RuntimeError):
    """Raised by ``State.check_host_updates()`` when the launcher asked for a resize (``B200DP_RESIZE_FILE`` exists)."""

    def __init__(self, skip_sync: bool = False):
        super().__init__("hosts updated")
        self.skip_sync = skip_sync


def _atomic_write(path: str, blob: bytes) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(blob)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


class State:
    """Named python values (ints, floats, small objects, tensors) that survive a failure.  ``ObjectState`` in Horovod."""

    def __init__(self, directory: Optional[str] = None, **kwargs: Any):
        self._directory = directory if directory is not None else os.environ.get("B200DP_ELASTIC_DIR")
        self._names: List[str] = list(kwargs)
        for k, v in kwargs.items():
            setattr(self, k, v)
        self._saved: Optional[Dict[str, Any]] = None
        self._reset_callbacks: List[Callable[[], None]] = []
        self.commits = 0
        self.save()

    # -- Horovod surface -----------------------------------------------------------------------------------------
    def register_reset_callbacks(self, callbacks) -> None:
        self._reset_callbacks.extend(callbacks)

    def on_reset(self) -> None:
        for cb in self._reset_callbacks:
            cb()

    def save(self) -> None:
        """Snapshot in memory (no persistence, no host check)."""
        self._saved = self._capture()

    def commit(self) -> None:
        """Snapshot, persist on rank 0 when a directory is configured, then look for a resize request."""
        self.save()
        self.commits += 1
        if self._directory and (not _w.is_initialized() or _w.rank() == 0):
            os.makedirs(self._directory, exist_ok=True)
            _atomic_write(self._path(), pickle.dumps({"commits": self.commits, "state": self._saved}, protocol=4))
        self.check_host_updates()

    def restore(self) -> None:
        if self._saved is not None:
            self._apply(self._saved)

    def sync(self) -> None:
        """Make every rank equal to rank 0 (after a (re)start: rank 0 may have loaded a persisted commit)."""
        if _w.is_initialized() and _w.size() > 1:
            self._apply(C.broadcast_object(self._capture(), 0))
        self._sync_counters()
        self.save()

    def _sync_counters(self) -> None:
        if _w.is_initialized() and _w.size() > 1:
            self.commits = int(C.broadcast_object(self.commits, 0))

    def check_host_updates(self) -> None:
        f = os.environ.get("B200DP_RESIZE_FILE")
        if f and os.path.exists(f):
            raise HostsUpdatedInterrupt()

    # -- persistence ---------------------------------------------------------------------------------------------
    def _path(self) -> str:
        return os.path.join(self._directory, "elastic_state.pkl")

    def load_persisted(self) -> bool:
        """Rank 0: adopt the last persisted commit, if any.  Returns whether one was found."""
        if not self._directory or not os.path.exists(self._path()):
            return False
        with open(self._path(), "rb") as f:
            blob = pickle.load(f)
        self._apply(blob["state"])
        self.commits = int(blob.get("commits", 0))
        self.save()
        return True

    # -- what a snapshot is --------------------------------------------------------------------------------------
    def _capture(self) -> Dict[str, Any]:
        out = {}
        for k in self._names:
            v = getattr(self, k)
            out[k] = v.detach().cpu().clone() if isinstance(v, torch.Tensor) else copy.deepcopy(v)
        return out

    def _apply(self, snap: Dict[str, Any]) -> None:
        for k in self._names:
            if k not in snap:
                continue
            cur = getattr(self, k, None)
            if isinstance(cur, torch.Tensor) and isinstance(snap[k], torch.Tensor) and cur.shape == snap[k].shape:
                cur.detach().copy_(snap[k].to(cur.device))
            else:
                setattr(self, k, copy.deepcopy(snap[k]))


ObjectState = State


class EngineState(State):
    """Training state of a :class:`b200dp.train.engine.Engine` (weights, Adam moments, step, seed) plus named values."""

    def __init__(self, engine, directory: Optional[str] = None, **kwargs: Any):
        self.engine = engine
        super().__init__(directory, **kwargs)

    def _capture(self) -> Dict[str, Any]:
        snap = super()._capture()
        st = self.engine.state()
        snap["__engine__"] = {"tensors": {k: v.clone() for k, v in st["tensors"].items()}, "meta": copy.deepcopy(st["meta"])}
        return snap

    def _apply(self, snap: Dict[str, Any]) -> None:
        super()._apply(snap)
        if "__engine__" in snap:
            self.engine.load_state(snap["__engine__"]["tensors"], snap["__engine__"]["meta"])

    def sync(self) -> None:
        self.engine.broadcast_state(0)
        if _w.is_initialized() and _w.size() > 1:
            State._apply(self, C.broadcast_object(State._capture(self), 0))
        self._sync_counters()
        self.save()


class TorchState(State):
    """``hvd.elastic.TorchState(model, optimizer, epoch=0, ...)``: a torch module and optimizer plus named values."""

    def __init__(self, model: Optional[torch.nn.Module] = None, optimizer: Optional[torch.optim.Optimizer] = None,
                 directory: Optional[str] = None, **kwargs: Any):
        self.model, self.optimizer = model, optimizer
        super().__init__(directory, **kwargs)

    def _capture(self) -> Dict[str, Any]:
        snap = super()._capture()
        if self.model is not None:
            snap["__model__"] = {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}
        if self.optimizer is not None:
            snap["__optimizer__"] = copy.deepcopy(self.optimizer.state_dict())
        return snap

    def _apply(self, snap: Dict[str, Any]) -> None:
        super()._apply(snap)
        if self.model is not None and "__model__" in snap:
            self.model.load_state_dict(snap["__model__"])
        if self.optimizer is not None and "__optimizer__" in snap:
            self.optimizer.load_state_dict(snap["__optimizer__"])

    def sync(self) -> None:
        if _w.is_initialized() and _w.size() > 1:
            if self.model is not None:
                C.broadcast_parameters(self.model.state_dict(), 0)
            if self.optimizer is not None:
                C.broadcast_optimizer_state(self.optimizer, 0)
            State._apply(self, C.broadcast_object(State._capture(self), 0))
        self._sync_counters()
        self.save()


def run(func: Callable) -> Callable:
    """Decorator: ``func(state, *args, **kwargs)`` starts from the last commit.

    On entry rank 0 adopts a persisted commit if one exists (this incarnation is a relaunch), everything is synced from
    rank 0 and the reset callbacks run when the world size differs from the one the commit was taken in.  A
    :class:`HostsUpdatedInterrupt` (resize requested at a commit point) exits with code 75 so the supervisor relaunches
    with the new size; a :class:`HorovodInternalError` restores the last commit and re-raises for the same reason."""

    @functools.wraps(func)
    def wrapper(state: State, *args, **kwargs):
        log = get_logger()
        me = _w.rank() if _w.is_initialized() else 0
        resumed = state.load_persisted() if me == 0 else False
        state.sync()
        if resumed:
            log.info("elastic_resume", commits=state.commits, restart=os.environ.get("B200DP_RESTART_COUNT", "0"))
        if os.environ.get("B200DP_RESTART_COUNT", "0") != "0":
            state.on_reset()
        try:
            return func(state, *args, **kwargs)
        except HostsUpdatedInterrupt:
            log.warning("elastic_resize_requested", commits=state.commits)
            raise SystemExit(75)
        except HorovodInternalError as e:
            log.error("elastic_failure", error=str(e))
            state.restore()
            raise
    return wrapper
