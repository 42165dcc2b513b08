"""Rank-0 checkpointing: periodic + at-exit saves, ``latest`` pointer, keep-last-K,
auto-resume, best-only epoch checkpoints and final model export.

Reference behaviour being reproduced:

* ``checkpoint_dir='./checkpoints' if hvd.rank() == 0 else None`` — only worker 0
  writes (``horovod/tensorflow_mnist.py:157-159``);
* ``MonitoredTrainingSession`` restores the latest checkpoint at start, saves
  every 600 s and at exit; weights + Adam slots + ``global_step``
  (``horovod/tensorflow_mnist.py:162-167``); other ranks get the restored state
  from the rank-0 broadcast (``:139-143``);
* Keras ``ModelCheckpoint('./checkpoints/mnist-{epoch}.h5', save_best_only=True)``
  and ``model.save('./final_model')`` on rank 0
  (``horovod/tensorflow_mnist_gpu.py:160-163,190-191``).

Format: ``<dir>/ckpt-<step>.safetensors`` (tensors) with the JSON metadata in the
safetensors header; written to a temp name, fsync'ed, renamed (atomic on POSIX);
``<dir>/latest`` holds the newest file name and is replaced the same way.
"""
from __future__ import annotations

import json
import os
import re
import time
from typing import Any, Dict, List, Optional, Tuple

import torch
from safetensors.torch import load_file, save_file
from safetensors import safe_open

_CKPT_RE = re.compile(r"^ckpt-(\d+)\.safetensors$")


def _fsync_dir(path: str) -> None:
    try:
        fd = os.open(path, os.O_RDONLY)
        try:
            os.fsync(fd)
        finally:
            os.close(fd)
    except OSError:
        pass


def atomic_write_bytes(path: str, data: bytes) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(data)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    _fsync_dir(os.path.dirname(path) or ".")


def save_tensors_atomic(path: str, tensors: Dict[str, torch.Tensor], meta: Dict[str, Any]) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    cpu = {k: v.detach().to("cpu").contiguous() for k, v in tensors.items()}
    save_file(cpu, tmp, metadata={"b200dp": json.dumps(meta, sort_keys=True)})
    with open(tmp, "rb") as f:
        os.fsync(f.fileno())
    os.replace(tmp, path)
    _fsync_dir(os.path.dirname(path) or ".")


def load_tensors(path: str) -> Tuple[Dict[str, torch.Tensor], Dict[str, Any]]:
    tensors = load_file(path)
    with safe_open(path, framework="pt") as f:
        md = f.metadata() or {}
    meta = json.loads(md.get("b200dp", "{}"))
    return tensors, meta


class CheckpointManager:
    """Owns one checkpoint directory.  Construct it with ``directory=None`` on
    non-zero ranks — every method is then a no-op, exactly the reference's
    ``checkpoint_dir=None`` convention."""

    def __init__(self, directory: Optional[str], *, keep: int = 5, save_secs: Optional[float] = 600.0,
                 save_steps: Optional[int] = None):
        self.dir = directory
        self.keep = keep
        self.save_secs = save_secs
        self.save_steps = save_steps
        self._last_save_t = time.monotonic()
        self._last_save_step: Optional[int] = None
        if self.dir:
            os.makedirs(self.dir, exist_ok=True)

    @property
    def enabled(self) -> bool:
        return bool(self.dir)

    # -- discovery -----------------------------------------------------------
    def list(self) -> List[Tuple[int, str]]:
        if not self.enabled or not os.path.isdir(self.dir):
            return []
        out = []
        for fn in os.listdir(self.dir):
            m = _CKPT_RE.match(fn)
            if m:
                out.append((int(m.group(1)), os.path.join(self.dir, fn)))
        return sorted(out)

    def latest(self) -> Optional[str]:
        """Path of the newest complete checkpoint (pointer file first, directory
        scan as fallback so a lost pointer never loses the run)."""
        if not self.enabled:
            return None
        ptr = os.path.join(self.dir, "latest")
        if os.path.exists(ptr):
            name = open(ptr).read().strip()
            p = os.path.join(self.dir, name)
            if name and os.path.exists(p):
                return p
        found = self.list()
        return found[-1][1] if found else None

    # -- save / load ---------------------------------------------------------
    def save(self, state: Dict[str, Any], step: int) -> Optional[str]:
        """``state``: ``{"tensors": {name: Tensor}, "meta": {...}}``."""
        if not self.enabled:
            return None
        meta = dict(state.get("meta", {}))
        meta["step"] = int(step)
        meta["saved_at"] = time.time()
        path = os.path.join(self.dir, f"ckpt-{int(step)}.safetensors")
        save_tensors_atomic(path, state["tensors"], meta)
        atomic_write_bytes(os.path.join(self.dir, "latest"), os.path.basename(path).encode())
        self._last_save_t = time.monotonic()
        self._last_save_step = int(step)
        self._prune()
        return path

    def _prune(self) -> None:
        if self.keep and self.keep > 0:
            found = self.list()
            for _, p in found[:-self.keep]:
                try:
                    os.remove(p)
                except OSError:
                    pass

    def restore(self) -> Optional[Dict[str, Any]]:
        """Load the newest checkpoint that can be read.  Writes are atomic, so a half-written file never carries the
        final name; this guards against the rest (bit rot or truncation on shared storage): an unreadable checkpoint is
        reported and the next older one is tried, so one bad file costs ``save_secs`` of work instead of the run."""
        first = self.latest()
        if first is None:
            return None
        candidates = [first] + [p for _, p in reversed(self.list()) if p != first]
        last_err: Optional[Exception] = None
        for p in candidates:
            try:
                tensors, meta = load_tensors(p)
            except Exception as e:                     # safetensors raises several types for damaged files
                last_err = e
                from ..utils.logging import get_logger
                get_logger().warning("checkpoint_unreadable", path=p, error=repr(e))
                continue
            return {"tensors": tensors, "meta": meta, "path": p}
        raise RuntimeError(f"no readable checkpoint in {self.dir}: {last_err!r}")

    def due(self, step: int) -> bool:
        if not self.enabled or self._last_save_step == step:
            return False
        if self.save_steps and step > 0 and step % self.save_steps == 0:
            return True
        if self.save_secs is not None and time.monotonic() - self._last_save_t >= self.save_secs:
            return True
        return False

    def maybe_save(self, state_fn, step: int) -> Optional[str]:
        if self.due(step):
            return self.save(state_fn(), step)
        return None

    # -- asynchronous saves ---------------------------------------------------
    def save_async(self, pending, step: int) -> None:
        """``pending``: a handle whose ``result()`` returns the state dict (``B200Engine.state_async``).  The file is
        written by a background thread — the training thread does not wait for the device-to-host drain, the
        serialisation or the fsync.  One save is in flight at a time; ``wait()`` joins it."""
        if not self.enabled:
            return
        self.wait()
        self._last_save_t = time.monotonic()
        self._last_save_step = int(step)

        def run():
            try:
                self._async_path = self.save(pending.result(), step)
            except Exception as e:          # surfaced by wait()
                self._async_err = e
        import threading
        self._async_err = None
        self._writer = threading.Thread(target=run, name="b200dp-ckpt-writer", daemon=False)
        self._writer.start()

    def wait(self) -> Optional[str]:
        """Join the save in flight (if any); re-raises its error."""
        w = getattr(self, "_writer", None)
        if w is not None:
            w.join()
            self._writer = None
            err = getattr(self, "_async_err", None)
            if err is not None:
                self._async_err = None
                raise err
            return getattr(self, "_async_path", None)
        return None

    def maybe_save_async(self, pending_fn, step: int) -> bool:
        if self.due(step):
            self.save_async(pending_fn(), step)
            return True
        return False


def export_model(directory: str, tensors: Dict[str, torch.Tensor], meta: Dict[str, Any]) -> str:
    """Final inference export (reference: ``model.save('./final_model')``):
    a directory with the weights (no optimizer state) and a JSON description."""
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, "model.safetensors")
    save_tensors_atomic(path, tensors, meta)
    atomic_write_bytes(os.path.join(directory, "model.json"),
                       json.dumps(meta, indent=1, sort_keys=True).encode())
    return path
