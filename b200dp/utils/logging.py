"""Per-rank structured stdout logging that Promtail/Loki can ingest.

The reference logs free text: TF INFO lines every 10 steps on every rank
(``horovod/tensorflow_mnist.py:27,148-149``), a Keras progress bar on rank 0
(``horovod/tensorflow_mnist_gpu.py:181``); ``mpirun`` funnels all of it to the
launcher pod's stdout, where Promtail tails it into Loki (``deploy_stack.sh:25-31``).

Here every rank writes ONE JSON object per line (so a Promtail
``pipeline_stages: - json`` stage can lift ``rank``, ``step``, ``loss`` into
labels/fields — see ``deploy/promtail-pipeline.yaml``) and, optionally, the
TF-style text line next to it for eyes and for greps written against the reference.
"""
from __future__ import annotations

import json
import os
import socket
import sys
import threading
import time
from typing import Any, Dict, Optional, TextIO

_LEVELS = {"DEBUG": 10, "INFO": 20, "WARNING": 30, "ERROR": 40}


class JsonLogger:
    def __init__(self, rank: int = 0, size: int = 1, local_rank: int = 0, *, stream: Optional[TextIO] = None,
                 file: Optional[str] = None, level: Optional[str] = None, tf_style: Optional[bool] = None):
        self.rank, self.size, self.local_rank = rank, size, local_rank
        self.host = socket.gethostname()
        self.stream = stream if stream is not None else sys.stdout
        lvl = level or os.environ.get("B200DP_LOG_LEVEL") or os.environ.get("HOROVOD_LOG_LEVEL") or "INFO"
        self.level = _LEVELS.get(lvl.upper(), 20)
        if tf_style is None:
            tf_style = os.environ.get("B200DP_LOG_TF_STYLE", "1") == "1"
        self.tf_style = tf_style
        self._file = open(file, "a", buffering=1) if file else None
        self._lock = threading.Lock()
        self._last_step_t: Optional[float] = None

    def log(self, event: str, level: str = "INFO", **fields: Any) -> Dict[str, Any]:
        rec: Dict[str, Any] = {
            "ts": round(time.time(), 6), "level": level, "event": event,
            "rank": self.rank, "local_rank": self.local_rank, "size": self.size, "host": self.host,
        }
        for k, v in fields.items():
            if hasattr(v, "item"):
                try:
                    v = v.item()
                except Exception:
                    v = str(v)
            rec[k] = v
        if _LEVELS.get(level, 20) < self.level:
            return rec
        line = json.dumps(rec, separators=(",", ":"), allow_nan=True, default=str)
        with self._lock:
            self.stream.write(line + "\n")
            self.stream.flush()
            if self._file:
                self._file.write(line + "\n")
        return rec

    def info(self, event: str, **f):
        return self.log(event, "INFO", **f)

    def warning(self, event: str, **f):
        return self.log(event, "WARNING", **f)

    def error(self, event: str, **f):
        return self.log(event, "ERROR", **f)

    def step(self, step: int, loss: float, **fields: Any) -> None:
        """The every-N-steps training line: JSON plus TF's
        ``INFO:tensorflow:loss = …, step = … (… sec)`` text form on stderr."""
        now = time.time()
        dt = None if self._last_step_t is None else now - self._last_step_t
        prev = getattr(self, "_last_step_n", None)
        self._last_step_t, self._last_step_n = now, int(step)
        if dt is not None and prev is not None and int(step) > prev:
            # wall time per step since the previous line (the reference prints the interval: "(x.xxx sec)")
            fields.setdefault("step_ms", round(dt * 1e3 / (int(step) - prev), 4))
            if "batch" in fields:
                fields.setdefault("images_per_sec", round(fields["batch"] * self.size * (int(step) - prev) / max(dt, 1e-9), 1))
        self.info("train_step", step=int(step), loss=float(loss), **fields)
        if self.tf_style and self.level <= 20:
            tail = f" ({dt:.3f} sec)" if dt is not None else ""
            sys.stderr.write(f"INFO:tensorflow:loss = {float(loss):.7g}, step = {int(step)}{tail}\n")
            sys.stderr.flush()

    def close(self):
        if self._file:
            self._file.close()
            self._file = None


_default: Optional[JsonLogger] = None


def get_logger() -> JsonLogger:
    global _default
    if _default is None:
        try:
            from ..runtime import world as _w
            w = _w.world()
            _default = JsonLogger(w.rank, w.size, w.local_rank, file=os.environ.get("B200DP_LOG_FILE"))
        except Exception:
            _default = JsonLogger()
    return _default


def reset_logger() -> None:
    global _default
    if _default is not None:
        _default.close()
    _default = None
