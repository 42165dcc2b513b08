"""Prometheus metrics for a training job (optional; ``B200DP_METRICS_PORT``).

The reference's observability is log shipping only — Promtail tails pod stdout into Loki and Grafana queries it
(``deploy_stack.sh:22-31``); it defines no Prometheus metrics.  Numbers that are awkward to rebuild from log
lines (a current step rate, the age of the last step, restart counts) are exposed here as gauges/counters on a plain
``/metrics`` endpoint per rank: port ``B200DP_METRICS_PORT + local_rank`` so the ranks of one host do not collide.  The
MPIJob manifests annotate the worker pods for scraping (``deploy/b200dp-mnist.yaml``).

The exporter never forces a device synchronisation of its own: it reads the loss only when the hook's cadence says so
(the same cadence as the logging hook), so the training loop stays free of per-step host round trips.
"""
from __future__ import annotations

import os
import time
from typing import Any, Dict, Optional

try:                                         # prometheus_client is in the image; the package must import without it
    import prometheus_client as _pc
except Exception:                            # pragma: no cover
    _pc = None


class MetricsExporter:
    """Gauges and counters of one rank.  ``registry=None`` uses a private registry (tests, several engines per process)."""

    def __init__(self, rank: int = 0, size: int = 1, *, labels: Optional[Dict[str, str]] = None, registry: Any = None):
        if _pc is None:
            raise RuntimeError("prometheus_client is not installed")
        self.registry = registry if registry is not None else _pc.CollectorRegistry()
        self.rank, self.size = rank, size
        base = {"rank": str(rank)}
        base.update(labels or {})
        self._labels = base
        names = list(base)
        g = lambda n, d: _pc.Gauge(n, d, names, registry=self.registry).labels(**base)
        c = lambda n, d: _pc.Counter(n, d, names, registry=self.registry).labels(**base)
        self.step = g("b200dp_global_step", "last completed training step")
        self.loss = g("b200dp_loss", "training loss at the last sampled step")
        self.accuracy = g("b200dp_accuracy", "training accuracy at the last sampled step")
        self.images_per_second = g("b200dp_images_per_second", "whole-job training throughput over the last sampling interval")
        self.step_seconds = g("b200dp_step_seconds", "mean wall time per step over the last sampling interval")
        self.last_step_unix = g("b200dp_last_step_timestamp_seconds", "unix time of the last sampled step (alert on its age: stall)")
        self.world_size = g("b200dp_world_size", "ranks in the job")
        self.restarts = g("b200dp_restart_count", "how many times the launcher has restarted this job")
        self.checkpoints = c("b200dp_checkpoints_total", "checkpoints written by this rank")
        self.steps_total = c("b200dp_steps_total", "training steps run by this rank in this incarnation")
        self.world_size.set(size)
        self.restarts.set(float(os.environ.get("B200DP_RESTART_COUNT", "0")))
        self._t: Optional[float] = None
        self._s: Optional[int] = None
        self._server = None

    # ------------------------------------------------------------------------------------------------
    def serve(self, port: int, addr: str = "0.0.0.0") -> int:
        """Start the ``/metrics`` endpoint; returns the port actually bound (``port=0`` picks a free one)."""
        res = _pc.start_http_server(port, addr=addr, registry=self.registry)
        self._server = res
        try:                                  # prometheus_client >= 0.17 returns (server, thread)
            return int(res[0].server_port)
        except Exception:
            return port

    def close(self) -> None:
        if self._server is not None:
            try:
                self._server[0].shutdown()
                self._server[0].server_close()
            except Exception:
                pass
            self._server = None

    def observe_step(self, global_step: int, batch: int, loss: Optional[float] = None, accuracy: Optional[float] = None) -> None:
        """Call at the sampling cadence (not every step): rates are computed between two calls."""
        now = time.monotonic()
        if self._t is not None and global_step > self._s:
            dt = max(now - self._t, 1e-9)
            n = global_step - self._s
            self.step_seconds.set(dt / n)
            self.images_per_second.set(n * batch * self.size / dt)
            self.steps_total.inc(n)
        self._t, self._s = now, global_step
        self.step.set(global_step)
        self.last_step_unix.set(time.time())
        if loss is not None:
            self.loss.set(loss)
        if accuracy is not None:
            self.accuracy.set(accuracy)

    def text(self) -> str:
        return _pc.generate_latest(self.registry).decode()


_exporter: Optional[MetricsExporter] = None


def exporter_from_env(rank: int, size: int, local_rank: int = 0) -> Optional[MetricsExporter]:
    """The process-wide exporter, created (and served) on first use when ``B200DP_METRICS_PORT`` is set."""
    global _exporter
    if _exporter is not None:
        return _exporter
    port = os.environ.get("B200DP_METRICS_PORT")
    if not port or _pc is None:
        return None
    _exporter = MetricsExporter(rank, size)
    _exporter.serve(int(port) + local_rank)
    return _exporter
