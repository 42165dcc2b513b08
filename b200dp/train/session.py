"""TF1-style monitored training loop with hooks.

Equivalent of ``tf.train.MonitoredTrainingSession(checkpoint_dir, hooks, config)``
plus the three hooks the reference installs (``horovod/tensorflow_mnist.py:138-171``):

* :class:`BroadcastGlobalVariablesHook` — after session creation (i.e. after init
  or checkpoint restore on rank 0) broadcast every variable from the root;
* :class:`StopAtStepHook` — ``last_step = num_steps // size``;
* :class:`LoggingTensorHook` — ``step``/``loss`` every N iterations on EVERY rank.

With a ``checkpoint_dir`` (rank 0) the session also restores the latest checkpoint
on entry, saves every ``save_checkpoint_secs`` (600 s default, TF's) and at exit,
and reports ``global_step/sec`` every 100 steps — what TF's chief session adds
implicitly.  "Closes when done or an error occurs" (reference comment ``:162-164``):
``__exit__`` saves on clean exit only and always releases resources.
"""
from __future__ import annotations

import time
from typing import Any, Dict, Iterable, List, Optional

from ..ckpt.manager import CheckpointManager
from ..runtime import world as _w
from ..utils.logging import get_logger
from .engine import Engine, StepResult


class SessionRunHook:
    def begin(self, session: "MonitoredTrainingSession") -> None: ...
    def after_create_session(self, session: "MonitoredTrainingSession") -> None: ...
    def before_run(self, session: "MonitoredTrainingSession") -> None: ...
    def after_run(self, session: "MonitoredTrainingSession", result: StepResult) -> None: ...
    def end(self, session: "MonitoredTrainingSession") -> None: ...


class BroadcastGlobalVariablesHook(SessionRunHook):
    def __init__(self, root_rank: int = 0):
        self.root_rank = root_rank

    def after_create_session(self, session):
        session.engine.broadcast_state(self.root_rank)


class StopAtStepHook(SessionRunHook):
    def __init__(self, num_steps: Optional[int] = None, last_step: Optional[int] = None):
        if (num_steps is None) == (last_step is None):
            raise ValueError("exactly one of num_steps / last_step")
        self._num_steps, self._last_step = num_steps, last_step

    def after_create_session(self, session):
        if self._last_step is None:
            self._last_step = session.engine.global_step + self._num_steps
        if session.engine.global_step >= self._last_step:
            session.request_stop()

    def after_run(self, session, result):
        if session.engine.global_step >= self._last_step:
            session.request_stop()


class LoggingTensorHook(SessionRunHook):
    """Logs ``step`` and ``loss`` every ``every_n_iter`` local iterations (first one
    included, as TF does)."""

    def __init__(self, tensors: Optional[Iterable[str]] = None, every_n_iter: int = 10):
        self.every = max(1, every_n_iter)
        self.tensors = tuple(tensors) if tensors else ("step", "loss")
        self._iter = 0

    def before_run(self, session):
        # the step that will be logged also samples the device time of its collectives (engines that can: CUDA events
        # around the comm kernels of one eagerly launched step)
        if self._iter % self.every == 0 and hasattr(session.engine, "sample_comm_timing"):
            session.engine.sample_comm_timing = True

    def after_run(self, session, result):
        if self._iter % self.every == 0:
            extra: Dict[str, Any] = {"batch": int(result.n)}
            if "accuracy" in self.tensors:
                extra["accuracy"] = result.accuracy()
            eng = session.engine
            if getattr(eng, "adam", None) is not None:
                extra["lr"] = float(eng.adam.lr)
            ms = getattr(eng, "last_comm_ms", None)
            ms = ms() if callable(ms) else ms
            if ms is not None:
                extra["allreduce_ms"] = round(float(ms), 4)
            get_logger().step(eng.global_step, result.loss_value(), **extra)
        self._iter += 1


class StepCounterHook(SessionRunHook):
    """``global_step/sec`` every N steps — the only throughput signal the reference emits."""

    def __init__(self, every_n_steps: int = 100, batch_size: Optional[int] = None):
        self.every = every_n_steps
        self.batch_size = batch_size
        self._t = None
        self._s = None

    def after_run(self, session, result):
        s = session.engine.global_step
        if self._t is None:
            self._t, self._s = time.monotonic(), s
            return
        if s - self._s >= self.every:
            now = time.monotonic()
            rate = (s - self._s) / max(now - self._t, 1e-9)
            f = {"global_step_per_sec": round(rate, 3), "step": s}
            if result.n:
                f["images_per_sec"] = round(rate * result.n * (_w.size() if _w.is_initialized() else 1), 1)
            get_logger().info("step_rate", **f)
            self._t, self._s = now, s


class CheckpointSaverHook(SessionRunHook):
    def __init__(self, manager: CheckpointManager):
        self.manager = manager

    @staticmethod
    def _count(session) -> None:
        if session.metrics is not None:
            session.metrics.checkpoints.inc()

    def after_run(self, session, result):
        eng = session.engine
        if hasattr(eng, "state_async"):
            # periodic saves do not stall the job: device-side snapshot now, drain + file write on a background thread
            if self.manager.maybe_save_async(eng.state_async, eng.global_step):
                get_logger().info("checkpoint_started", step=eng.global_step)
                self._count(session)
            return
        p = self.manager.maybe_save(eng.state, eng.global_step)
        if p:
            get_logger().info("checkpoint_saved", path=p, step=eng.global_step)
            self._count(session)

    def end(self, session):
        if self.manager.enabled:
            self.manager.wait()
            p = self.manager.save(session.engine.state(), session.engine.global_step)
            get_logger().info("checkpoint_saved", path=p, step=session.engine.global_step, final=True)
            self._count(session)


class MetricsHook(SessionRunHook):
    """Feeds the Prometheus exporter (:mod:`b200dp.utils.metrics`) every ``every_n_iter`` iterations — the cadence at
    which the loss is read back anyway, so the exporter adds no device synchronisation of its own."""

    def __init__(self, exporter, every_n_iter: int = 10):
        self.exporter = exporter
        self.every = max(1, every_n_iter)
        self._iter = 0

    def after_run(self, session, result):
        if self._iter % self.every == 0:
            self.exporter.observe_step(session.engine.global_step, result.n, loss=result.loss_value(), accuracy=result.accuracy())
        self._iter += 1


class FaultInjectionHook(SessionRunHook):
    """Test hook: ``B200DP_FAULT="rank:step"`` makes that rank die at that step
    (exercises the launcher's fail-fast teardown and restart-from-checkpoint)."""

    def __init__(self, spec: Optional[str] = None):
        import os
        spec = spec if spec is not None else os.environ.get("B200DP_FAULT", "")
        self.rank = self.step = None
        if spec:
            r, s = spec.split(":")
            self.rank, self.step = int(r), int(s)

    def after_run(self, session, result):
        import os
        if self.rank is None:
            return
        me = _w.rank() if _w.is_initialized() else 0
        if me == self.rank and session.engine.global_step == self.step and os.environ.get("B200DP_RESTART_COUNT", "0") == "0":
            get_logger().error("fault_injected", step=self.step)
            os._exit(17)


class MonitoredTrainingSession:
    def __init__(self, engine: Engine, checkpoint_dir: Optional[str] = None,
                 hooks: Optional[List[SessionRunHook]] = None, *, save_checkpoint_secs: Optional[float] = 600.0,
                 save_checkpoint_steps: Optional[int] = None, keep_checkpoint_max: int = 5,
                 log_step_count_steps: Optional[int] = 100, config: Any = None):
        self.engine = engine
        self.hooks: List[SessionRunHook] = list(hooks or [])
        self.manager = CheckpointManager(checkpoint_dir, keep=keep_checkpoint_max,
                                         save_secs=save_checkpoint_secs, save_steps=save_checkpoint_steps)
        if self.manager.enabled:
            self.hooks.append(CheckpointSaverHook(self.manager))
            if log_step_count_steps:
                self.hooks.append(StepCounterHook(log_step_count_steps))
        self.hooks.append(FaultInjectionHook())
        # B200DP_METRICS_PORT=<port>: a Prometheus /metrics endpoint per rank (port + local_rank)
        from ..utils.metrics import exporter_from_env
        me = (_w.rank(), _w.size(), _w.local_rank()) if _w.is_initialized() else (0, 1, 0)
        self.metrics = exporter_from_env(*me)
        if self.metrics is not None and not any(isinstance(h, MetricsHook) for h in self.hooks):
            self.hooks.append(MetricsHook(self.metrics))
        self._stop = False
        self._entered = False
        self.last_result: Optional[StepResult] = None
        self._timeline = None

    # -------------------------------------------------------------------------
    def __enter__(self) -> "MonitoredTrainingSession":
        log = get_logger()
        for h in self.hooks:
            h.begin(self)
        if self.manager.enabled:
            restored = self.manager.restore()
            if restored is not None:
                self.engine.load_state(restored["tensors"], restored["meta"])
                log.info("checkpoint_restored", path=restored["path"], step=self.engine.global_step)
        for h in self.hooks:
            h.after_create_session(self)
        # B200DP_TIMELINE=<path> (Horovod: HOROVOD_TIMELINE): Chrome-trace of every kernel of the run, one file per rank.
        # Tracing needs Python-visible launches, so the step runs eagerly instead of from its CUDA graph.
        import os
        tl_path = os.environ.get("B200DP_TIMELINE") or os.environ.get("HOROVOD_TIMELINE")
        if tl_path and hasattr(self.engine, "use_graph"):
            from ..utils.timeline import Timeline
            self.engine.use_graph = False
            self._timeline = (Timeline(), tl_path)
            self._timeline[0].__enter__()
        elif tl_path and hasattr(self.engine, "host_timeline"):
            # CPU/gloo engine: host-side spans (forward, backward, every collective, the optimizer)
            from ..utils.timeline import HostTimeline
            self.engine.host_timeline = HostTimeline(_w.rank() if _w.is_initialized() else 0)
            self._timeline = (self.engine.host_timeline, tl_path)
        self._entered = True
        return self

    def __exit__(self, exc_type, exc, tb) -> bool:
        if self._timeline is not None:
            tl, path = self._timeline
            self._timeline = None
            if hasattr(tl, "__exit__"):
                tl.__exit__(None, None, None)
            else:
                self.engine.host_timeline = None
            r = _w.rank() if _w.is_initialized() else 0
            out = path if path.endswith(".json") and r == 0 else f"{path}.rank{r}.json"
            try:
                tl.save_chrome_trace(out)
                get_logger().info("timeline_saved", path=out, kernels=len(tl.records))
            except OSError as e:
                get_logger().warning("timeline_save_failed", error=str(e))
        try:
            if exc_type is None:
                # every rank leaves the loop at the same global step: a collective point.  Sharded optimizer state is
                # gathered everywhere before rank 0's final save, which then needs no peer (peers may shut down first).
                self.engine.finalize()
                for h in self.hooks:
                    h.end(self)
            else:
                get_logger().error("session_error", error=repr(exc))
        finally:
            self._entered = False
        return False

    def should_stop(self) -> bool:
        return self._stop

    def request_stop(self) -> None:
        self._stop = True

    def run(self, x, y) -> StepResult:
        """One synchronous training step on this rank's minibatch."""
        if self._stop:
            raise RuntimeError("run() called after should_stop()")
        for h in self.hooks:
            h.before_run(self)
        res = self.engine.train_step(x, y)
        self.last_result = res
        for h in self.hooks:
            h.after_run(self, res)
        return res
