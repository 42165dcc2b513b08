"""Host-side stall watchdog — the counterpart of Horovod's stall inspector ([upstream]: warns when ranks have not
agreed on a tensor for 60 s).  With a static bucket plan there is nothing to negotiate, so a stall can only mean
a rank stopped stepping; the device side already bounds every flag wait (``csrc/comm.cu``: timeout -> diagnostic ->
trap).  This thread covers the host side: if ``beat()`` has not been called for ``warn_secs`` it logs which step
the rank is stuck at (every ``warn_secs``), and after ``abort_secs`` (optional) it terminates the process so the
launcher's fail-fast teardown and restart-from-checkpoint take over.
"""
from __future__ import annotations

import os
import threading
import time
from typing import Optional

from .logging import get_logger


class StallWatchdog:
    def __init__(self, warn_secs: float = 60.0, abort_secs: Optional[float] = None, poll_secs: float = 1.0):
        self.warn_secs, self.abort_secs, self.poll_secs = warn_secs, abort_secs, poll_secs
        self._last = time.monotonic()
        self._step = 0
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.warnings = 0

    def beat(self, step: int) -> None:
        self._last = time.monotonic()
        self._step = step

    def start(self) -> "StallWatchdog":
        self._thread = threading.Thread(target=self._run, name="b200dp-watchdog", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2 * self.poll_secs + 1)

    def _run(self) -> None:
        next_warn = self.warn_secs
        while not self._stop.wait(self.poll_secs):
            idle = time.monotonic() - self._last
            if idle < self.warn_secs:
                next_warn = self.warn_secs
                continue
            if idle >= next_warn:
                self.warnings += 1
                get_logger().warning("stall", step=self._step, idle_sec=round(idle, 1),
                                     hint="a peer may have died or a collective is waiting; see the launcher log")
                next_warn += self.warn_secs
            if self.abort_secs is not None and idle >= self.abort_secs:
                get_logger().error("stall_abort", step=self._step, idle_sec=round(idle, 1))
                os._exit(86)


class StallWatchdogHook:
    """Session hook form: beats once per training step."""

    def __init__(self, warn_secs: float = 60.0, abort_secs: Optional[float] = None):
        self.wd = StallWatchdog(warn_secs, abort_secs)

    def begin(self, session):
        self.wd.start()

    def after_create_session(self, session):
        self.wd.beat(session.engine.global_step)

    def before_run(self, session):
        pass

    def after_run(self, session, result):
        self.wd.beat(session.engine.global_step)

    def end(self, session):
        self.wd.stop()
