"""Kernel timeline from CUDA events (the counterpart of ``HOROVOD_TIMELINE``, which the reference
never enables — ``horovod/tensorflow-mnist.yaml:27-30`` forwards no HOROVOD_* variable).

``with Timeline() as tl:`` wraps every ``b200dp.ops.kernels`` entry point so that each call is
bracketed by two CUDA events on its launch stream; ``tl.summary()`` gives per-kernel device time,
``tl.save_chrome_trace(path)`` writes a ``chrome://tracing`` / Perfetto JSON (one track per stream).
Events cost ~1 µs each and serialise nothing, so overlap between streams stays visible.  Use it on
the eager path (``use_graph=False``): graph replays launch no Python-visible kernels.
"""
from __future__ import annotations

import json
from collections import defaultdict
from typing import Dict, List, Tuple

import torch

from ..ops import kernels as K

_WRAPPED = ["conv1_pool_fwd", "conv1_wgrad", "head_fwd_bwd", "head_wgrad", "unpool_relu_bwd", "maxpool_fwd", "conv_direct_fwd",
            "conv_direct_dgrad", "conv_direct_wgrad", "gemm_simple", "cast_f32_to_bf16", "bias_relu_bf16", "adam_local",
            "gemm_tc", "gemm_tc_reduce_scatter", "conv_tc_fwd", "conv_tc_dgrad", "conv_tc_wgrad", "conv1_valid_fwd",
            "conv1_valid_wgrad", "dropout_bf16"]


class Timeline:
    def __init__(self):
        self.records: List[Tuple[str, int, torch.cuda.Event, torch.cuda.Event]] = []
        self._orig = {}
        self._t0 = None

    def _wrap(self, name, fn):
        def inner(*a, **kw):
            st = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            r = fn(*a, **kw)
            e1.record(st)
            self.records.append((name, st.cuda_stream, e0, e1))
            return r
        return inner

    def __enter__(self):
        self._t0 = torch.cuda.Event(enable_timing=True)
        self._t0.record()
        for n in _WRAPPED:
            self._orig[n] = getattr(K, n)
            setattr(K, n, self._wrap(n, self._orig[n]))
        from ..parallel.arena import SymmetricArena
        self._orig["fused"] = SymmetricArena.fused_allreduce_adam
        tl = self

        def fused(arena, **kw):
            st = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            r = tl._orig["fused"](arena, **kw)
            e1.record(st)
            tl.records.append((f"fused_allreduce_adam[ch{kw.get('channel')}]", st.cuda_stream, e0, e1))
            return r
        SymmetricArena.fused_allreduce_adam = fused
        return self

    def __exit__(self, *exc):
        from ..parallel.arena import SymmetricArena
        SymmetricArena.fused_allreduce_adam = self._orig.pop("fused")
        for n, f in self._orig.items():
            setattr(K, n, f)
        self._orig.clear()
        torch.cuda.synchronize()
        return False

    def events(self):
        out = []
        for name, stream, e0, e1 in self.records:
            out.append({"name": name, "stream": stream, "start_us": self._t0.elapsed_time(e0) * 1e3, "dur_us": e0.elapsed_time(e1) * 1e3})
        return out

    def summary(self) -> Dict[str, Dict[str, float]]:
        agg = defaultdict(list)
        for ev in self.events():
            agg[ev["name"]].append(ev["dur_us"])
        return {k: {"calls": len(v), "mean_us": sum(v) / len(v), "min_us": min(v), "total_us": sum(v)} for k, v in agg.items()}

    def save_chrome_trace(self, path: str) -> None:
        streams = {}
        tr = []
        for ev in self.events():
            tid = streams.setdefault(ev["stream"], len(streams))
            tr.append({"name": ev["name"], "ph": "X", "ts": ev["start_us"], "dur": ev["dur_us"], "pid": 0, "tid": tid, "cat": "kernel"})
        with open(path, "w") as f:
            json.dump({"traceEvents": tr, "displayTimeUnit": "ns"}, f)


class HostTimeline:
    """Host-side spans (``time.perf_counter``) for the CPU/gloo engine: forward, backward, each gradient allreduce, the
    optimizer.  This is what Horovod's timeline shows per tensor (NEGOTIATE / ALLREDUCE phases); there is no negotiation
    here, so a collective is one span.  Same Chrome-trace output as :class:`Timeline`; ``pid`` is the rank."""

    def __init__(self, rank: int = 0):
        import time
        self._now = time.perf_counter
        self._t0 = self._now()
        self.rank = rank
        self.records: List[Tuple[str, str, float, float]] = []

    def span(self, name: str, cat: str = "host"):
        tl = self

        class _Span:
            def __enter__(self_inner):
                self_inner.t = tl._now()
                return self_inner

            def __exit__(self_inner, *exc):
                tl.records.append((name, cat, (self_inner.t - tl._t0) * 1e6, (tl._now() - self_inner.t) * 1e6))
                return False
        return _Span()

    def summary(self) -> Dict[str, Dict[str, float]]:
        agg = defaultdict(list)
        for name, _, _, dur in self.records:
            agg[name].append(dur)
        return {k: {"calls": len(v), "mean_us": sum(v) / len(v), "min_us": min(v), "total_us": sum(v)} for k, v in agg.items()}

    def save_chrome_trace(self, path: str) -> None:
        cats: Dict[str, int] = {}
        tr = [{"name": n, "ph": "X", "ts": ts, "dur": dur, "pid": self.rank, "tid": cats.setdefault(c, len(cats)), "cat": c}
              for n, c, ts, dur in self.records]
        with open(path, "w") as f:
            json.dump({"traceEvents": tr, "displayTimeUnit": "ns"}, f)
