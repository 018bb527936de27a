"""Tracing / profiling hooks (SURVEY §5 "Tracing / profiling").

The reference has four ad-hoc mechanisms: ``TimingReport`` scopes (``DGraph/utils/TimingReport.py:44-75``), manual CUDA events
(``experiments/OGB/main.py:134-157``), one ``nvtx.annotate`` (``GraphCast/microbenchmark_graphcast.py:126``) and a
``torch.profiler`` pass under ``--benchmark`` (``train_graphcast.py:161-169``).  Here they are one module:

* :func:`nvtx_range` -- NVTX push/pop (no-op without CUDA) so Nsight tools show model stages;
  ``DGRAPH_B200_NVTX=1`` also wraps every :class:`~dgraph_b200.utils.timing.TimingReport` scope.
* :class:`KernelTrace` -- per-launch trace of *our* native kernels (name, host time, optional CUDA-event duration),
  fed by the post-launch hook of ``ops/_native.check``; ``summary()`` aggregates per kernel, ``to_chrome_trace()``
  writes a ``chrome://tracing`` / Perfetto JSON.
* :func:`profile_step` -- ``torch.profiler`` wrapper used by the trainers' ``--benchmark`` flag.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from collections import defaultdict
from typing import Callable, Dict, List, Optional

import torch

from dgraph_b200.ops import _native as N


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


def nvtx_enabled() -> bool:
    return os.environ.get("DGRAPH_B200_NVTX", "0") == "1"


class KernelTrace:
    """``with KernelTrace(cuda_events=True) as tr: step()`` then ``tr.summary()``.

    With ``cuda_events`` every launch is bracketed by the previous launch's event, so ``gpu_ms`` is the stream time between
    consecutive native launches (includes any ATen kernels in between -- it is a timeline, not an isolated kernel time; use
    ncu for that).  Event recording perturbs nothing on the device but costs ~2 us of host time per launch.
    """

    def __init__(self, cuda_events: bool = False):
        self.cuda_events = cuda_events and torch.cuda.is_available()
        self.records: List[dict] = []
        self._events: List = []
        self._t0 = 0.0

    def _hook(self, what: str):
        rec = {"name": what, "host_us": (time.perf_counter() - self._t0) * 1e6}
        if self.cuda_events:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._events.append(ev)
        self.records.append(rec)

    def __enter__(self):
        self._t0 = time.perf_counter()
        if self.cuda_events:
            self._start = torch.cuda.Event(enable_timing=True)
            self._start.record()
        N._post_launch_hooks.append(self._hook)
        return self

    def __exit__(self, *exc):
        N._post_launch_hooks.remove(self._hook)
        if self.cuda_events:
            torch.cuda.synchronize()
            prev = self._start
            t = 0.0
            for rec, ev in zip(self.records, self._events):
                d = prev.elapsed_time(ev)
                rec["gpu_start_ms"], rec["gpu_ms"] = t, d
                t += d
                prev = ev
        return False

    def summary(self) -> Dict[str, dict]:
        agg: Dict[str, dict] = defaultdict(lambda: {"launches": 0, "gpu_ms": 0.0})
        for r in self.records:
            a = agg[r["name"]]
            a["launches"] += 1
            a["gpu_ms"] += r.get("gpu_ms", 0.0)
        return dict(sorted(agg.items(), key=lambda kv: -kv[1]["gpu_ms"]))

    def to_chrome_trace(self, path: str, pid: int = 0):
        ev = []
        for r in self.records:
            if "gpu_ms" in r:
                ev.append({"name": r["name"], "ph": "X", "pid": pid, "tid": "gpu", "ts": r["gpu_start_ms"] * 1e3,
                           "dur": r["gpu_ms"] * 1e3})
            ev.append({"name": r["name"], "ph": "i", "pid": pid, "tid": "host", "ts": r["host_us"], "s": "t"})
        with open(path, "w") as f:
            json.dump({"traceEvents": ev}, f)


def profile_step(fn: Callable, out: Optional[str] = None, with_cuda: Optional[bool] = None, row_limit: int = 25) -> str:
    """One call of ``fn`` under ``torch.profiler`` (CPU + memory, plus CUDA when available); returns the table and optionally
    writes a chrome trace.  Numbers taken under the profiler are diagnostics, never benchmark values."""
    from torch.profiler import ProfilerActivity, profile

    acts = [ProfilerActivity.CPU]
    if with_cuda if with_cuda is not None else torch.cuda.is_available():
        acts.append(ProfilerActivity.CUDA)
    with profile(activities=acts, profile_memory=True, record_shapes=False) as prof:
        fn()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
    if out:
        prof.export_chrome_trace(out)
    key = "cuda_time_total" if len(acts) > 1 else "cpu_time_total"
    return prof.key_averages().table(sort_by=key, row_limit=row_limit)
