"""CUDA-event timing scopes.

Parity: reference ``DGraph/utils/TimingReport.py:5-84`` (static registry ``name -> [ms]``, ``init(comm)``,
``start`` / ``stop`` / ``add_time`` / context-manager form).  Difference: scopes are **disabled until
``TimingReport.init`` is called** and can be turned off (``TimingReport.enabled = False``), because every reference
scope costs two barriers and a device synchronise -- instrumentation must not sit on the hot path
by default.  On CPU tensors it falls back to ``time.perf_counter``.
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional

import torch


class TimingReport:
    _timers: Dict[str, List[float]] = {}
    _open: Dict[str, object] = {}
    _comm = None
    _is_initialized = False
    enabled = False
    sync_ranks = True  # barrier around scopes like the reference; set False for low-overhead local timing

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if TimingReport.enabled:
            TimingReport.start(self.name)
        return self

    def __exit__(self, *exc):
        if TimingReport.enabled:
            TimingReport.stop(self.name)
        return False

    @staticmethod
    def init(communicator, enabled: bool = True, sync_ranks: bool = True):
        TimingReport._comm = communicator
        TimingReport._is_initialized = True
        TimingReport.enabled = enabled
        TimingReport.sync_ranks = sync_ranks
        TimingReport._timers = {}

    @staticmethod
    def _barrier():
        if TimingReport.sync_ranks and TimingReport._comm is not None:
            TimingReport._comm.barrier()

    @staticmethod
    def start(name: str):
        TimingReport._barrier()
        if torch.cuda.is_available():
            if os.environ.get("DGRAPH_B200_NVTX", "0") == "1":
                torch.cuda.nvtx.range_push(name)
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            TimingReport._open[name] = ev
        else:
            TimingReport._open[name] = time.perf_counter()

    @staticmethod
    def stop(name: str):
        st = TimingReport._open.pop(name, None)
        if st is None:
            raise ValueError(f"Timer {name} has not been started")
        if torch.cuda.is_available():
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            if os.environ.get("DGRAPH_B200_NVTX", "0") == "1":
                torch.cuda.nvtx.range_pop()
            torch.cuda.synchronize()
            ms = st.elapsed_time(ev)
        else:
            ms = (time.perf_counter() - st) * 1e3
        TimingReport._timers.setdefault(name, []).append(ms)
        TimingReport._barrier()

    @staticmethod
    def add_time(name: str, elapsed_time: float):
        """Manually add one entry (milliseconds) under ``name``."""
        TimingReport._timers.setdefault(name, []).append(elapsed_time)

    @staticmethod
    def get_timers() -> Dict[str, List[float]]:
        return TimingReport._timers

    @staticmethod
    def dump(path: str):
        with open(path, "w") as f:
            json.dump(TimingReport._timers, f)

    @staticmethod
    def reset():
        TimingReport._timers = {}
        TimingReport._open = {}
