"""Halo rows still in flight.

``HaloExchange.augmented(x, pattern, defer=True)`` (peer-memory engine) launches the NVLink pull of the halo rows on a side
stream and returns immediately; the rows ``[n_local, n_aug)`` of the returned buffer are valid only after the CUDA event
registered here.  The caller promises to hand the buffer to a consumer that calls :func:`take` (``ops/graphconv.py`` runs the
local-row GEMM and the local-neighbour part of the aggregation first, then waits).  ``defer`` is opt-in per call: every other
consumer gets a fully populated buffer.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

_events: Dict[int, "torch.cuda.Event"] = {}


def put(t: torch.Tensor, ev) -> None:
    _events[t.data_ptr()] = ev


def take(t: torch.Tensor):
    return _events.pop(t.data_ptr(), None)


def wait_all() -> None:
    """Make the current stream wait for every outstanding pull (called before the next inter-rank barrier is enqueued)."""
    if _events:
        s = torch.cuda.current_stream()
        for ev in _events.values():
            s.wait_event(ev)
        _events.clear()
