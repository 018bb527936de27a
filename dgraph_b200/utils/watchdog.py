"""Failure detection (SURVEY §5 "Failure detection": the reference has no heartbeats, timeouts or retries).

* Device side: every peer-memory barrier spins a bounded number of iterations (``DGRAPH_B200_BARRIER_SPINS``) and raises a
  flag instead of hanging the GPU; ``PeerMemoryEngine.check_timeouts()`` turns it into an exception (csrc/peer.cu).
* Host side, this module:
  - :class:`StepWatchdog` -- a daemon thread that fires ``on_timeout`` (default: dump all Python stacks and ``os._exit``)
    when ``beat()`` has not been called for ``timeout_s``; a hung collective thus kills the rank instead of the job idling
    until the scheduler's limit, and the launcher (torchrun) tears the others down.
  - :class:`Heartbeat` -- per-rank heartbeat files under a shared directory; ``stale()`` lists ranks that stopped beating.
  - :func:`health_check` -- one all-reduce(MAX) of [non-finite loss, device barrier timeout, user flag]; every rank learns
    that *some* rank failed and can stop at the same step (checkpoint, then exit) instead of deadlocking in the next exchange.
  - :func:`inject_fault` -- ``DGRAPH_B200_FAULT_INJECT="nan:rank=1:step=3"`` / ``"stall:rank=0:step=2:seconds=5"`` for testing
    the above.
"""
from __future__ import annotations

import faulthandler
import os
import sys
import threading
import time
from pathlib import Path
from typing import Callable, List, Optional

import torch


class StepWatchdog:
    def __init__(self, timeout_s: float, on_timeout: Optional[Callable[[float], None]] = None, poll_s: float = 0.5):
        self.timeout_s, self.poll_s = float(timeout_s), float(poll_s)
        self.on_timeout = on_timeout or self._default
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name="dgraph-b200-watchdog", daemon=True)
        self.fired = False

    @staticmethod
    def _default(idle_s: float):
        sys.stderr.write(f"[dgraph_b200 watchdog] no progress for {idle_s:.0f} s -- dumping stacks and aborting this rank\n")
        faulthandler.dump_traceback(all_threads=True)
        os._exit(124)

    def _run(self):
        while not self._stop.wait(self.poll_s):
            idle = time.monotonic() - self._last
            if idle > self.timeout_s:
                self.fired = True
                self.on_timeout(idle)
                self._last = time.monotonic()

    def start(self):
        self._last = time.monotonic()
        self._thread.start()
        return self

    def beat(self):
        self._last = time.monotonic()

    def stop(self):
        self._stop.set()

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()
        return False


class Heartbeat:
    def __init__(self, directory: str, rank: int, world_size: int):
        self.dir, self.rank, self.world = Path(directory), rank, world_size
        self.dir.mkdir(parents=True, exist_ok=True)

    def _file(self, r: int) -> Path:
        return self.dir / f"heartbeat_{r}.txt"

    def beat(self, step: int):
        tmp = self._file(self.rank).with_suffix(".tmp")
        tmp.write_text(f"{step} {time.time()}\n")
        os.replace(tmp, self._file(self.rank))

    def read(self, r: int):
        try:
            s, t = self._file(r).read_text().split()
            return int(s), float(t)
        except (OSError, ValueError):
            return None

    def stale(self, max_age_s: float) -> List[int]:
        now = time.time()
        out = []
        for r in range(self.world):
            v = self.read(r)
            if v is None or now - v[1] > max_age_s:
                out.append(r)
        return out


def health_check(comm, loss: Optional[torch.Tensor] = None, failed: bool = False) -> bool:
    """True when every rank is healthy.  Costs one 3-element all-reduce; call it every N steps, not every step."""
    dev = loss.device if loss is not None else ("cuda" if torch.cuda.is_available() else "cpu")
    flags = torch.zeros(3, dtype=torch.float32, device=dev)
    if loss is not None:
        flags[0] = (~torch.isfinite(loss.detach().float())).any().float()
    eng = getattr(comm, "engine", None)
    tf = getattr(eng, "_timeout_flag", None)
    if tf is not None:
        flags[1] = tf.float()[0]
    flags[2] = float(failed)
    comm.all_reduce(flags, op="max", scope="world") if hasattr(comm, "get_num_replicas") else comm.all_reduce(flags, op="max")
    return not bool(flags.any())


def inject_fault(rank: int, step: int, loss: Optional[torch.Tensor] = None):
    """Testing aid, driven by ``DGRAPH_B200_FAULT_INJECT`` (see module docstring).  Returns the (possibly poisoned) loss."""
    spec = os.environ.get("DGRAPH_B200_FAULT_INJECT", "")
    if not spec:
        return loss
    kind, *kv = spec.split(":")
    opts = dict(x.split("=") for x in kv)
    if int(opts.get("rank", -1)) not in (-1, rank) or int(opts.get("step", -1)) not in (-1, step):
        return loss
    if kind == "nan" and loss is not None:
        return loss * float("nan")
    if kind == "stall":
        time.sleep(float(opts.get("seconds", 1)))
    if kind == "exit":
        os._exit(int(opts.get("code", 1)))
    return loss
