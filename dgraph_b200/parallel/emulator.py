"""Single-process world emulator (SURVEY §4 "what the new framework should add").

``run_world(W, fn)`` runs ``fn(comm)`` on W Python threads, each with a :class:`Communicator` bound to a
:class:`ThreadEngine`.  CPU autograd executes backward on the calling thread, so the distributed backward
passes (which issue collectives) stay rank-aligned.  The reference has no fake backend at all; its distributed
tests need 2 GPUs (``tests/test_comm_info.py:441-449``).
"""
from __future__ import annotations

import threading
import traceback
from typing import Callable, List

from dgraph_b200.parallel.collective_engine import ThreadEngine, _ThreadWorld


def run_world(world_size: int, fn: Callable, *args, timeout: float = 120.0, **kwargs) -> List:
    from dgraph_b200.communicator import Communicator

    world = _ThreadWorld(world_size)
    results: List = [None] * world_size
    errors: List = [None] * world_size

    def _target(r):
        try:
            comm = Communicator._from_engine(ThreadEngine(world, r), backend="emulated")
            results[r] = fn(comm, *args, **kwargs)
        except BaseException as e:  # noqa: BLE001 - propagate to the caller below
            errors[r] = (e, traceback.format_exc())
            world.barrier.abort()

    threads = [threading.Thread(target=_target, args=(r,), daemon=True) for r in range(world_size)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout)
    for r, err in enumerate(errors):
        if err is not None and not isinstance(err[0], threading.BrokenBarrierError):
            raise RuntimeError(f"emulated rank {r} failed:\n{err[1]}") from err[0]
    for r, err in enumerate(errors):
        if err is not None:
            raise RuntimeError(f"emulated rank {r} failed:\n{err[1]}") from err[0]
    if any(t.is_alive() for t in threads):
        raise TimeoutError("emulated world did not finish (deadlock?)")
    return results
