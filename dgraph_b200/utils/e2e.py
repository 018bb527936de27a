"""End-to-end step loop shared by both arms of ``bench.py`` (ours and the unmodified reference), so their numbers differ only in
what ``step`` does: every iteration copies that step's inputs from pinned host memory to the device and reads that step's loss
back on the host.

Copy scheme (identical for both arms): two device input buffers; the copy for step ``i + 1`` is issued on a side stream while step
``i`` computes; the loss of step ``i`` lands in pinned host memory through an asynchronous copy and is read on the host right after
step ``i + 1`` has been enqueued (the last one after the loop), so the host never idles the GPU and still sees every step's loss
inside the timed region.  Harness code only: no model, kernel or engine of either framework.
"""
from __future__ import annotations

from typing import Callable, List, Sequence

import torch


def make_e2e_loop(step: Callable, host_inputs: Sequence[torch.Tensor], device_buffers: List[Sequence[torch.Tensor]],
                  sink: List[float]):
    """``step(*device_inputs) -> loss tensor``.  ``device_buffers``: two tuples of device tensors shaped like ``host_inputs``
    (the caller allocates them, e.g. in symmetric memory).  Returns ``loop(k)``; the host-side loss values are appended to ``sink``."""
    assert len(device_buffers) == 2
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    free = [torch.cuda.Event(), torch.cuda.Event()]
    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ready = [torch.cuda.Event(), torch.cuda.Event()]

    def issue(i):
        b = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(free[b])
            for d, h in zip(device_buffers[b], host_inputs):
                d.copy_(h, non_blocking=True)
            ready[b].record(copy_stream)

    def loop(k):
        main = torch.cuda.current_stream()
        for b in range(2):
            free[b].record(main)
        issue(0)
        for i in range(k):
            b = i % 2
            if i + 1 < k:
                issue(i + 1)  # next step's inputs travel while this step computes
            main.wait_event(ready[b])
            loss = step(*device_buffers[b])
            free[b].record(main)
            loss_host[b].copy_(loss.detach().float(), non_blocking=True)
            loss_ready[b].record(main)
            if i > 0:  # host read of the previous step's loss, after this step has been enqueued
                loss_ready[1 - b].synchronize()
                sink.append(float(loss_host[1 - b]))
        loss_ready[(k - 1) % 2].synchronize()
        sink.append(float(loss_host[(k - 1) % 2]))

    return loop
