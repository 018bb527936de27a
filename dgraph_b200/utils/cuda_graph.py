"""Whole-training-step CUDA graphs ("CUDA streams and graphs instead of a tracing compiler").

A full-graph GNN step is the *same* launch sequence every epoch -- same plans, same shapes, same buffers -- so it is captured
once and replayed: one ``cudaGraphLaunch`` instead of ~75 Python -> ctypes -> ``<<<>>>`` round trips plus the autograd engine.
That matters when the step is launch-bound (ogbn-arxiv-sized graphs, the per-rank slice of a graph at 8 GPUs, GraphCast's
~600 small kernels), not for a 30 ms products epoch.  The reference has nothing comparable (no graphs, no fused optimizer;
SURVEY §5.8 last table row).

Requirements on the captured region (all met by this package's single-GPU path):
* no host synchronisation: index tensors instead of boolean masks, no ``.item()``;
* values that change per step live in device memory: ``FusedAdam(capturable=True)`` keeps ``t`` and the learning rate there;
* inputs are copied into static buffers before each replay; outputs are read from static tensors.
Multi-GPU: the peer-memory engine normally passes the barrier epoch as a kernel argument (frozen at capture time); with
``DGRAPH_B200_DEVICE_EPOCH=1`` (set before the communicator is created; `bench.py` does so for its multi-GPU runs) the epoch, like the
optimizer step, is read from and advanced in device memory, and a step with halo exchanges can be captured.  Every rank must
capture and replay the same sequence (the barriers inside the graph still wait for all peers).
"""
from __future__ import annotations

import sys
from typing import Callable, Optional, Sequence

import torch

from dgraph_b200.ops import _native as N


class GraphedTrainStep:
    """``step = GraphedTrainStep(model, opt, loss_fn, (x, y)); loss = step(x, y)`` -- ``loss_fn(model_out, *targets)``.

    ``forward_fn(model, *inputs)`` defaults to ``model(inputs[0])``-style calls being wrapped by the caller; pass a closure
    when the model needs extra (static) arguments such as the communication pattern.
    """

    def __init__(self, forward_loss: Callable[..., torch.Tensor], optimizer, example_inputs: Sequence[torch.Tensor],
                 warmup: int = 3, pool=None, static_inputs: Optional[Sequence[torch.Tensor]] = None):
        """``static_inputs``: use these tensors as the graph's input buffers instead of clones of ``example_inputs`` -- e.g. features
        that must live in symmetric memory (``HaloExchange.symmetric_rows``) so that the captured exchange reads them in place."""
        assert torch.cuda.is_available(), "CUDA graphs need a GPU"
        assert getattr(optimizer, "capturable", False), "use FusedAdam(..., capturable=True)"
        self.opt = optimizer
        # The graph replays raw addresses.  Everything the captured region reads besides its inputs -- label / index tensors, plans,
        # the model -- is reachable from this closure and must stay allocated for as long as the graph can be replayed (a caller that
        # built the closure in a function and kept only the step would otherwise free them, and replays would read recycled memory).
        self.forward_loss = forward_loss
        if static_inputs is not None:
            self.static_in = list(static_inputs)
            for dst, src in zip(self.static_in, example_inputs):
                if dst.data_ptr() != src.data_ptr():
                    dst.copy_(src)
        else:
            self.static_in = [t.clone() for t in example_inputs]
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up off the capture stream: lazy inits, TMA descriptor entry point, autotune-free
            for _ in range(max(1, warmup)):
                optimizer.zero_grad()
                forward_loss(*self.static_in).backward()
                optimizer.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.warmup_steps = max(1, warmup)
        l0 = N.launches()
        with torch.cuda.graph(self.graph, pool=pool):
            optimizer.zero_grad()
            self.static_loss = forward_loss(*self.static_in)
            self.static_loss.backward()
            optimizer.step()
        self.launches_per_replay = N.launches() - l0

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        for dst, src in zip(self.static_in, inputs):
            if src is not dst and src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.opt.sync_hyper()
        self.graph.replay()
        N.count_launch(self.launches_per_replay)  # keep the launch accounting of bench.py truthful under replay
        return self.static_loss


def capture_or_eager(make: Callable[[], GraphedTrainStep], world: int = 1) -> Optional[GraphedTrainStep]:
    """``make()`` builds a :class:`GraphedTrainStep`.  When the capture raises on any rank, EVERY rank gets ``None`` and keeps
    stepping eagerly (the ranks agree through one small all-reduce): nothing ran on the device during the failed capture, so the
    barrier epochs in device memory are still in step.  The caller restores the optimizer snapshot either way."""
    err = None
    try:
        g = make()
    except Exception as e:  # noqa: BLE001 - reported below; eager stepping is always available
        g, err = None, e
    ok = torch.tensor([0 if g is None else 1], dtype=torch.int32, device="cuda")
    if world > 1:
        import torch.distributed as dist

        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok) == 1:
        return g
    print(f"CUDA-graph capture failed{'' if err is None else f' ({type(err).__name__}: {err})'}; stepping eagerly", file=sys.stderr)
    return None
