"""Timing protocol shared by every `bench.py --config` workload and by both arms (ours / reference): W untimed warm-up steps, K
timed steps between barrier + synchronize, CUDA events, max over ranks; clocks sampled during the timed region; an end-to-end arm
whose every step copies the step's inputs from pinned host memory and reads the loss back (`dgraph_b200/utils/e2e.py`, the same
loop for both arms).  Harness code only."""
from __future__ import annotations

import json
import time
from dataclasses import dataclass, field
from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist


@dataclass
class Workload:
    metric: str
    unit: str
    work_per_step: float  # value = work_per_step / seconds_per_step   (0 -> the metric is the step time itself, in ms)
    config: dict
    host_inputs: Sequence[torch.Tensor]  # pinned host tensors: one step's inputs
    alloc_inputs: Callable  # (tag) -> tuple of device tensors shaped like host_inputs
    step: Callable  # (*device_inputs) -> loss tensor; does zero_grad + forward + loss + backward + optimizer step
    dtype: str = "bf16"
    scaling: str = "strong"
    launches: Callable = lambda: 0  # number of OUR native kernel launches so far
    extra: dict = field(default_factory=dict)
    finalize: Callable = lambda: None
    loss_weight: Optional[float] = None  # local loss * weight summed over ranks / summed weights = global loss


def run_workload(wl: Workload, args, rank: int, world: int, dev, impl: str, out, clock_sampler=None) -> None:
    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(k)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        sync()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    dev_in = wl.alloc_inputs(0)
    for d, h in zip(dev_in, wl.host_inputs):
        d.copy_(h, non_blocking=True)
    losses = []

    def loop_dev(k):
        for _ in range(k):
            losses.append(wl.step(*dev_in))

    loop_dev(args.warmup)
    if clock_sampler is not None and rank == 0:
        clock_sampler.start()
    l0 = wl.launches()
    ms_step = timed(loop_dev, args.steps) / args.steps
    n_launch = wl.launches() - l0
    clocks = clock_sampler.stop() if (clock_sampler is not None and rank == 0) else None

    def value_of(ms):
        return wl.work_per_step / (ms * 1e-3) if wl.work_per_step else ms

    e2e = None
    if not args.no_e2e:
        from dgraph_b200.utils.e2e import make_e2e_loop

        sink = []
        loop_e2e = make_e2e_loop(wl.step, tuple(wl.host_inputs), [dev_in, wl.alloc_inputs(1)], sink)
        loop_e2e(max(2, min(args.warmup, 3)))
        ms_e2e = timed(loop_e2e, args.steps) / args.steps
        e2e = {"value": value_of(ms_e2e), "unit": wl.unit, "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": int(sum(h.numel() * h.element_size() for h in wl.host_inputs)), "d2h_bytes_per_step": 4,
               "copy_scheme": "double-buffered H2D on a side stream; loss read on the host one step later (every step, inside the "
                              "timed region); identical in both arms"}
    w = float(wl.loss_weight if wl.loss_weight is not None else 1.0)
    gl = torch.tensor([float(losses[0].detach()) * w, float(losses[-1].detach()) * w, w], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(gl)
    if rank == 0:
        line = {"metric": wl.metric, "value": value_of(ms_step), "unit": wl.unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": bool(wl.work_per_step), "scaling": wl.scaling,
                "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic", "impl": impl, "config": wl.config, "e2e": e2e,
                "gpu_launches": n_launch, "clocks": clocks, "loss_first": float(gl[0] / gl[2]), "final_loss": float(gl[1] / gl[2])}
        line.update(wl.extra)
        print(json.dumps(line), file=out, flush=True)
    wl.finalize()


def median_ms(fn, iters: int, warm: int = 10) -> float:
    """Median CUDA-event time of ``fn`` (ms): ``warm`` untimed + ``iters`` timed calls, each call bracketed by its own events."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def now() -> float:
    return time.perf_counter()
