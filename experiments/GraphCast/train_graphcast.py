"""GraphCast training entry point (parity: reference ``experiments/GraphCast/train_graphcast.py:29-173``).

``torchrun --nnodes 1 --nproc-per-node N experiments/GraphCast/train_graphcast.py --is_distributed --use_synthetic_data``

Same flags as the reference; adds ``--iters``, ``--mesh_level``, ``--lat``, ``--lon``, ``--resume``.  Checkpoints are plain
``state_dict`` files ``model_{iter}.pth`` (reference format) plus a ``.train_state`` side file for resume.
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

import torch  # noqa: E402

from dgraph_b200.utils.cli import run_cli  # noqa: E402


def loss_fn(pred, target, ranks_per_sample, world):
    """MSE scaled so that summing over the ranks of one sample and averaging over replicas gives the global mean
    (``train_graphcast.py:29-34``)."""
    return torch.nn.functional.mse_loss(pred.float(), target.float()) * (ranks_per_sample / world)


def compute_loss(ground_truth, prediction, comm):
    """The reference's signature (``train_graphcast.py:29-34``): mean squared error per sample, summed over the batch, scaled by
    ``ranks_per_sample / world_size`` so that a sum over all ranks averages over the samples in flight.  Accepts ``[N, C]`` (one
    sample) or ``[B, N, C]``."""
    d = (ground_truth.float() - prediction.float()).pow(2)
    d = d.unsqueeze(0) if d.dim() == 2 else d
    per_sample = d.mean(dim=(1, 2)).sum()
    rps = getattr(comm, "_ranks_per_sample", None)
    if rps is None:
        eng = getattr(comm, "engine", None)
        rps = getattr(eng, "ranks_per_graph", comm.get_world_size())
    world = comm.get_global_world_size() if hasattr(comm, "get_global_world_size") else comm.get_world_size()
    return per_sample * rps / world


def main(batch_size: int = 1, is_distributed: bool = False, test_run: bool = False, use_synthetic_data: bool = True,
         benchmark: bool = False, backend: str = "single", procs_per_graph: int = 2, iters: int = 10, mesh_level: int = 6,
         lat: int = 721, lon: int = 1440, hidden_dim: int = 128, save_dir: str = "", resume: str = "", placement: str = "bands",
         cuda_graph: bool = False):
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset
    from dgraph_b200.models.graphcast import Config, DGraphCast
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    assert backend in ("single", "nccl", "nvshmem", "mpi")
    assert batch_size == 1, "full-graph training: one sample per step (as the reference)"
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    if cuda_graph and world_env > 1:  # device-resident barrier epochs make steps with halo exchanges capturable
        os.environ["DGRAPH_B200_DEVICE_EPOCH"] = "1"
    hybrid = is_distributed and 0 < procs_per_graph < world_env and world_env % procs_per_graph == 0
    # hybrid data-parallel x graph-partition: `procs_per_graph` ranks share one sample's graph, replicas take different samples
    comm, rank, world, dev = init_distributed("nccl" if backend == "single" else backend,
                                              **({"ranks_per_graph": procs_per_graph} if hybrid else {}))
    g_rank, g_world = comm.get_rank(), comm.get_world_size()
    cfg = Config()
    if test_run:
        mesh_level, lat, lon, iters = min(mesh_level, 3), min(lat, 46), min(lon, 90), min(iters, 3)
    cfg.model.mesh_level, cfg.model.hidden_dim, cfg.data.latlon_res = mesh_level, hidden_dim, (lat, lon)
    cdt = torch.bfloat16 if dev.type == "cuda" else torch.float32
    gen = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level, multi=cfg.model.multimesh, placement=placement)
    sg = gen.shard(g_rank, g_world).to(dev, cdt)
    data = SyntheticWeatherDataset(cfg.model.input_grid_dim, cfg.data.latlon_res, num_samples=max(iters, 2),
                                   grid_rows=sg.grid_rows_global.cpu())
    torch.manual_seed(0)
    model = DGraphCast(cfg, comm).to(dev)
    use_graph = bool(cuda_graph) and dev.type == "cuda" and (comm.get_world_size() == 1 or getattr(comm.engine, "device_epoch", False))
    opt = FusedAdam(model.parameters(), lr=cfg.training.lr, comm=comm, capturable=use_graph)
    tc = cfg.training
    sched = torch.optim.lr_scheduler.SequentialLR(opt, [
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-3, end_factor=1.0, total_iters=tc.num_iters_step1),
        torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=tc.num_iters_step2, eta_min=0.0),
        torch.optim.lr_scheduler.LambdaLR(opt, lambda _e: tc.lr_step3 / tc.lr)],
        milestones=[tc.num_iters_step1, tc.num_iters_step1 + tc.num_iters_step2])
    start = load_checkpoint(resume, model, opt, sched, map_location=dev) if resume else 0
    from dgraph_b200.utils.dist_utils import CommAwareDistributedSampler

    sampler = CommAwareDistributedSampler(data, comm, shuffle=False)  # every rank of a replica sees the same sample
    order = list(sampler)
    times = []
    graphed = None
    if use_graph:
        # ~600 small launches per step (far more CPU time than GPU time once the graph is split over 8 GPUs) become one
        # cudaGraphLaunch.  Capturing runs two eager warm-up steps on the first sample.
        from dgraph_b200.utils.cuda_graph import GraphedTrainStep

        first = data[order[start % len(order)]]
        graphed = GraphedTrainStep(lambda a, b: loss_fn(model(a, sg), b, world, world), opt,
                                   (first["invar"].to(dev, cdt), first["outvar"].to(dev, cdt)), warmup=2)
    for it in range(start, start + iters):
        item = data[order[it % len(order)]]
        x, y = item["invar"].to(dev, cdt), item["outvar"].to(dev, cdt)
        comm.barrier()
        t0 = time.perf_counter()
        if graphed is not None:
            loss = graphed(x, y)
        else:
            opt.zero_grad()
            loss = loss_fn(model(x, sg), y, world, world)  # factor 1: FusedAdam already averages over the whole world
            loss.backward()
            opt.step()
        sched.step()
        lv = float(loss.detach())
        times.append((time.perf_counter() - t0) * 1e3)
        if rank == 0:
            print(f"iter {it:5d} | loss {lv:.6f} | {times[-1]:.1f} ms", flush=True)
        if save_dir and (it + 1) % tc.save_freq == 0:
            save_checkpoint(os.path.join(save_dir, f"model_{it}.pth"), model, opt, sched, step=it + 1, rank=rank)
    if benchmark and rank == 0:
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if dev.type == "cuda" else []),
                     profile_memory=True) as prof:
            with torch.no_grad():
                model(x, sg)
        print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=15))
    if rank == 0 and len(times) > 1:
        print(f"mean step (excl. first): {sum(times[1:]) / (len(times) - 1):.2f} ms", flush=True)
    cleanup(comm)


if __name__ == "__main__":
    run_cli(main)
