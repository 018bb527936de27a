"""Halo gather / scatter micro-benchmark (parity: reference ``experiments/Benchmarks/TestNCCL.py:23-290``).

``torchrun --nproc-per-node N experiments/Benchmarks/TestNCCL.py --message_size 128 --num_iters 1000 --log_dir logs``

Protocol of the reference: all-pairs pattern, per-iteration CUDA-event times, first 10 iterations dropped, mean +- std;
``.npy`` per rank.  Adds ``--sweep``: halo bandwidth sweep (rows x feature width) comparing the peer-memory pull engine with
the NCCL all-to-all engine on identical splits; reports GB/s (max over ranks) and the fraction of the 770 GB/s per-direction
NVLink peer-copy roofline (B200_PROFILING.md).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from graph_utils import (all_pairs_graph, get_nccl_gather_benchmark_data, get_nccl_scatter_benchmark_data,  # noqa: E402
                         halo_rows_graph)


def _time(fn, iters, dev):
    ts = []
    for _ in range(iters):
        if dev.type == "cuda":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        else:
            import time

            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e3)
    return np.asarray(ts)


def run(comm, graph, F, iters, dev, engine_name):
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    rank, world = comm.get_rank(), comm.get_world_size()
    pat = build_communication_pattern(graph.edge_list.to(dev), graph.partitioning.to(dev), rank, world)
    dt = torch.bfloat16 if dev.type == "cuda" else torch.float32
    x = torch.randn(pat.num_local_vertices, F, device=dev).to(dt).requires_grad_(True)
    ex = HaloExchange(comm)
    gather = _time(lambda: ex.augmented(x, pat), iters, dev)

    def fwd_bwd():
        a = ex.augmented(x, pat)
        a.backward(torch.ones_like(a))

    both = _time(fwd_bwd, iters, dev)
    t = torch.tensor([gather[10:].mean(), both[10:].mean()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    in_bytes = pat.num_halo_vertices * F * x.element_size()
    out_bytes = int(pat.send_offset[-1]) * F * x.element_size()
    b = torch.tensor([max(in_bytes, out_bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(b, op=dist.ReduceOp.MAX)
    return {"engine": engine_name, "F": F, "halo_rows": pat.num_halo_vertices, "gather_ms": float(t[0]),
            "gather_scatter_ms": float(t[1]), "gather_GBps": float(b[0]) / float(t[0]) / 1e6,
            "frac_of_770GBps": float(b[0]) / float(t[0]) / 1e6 / 770.0, "per_iter": gather}


def run_legacy(comm, F, iters, dev, use_cache):
    """The reference's own protocol: all-pairs pattern (one vertex per rank), vector-addressed ``comm.gather`` /
    ``comm.scatter`` with ``indices + rank_mappings``, per-iteration times.  ``use_cache=False`` drops the cached plan before
    every call, i.e. pays the two metadata collectives each time (the reference's "no cache" arm)."""
    rank, world = comm.get_rank(), comm.get_world_size()
    gd = get_nccl_gather_benchmark_data(F, world, dev)
    sd = get_nccl_scatter_benchmark_data(F, world, dev)
    x = gd.vertex_data[rank:rank + 1].unsqueeze(0).contiguous()
    g_map = torch.cat([gd.edge_rank_placement, gd.edge_src_rank], 0)
    s_map = torch.cat([sd.edge_rank_placement, sd.edge_dest_rank], 0)
    mine = sd.edge_rank_placement[0] == rank
    vals = sd.vertex_data[:, mine].contiguous()

    def drop():
        if not use_cache and hasattr(comm.engine, "_legacy_plans"):
            comm.engine._legacy_plans.clear()

    def gather():
        drop()
        comm.gather(x, gd.edge_indices, g_map)

    def scatter():
        drop()
        comm.scatter(vals, sd.edge_indices, s_map, sd.num_local_vertices)

    return _time(gather, iters, dev), _time(scatter, iters, dev)


class NCCLBenchmark:
    """The reference's benchmark object (``TestNCCL.py:23-111``): per-iteration device times of the vector-addressed
    ``comm.gather`` / ``comm.scatter`` (``indices`` + stacked ``rank_mappings``).  ``cache=False`` drops the engine's plan cache
    before every call (the reference's "without cache" arm); any other value keeps it."""

    def __init__(self, comm, world_size: int, rank: int):
        self.comm, self.world_size, self.rank = comm, world_size, rank

    def comm_stats_gather(self, edge_rank_placement, edge_src_rank):
        """Messages this rank receives (its edges whose source lives elsewhere) and sends (others' edges whose source is here)."""
        mine = edge_rank_placement == self.rank
        return {"local_comm_rev": (edge_src_rank[mine] != self.rank).sum(), "local_comm_send": (edge_src_rank[~mine] == self.rank).sum()}

    comm_stats_scatter = comm_stats_gather

    def _device(self, t):
        return t.device

    def print(self, message):
        if self.rank == 0:
            print(message, flush=True)

    def save_np(self, np_array, filename, rank_to_save: int = 0):
        if self.rank == rank_to_save:
            np.save(filename, np_array)

    def _drop_cache(self, cache):
        if cache is False and hasattr(self.comm.engine, "_legacy_plans"):
            self.comm.engine._legacy_plans.clear()

    def benchmark_gather(self, data, edge_rank_placement, edge_src_rank, edge_indices, cache=None, num_iters: int = 1000):
        maps = torch.cat([edge_rank_placement, edge_src_rank], dim=0)

        def one():
            self._drop_cache(cache)
            self.comm.gather(data, edge_indices, maps)

        self.comm.barrier()
        return _time(one, num_iters, self._device(data))

    def benchmark_scatter(self, data, edge_rank_placement, edge_dest_rank, edge_indices, num_local_vertices, cache=None,
                          num_iters: int = 1000):
        maps = torch.cat([edge_rank_placement, edge_dest_rank], dim=0)

        def one():
            self._drop_cache(cache)
            self.comm.scatter(data, edge_indices, maps, num_local_vertices)

        self.comm.barrier()
        return _time(one, num_iters, self._device(data))


def run_gather_benchmark(benchmark: NCCLBenchmark, num_iters: int, graph_data, cache=None):
    """Times ``comm.gather`` on a ``GatherGraphData`` (``graph_utils.py``): this rank's vertex rows as ``[1, n_local, F]``."""
    F = graph_data.vertex_data.shape[-1]
    local = graph_data.vertex_data[graph_data.vertex_rank_mapping == benchmark.rank, :].view(1, -1, F).contiguous()
    return benchmark.benchmark_gather(local, graph_data.edge_rank_placement, graph_data.edge_src_rank, graph_data.edge_indices,
                                      num_iters=num_iters, cache=cache)


def run_scatter_benchmark(benchmark: NCCLBenchmark, num_iters: int, graph_data, cache=None):
    """Times ``comm.scatter`` on a ``ScatterGraphData``: this rank's edge rows as ``[1, e_local, F]``."""
    F = graph_data.vertex_data.shape[-1]
    local = graph_data.vertex_data[graph_data.data_rank_mapping == benchmark.rank].view(1, -1, F).contiguous()
    return benchmark.benchmark_scatter(local, graph_data.edge_rank_placement, graph_data.edge_dest_rank, graph_data.edge_indices,
                                       graph_data.num_local_vertices, num_iters=num_iters, cache=cache)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--message_size", type=int, default=128)
    ap.add_argument("--benchmark_cache", action="store_true")
    ap.add_argument("--num_iters", type=int, default=1000)
    ap.add_argument("--log_dir", type=str, default="logs")
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--backend", type=str, default="nccl")
    a = ap.parse_args()
    comm, rank, world, dev = init_distributed(a.backend)
    os.makedirs(a.log_dir, exist_ok=True)
    name = type(comm.engine).__name__
    results = []
    if not a.sweep and world > 1:  # the reference's legacy-API arm (with the plan cache, and without it on --benchmark_cache)
        for cached in ([True, False] if a.benchmark_cache else [True]):
            g_t, s_t = run_legacy(comm, a.message_size, min(a.num_iters, 1000), dev, cached)
            tag = "cache" if cached else "nocache"
            np.save(os.path.join(a.log_dir, f"legacy_gather_{tag}_{name}_rank{rank}.npy"), g_t)
            np.save(os.path.join(a.log_dir, f"legacy_scatter_{tag}_{name}_rank{rank}.npy"), s_t)
            results.append({"engine": name, "api": f"legacy ({tag})", "F": a.message_size,
                            "gather_ms": float(g_t[10:].mean()) if len(g_t) > 10 else float(g_t.mean()),
                            "scatter_ms": float(s_t[10:].mean()) if len(s_t) > 10 else float(s_t.mean())})
    if not a.sweep:
        r = run(comm, all_pairs_graph(world), a.message_size, a.num_iters, dev, name)
        np.save(os.path.join(a.log_dir, f"gather_times_{name}_rank{rank}.npy"), r.pop("per_iter"))
        results.append(r)
    else:
        for rows in (1_000, 10_000, 100_000, 1_000_000):
            for F in (64, 256, 768):
                per_peer = max(rows // max(world - 1, 1), 1)
                g = halo_rows_graph(world, max(per_peer, 1024), per_peer)  # local rows kept small: the timed exchange is dominated by the pull
                r = run(comm, g, F, min(a.num_iters, 60), dev, name)
                r.pop("per_iter")
                results.append(r)
    if rank == 0:
        for r in results:
            print(json.dumps(r), flush=True)
        with open(os.path.join(a.log_dir, f"halo_bench_{name}_world{world}.json"), "w") as f:
            json.dump(results, f)
    cleanup(comm)


if __name__ == "__main__":
    main()
