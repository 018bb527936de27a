"""Mesh edge-block micro-benchmark (parity: reference ``experiments/GraphCast/microbenchmark_graphcast.py:19-253``): hidden 512,
100 iterations / 10 warm-up on the multimesh, "comm" (halo exchange of the node features) vs "process" (edge MLP + residual)
in ms, saved as ``.npy``.   ``torchrun --nproc-per-node 4 microbenchmark_graphcast.py``"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402

from dgraph_b200.utils.cli import run_cli  # noqa: E402


def load_data(rank, world_size, mesh_level: int = 6, grid=(181, 360), device="cpu", dtype=torch.float32):
    """This rank's shard of the static GraphCast graph (the reference loads its dataset object for the same purpose,
    ``microbenchmark_graphcast.py:19-60``)."""
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator

    return DistributedGraphCastGraphGenerator(tuple(grid), mesh_level).shard(rank, world_size).to(device, dtype)


def _timed(fn, dev):
    if dev.type == "cuda":
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        return out, a.elapsed_time(b)
    import time

    t0 = time.perf_counter()
    out = fn()
    return out, (time.perf_counter() - t0) * 1e3


def time_mesh_edge_block(local_vertex_features, local_edge_features, static_graph, comm, iters: int = 100, warmup: int = 10,
                         hidden_dim: int = 512):
    """Per-iteration milliseconds of one mesh edge block: ``comm`` = halo exchange of the vertex features, ``process`` = gathered
    edge MLP + LayerNorm + residual (``microbenchmark_graphcast.py:63-181``)."""
    from dgraph_b200.models.graphcast import MeshEdgeBlock
    from dgraph_b200.parallel import HaloExchange

    x, e, sg = local_vertex_features, local_edge_features, static_graph
    dev = x.device
    block = MeshEdgeBlock(hidden_dim, hidden_dim, hidden_dim, hidden_dim, comm, hidden_dim).to(dev)
    ex = HaloExchange(comm)
    pat = sg.patterns.mesh
    comm_ms, proc_ms = [], []
    with torch.no_grad():
        for it in range(iters + warmup):
            aug, t1 = _timed(lambda: ex.augmented(x, pat), dev)
            _, t2 = _timed(lambda: block(x, aug, e, sg.mesh_graph_src_indices, sg.mesh_graph_dst_indices), dev)
            if it >= warmup:
                comm_ms.append(t1)
                proc_ms.append(t2)
    return np.asarray(comm_ms), np.asarray(proc_ms)


def main(hidden: int = 512, iters: int = 100, warmup: int = 10, mesh_level: int = 6, log_dir: str = "logs"):
    comm, rank, world, dev = init_distributed("nccl")
    cdt = torch.bfloat16 if dev.type == "cuda" else torch.float32
    sg = load_data(rank, world, mesh_level, device=dev, dtype=cdt)
    n, E = sg.num_local_mesh, int(sg.mesh_graph_src_indices.numel())
    x = torch.randn(n, hidden, device=dev).to(cdt)
    e = torch.randn(E, hidden, device=dev).to(cdt)
    comm_ms, proc_ms = time_mesh_edge_block(x, e, sg, comm, iters, warmup, hidden)
    os.makedirs(log_dir, exist_ok=True)
    np.save(os.path.join(log_dir, f"graphcast_mesh_edge_comm_rank{rank}.npy"), comm_ms)
    np.save(os.path.join(log_dir, f"graphcast_mesh_edge_process_rank{rank}.npy"), proc_ms)
    if rank == 0:
        print(f"mesh edge block hidden {hidden}, {E} local edges, world {world}: comm {np.mean(comm_ms):.3f} +- {np.std(comm_ms):.3f} ms, "
              f"process {np.mean(proc_ms):.3f} +- {np.std(proc_ms):.3f} ms")
    cleanup(comm)


if __name__ == "__main__":
    run_cli(main)
