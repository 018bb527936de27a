"""GraphSAGE on an ogbn-papers100M-shaped graph (111 M nodes, 1.6 B edges, 128-d, 172 classes) partitioned over the ranks --
BASELINE.json config 3.  Every rank generates ONLY its own edge shard (the reference needs the global edge list on every
rank, 25 GB at this size) and the halo plan is built from shards (`build_communication_pattern_from_shard`).

torchrun --nproc-per-node 8 benchmarks/bench_sage_papers.py [--scale 1.0] [--steps 5]
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "experiments"))
from _common import cleanup, init_distributed  # noqa: E402

from dgraph_b200.data.synthetic import SHAPES  # noqa: E402
from dgraph_b200.models.sage import CommAwareSAGE  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402
from dgraph_b200.parallel import HaloExchange, build_communication_pattern_from_shard  # noqa: E402


def shard_edges(V, E, rank, world, dev, cut=0.1, window=8192, seed=0):
    """In-neighbour lists of the vertices this rank owns: [E_loc, 2] (owned vertex, neighbour), global ids."""
    lo, hi = V * rank // world, V * (rank + 1) // world
    n = hi - lo
    e_loc = E * n // V
    g = torch.Generator(device=dev).manual_seed(seed * 1000 + rank)
    u = torch.randint(lo, hi, (e_loc,), generator=g, device=dev)
    off = torch.randint(1, window + 1, (e_loc,), generator=g, device=dev) * (torch.randint(0, 2, (e_loc,), generator=g, device=dev) * 2 - 1)
    far = torch.rand(e_loc, generator=g, device=dev) < cut
    v = torch.where(far, torch.randint(0, V, (e_loc,), generator=g, device=dev), (u + off) % V)
    return torch.stack([u, v], 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--hidden", type=int, default=256)
    ap.add_argument("--cut", type=float, default=0.05, help="fraction of edges whose neighbour is drawn uniformly (long range)")
    a = ap.parse_args()
    comm, rank, world, dev = init_distributed("nccl")
    V0, E0, F, C = SHAPES["ogbn-papers100M"]
    V, E = int(V0 * a.scale), int(E0 * a.scale)
    t0 = time.time()
    edges = shard_edges(V, E, rank, world, dev, cut=a.cut)
    offsets = torch.tensor([V * r // world for r in range(world + 1)])
    pat = build_communication_pattern_from_shard(edges, offsets, rank, world)
    del edges
    pat.kernel_plan(dev)
    n = pat.num_local_vertices
    cdt = torch.bfloat16 if dev.type == "cuda" else torch.float32
    x = torch.randn(n, F, device=dev).to(cdt)
    y = torch.randint(0, C, (n,), device=dev)
    setup_s = time.time() - t0
    model = CommAwareSAGE(F, a.hidden, C, HaloExchange(comm), comm, compute_dtype=cdt).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-3, comm=comm)
    idx = torch.arange(0, n, 100, device=dev)  # ~1 % labelled vertices (papers100M: 1.1 %)

    def step():
        opt.zero_grad()
        out = model(x, pat)
        loss = torch.nn.functional.cross_entropy(out.index_select(0, idx).float(), y.index_select(0, idx))
        loss.backward()
        opt.step()
        return loss

    for _ in range(a.warmup):
        step()
    comm.barrier()
    if dev.type == "cuda":
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t1 = time.perf_counter()
    for _ in range(a.steps):
        loss = step()
    if dev.type == "cuda":
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
    else:
        ms = (time.perf_counter() - t1) * 1e3 / a.steps
    t = torch.tensor([ms, float(pat.num_halo_vertices), float(pat.num_local_edges)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"model": "GraphSAGE 2 layers hidden %d" % a.hidden, "graph": f"papers100M-shaped V={V} E~{E} F={F}",
                          "cut": a.cut, "n_gpus": world, "ms_per_epoch": float(t[0]), "edges_per_s": E / (float(t[0]) * 1e-3),
                          "max_halo_rows": int(t[1]), "max_local_edges": int(t[2]), "setup_s": round(setup_s, 1),
                          "loss": float(loss.detach()), "engine": type(comm.engine).__name__,
                          "mem_GB": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1) if dev.type == "cuda" else None}),
              flush=True)
    cleanup(comm)


if __name__ == "__main__":
    main()
