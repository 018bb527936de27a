"""Edge-cut / balance / time of the partitioners (CPU only): `python benchmarks/bench_partition.py > profiles/partition_quality.md`."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.data import make_graph  # noqa: E402
from dgraph_b200.data.partition import bfs_blocks, contiguous_blocks, edge_cut_fraction, hashed, multilevel  # noqa: E402


def grid(n):
    idx = torch.arange(n * n).view(n, n)
    e = torch.cat([torch.stack([idx[:, :-1].reshape(-1), idx[:, 1:].reshape(-1)], 1),
                   torch.stack([idx[:-1].reshape(-1), idx[1:].reshape(-1)], 1)])
    perm = torch.randperm(n * n, generator=torch.Generator().manual_seed(0))
    return perm[e], n * n


def main():
    cases = [("2-D grid 300 x 300, shuffled ids", *grid(300))]
    g = make_graph("ogbn-arxiv", seed=0)
    cases.append(("ogbn-arxiv-shaped synthetic (V=169,343, 2.33 M directed edges, 10 % long-range)", g.edge_list, g.num_nodes))
    g = make_graph("ogbn-products", seed=0, scale=0.05)
    cases.append((f"ogbn-products-shaped synthetic at 5 % scale (V={g.num_nodes:,}, {g.edge_list.shape[0]:,} directed edges)",
                  g.edge_list, g.num_nodes))
    print("| graph | parts | method | cut fraction | max part / ideal | seconds |")
    print("|---|---|---|---|---|---|")
    for name, e, V in cases:
        for W in (2, 8):
            for mname, fn in (("hash", lambda: hashed(V, W)), ("contiguous ids", lambda: contiguous_blocks(V, W)),
                              ("BFS blocks (PyTorch)", lambda: bfs_blocks(e, V, W)),
                              ("multilevel (native)", lambda: multilevel(e, V, W))):
                t = time.time()
                p = fn()
                dt = time.time() - t
                sizes = torch.bincount(p, minlength=W)
                print(f"| {name} | {W} | {mname} | {edge_cut_fraction(e, p):.4f} | {float(sizes.max()) / (V / W):.3f} | {dt:.2f} |")


if __name__ == "__main__":
    main()
