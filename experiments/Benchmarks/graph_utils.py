"""Benchmark graphs (parity: reference ``experiments/Benchmarks/graph_utils.py:7-230``): the all-pairs pattern -- every rank
owns ``verts_per_rank`` vertices and has one edge to every vertex of every other rank -- plus row-count driven halo patterns
for the bandwidth sweep."""
from dataclasses import dataclass

import torch


@dataclass
class GatherGraphData:
    edge_list: torch.Tensor  # [E, 2] global (central, neighbour)
    partitioning: torch.Tensor  # [V]
    num_vertices: int


ScatterGraphData = GatherGraphData


def all_pairs_graph(world_size: int, verts_per_rank: int = 1) -> GatherGraphData:
    V = world_size * verts_per_rank
    part = torch.arange(V) // verts_per_rank
    u, v = torch.meshgrid(torch.arange(V), torch.arange(V), indexing="ij")
    m = part[u] != part[v]
    return GatherGraphData(torch.stack([u[m], v[m]], 1), part, V)


def halo_rows_graph(world_size: int, local_rows: int, halo_rows_per_peer: int, seed: int = 0) -> GatherGraphData:
    """Every rank pulls ``halo_rows_per_peer`` distinct rows from every other rank (one edge per pulled row)."""
    V = world_size * local_rows
    part = torch.arange(V) // local_rows
    g = torch.Generator().manual_seed(seed)
    es = []
    for r in range(world_size):
        for q in range(world_size):
            if q == r:
                continue
            nb = torch.randperm(local_rows, generator=g)[:halo_rows_per_peer] + q * local_rows
            ce = torch.randint(0, local_rows, (nb.numel(),), generator=g) + r * local_rows
            es.append(torch.stack([ce, nb], 1))
    return GatherGraphData(torch.cat(es), part, V)
