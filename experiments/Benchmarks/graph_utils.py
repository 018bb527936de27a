"""Benchmark graphs (parity: reference ``experiments/Benchmarks/graph_utils.py:7-230``): the all-pairs pattern -- every rank
owns ``verts_per_rank`` vertices and has one edge to every vertex of every other rank -- plus row-count driven halo patterns
for the bandwidth sweep."""
from dataclasses import dataclass

import torch


@dataclass
class PatternBenchGraph:
    """Edge list + placement for the halo-plan flavour of the benchmarks (``TestNCCL.py --sweep``)."""
    edge_list: torch.Tensor  # [E, 2] global (central, neighbour)
    partitioning: torch.Tensor  # [V]
    num_vertices: int


def all_pairs_graph(world_size: int, verts_per_rank: int = 1) -> PatternBenchGraph:
    V = world_size * verts_per_rank
    part = torch.arange(V) // verts_per_rank
    u, v = torch.meshgrid(torch.arange(V), torch.arange(V), indexing="ij")
    m = part[u] != part[v]
    return PatternBenchGraph(torch.stack([u[m], v[m]], 1), part, V)


def halo_rows_graph(world_size: int, local_rows: int, halo_rows_per_peer: int, seed: int = 0) -> PatternBenchGraph:
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
    return PatternBenchGraph(torch.cat(es), part, V)


# ---------------------------------------------------------------------------------------------------------------------------
# The reference's data holders and generators for the vector-addressed (legacy API) benchmarks: one vertex per rank, one edge from
# every rank to every other rank ("all pairs").  Tensors follow the layouts of `comm.gather(x, indices, rank_mappings)` /
# `comm.scatter(x, indices, rank_mappings, output_size)`.
# ---------------------------------------------------------------------------------------------------------------------------
@dataclass
class GatherGraphData:
    vertex_data: torch.Tensor          # [W, F]: row r is the vertex owned by rank r
    vertex_rank_mapping: torch.Tensor  # [W]
    edge_rank_placement: torch.Tensor  # [1, E]: rank that holds each edge
    edge_src_rank: torch.Tensor        # [1, E]: rank that owns the vertex each edge reads
    edge_indices: torch.Tensor         # [1, E]: global vertex id each edge reads


@dataclass
class ScatterGraphData:
    vertex_data: torch.Tensor          # [1, E, F]: one row per edge
    data_rank_mapping: torch.Tensor    # [1, E]
    edge_rank_placement: torch.Tensor  # [1, E]
    edge_dest_rank: torch.Tensor       # [1, E]: rank that owns the vertex each edge adds to
    edge_indices: torch.Tensor         # [1, E]: global vertex id each edge adds to
    num_local_vertices: int


def _all_pairs(world_size: int, ranks_per_node=None, intra_node_only=False):
    assert ranks_per_node is not None or not intra_node_only, "ranks_per_node must be provided if intra_node_only is True"
    holder, target = [], []
    for i in range(world_size):
        for j in range(world_size):
            if i == j or (intra_node_only and i // ranks_per_node != j // ranks_per_node):
                continue
            holder.append(i)
            target.append(j)
    return torch.tensor(holder, dtype=torch.long).unsqueeze(0), torch.tensor(target, dtype=torch.long).unsqueeze(0)


def get_nccl_gather_benchmark_data(sample_size, world_size, device, ranks_per_node=None, intra_node_only=False):
    g = torch.Generator().manual_seed(0)
    holder, target = _all_pairs(world_size, ranks_per_node, intra_node_only)
    return GatherGraphData(vertex_data=torch.randn(world_size, sample_size, generator=g).to(device),
                           vertex_rank_mapping=torch.arange(world_size, device=device),
                           edge_rank_placement=holder.to(device), edge_src_rank=target.to(device),
                           edge_indices=target.clone().to(device))


def get_nccl_scatter_benchmark_data(sample_size, world_size, device, ranks_per_node=None, intra_node_only=False):
    g = torch.Generator().manual_seed(0)
    holder, target = _all_pairs(world_size, ranks_per_node, intra_node_only)
    E = holder.shape[1]
    return ScatterGraphData(vertex_data=torch.randn(1, E, sample_size, generator=g).to(device),
                            data_rank_mapping=holder.clone().to(device), edge_rank_placement=holder.to(device),
                            edge_dest_rank=target.to(device), edge_indices=target.clone().to(device), num_local_vertices=1)


def _local_share(full, rank: int):
    """Cut a rank's share out of an all-ranks record: its vertex rows as ``[1, n_local, F]`` and its edges as ``[1, E_local]``."""
    mine = full.edge_rank_placement[0] == rank
    return mine


def get_nvshmem_gather_benchmark_data(sample_size, rank, world_size, device, ranks_per_node=None, intra_node_only=False):
    """One-sided convention of the reference (``graph_utils.py:121-166``): the record holds THIS rank's vertex rows
    ``[1, n_local, F]`` and the owner ranks / global ids of THIS rank's edges ``[1, E_local]``; ``edge_rank_placement`` stays the
    all-ranks tensor."""
    full = get_nccl_gather_benchmark_data(sample_size, world_size, device, ranks_per_node, intra_node_only)
    mine = _local_share(full, rank)
    F = full.vertex_data.shape[-1]
    return GatherGraphData(vertex_data=full.vertex_data[full.vertex_rank_mapping == rank].view(1, -1, F),
                           vertex_rank_mapping=full.vertex_rank_mapping, edge_rank_placement=full.edge_rank_placement,
                           edge_src_rank=full.edge_src_rank[:, mine].contiguous(), edge_indices=full.edge_indices[:, mine].contiguous())


def get_nvshmem_scatter_benchmark_data(sample_size, rank, world_size, device, ranks_per_node=None, intra_node_only=False):
    """Scatter counterpart (``graph_utils.py:169-230``): this rank's edge rows ``[1, E_local, F]`` with their destinations."""
    full = get_nccl_scatter_benchmark_data(sample_size, world_size, device, ranks_per_node, intra_node_only)
    mine = _local_share(full, rank)
    return ScatterGraphData(vertex_data=full.vertex_data[:, mine].contiguous(), data_rank_mapping=full.data_rank_mapping,
                            edge_rank_placement=full.edge_rank_placement, edge_dest_rank=full.edge_dest_rank[:, mine].contiguous(),
                            edge_indices=full.edge_indices[:, mine].contiguous(), num_local_vertices=full.num_local_vertices)


def safe_create_dir(directory, rank):
    """Rank 0 creates the directory (no error when it exists)."""
    import os

    if rank == 0:
        os.makedirs(directory, exist_ok=True)
