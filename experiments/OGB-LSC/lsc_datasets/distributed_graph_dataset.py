"""Reference file name (``lsc_datasets/distributed_graph_dataset.py:24-489``)."""
from . import DistributedHeteroGraphDataset  # noqa: F401
from dgraph_b200.data.hetero import NODE_TYPES, RELATIONS, node_offsets  # noqa: F401

import torch  # noqa: E402


def get_vertex_offsets(num_vertices: int, world_size: int) -> torch.Tensor:
    """``[W + 1]`` contiguous vertex ranges of ``ceil(V / W)`` vertices each, the last ones clamped to ``V`` -- the reference's split
    (``distributed_graph_dataset.py:40-46``).  ``dgraph_b200.data.hetero.node_offsets`` spreads the remainder evenly instead."""
    per = (num_vertices + world_size - 1) // world_size
    return (torch.arange(world_size + 1, dtype=torch.long) * per).clamp_(max=num_vertices)


def get_rank_mappings(num_vertices: int, world_size: int, rank: int):
    """``(owner rank of every vertex [V], number of vertices of `rank`)``: blocks of ``ceil(V / W)``, the last rank takes what is
    left (``distributed_graph_dataset.py:24-37``)."""
    per = max((num_vertices + world_size - 1) // world_size, 1)
    owner = (torch.arange(num_vertices, dtype=torch.long) // per).clamp_(max=world_size - 1)
    return owner, int((owner == rank).sum())


def edge_mapping_from_vertex_mapping(edge_index: torch.Tensor, src_rank_mappings: torch.Tensor, dst_rank_mappings: torch.Tensor):
    """Owner ranks of both endpoints of every directed edge ``edge_index[0] -> edge_index[1]``; source and destination types have
    their own placements in a heterogeneous graph."""
    return src_rank_mappings[edge_index[0]], dst_rank_mappings[edge_index[1]]
