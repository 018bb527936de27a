"""Rank-sorting of a homogeneous graph (parity: reference ``DGraph/data/preprocess.py:6-114``).

``process_homogenous_data`` sorts nodes by owner rank, renumbers edges with the **inverse** permutation (the reference
reuses ``argsort(placement)`` as its own inverse), sorts edges by source rank and returns a
:class:`DistributedGraph`."""
from __future__ import annotations

from typing import Optional

import torch

from dgraph_b200.data.graph import DistributedGraph


def node_renumbering(node_rank_placement: torch.Tensor):
    """``(renumbered_nodes, contiguous_rank_mapping)`` as in the reference (``preprocess.py:6-11``): ``renumbered_nodes`` is the
    stable sort permutation by owner rank (position ``i`` of the new numbering holds old node ``renumbered_nodes[i]``) and
    ``contiguous_rank_mapping`` the sorted placement.  Edges need the inverse permutation: :func:`inverse_permutation`."""
    contiguous_rank_mapping, renumbered_nodes = torch.sort(node_rank_placement, stable=True)
    return renumbered_nodes, contiguous_rank_mapping


def inverse_permutation(new_to_old: torch.Tensor) -> torch.Tensor:
    old_to_new = torch.empty_like(new_to_old)
    old_to_new[new_to_old] = torch.arange(new_to_old.numel(), device=new_to_old.device)
    return old_to_new


def edge_renumbering(edge_indices: torch.Tensor, renumbered_nodes: torch.Tensor, vertex_mapping: torch.Tensor,
                     edge_features: Optional[torch.Tensor] = None):
    """``(edge_index [2, E], source ranks, destination ranks, edge_features)`` with the endpoints renumbered and the edges
    ordered by the rank of their source (``preprocess.py:14-40``).  ``renumbered_nodes`` maps an OLD node id to its new id
    (the inverse of the sort permutation; the reference passes the permutation itself); ``vertex_mapping`` is the
    placement in the new numbering."""
    e = renumbered_nodes[edge_indices]
    src_rank = vertex_mapping[e[0]]
    order = torch.argsort(src_rank, stable=True)
    e = e[:, order]
    return e, src_rank[order], vertex_mapping[e[1]], None if edge_features is None else edge_features[order]


def process_homogenous_data(graph_data: dict, labels: torch.Tensor, rank: Optional[int] = None, world_Size: Optional[int] = None,
                            split_idx: Optional[dict] = None, node_rank_placement: Optional[torch.Tensor] = None, *args,
                            world_size: Optional[int] = None, **kwargs) -> DistributedGraph:
    """The reference's argument order (``preprocess.py:43-52``): ``graph_data`` = ``{"node_feat": [V, F], "edge_index": [2, E],
    "num_nodes": V}`` (OGB layout), ``labels``, ``rank`` (unused: the result holds every rank's part), ``world_Size``,
    ``split_idx``, ``node_rank_placement`` (round robin when absent).  ``world_size=`` is accepted as a keyword too."""
    W = world_Size if world_Size is not None else world_size
    assert W is not None, "world size missing"
    assert "node_feat" in graph_data and "edge_index" in graph_data, "node_feat / edge_index not found"
    x = torch.as_tensor(graph_data["node_feat"])
    ei = torch.as_tensor(graph_data["edge_index"]).long()
    V = int(graph_data.get("num_nodes", x.shape[0]))
    if node_rank_placement is None:
        node_rank_placement = torch.arange(V) % W
    new_to_old, placement = node_renumbering(node_rank_placement)
    old_to_new = inverse_permutation(new_to_old)
    ef = graph_data.get("edge_feat")
    e, e_loc, e_dst, ef = edge_renumbering(ei, old_to_new, placement, None if ef is None else torch.as_tensor(ef))
    masks = {}
    for k_in, k_out in (("train", "train_mask"), ("valid", "val_mask"), ("test", "test_mask")):
        masks[k_out] = None if not split_idx or k_in not in split_idx else torch.sort(old_to_new[torch.as_tensor(split_idx[k_in]).long()]).values
    return DistributedGraph(node_features=x[new_to_old], edge_index=e, labels=torch.as_tensor(labels).reshape(-1)[new_to_old],
                            node_loc=placement, edge_loc=e_loc, edge_dest_rank_mapping=e_dst, num_nodes=V,
                            num_edges=int(e.shape[1]), world_size=W, edge_features=ef, **masks)
