"""Rank-sorting of a homogeneous graph (parity: reference ``DGraph/data/preprocess.py:6-114``).

``process_homogenous_data`` sorts nodes by owner rank, renumbers edges with the **inverse** permutation (the reference
reuses ``argsort(placement)`` as its own inverse), sorts edges by source rank and returns a
:class:`DistributedGraph`."""
from __future__ import annotations

from typing import Optional

import torch

from dgraph_b200.data.graph import DistributedGraph


def node_renumbering(node_rank_placement: torch.Tensor):
    """``(new_to_old, old_to_new)`` for a stable sort of nodes by owner rank."""
    new_to_old = torch.argsort(node_rank_placement, stable=True)
    old_to_new = torch.empty_like(new_to_old)
    old_to_new[new_to_old] = torch.arange(new_to_old.numel(), device=new_to_old.device)
    return new_to_old, old_to_new


def edge_renumbering(edge_index: torch.Tensor, old_to_new: torch.Tensor, sorted_placement: torch.Tensor):
    """Renumber ``[2, E]`` edges and order them by the rank of their source."""
    e = old_to_new[edge_index]
    src_rank = sorted_placement[e[0]]
    order = torch.argsort(src_rank, stable=True)
    e = e[:, order]
    return e, src_rank[order], sorted_placement[e[1]], order


def process_homogenous_data(graph_data: dict, labels: torch.Tensor, world_size: int, split_idx: Optional[dict] = None,
                            node_rank_placement: Optional[torch.Tensor] = None, **_unused) -> DistributedGraph:
    """``graph_data``: ``{"node_feat": [V, F], "edge_index": [2, E], "num_nodes": V}`` (OGB layout)."""
    x = torch.as_tensor(graph_data["node_feat"])
    ei = torch.as_tensor(graph_data["edge_index"]).long()
    V = int(graph_data.get("num_nodes", x.shape[0]))
    if node_rank_placement is None:
        node_rank_placement = torch.arange(V) % world_size
    new_to_old, old_to_new = node_renumbering(node_rank_placement)
    placement = node_rank_placement[new_to_old]
    e, e_loc, e_dst, order = edge_renumbering(ei, old_to_new, placement)
    masks = {}
    for k_in, k_out in (("train", "train_mask"), ("valid", "val_mask"), ("test", "test_mask")):
        masks[k_out] = None if not split_idx or k_in not in split_idx else torch.sort(old_to_new[torch.as_tensor(split_idx[k_in]).long()]).values
    ef = graph_data.get("edge_feat")
    return DistributedGraph(node_features=x[new_to_old], edge_index=e, labels=torch.as_tensor(labels).reshape(-1)[new_to_old],
                            node_loc=placement, edge_loc=e_loc, edge_dest_rank_mapping=e_dst, num_nodes=V,
                            num_edges=int(e.shape[1]), world_size=world_size,
                            edge_features=None if ef is None else torch.as_tensor(ef)[order], **masks)
