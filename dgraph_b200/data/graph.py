"""Globally replicated, rank-sorted graph container.

Parity: reference ``DGraph/data/graph.py:24-288`` (``DistributedGraph`` accessors, ``get_round_robin_node_rank_map``).
Local index masks are derived from the true per-rank node ranges (the reference's ``% nodes_per_rank`` shortcut is only
right for equal shards).
"""
from __future__ import annotations

from typing import Optional

import torch


class DistributedGraph:
    def __init__(self, node_features: torch.Tensor, edge_index: torch.Tensor, labels: torch.Tensor, node_loc: torch.Tensor,
                 edge_loc: torch.Tensor, edge_dest_rank_mapping: torch.Tensor, num_nodes: int, num_edges: int,
                 world_size: int, edge_features: Optional[torch.Tensor] = None, train_mask: Optional[torch.Tensor] = None,
                 val_mask: Optional[torch.Tensor] = None, test_mask: Optional[torch.Tensor] = None,
                 graph_labels: Optional[torch.Tensor] = None):
        assert node_features.dim() == 2, "Invalid node features shape. Expect 2D tensor"
        assert edge_index.dim() == 2 and edge_index.shape[0] == 2, "Invalid edge index shape. Expect (2, num_edges)"
        assert node_loc.dim() == 1 and node_loc.shape[0] == num_nodes, "Invalid node_loc shape"
        assert edge_loc.dim() == 1 and edge_loc.shape[0] == num_edges, "Invalid edge_loc shape"
        assert node_features.shape[0] == num_nodes and edge_index.shape[1] == num_edges
        self.num_nodes, self.num_edges, self.world_size = num_nodes, num_edges, world_size
        self.node_features, self.edge_index, self.edge_features, self.labels = node_features, edge_index, edge_features, labels
        self.train_mask, self.val_mask, self.test_mask, self.graph_labels = train_mask, val_mask, test_mask, graph_labels
        self.node_loc, self.edge_loc, self.edge_dest_rank_mapping = node_loc, edge_loc, edge_dest_rank_mapping
        self._nodes_per_rank = torch.bincount(node_loc, minlength=world_size)
        self._edges_per_rank = torch.bincount(edge_loc, minlength=world_size)
        self.max_node_per_rank = int(self._nodes_per_rank.max())
        self.max_edge_per_rank = int(self._edges_per_rank.max()) if num_edges else 0
        self.rank_mappings = torch.stack([edge_loc, edge_dest_rank_mapping], dim=0)

    def get_nodes_per_rank(self):
        return self._nodes_per_rank

    def get_edges_per_rank(self):
        return self._edges_per_rank

    def get_max_node_per_rank(self) -> int:
        return self.max_node_per_rank

    def get_max_edge_per_rank(self) -> int:
        return self.max_edge_per_rank

    def get_local_node_features(self, rank):
        return self.node_features[self.node_loc == rank]

    def get_global_node_features(self):
        return self.node_features

    def get_local_edge_indices(self, rank):
        return self.edge_index[:, self.edge_loc == rank]

    def get_global_edge_indices(self):
        return self.edge_index

    def get_global_rank_mappings(self):
        return self.rank_mappings

    def get_local_rank_mappings(self, rank):
        return self.rank_mappings[:, self.edge_loc == rank]

    def get_local_labels(self, rank):
        return self.labels[self.node_loc == rank]

    def get_global_labels(self):
        return self.labels

    def get_local_mask(self, mask: str, rank):
        """Index mask (``train`` / ``val`` / ``test``) restricted to this rank's contiguous node range, as local indices."""
        m = {"train": self.train_mask, "val": self.val_mask, "test": self.test_mask}.get(mask, "bad")
        if isinstance(m, str):
            raise ValueError(f"Invalid mask {mask}")
        assert m is not None, f"{mask} mask not found"
        start = int(self._nodes_per_rank[:rank].sum())
        end = start + int(self._nodes_per_rank[rank])
        m = m.long()
        return m[(m >= start) & (m < end)] - start

    def _get_index_to_rank_mapping(self, _indices):
        return self.node_loc[_indices.long()]

    def get_sender_receiver_ranks(self):
        return self.edge_loc, self.edge_dest_rank_mapping


def get_round_robin_node_rank_map(num_nodes: int, world_size: int) -> torch.Tensor:
    """``node i -> rank i % world_size`` (``graph.py:270-288``, vectorised)."""
    return torch.arange(num_nodes) % world_size
