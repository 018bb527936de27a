"""Reference file name (``experiments/OGB/ogb_comm_dataset.py:54-145``): ``ds[0] -> (local_feats, local_labels, comm_pattern)``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.data.ogbn_datasets import DGraphOGBDataset  # noqa: F401,E402

__all__ = ["DGraphOGBDataset"]


def _build_local_split_masks(node_rank_placement, split_idx: dict, rank: int) -> dict:
    """Global split index lists (``{"train": ids, "valid": ids, "test": ids}``) -> boolean masks over this rank's vertices, in
    the order of ``torch.where(node_rank_placement == rank)`` (``ogb_comm_dataset.py:14-39``)."""
    import torch

    mine = torch.where(node_rank_placement == rank)[0]
    out = {}
    for name, ids in split_idx.items():
        member = torch.zeros(node_rank_placement.shape[0], dtype=torch.bool)
        member[torch.as_tensor(ids).long()] = True
        out[name] = member[mine]
    return out


def generate_communication_pattern(edge_index, node_rank_placement, rank: int, world_size: int):
    """Halo plan of one rank from the global ``[E, 2]`` edge list and the vertex placement (``ogb_comm_dataset.py:42-51``)."""
    from dgraph_b200.parallel import build_communication_pattern

    return build_communication_pattern(edge_index, node_rank_placement, rank, world_size)
