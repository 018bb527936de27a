"""Reference file name (``lsc_datasets/MAG240M_dataset.py:31-340``): reads the processed MAG240M files from disk, no ``ogb``."""
from . import DGraph_MAG240M_Dataset  # noqa: F401
from dgraph_b200.data.ogb_raw import MAG240MFiles, find_mag240m  # noqa: F401

from typing import Optional  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402


def get_col_slice(x, start_row_idx: int, end_row_idx: int, start_col_idx: int, end_col_idx: int, chunk: int = 100_000) -> np.ndarray:
    """Rows ``[start_row, end_row)`` x columns ``[start_col, end_col)`` of a (memory-mapped) matrix, read in row chunks so that only
    one chunk of the full-width rows is resident at a time."""
    out = np.empty((end_row_idx - start_row_idx, end_col_idx - start_col_idx), dtype=x.dtype)
    for i in range(start_row_idx, end_row_idx, chunk):
        j = min(i + chunk, end_row_idx)
        out[i - start_row_idx:j - start_row_idx] = x[i:j, start_col_idx:end_col_idx]
    return out


def save_col_slice(x_src, x_dst, start_row_idx: int, end_row_idx: int, start_col_idx: int, end_col_idx: int,
                   chunk: int = 100_000) -> None:
    """Inverse of :func:`get_col_slice`: writes ``x_src`` into that window of ``x_dst`` chunk by chunk."""
    assert x_src.shape == (end_row_idx - start_row_idx, end_col_idx - start_col_idx)
    for i in range(0, end_row_idx - start_row_idx, chunk):
        j = min(i + chunk, end_row_idx - start_row_idx)
        x_dst[start_row_idx + i:start_row_idx + j, start_col_idx:end_col_idx] = x_src[i:j]


def get_edge_mappings(src_indices: torch.Tensor, dst_indices: torch.Tensor, rank_mappings: torch.Tensor) -> torch.Tensor:
    """Rank that owns each edge: the owner of its destination vertex (where the messages are accumulated)."""
    return rank_mappings[dst_indices]


def _generate_features_from_paper_features(out, num_nodes: int, num_papers: int, paper_feat, edge_index, num_features: int,
                                           dim_chunk_size: int = 64) -> None:
    """Features of authors / institutions as the mean of their neighbours' features (``MAG240M_dataset.py:65-102``, there a
    ``torch_sparse`` SpMM): ``out[v] = mean_{(v, p) in edge_index} paper_feat[p]``, 64 columns at a time
    (``dgraph_b200.data.ogb_raw.mean_neighbor_features``: a CSR segment mean, no ``torch_sparse``)."""
    from dgraph_b200.data.ogb_raw import mean_neighbor_features

    mean_neighbor_features(out, num_nodes, num_papers, paper_feat, edge_index, num_features, dim_chunk_size)


def load_or_generate_vertex_rank_mask(rank_mapping: Optional[torch.Tensor], num_vertices: int, world_size: int, rank: int):
    """``(boolean mask of the vertices `rank` owns, their number)``; without a placement the contiguous block split is used."""
    from .distributed_graph_dataset import get_rank_mappings

    if rank_mapping is None:
        rank_mapping, _ = get_rank_mappings(num_vertices, world_size, rank)
    mask = rank_mapping == rank
    return mask, int(mask.sum())
