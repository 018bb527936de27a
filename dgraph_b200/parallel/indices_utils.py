"""Per-peer message counts and placements for the legacy ``indices + rank map`` API.

Parity: reference ``DGraph/distributed/nccl/_indices_utils.py:5-245`` (same function names; the legacy gather / scatter functions
that should import them do not).  A *message* is one edge whose data lives on ``src_ranks[e]`` and is needed on
``dest_ranks[e]``; only edges with different ranks are communicated.  All functions are plain tensor code, device-agnostic.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch


def _count_per_rank(ranks: torch.Tensor, world_size: int) -> torch.Tensor:
    return torch.bincount(ranks.reshape(-1).long(), minlength=world_size).long()


def _get_send_comm_vector(comm_senders, comm_receivers, rank: int, world_size: int) -> torch.Tensor:
    """``[W]``: messages this rank sends to every rank."""
    return _count_per_rank(comm_receivers[comm_senders == rank], world_size)


def _get_recv_comm_vector(comm_senders, comm_receivers, rank: int, world_size: int) -> torch.Tensor:
    """``[W]``: messages this rank receives from every rank."""
    return _count_per_rank(comm_senders[comm_receivers == rank], world_size)


def _get_send_recv_comm_vectors(src_ranks, dest_ranks, rank: int, world_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(send [W], recv [W])`` over the cross-rank messages only."""
    cross = src_ranks != dest_ranks
    s, d = src_ranks[cross], dest_ranks[cross]
    return _get_send_comm_vector(s, d, rank, world_size), _get_recv_comm_vector(s, d, rank, world_size)


def _get_local_send_placement(send_comm_vector, indices, src_ranks, dest_ranks, rank: int, num_src_rows: int) -> Dict[int, torch.Tensor]:
    """For every peer that gets something: the local rows (``index % num_src_rows``) to ship, in edge order."""
    out: Dict[int, torch.Tensor] = {}
    idx = indices.reshape(-1)
    s, d = src_ranks.reshape(-1), dest_ranks.reshape(-1)
    for peer, n in enumerate(send_comm_vector.tolist()):
        if n and peer != rank:
            out[peer] = idx[(s == rank) & (d == peer)] % num_src_rows
    return out


def _get_local_unique_recv_placement(indices, src_ranks, receive_from_mask, num_local_output_rows: int, rank: int,
                                     world_size: int) -> Dict[int, torch.Tensor]:
    """Scatter with local pre-reduction: every sender ships one row per *distinct* target, so the receiver's placement for a
    sender is the set of distinct local rows (``index % num_local_output_rows``) that sender addresses."""
    out: Dict[int, torch.Tensor] = {}
    idx, s, m = indices.reshape(-1), src_ranks.reshape(-1), receive_from_mask.reshape(-1)
    for sender in range(world_size):
        if sender == rank:
            continue
        sel = m & (s == sender)
        if bool(sel.any()):
            out[sender] = torch.unique(idx[sel] % num_local_output_rows)
    return out


def _get_local_recv_placement(recv_comm_vector, local_rank_mapping, rank: int) -> Dict[int, torch.Tensor]:
    """Positions (``argwhere``) of the local edges whose data comes from each sending peer."""
    out: Dict[int, torch.Tensor] = {}
    for peer, n in enumerate(recv_comm_vector.tolist()):
        if n and peer != rank:
            out[peer] = torch.argwhere(local_rank_mapping == peer)
    return out


def _allocate_local_recv_buffers(recv_comm_vector, num_features: int, rank: int, device) -> Dict[int, torch.Tensor]:
    """One zeroed ``[1, n_peer, F]`` buffer per sending peer."""
    return {peer: torch.zeros(1, int(n), num_features, device=device)
            for peer, n in enumerate(recv_comm_vector.tolist()) if n and peer != rank}


def _get_local_recv_buffer_w_placement(recv_comm_vector, local_rank_mapping, num_features: int, rank: int, device):
    """``(buffers, placements)``: the two functions above in one call."""
    return (_allocate_local_recv_buffers(recv_comm_vector, num_features, rank, device),
            _get_local_recv_placement(recv_comm_vector, local_rank_mapping, rank))


def _generate_local_rank_mapping(_global_rank_mapping: torch.Tensor, world_size: int) -> torch.Tensor:
    """Rank of every position under an even block split of a length-``n`` array into ``world_size`` blocks (the last block is
    shorter when ``n`` is not a multiple): ``[0, 0, ..., 1, 1, ..., W-1, ...]`` truncated to ``n`` entries."""
    n = int(_global_rank_mapping.reshape(-1).shape[0])
    per = (n + world_size - 1) // world_size if n else 0
    full = torch.arange(world_size, device=_global_rank_mapping.device).repeat_interleave(max(per, 1))
    return full[:n].long()
