"""Rank-local index helpers and kernels.

Parity: reference ``DGraph/distributed/RankLocalOps.py:35-327`` (``OptimizedLocalScatterGather``,
``OptimizedLocalScatterSumGather``, ``OptimizedRankLocalMaskedGather``, ``RankLocalMaskedGather`` / ``Scatter``,
``RankLocalRenumbering``, ``RankLocalRenumberingWithMapping``, ``LocalAggregateWithRemapping``) on top of
``csrc/rank_local.cu`` (the ``torch_local`` equivalent).  Tensors use the reference's ``[1, N, F]`` layout (2-D accepted).
The torch fallbacks index the row dimension (the reference's fallback indexes the batch dimension) and the
scatter-sum fallback is ``index_add_`` rather than an O(E) Python loop.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from dgraph_b200.ops import _native as N


def _rows(t):
    return t[0] if t.dim() == 3 else t


def _native_ok(src, out):
    rb = src.shape[-1] * src.element_size()
    return (src.is_cuda and out.is_cuda and src.dtype == out.dtype and src.dtype in (torch.float32, torch.bfloat16)
            and rb % 16 == 0 and _rows(src).stride(-1) == 1 and _rows(out).stride(-1) == 1
            and (_rows(src).stride(0) * src.element_size()) % 16 == 0 and (_rows(out).stride(0) * out.element_size()) % 16 == 0)


def _launch(out, src, dst_idx, src_idx, placement, rank, mode):
    o, s = _rows(out), _rows(src)
    n = int(dst_idx.numel() if dst_idx is not None else (src_idx.numel() if src_idx is not None else s.shape[0]))
    es = s.element_size()
    li = lambda t: None if t is None else t.to(torch.int64).contiguous()  # noqa: E731
    d, si, pl = li(dst_idx), li(src_idx), li(placement)
    N.check(N.lib().dgb_rows_indexed(N.ptr(o), N.i64(o.stride(0) * es), N.ptr(s), N.i64(s.stride(0) * es), N.ptr(d), N.ptr(si),
                                     N.ptr(pl), N.i32(rank), N.i64(n), N.i32(s.shape[-1] * es), N.i32(es), N.i32(mode),
                                     N.stream()), "dgb_rows_indexed")
    N.count_launch()
    return out


def OptimizedLocalScatterGather(src: torch.Tensor, src_indices: torch.Tensor, dst_indices: torch.Tensor, output: torch.Tensor):
    """``output[:, dst_indices[i]] = src[:, src_indices[i]]``."""
    if _native_ok(src, output):
        return _launch(output, src, dst_indices, src_indices, None, 0, 0)
    _rows(output)[dst_indices.long()] = _rows(src)[src_indices.long()]
    return output


def OptimizedLocalScatterSumGather(src: torch.Tensor, src_indices: torch.Tensor, dst_indices: torch.Tensor, output: torch.Tensor):
    """``output[:, dst_indices[i]] += src[:, src_indices[i]]`` (duplicates in ``dst_indices`` accumulate)."""
    if _native_ok(src, output):
        return _launch(output, src, dst_indices, src_indices, None, 0, 1)
    _rows(output).index_add_(0, dst_indices.long(), _rows(src)[src_indices.long()])
    return output


def OptimizedRankLocalMaskedGather(src: torch.Tensor, indices: torch.Tensor, rank_mapping: torch.Tensor, output: torch.Tensor, rank: int):
    """``output[:, i] = src[:, indices[i]]`` for the entries with ``rank_mapping[i] == rank``; other rows are untouched."""
    idx, rm = indices.reshape(-1), rank_mapping.reshape(-1)
    if _native_ok(src, output):
        return _launch(output, src, None, idx, rm, rank, 0)
    m = rm == rank
    _rows(output)[m] = _rows(src)[idx[m].long()]
    return output


def RankLocalMaskedGather(_src: torch.Tensor, indices: torch.Tensor, rank_mapping: torch.Tensor, rank: int) -> torch.Tensor:
    """Rows of ``_src`` selected by the indices whose mapping equals ``rank`` (compact result)."""
    m = rank_mapping.reshape(-1) == rank
    return _rows(_src).index_select(0, indices.reshape(-1)[m].long()).unsqueeze(0)


def RankLocalMaskedScatter(_src: torch.Tensor, _output: torch.Tensor, local_indices_slice: torch.Tensor,
                           local_dest_ranks: torch.Tensor, rank: int):
    """``_output[:, local_indices_slice[i] % rows] += _src[:, i]`` for entries with ``local_dest_ranks[i] == rank`` (plain sum, no
    ReLU; indices may be global ids of a contiguous even split, hence the modulo as in ``RankLocalOps.py:220-245``)."""
    rows = _rows(_output).shape[0]
    idx, rm = local_indices_slice.reshape(-1), local_dest_ranks.reshape(-1)
    if rows:
        idx = idx % rows  # identity for local ids; no host synchronisation to find out
    if _native_ok(_src, _output):
        return _launch(_output, _src, idx, None, rm, rank, 1)
    m = rm == rank
    _rows(_output).index_add_(0, idx[m].long(), _rows(_src)[m])
    return _output


def RankLocalRenumbering(_indices: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(unique_sorted, renumbered)`` with ``unique_sorted[renumbered] == _indices``."""
    u, inv = torch.unique(_indices, return_inverse=True)
    return u, inv


def RankLocalRenumberingWithMapping(_indices: torch.Tensor, rank_mapping: torch.Tensor):
    """``(renumbered, unique, rank of every unique index)``."""
    u, inv = torch.unique(_indices, return_inverse=True)
    rm = torch.zeros_like(u)
    rm.scatter_(0, inv.reshape(-1), rank_mapping.reshape(-1).to(u.dtype))
    return inv, u, rm


def LocalAggregateWithRemapping(global_data: torch.Tensor, global_indices: torch.Tensor, global_mapping=None, num_features=None,
                                device=None):
    """Pre-reduce rows that target the same global index (``RankLocalOps.py:287-327``):
    ``(summed_rows [1, U, F], rank of every distinct index [U])`` in ascending index order.  Without ``global_mapping`` -- the
    short form ``(data, indices, num_features)`` -- the second result is the list of distinct indices itself."""
    if global_mapping is not None and not torch.is_tensor(global_mapping):  # short form: third argument is the feature count
        global_mapping, num_features, device = None, global_mapping, (num_features if device is None else device)
    idx = global_indices.reshape(-1)
    u, inv = torch.unique(idx, return_inverse=True)
    rows = _rows(global_data)
    F = int(num_features) if num_features is not None else rows.shape[1]
    out = torch.zeros((u.numel(), F), dtype=rows.dtype, device=device or rows.device)
    out.index_add_(0, inv.to(out.device), rows.to(out.device))
    if global_mapping is None:
        return out.unsqueeze(0), u
    owner = torch.zeros_like(u)
    owner.scatter_(0, inv, global_mapping.reshape(-1).to(u.dtype))
    return out.unsqueeze(0), owner


def RankLocalReNumbering(_indices: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(renumbered, unique)``: every index replaced by its position among the sorted distinct values
    (reference ``RankLocalOps.py:248-258``, there an O(U) Python loop)."""
    u, inv = torch.unique(_indices, return_inverse=True)
    return inv.reshape(_indices.shape), u


def RankLocalGather(_src: torch.Tensor, indices: torch.Tensor, rank_mapping: torch.Tensor, rank: int) -> torch.Tensor:
    """Rows ``_src[indices[i]]`` for the entries with ``rank_mapping[i] == rank`` (compact, 2-D source)."""
    sel = indices.reshape(-1)[rank_mapping.reshape(-1) == rank].long()
    return _rows(_src).index_select(0, sel)


OutOfPlaceRankLocalMaskedGather = RankLocalGather  # the reference defines the same body under both names
