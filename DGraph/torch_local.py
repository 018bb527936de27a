"""``DGraph.torch_local`` -- the four launchers of the reference's pybind11 module (``csrc/torch_local_bindings.cpp:20-26``,
``include/torch_local.hpp:4-40``) with the same positional signatures, backed by ``dgb_rows_indexed`` (``csrc/rank_local.cu``)
on CUDA tensors and by index ops elsewhere.  Tensors are ``[batch, rows, cols]`` with batch 1 (or ``[rows, cols]``); the size
arguments are accepted for compatibility and checked against the tensors."""
from __future__ import annotations

import torch

from dgraph_b200.parallel import rank_local as _rl


def _check(t: torch.Tensor, rows: int, cols: int, what: str):
    r = t[0] if t.dim() == 3 else t
    assert t.dim() == 2 or t.shape[0] == 1, f"{what}: batch must be 1"
    assert r.shape[-1] == cols, f"{what}: has {r.shape[-1]} columns, num_cols={cols}"
    assert r.shape[0] >= rows, f"{what}: has {r.shape[0]} rows, expected at least {rows}"


def local_masked_gather(input, indices, rank_local_placement, output, num_batches, num_values_rows, num_cols, num_output_rows,
                        rank):
    """``output[b, r, :] = input[b, indices[r], :]`` where ``rank_local_placement[r] == rank`` (other rows untouched)."""
    _check(input, num_values_rows, num_cols, "input")
    _check(output, num_output_rows, num_cols, "output")
    return _rl.OptimizedRankLocalMaskedGather(input, indices, rank_local_placement, output, rank)


def local_masked_scatter(input, indices, rank_local_placement, output, num_batches, num_values_rows, num_cols, num_output_rows,
                         rank):
    """``output[b, indices[r], :] += input[b, r, :]`` where ``rank_local_placement[r] == rank``.  The reference kernel also
    applies a ReLU to the scattered values (an accident of its dead fused kernel); a plain sum is used here."""
    _check(input, num_values_rows, num_cols, "input")
    _check(output, num_output_rows, num_cols, "output")
    return _rl.RankLocalMaskedScatter(input, output, indices, rank_local_placement, rank)


def local_masked_scatter_gather(input, indices, input_mask, output, num_batches, num_values_rows, num_cols, num_output_rows):
    """``output[b, input_mask[i], :] = input[b, indices[i], :]`` for ``i < num_values_rows``."""
    _check(output, num_output_rows, num_cols, "output")
    return _rl.OptimizedLocalScatterGather(input, indices.reshape(-1)[:num_values_rows], input_mask.reshape(-1)[:num_values_rows],
                                           output)


def local_masked_scatter_add_gather(input, indices, input_mask, output, num_batches, num_values_rows, num_cols,
                                    num_output_rows):
    """``output[b, input_mask[i], :] += input[b, indices[i], :]`` (duplicates accumulate)."""
    _check(output, num_output_rows, num_cols, "output")
    return _rl.OptimizedLocalScatterSumGather(input, indices.reshape(-1)[:num_values_rows],
                                              input_mask.reshape(-1)[:num_values_rows], output)
