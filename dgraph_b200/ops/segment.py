"""Segment aggregation ops (CSR walk, atomic-free) with autograd.

* :func:`graphconv_aggregate` -- ``out[i] = sum_{(i,j) in E} relu(P[i] + Q[j])``: the whole message +
  aggregation stage of the reference ``GraphConvLayer`` (``experiments/OGB/GCN.py:43-65``) after splitting the
  first Linear over its ``cat`` input (``W [x_i; x_j] = W_i x_i + W_j x_j``).  Nothing of size ``[E, *]`` is ever
  materialised; backward recomputes the ReLU mask from ``P``/``Q``.
* :func:`segment_sum` / :func:`segment_mean` -- GraphSAGE / GCN neighbourhood aggregation
  (``out[i] = (1/deg) sum_j x[j]``), and the building block of distributed scatter-sum.

CUDA bf16 tensors run the sm_100a kernels in ``csrc/segment_ops.cu``; everything else (CPU, fp32/fp64) runs the
plain PyTorch formulation below, which is also the numerics oracle in the tests.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _sparsity
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan


# Save 1 bit per (edge, column) in the forward so the backward gathers one row per edge instead of two (E * H / 8 bytes).
SAVE_RELU_MASK = os.environ.get("DGRAPH_B200_RELU_MASK", "1") == "1"
# Where the random 32-byte mask accesses go: "csc" = the forward kernel scatters its mask rows to the CSC slot and the backward
# walk streams them; "csr" (default) = forward streams, backward gathers through csc_to_csr.  Measured on the products-shaped
# graph (B200): csr 4.09 ms forward + 4.38 ms backward, csc 4.75 + 4.26 -- the scattered byte stores cost the issue-bound forward
# kernel more than the streamed reads save the latency-bound backward kernel.
MASK_IN_CSC_ORDER = os.environ.get("DGRAPH_B200_MASK_ORDER", "csr") == "csc"


# ------------------------------------------------------------------------------------------------
# PyTorch reference formulations (oracle + CPU path)
# ------------------------------------------------------------------------------------------------
def _coo(kp: GraphKernelPlan):
    """(central row, neighbour row) per CSR slot."""
    key = "_coo_rows"
    c = getattr(kp, key, None)
    if c is None:
        deg = (kp.rowptr[1:] - kp.rowptr[:-1]).long()
        c = torch.repeat_interleave(torch.arange(kp.num_local, device=kp.rowptr.device), deg)
        setattr(kp, key, c)
    return c, kp.col.long()


def graphconv_aggregate_ref(P, Qaug, kp: GraphKernelPlan):
    r, c = _coo(kp)
    m = torch.relu(P.index_select(0, r) + Qaug.index_select(0, c))
    return torch.zeros_like(P).index_add_(0, r, m)


def segment_sum_ref(xaug, kp: GraphKernelPlan, mean: bool = False, slot_w=None):
    r, c = _coo(kp)
    m = xaug.index_select(0, c)
    if slot_w is not None:
        m = m * slot_w.to(m.dtype).unsqueeze(1)
    out = torch.zeros((kp.num_local, xaug.shape[1]), dtype=xaug.dtype, device=xaug.device).index_add_(0, r, m)
    if mean:
        deg = (kp.rowptr[1:] - kp.rowptr[:-1]).clamp(min=1).to(out.dtype)
        out = out / deg.unsqueeze(1)
    return out


# ------------------------------------------------------------------------------------------------
# native launch helpers
# ------------------------------------------------------------------------------------------------
def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensor expected"
    return t.stride(0)


def _native_seg_sum(ptr, idx, n_rows, src, slot_w, mean, peer=None, out=None, out2=None, out_split=0):
    """``peer = (pointer table [W], W, shift, ld_peer)``: ``idx`` is a ``GraphKernelPlan.col_peer(shift)`` array and negative entries are
    read from the owner rank inside the kernel.  ``out2``: rows ``>= out_split`` go to ``out2[r - out_split]`` (a symmetric slot)."""
    H = src.shape[1]
    if out is None:
        out = torch.empty((n_rows if out2 is None else out_split, H), dtype=torch.bfloat16, device=src.device)
    table, world, shift, ld_peer = (None, 0, 0, 0) if peer is None else peer
    N.check(N.lib().dgb_seg_sum(N.ptr(ptr), N.ptr(idx), N.i64(n_rows), N.i32(H), N.ptr(src), N.i64(_ld(src)),
                                N.ptr(slot_w), N.i32(1 if mean else 0), N.ptr(out), N.i64(_ld(out)), N.ptr(table), N.i32(world),
                                N.i32(shift), N.i64(ld_peer), N.ptr(out2), N.i64(0 if out2 is None else _ld(out2)),
                                N.i64(out_split), N.stream()),
            "dgb_seg_sum")
    N.count_launch()
    return out


def _mean_bwd_weights(kp: GraphKernelPlan):
    w = getattr(kp, "_csc_invdeg", None)
    if w is None:
        deg = (kp.rowptr[1:] - kp.rowptr[:-1]).clamp(min=1).float()
        w = (1.0 / deg)[kp.row.long()].contiguous()
        setattr(kp, "_csc_invdeg", w)
    return w


class _SegSum(Function):
    @staticmethod
    def forward(ctx, xaug, kp, mean):
        ctx.kp, ctx.mean = kp, mean
        return _native_seg_sum(kp.rowptr, kp.col, kp.num_local, xaug, None, mean)

    @staticmethod
    def backward(ctx, g):
        kp = ctx.kp
        w = _mean_bwd_weights(kp) if ctx.mean else None
        gx = _native_seg_sum(kp.colptr, kp.row, kp.num_aug, g.contiguous(), w, False)
        return gx, None, None


def segment_sum(xaug: torch.Tensor, kp: GraphKernelPlan, mean: bool = False) -> torch.Tensor:
    """``out[i] = sum_{j in N(i)} xaug[j]`` (``mean=True`` divides by the degree).  ``xaug = cat([local, halo])``."""
    if xaug.is_cuda and xaug.dtype == torch.bfloat16:
        N.lib()
        return _SegSum.apply(xaug if xaug.stride(1) == 1 else xaug.contiguous(), kp, mean)
    return segment_sum_ref(xaug, kp, mean)


def segment_mean(xaug, kp):
    return segment_sum(xaug, kp, mean=True)


def _check_degree(kp: GraphKernelPlan) -> None:
    from dgraph_b200.parallel.kernel_plan import MAX_COUNTED_DEGREE

    if kp.max_row_degree() > MAX_COUNTED_DEGREE:
        raise RuntimeError(
            f"fused GraphConv aggregation: a vertex has {kp.max_row_degree()} neighbours, more than the {MAX_COUNTED_DEGREE} its "
            "16-bit per-column counters of active edges can hold (GraphConvLayer falls back to the edge-materialising formulation "
            "for such graphs)")


class _GraphConvAgg(Function):
    @staticmethod
    def forward(ctx, P, Qaug, kp: GraphKernelPlan):
        H = P.shape[1]
        n = kp.num_local
        out = torch.empty((n, H), dtype=torch.bfloat16, device=P.device)
        need_p = P.requires_grad
        need_q = Qaug.requires_grad
        cnt = torch.empty((n, H), dtype=torch.uint16, device=P.device) if need_p else None
        # ReLU bit masks (1 byte per edge and 8 columns = E * H / 8 bytes) let backward gather ONE row per edge
        mask = torch.empty((kp.num_edges, H // 8), dtype=torch.uint8, device=P.device) if (need_q and SAVE_RELU_MASK) else None
        smap = kp.csr_to_csc() if (mask is not None and MASK_IN_CSC_ORDER) else None
        N.check(N.lib().dgb_seg_add_relu_fwd_ex(
            N.ptr(kp.rowptr), N.ptr(None), N.ptr(kp.col), N.i64(n), N.i32(H), N.ptr(P), N.i64(_ld(P)), N.ptr(Qaug),
            N.i64(_ld(Qaug)), N.ptr(out), N.i64(H), N.ptr(cnt), N.i64(H), N.ptr(mask), N.ptr(smap), N.i32(0),
            N.ptr(None), N.i32(0), N.i32(0), N.i64(0), N.stream()), "dgb_seg_add_relu_fwd")
        ctx.mask_csc = smap is not None
        N.count_launch()
        if mask is not None:
            ctx.save_for_backward(cnt, mask)
        else:
            ctx.save_for_backward(cnt, P, Qaug)
        ctx.has_mask = mask is not None
        ctx.kp = kp
        return out

    @staticmethod
    def backward(ctx, g):
        kp = ctx.kp
        # incoming gradient zero outside a known set of rows (loss over the labelled vertices): walk only the edges that start there
        hint = _sparsity.take(g) if (ctx.has_mask and not ctx.mask_csc) else None
        g = g.contiguous()
        H = g.shape[1]
        dP = dQ = None
        if ctx.has_mask:
            cnt, mask = ctx.saved_tensors
        else:
            cnt, P, Qaug = ctx.saved_tensors
        if ctx.needs_input_grad[0]:
            dP = torch.empty_like(g)
            N.check(N.lib().dgb_mul_cnt(N.ptr(g), N.i64(H), N.ptr(cnt), N.i64(H), N.ptr(dP), N.i64(H),
                                        N.i64(g.shape[0]), N.i32(H), N.stream()), "dgb_mul_cnt")
            N.count_launch()
        if ctx.needs_input_grad[1] and ctx.has_mask:
            dQ = torch.empty((kp.num_aug, H), dtype=torch.bfloat16, device=g.device)
            colptr, row, slot_map = (kp.colptr, kp.row, kp.csc_to_csr) if hint is None else kp.restricted_csc(hint[0])
            N.check(N.lib().dgb_seg_masked_sum(
                N.ptr(colptr), N.ptr(None), N.ptr(row), N.ptr(None if ctx.mask_csc else slot_map), N.ptr(mask),
                N.i64(kp.num_aug), N.i32(H), N.ptr(g),
                N.i64(H), N.ptr(dQ), N.i64(H), N.ptr(None), N.i64(0), N.ptr(None), N.i64(0), N.i64(0), N.ptr(None), N.i64(0),
                N.i64(0), N.i32(0), N.stream()),
                "dgb_seg_masked_sum")
            N.count_launch()
        elif ctx.needs_input_grad[1]:
            dQ = torch.empty((kp.num_aug, H), dtype=torch.bfloat16, device=g.device)
            N.check(N.lib().dgb_seg_add_relu_bwd(
                N.ptr(kp.colptr), N.ptr(kp.row), N.i64(kp.num_aug), N.i32(H), N.ptr(g), N.i64(H), N.ptr(P),
                N.i64(_ld(P)), N.ptr(Qaug), N.i64(_ld(Qaug)), N.ptr(dQ), N.i64(H), N.stream()),
                "dgb_seg_add_relu_bwd")
            N.count_launch()
        return dP, dQ, None


def graphconv_aggregate(P: torch.Tensor, Qaug: torch.Tensor, kp: GraphKernelPlan) -> torch.Tensor:
    """``out[i] = sum_{(i,j)} relu(P[i] + Qaug[j])`` for the local rows ``i``; ``Qaug`` has local + halo rows."""
    if P.is_cuda and P.dtype == torch.bfloat16:
        N.lib()
        assert kp.rowptr.dtype == torch.int32, "per-rank edge count must stay below 2^31"
        _check_degree(kp)
        if P.stride(1) != 1:
            P = P.contiguous()
        if Qaug.stride(1) != 1:
            Qaug = Qaug.contiguous()
        return _GraphConvAgg.apply(P, Qaug, kp)
    return graphconv_aggregate_ref(P, Qaug, kp)
