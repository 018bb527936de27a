"""Fused GraphConv layer: ``out_i = sum_j ReLU(W [x_i ; x_j] + b)`` as ONE autograd node.

Forward  : one tcgen05 GEMM ``PQ = x_aug [W_i ; W_j]^T (+ b on the P half)`` -> one fused gather-add-ReLU-segment-sum.
Backward : ``dP = g * cnt`` and ``dQ = masked CSC sum`` written straight into the two halves of one ``[n_aug, 2H]``
           buffer -> one GEMM for ``dX`` (skipped when the input needs no gradient), one split-K tcgen05 GEMM for ``dW``, one column-sum
           for ``db``.  No ``[E, *]`` tensor, no ``cat``/slice copies, no gradient-accumulation adds.

Semantics: reference ``GraphConvLayer.forward`` (``experiments/OGB/GCN.py:34-67``).
"""
from __future__ import annotations

import os

import torch
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w, gemm_bf16_tn
from dgraph_b200.ops.reduce import colsum
from dgraph_b200.ops import segment as S
from dgraph_b200.ops import _sparsity
from dgraph_b200.ops import _pending


class _GraphConvFused(Function):
    @staticmethod
    def forward(ctx, x_aug, weight, bias, kp, n_local, f_in):
        H = weight.shape[0]
        Fp = x_aug.shape[1]
        dev = x_aug.device
        n_aug = x_aug.shape[0]
        # [W_i ; W_j] stacked along the output dimension, zero-padded to the (16-byte aligned) feature width
        wcat = torch.zeros((2 * H, Fp), dtype=torch.bfloat16, device=dev)
        wcat[:H, :f_in] = weight[:, :f_in]
        wcat[H:, :f_in] = weight[:, f_in:2 * f_in]
        bcat = torch.zeros(2 * H, dtype=torch.float32, device=dev)
        bcat[:H] = bias
        # local rows need P and Q (one GEMM, N = 2H); halo rows only Q (N = H) written into the same [n_aug, 2H] buffer
        PQ = torch.empty((n_aug, 2 * H), dtype=torch.bfloat16, device=dev)
        pull_done = _pending.take(x_aug)  # halo rows still arriving over NVLink (HaloExchange.augmented(defer=True))
        gemm_bf16_tn(x_aug[:n_local], wcat, bcat, None, out=PQ[:n_local])
        P, Q = PQ[:n_local, :H], PQ[:, H:]
        need_w = weight.requires_grad or x_aug.requires_grad
        out = torch.empty((n_local, H), dtype=torch.bfloat16, device=dev)
        cnt = torch.empty((n_local, H), dtype=torch.uint16, device=dev) if need_w else None
        mask = torch.empty((kp.num_edges, H // 8), dtype=torch.uint8, device=dev) if need_w else None
        smap = kp.csr_to_csc() if (mask is not None and S.MASK_IN_CSC_ORDER) else None

        def aggregate(ptr, ptr_end, accumulate):
            N.check(N.lib().dgb_seg_add_relu_fwd_ex(
                N.ptr(ptr), N.ptr(ptr_end), N.ptr(kp.col), N.i64(n_local), N.i32(H), N.ptr(P), N.i64(2 * H), N.ptr(Q),
                N.i64(2 * H), N.ptr(out), N.i64(H), N.ptr(cnt), N.i64(H), N.ptr(mask), N.ptr(smap), N.i32(accumulate),
                N.i64(n_aug), N.ptr(None), N.ptr(None), N.ptr(None), N.i64(0), N.stream()), "dgb_seg_add_relu_fwd")
            N.count_launch()

        # DGRAPH_B200_OVERLAP (opt-in, default "0" = pull on the main stream): "gemm" overlaps only the local-row GEMM above with
        # the pull (measured: no gain at 2 GPUs, 13.08 vs 13.10 ms/epoch -- both sides are HBM-bound); "split" also splits the
        # aggregation (measured slower at 2 GPUs, 14.4 ms: the second pass re-reads P / out / cnt of almost every row).
        split = (pull_done is not None and n_aug > n_local and kp.num_nbr_local == n_local
                 and os.environ.get("DGRAPH_B200_OVERLAP", "0") == "split")
        if split:
            # pass 1 touches only local neighbours (Q rows < n_local) and overlaps the pull; pass 2 adds the halo neighbours
            mid = kp.rowmid()
            aggregate(kp.rowptr, mid, 0)
            torch.cuda.current_stream().wait_event(pull_done)
            gemm_bf16_tn(x_aug[n_local:], wcat[H:], None, None, out=PQ[n_local:, H:])
            aggregate(mid, kp.rowptr[1:], 1)
        else:
            if pull_done is not None:
                torch.cuda.current_stream().wait_event(pull_done)
            if n_aug > n_local:
                gemm_bf16_tn(x_aug[n_local:], wcat[H:], None, None, out=PQ[n_local:, H:])
            aggregate(kp.rowptr, None, 0)
        ctx.mask_csc = smap is not None
        ctx.save_for_backward(x_aug, wcat, cnt, mask)
        ctx.kp, ctx.n_local, ctx.f_in, ctx.H = kp, n_local, f_in, H
        return out

    @staticmethod
    def backward(ctx, g):
        x_aug, wcat, cnt, mask = ctx.saved_tensors
        kp, n, f_in, H = ctx.kp, ctx.n_local, ctx.f_in, ctx.H
        hint = _sparsity.take(g)
        if ctx.mask_csc:
            hint = None  # the restricted walk addresses the mask through csc_to_csr
        g = g.contiguous()
        # g zero outside rows `idx` (loss over the labelled vertices): walk only the edges that start there
        colptr, row, slot_map = (kp.colptr, kp.row, kp.csc_to_csr) if hint is None else kp.restricted_csc(hint[0])
        n_aug = x_aug.shape[0]
        dev = g.device
        # [dP | dQ] for the local rows, dQ only for the halo rows (their P half is never read)
        dPQ = torch.empty((n_aug, 2 * H), dtype=torch.bfloat16, device=dev)
        lib = N.lib()
        dQ = dPQ[:, H:]
        # one CSC walk writes dQ (masked sums) for all rows and dP = g * cnt for the local rows
        N.check(lib.dgb_seg_masked_sum(N.ptr(colptr), N.ptr(row), N.ptr(None if ctx.mask_csc else slot_map), N.ptr(mask),
                                       N.i64(n_aug),
                                       N.i32(H), N.ptr(g), N.i64(H), N.ptr(dQ), N.i64(2 * H), N.ptr(cnt), N.i64(H),
                                       N.ptr(dPQ), N.i64(2 * H), N.i64(n), N.stream()),
                "dgb_seg_masked_sum")
        N.count_launch(1)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((n_aug, x_aug.shape[1]), dtype=torch.bfloat16, device=dev)
            wt = wcat.t().contiguous()  # [Fp, 2H]
            gemm_bf16_tn(dPQ[:n], wt, None, None, out=dx[:n])
            if n_aug > n:
                gemm_bf16_tn(dQ[n:], wt[:, H:], None, None, out=dx[n:])
        if ctx.needs_input_grad[1]:
            dw_i = gemm_bf16_nn_grad_w(dPQ[:n, :H], x_aug[:n])  # [H, Fp] fp32, split-K tcgen05 kernel
            dw_j = gemm_bf16_nn_grad_w(dQ, x_aug)
            dw = torch.cat([dw_i[:, :f_in], dw_j[:, :f_in]], dim=1)
        if ctx.needs_input_grad[2]:
            db = colsum(dPQ[:n, :H])
        return dx, dw, db, None, None, None


def graphconv_fused(x_aug, weight, bias, kp, n_local: int, f_in: int):
    """``x_aug`` bf16 ``[n_local + n_halo, F_pad]`` (extra columns zero); ``weight`` fp32 ``[H, 2 * f_in]``; ``bias`` fp32."""
    N.lib()
    assert kp.rowptr.dtype == torch.int32
    S._check_degree(kp)
    if x_aug.stride(1) != 1 or x_aug.stride(0) % 8:
        x_aug = x_aug.contiguous()
    return _GraphConvFused.apply(x_aug, weight, bias, kp, n_local, f_in)
