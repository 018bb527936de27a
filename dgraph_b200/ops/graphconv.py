"""Fused GraphConv layer: ``out_i = sum_j ReLU(W [x_i ; x_j] + b)`` as ONE autograd node, local or distributed.

Single rank / features already exchanged (``_GraphConvFused``):
  forward  : one tcgen05 GEMM ``PQ = x_aug [W_i ; W_j]^T (+ b on the P half)`` -> one fused gather-add-ReLU-segment-sum.
  backward : ``dP = g * cnt`` and ``dQ = masked CSC sum`` written straight into the two halves of one ``[n_aug, 2H]`` buffer ->
             one GEMM for ``dX`` (skipped when the input needs no gradient), one split-K tcgen05 GEMM for ``dW``, one column-sum
             for ``db``.

Distributed, exchange fused into the layer (``_GraphConvPullQ``; the default on the peer-memory engine when the input needs a
gradient or is at least as wide as the layer): every rank projects only the rows it owns and the *projected* rows travel.
  forward  : the GEMM writes ``[P | Q]`` of the owned rows straight into the rank's symmetric slot (no publish copy) -> flag
             barrier -> the aggregation kernel reads the ``Q`` rows of remote neighbours from their owners over NVLink in the same
             batch of loads as the local ones (``pull="kernel"``: no halo buffer, no exchange kernel), or -- when every halo row is
             re-read by several edges -- ONE pull kernel fills the halo rows of the slot first (``pull="buffer"``).
  backward : the CSC walk writes ``dQ`` of the halo rows straight into a second symmetric slot -> flag barrier -> the owners
             pull-reduce them *in place* into their ``dQ`` rows (deterministic order) -> ``dX`` / ``dW`` / ``db`` on owned rows only.
  Halo rows therefore cost no GEMM in either direction, nothing is copied into or out of the slots, and no ``[E, *]`` or
  ``cat([x, halo])`` tensor exists.

Semantics: reference ``GraphConvLayer.forward`` (``experiments/OGB/GCN.py:34-67``) behind ``HaloExchange`` (``haloExchange.py:121-139``).
"""
from __future__ import annotations

import os

import torch
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _pending, _sparsity, _wcache
from dgraph_b200.ops import segment as S
from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w, gemm_bf16_tn
from dgraph_b200.ops.reduce import colsum


def packed_weights(weight: torch.Tensor, bias, f_in: int, Fp: int):
    """``(wcat [2H, Fp] bf16, wcat_t [Fp, 2H] bf16, bcat [2H] fp32)`` of ``weight = [W_i | W_j]`` (``[H, 2 * f_in]``): one
    ``pack_graphconv_kernel`` launch per layer and optimizer step (``ops/_wcache.py``)."""
    H = weight.shape[0]

    def build():
        dev = weight.device
        w = weight.detach()
        if w.dtype != torch.float32 or not w.is_contiguous():
            w = w.float().contiguous()
        b = None if bias is None else bias.detach().float().contiguous()
        wcat = torch.empty((2 * H, Fp), dtype=torch.bfloat16, device=dev)
        wcat_t = torch.empty((Fp, 2 * H), dtype=torch.bfloat16, device=dev)
        bcat = torch.empty(2 * H, dtype=torch.float32, device=dev)
        N.check(N.lib().dgb_pack_graphconv_weights(N.ptr(w), N.ptr(b), N.i32(H), N.i32(f_in), N.i32(Fp), N.ptr(wcat),
                                                   N.ptr(wcat_t), N.ptr(bcat), N.stream()), "dgb_pack_graphconv_weights")
        N.count_launch()
        return wcat, wcat_t, bcat

    return _wcache.get(weight, ("graphconv", f_in, Fp, None if bias is None else (bias._version, bias.data_ptr())), build)


def _aggregate(kp, ptr, ptr_end, idx, n_local, H, P, ld_p, Q, ld_q, out, cnt, mask, smap, accumulate, peer):
    table, world, shift, ld_peer = (None, 0, 0, 0) if peer is None else peer
    N.check(N.lib().dgb_seg_add_relu_fwd_ex(
        N.ptr(ptr), N.ptr(ptr_end), N.ptr(idx), N.i64(n_local), N.i32(H), N.ptr(P), N.i64(ld_p), N.ptr(Q), N.i64(ld_q),
        N.ptr(out), N.i64(H), N.ptr(cnt), N.i64(H), N.ptr(mask), N.ptr(smap), N.i32(accumulate), N.ptr(table), N.i32(world),
        N.i32(shift), N.i64(ld_peer), N.stream()), "dgb_seg_add_relu_fwd")
    N.count_launch()


def _aggregate_rows(kp, idx, n_local, H, P, ld_p, Q, ld_q, out, cnt, mask, smap, peer):
    """The whole forward aggregation: one launch, plus -- on graphs with hub vertices -- a second small launch over the chunks of the
    over-long rows (``GraphKernelPlan.hub_plan``) whose partial sums / counts are added up here."""
    hp = kp.hub_plan("csr")
    if hp is None:
        return _aggregate(kp, kp.rowptr, None, idx, n_local, H, P, ld_p, Q, ld_q, out, cnt, mask, smap, 0, peer)
    _aggregate(kp, kp.rowptr, hp.main_end, idx, n_local, H, P, ld_p, Q, ld_q, out, cnt, mask, smap, 0, peer)  # hub rows: empty -> zeros
    nv = hp.vrow.numel()
    Pv = P.index_select(0, hp.vrow).contiguous()
    part = torch.empty((nv, H), dtype=torch.bfloat16, device=P.device)
    pcnt = torch.empty((nv, H), dtype=torch.uint16, device=P.device) if cnt is not None else None
    _aggregate(kp, hp.vbeg, hp.vend, idx, nv, H, Pv, H, Q, ld_q, part, pcnt, mask, smap, 0, peer)
    nh = hp.hubs.numel()
    out.index_copy_(0, hp.hubs, torch.zeros((nh, H), dtype=torch.float32, device=P.device).index_add_(0, hp.seg, part.float())
                    .to(torch.bfloat16))
    if cnt is not None:
        tot = torch.zeros((nh, H), dtype=torch.int32, device=P.device).index_add_(0, hp.seg, pcnt.view(torch.int16).to(torch.int32) & 0xFFFF)
        cnt.view(torch.int16).index_copy_(0, hp.hubs, tot.to(torch.int16))  # degree <= 65535 is checked: the sum fits 16 bits


def _masked_sum_cols(kp, full_walk, colptr, row, slot_map, mask, n_aug, H, g, dQ, ld_dq, cnt, dP, ld_dp, dp_rows, dQ2=None, ld_dq2=0,
                     split=0, skip_empty=False):
    """The whole backward CSC walk; with hub columns (``full_walk`` only: a walk restricted to the labelled rows has no long
    columns) their chunks run as rows of a second small launch and are added up here."""
    hp = kp.hub_plan("csc") if full_walk else None
    if hp is None:
        return _masked_sum(colptr, row, slot_map, mask, n_aug, H, g, dQ, ld_dq, cnt, dP, ld_dp, dp_rows, dQ2, ld_dq2, split, skip_empty)
    _masked_sum(colptr, row, slot_map, mask, n_aug, H, g, dQ, ld_dq, cnt, dP, ld_dp, dp_rows, dQ2, ld_dq2, split, skip_empty,
                colptr_end=hp.main_end)
    nv = hp.vrow.numel()
    part = torch.empty((nv, H), dtype=torch.bfloat16, device=g.device)
    _masked_sum(hp.vbeg, row, slot_map, mask, nv, H, g, part, H, None, None, 0, 0, colptr_end=hp.vend)
    tot = torch.zeros((hp.hubs.numel(), H), dtype=torch.float32, device=g.device).index_add_(0, hp.seg, part.float()).to(torch.bfloat16)
    if dQ2 is None:
        dQ.index_copy_(0, hp.hubs, tot)
    else:  # hub columns that are halo rows live in the symmetric slot
        loc = hp.hubs < split
        if bool(loc.any()):
            dQ.index_copy_(0, hp.hubs[loc], tot[loc])
        if bool((~loc).any()):
            dQ2.index_copy_(0, hp.hubs[~loc] - split, tot[~loc])


def _masked_sum(colptr, row, slot_map, mask, n_aug, H, g, dQ, ld_dq, cnt, dP, ld_dp, dp_rows, dQ2=None, ld_dq2=0, split=0,
                skip_empty=False, colptr_end=None):
    N.check(N.lib().dgb_seg_masked_sum(N.ptr(colptr), N.ptr(colptr_end), N.ptr(row), N.ptr(slot_map), N.ptr(mask), N.i64(n_aug), N.i32(H), N.ptr(g),
                                       N.i64(H), N.ptr(dQ), N.i64(ld_dq), N.ptr(cnt), N.i64(H), N.ptr(dP), N.i64(ld_dp),
                                       N.i64(dp_rows), N.ptr(dQ2), N.i64(ld_dq2), N.i64(split), N.i32(1 if skip_empty else 0),
                                       N.stream()),
            "dgb_seg_masked_sum")
    N.count_launch()


class _GraphConvFused(Function):
    @staticmethod
    def forward(ctx, x_aug, weight, bias, kp, n_local, f_in):
        H = weight.shape[0]
        Fp = x_aug.shape[1]
        dev = x_aug.device
        n_aug = x_aug.shape[0]
        wcat, wcat_t, bcat = packed_weights(weight, bias, f_in, Fp)
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
            if ptr is kp.rowptr and ptr_end is None and not accumulate:
                return _aggregate_rows(kp, kp.col, n_local, H, P, 2 * H, Q, 2 * H, out, cnt, mask, smap, None)
            _aggregate(kp, ptr, ptr_end, kp.col, n_local, H, P, 2 * H, Q, 2 * H, out, cnt, mask, smap, accumulate, None)

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
        ctx.save_for_backward(x_aug, wcat_t, cnt, mask)
        ctx.kp, ctx.n_local, ctx.f_in, ctx.H = kp, n_local, f_in, H
        return out

    @staticmethod
    def backward(ctx, g):
        x_aug, wcat_t, cnt, mask = ctx.saved_tensors
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
        dQ = dPQ[:, H:]
        # one CSC walk writes dQ (masked sums) for all rows and dP = g * cnt for the local rows
        _masked_sum_cols(kp, hint is None, colptr, row, None if ctx.mask_csc else slot_map, mask, n_aug, H, g, dQ, 2 * H, cnt, dPQ,
                         2 * H, n)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((n_aug, x_aug.shape[1]), dtype=torch.bfloat16, device=dev)
            gemm_bf16_tn(dPQ[:n], wcat_t, None, None, out=dx[:n])
            if n_aug > n:
                gemm_bf16_tn(dQ[n:], wcat_t[:, H:], None, None, out=dx[n:])
        if ctx.needs_input_grad[1]:
            dw_i = gemm_bf16_nn_grad_w(dPQ[:n, :H], x_aug[:n])  # [H, Fp] fp32, split-K tcgen05 kernel
            dw_j = gemm_bf16_nn_grad_w(dQ, x_aug)
            dw = torch.cat([dw_i[:, :f_in], dw_j[:, :f_in]], dim=1)
        if ctx.needs_input_grad[2]:
            db = colsum(dPQ[:n, :H])
        return dx, dw, db, None, None, None


class _GraphConvPullQ(Function):
    """Distributed layer with the exchange inside: projected rows travel, pulled by the aggregation kernel (or one pull kernel)."""

    @staticmethod
    def forward(ctx, x, weight, bias, kp, f_in, engine, pat, tag):
        info = engine.pattern_info(pat)
        H = weight.shape[0]
        n, Fp = x.shape
        dev = x.device
        wcat, wcat_t, bcat = packed_weights(weight, bias, f_in, Fp)
        pull = engine.pull_mode(info, kp)
        # [P | Q] of the owned rows, produced in place in the symmetric slot; in "buffer" mode the halo rows' Q follows behind
        rows = info.max_aug if pull == "buffer" else info.max_nbr_local
        PQ, off = engine.slot_tensor(("pq", info.uid, tag, pull), (rows, 2 * H), torch.bfloat16)
        gemm_bf16_tn(x, wcat, bcat, None, out=PQ[:n])
        engine.barrier_on_stream()  # every rank's Q rows are complete and visible
        q_table = engine.arena.ptr_table(off + 2 * H)  # peers' Q halves start H bf16 into each row
        need_w = weight.requires_grad or x.requires_grad
        out = torch.empty((n, H), dtype=torch.bfloat16, device=dev)
        cnt = torch.empty((n, H), dtype=torch.uint16, device=dev) if need_w else None
        mask = torch.empty((kp.num_edges, H // 8), dtype=torch.uint8, device=dev) if need_w else None
        P, Q = PQ[:n, :H], PQ[:, H:]
        if pull == "buffer":
            if info.n_halo:
                N.check(N.lib().dgb_halo_pull(N.ptr(Q[n:]), N.i64(4 * H), N.i64(info.n_halo), N.i32(2 * H), N.ptr(q_table),
                                              N.i64(4 * H), N.ptr(info.halo_owner), N.ptr(info.halo_row), N.i64(info.rot),
                                              N.stream()), "dgb_halo_pull")
                N.count_launch()
            _aggregate_rows(kp, kp.col, n, H, P, 2 * H, Q, 2 * H, out, cnt, mask, None, None)
        else:
            _aggregate_rows(kp, kp.col_peer(info.peer_shift), n, H, P, 2 * H, Q, 2 * H, out, cnt, mask, None,
                            (q_table, engine.get_world_size(), info.peer_shift, 2 * H))
        engine.mark_read(off)
        ctx.save_for_backward(x, wcat_t, cnt, mask)
        ctx.kp, ctx.f_in, ctx.H, ctx.engine, ctx.pat, ctx.tag = kp, f_in, H, engine, pat, tag
        return out

    @staticmethod
    def backward(ctx, g):
        x, wcat_t, cnt, mask = ctx.saved_tensors
        kp, f_in, H, engine = ctx.kp, ctx.f_in, ctx.H, ctx.engine
        info = engine.pattern_info(ctx.pat)
        n = x.shape[0]
        hint = _sparsity.take(g)
        g = g.contiguous()
        colptr, row, slot_map = (kp.colptr, kp.row, kp.csc_to_csr) if hint is None else kp.restricted_csc(hint[0])
        dev = g.device
        dPQ = torch.empty((n, 2 * H), dtype=torch.bfloat16, device=dev)  # [dP | dQ] of the owned rows
        dQ = dPQ[:, H:]
        dQh, goff = engine.slot_tensor(("dqh", info.uid, ctx.tag), (info.max_halo, H), torch.bfloat16)
        # Structural sparsity travels with the exchange: when the incoming gradient is zero outside the labelled rows, a halo row
        # receives a gradient only if one of its edges starts at a labelled row.  The owners learn once per label set which of
        # their halo copies those are (engine.restricted_reduce_plan) and read only them; the walk does not even write the others.
        rplan = None if hint is None else engine.restricted_reduce_plan(info, kp, hint[0], colptr)
        # one CSC walk: dQ of owned rows -> dPQ, dQ of halo rows -> the symmetric slot, dP = g * cnt
        _masked_sum_cols(kp, hint is None, colptr, row, slot_map, mask, kp.num_aug, H, g, dQ, 2 * H, cnt, dPQ, 2 * H, n, dQh, H, n,
                         skip_empty=rplan is not None)
        engine.barrier_on_stream()
        engine.halo_grad_reduce_inplace(info, dQ, 2 * H, H, goff, H, plan=rplan)  # dQ[b] += sum over the holders of b's halo copies
        engine.mark_read(goff)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm_bf16_tn(dPQ, wcat_t, None, None)
        if ctx.needs_input_grad[1]:
            dwc = gemm_bf16_nn_grad_w(dPQ, x)  # [2H, Fp] fp32: both halves of the weight in one split-K GEMM
            dw = torch.cat([dwc[:H, :f_in], dwc[H:, :f_in]], dim=1)
        if ctx.needs_input_grad[2]:
            db = colsum(dPQ[:, :H])
        return dx, dw, db, None, None, None, None, None


def graphconv_fused(x_aug, weight, bias, kp, n_local: int, f_in: int):
    """``x_aug`` bf16 ``[n_local + n_halo, F_pad]`` (extra columns zero); ``weight`` fp32 ``[H, 2 * f_in]``; ``bias`` fp32."""
    N.lib()
    assert kp.rowptr.dtype == torch.int32
    S._check_degree(kp)
    if x_aug.stride(1) != 1 or x_aug.stride(0) % 8:
        x_aug = x_aug.contiguous()
    return _GraphConvFused.apply(x_aug, weight, bias, kp, n_local, f_in)


def graphconv_pull_q(x_local, weight, bias, kp, f_in: int, engine, pattern, tag):
    """Distributed form: ``x_local`` are the owned rows only; the layer exchanges what it needs itself (see module docstring).
    ``tag`` distinguishes the call sites that are live at the same time (one symmetric slot per layer)."""
    N.lib()
    assert kp.rowptr.dtype == torch.int32
    S._check_degree(kp)
    if x_local.stride(1) != 1 or x_local.stride(0) % 8:
        x_local = x_local.contiguous()
    return _GraphConvPullQ.apply(x_local, weight, bias, kp, f_in, engine, pattern, tag)
