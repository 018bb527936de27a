"""Segment softmax + attention-weighted aggregation (``csrc/softmax_ops.cu``), with autograd.

``out[i] = sum_{e: dst_e = i} softmax_e(LeakyReLU(s_i[i] + s_j[src_e])) * h_j[src_e]`` over the CSR of a
:class:`GraphKernelPlan` -- the whole attention stage of the reference GAT layer (``experiments/OGB-LSC/RGAT.py:147-165,
191-202``) in one kernel, max-stabilised (the reference's ``exp`` is not).
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops.segment import _coo
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan


def seg_softmax_aggregate_ref(s_i, s_j, h_j, kp: GraphKernelPlan, slope: float = 0.2):
    r, c = _coo(kp)
    sc = torch.nn.functional.leaky_relu(s_i.index_select(0, r) + s_j.index_select(0, c), slope)
    mx = torch.full((kp.num_local,), float("-inf"), dtype=sc.dtype, device=sc.device)
    mx = mx.scatter_reduce(0, r, sc, reduce="amax", include_self=True)
    e = torch.exp(sc - mx.index_select(0, r))
    den = torch.zeros(kp.num_local, dtype=sc.dtype, device=sc.device).index_add_(0, r, e)
    alpha = e / den.index_select(0, r)
    msg = h_j.index_select(0, c) * alpha.unsqueeze(1).to(h_j.dtype)
    return torch.zeros((kp.num_local, h_j.shape[1]), dtype=h_j.dtype, device=h_j.device).index_add_(0, r, msg)


class _SegSoftmaxAgg(Function):
    @staticmethod
    def forward(ctx, s_i, s_j, h_j, kp, slope):
        n, C = kp.num_local, h_j.shape[1]
        dev = h_j.device
        out = torch.empty((n, C), dtype=torch.bfloat16, device=dev)
        rmax = torch.empty(n, dtype=torch.float32, device=dev)
        rden = torch.empty(n, dtype=torch.float32, device=dev)
        N.check(N.lib().dgb_seg_softmax_fwd(N.ptr(kp.rowptr), N.ptr(kp.col), N.i64(n), N.i32(C), N.f32(slope), N.ptr(s_i),
                                            N.ptr(s_j), N.ptr(h_j), N.i64(h_j.stride(0)), N.ptr(out), N.i64(C), N.ptr(rmax),
                                            N.ptr(rden), N.stream()), "dgb_seg_softmax_fwd")
        N.count_launch()
        ctx.save_for_backward(s_i, s_j, h_j, out, rmax, rden)
        ctx.kp, ctx.slope = kp, slope
        return out

    @staticmethod
    def backward(ctx, g):
        s_i, s_j, h_j, out, rmax, rden = ctx.saved_tensors
        kp = ctx.kp
        g = g.contiguous()
        C = h_j.shape[1]
        dev = g.device
        dscore = torch.empty(max(kp.num_edges, 1), dtype=torch.float32, device=dev)
        ds_i = torch.empty(kp.num_local, dtype=torch.float32, device=dev)
        ds_j = torch.empty(kp.num_aug, dtype=torch.float32, device=dev)
        dh = torch.empty((kp.num_aug, C), dtype=torch.bfloat16, device=dev)
        N.check(N.lib().dgb_seg_softmax_bwd(
            N.ptr(kp.rowptr), N.ptr(kp.col), N.ptr(kp.colptr), N.ptr(kp.row), N.ptr(kp.csc_to_csr), N.i64(kp.num_local),
            N.i64(kp.num_aug), N.i32(C), N.f32(ctx.slope), N.ptr(s_i), N.ptr(s_j), N.ptr(h_j), N.i64(h_j.stride(0)),
            N.ptr(out), N.i64(C), N.ptr(rmax), N.ptr(rden), N.ptr(g), N.i64(C), N.ptr(dscore), N.ptr(ds_i), N.ptr(ds_j),
            N.ptr(dh), N.i64(C), N.stream()), "dgb_seg_softmax_bwd")
        N.count_launch(2)
        return ds_i, ds_j, dh, None, None


def seg_softmax_aggregate(s_i: torch.Tensor, s_j: torch.Tensor, h_j: torch.Tensor, kp: GraphKernelPlan,
                          slope: float = 0.2) -> torch.Tensor:
    """``s_i [n_dst]``, ``s_j [n_src_aug]`` fp32 logit halves; ``h_j [n_src_aug, C]`` neighbour features."""
    if h_j.is_cuda and h_j.dtype == torch.bfloat16 and h_j.shape[1] % 8 == 0:
        N.lib()
        if h_j.stride(1) != 1 or h_j.stride(0) % 8:
            h_j = h_j.contiguous()
        return _SegSoftmaxAgg.apply(s_i.float().contiguous(), s_j.float().contiguous(), h_j, kp, float(slope))
    return seg_softmax_aggregate_ref(s_i.to(h_j.dtype), s_j.to(h_j.dtype), h_j, kp, slope)
