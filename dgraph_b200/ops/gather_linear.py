"""``act(Linear(cat([x_0[idx_0], x_1[idx_1], x_2], 1)))`` without materialising the gather or the concatenation.

Forward on CUDA bf16: ONE tcgen05 gather-GEMM (``csrc/gemm_tcgen05.cu``, ``dgb_gather_gemm_bf16``): four producer warps
gather the rows of up to three sources straight into the 128-byte-swizzled A tiles, the weight streams in by TMA, bias +
activation are fused in the epilogue.  This is the first Linear of the reference's ``MeshEdgeBlock`` / ``MeshNodeBlock``
(``experiments/GraphCast/layers.py:136-143,205-216``) and of ``GraphConvLayer`` with edge features.

Backward (round 1): dense formulation -- ``d_cat = d_pre @ W`` with the tcgen05 GEMM, split and ``index_add`` back to
the sources; ``dW = d_pre^T @ cat`` needs the gathered operand once (library GEMM).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn.functional as F
from torch.autograd import Function

from dgraph_b200.ops import _native as N


def gather_linear_ref(srcs: Sequence[torch.Tensor], idxs: Sequence[Optional[torch.Tensor]], weight, bias, act=None):
    cat = torch.cat([s if i is None else s.index_select(0, i.long()) for s, i in zip(srcs, idxs)], dim=1)
    y = F.linear(cat, weight.to(cat.dtype), None if bias is None else bias.to(cat.dtype))
    if act == "silu":
        y = F.silu(y)
    elif act == "relu":
        y = torch.relu(y)
    return y


class _GatherLinear(Function):
    @staticmethod
    def forward(ctx, weight, bias, act, n_src, scatter_plans, *args):
        from dgraph_b200.ops.gemm_native import gather_gemm_bf16, pack_segment_weight

        srcs, idxs = list(args[:n_src]), list(args[n_src:])
        ks = [int(s.shape[1]) for s in srcs]
        M = int(idxs[0].shape[0]) if idxs[0] is not None else int(srcs[0].shape[0])
        for s, i in zip(srcs, idxs):
            M = int(i.shape[0]) if i is not None else int(s.shape[0])
        wp = pack_segment_weight(weight, ks)
        # ReLU runs in the GEMM epilogue and its gradient needs only the output; SiLU's gradient needs the pre-activation, so
        # that one is stored and the activation is a second kernel
        pre = gather_gemm_bf16(srcs, idxs, wp, bias, M, ks, act="relu" if act == "relu" else None)
        ctx.act, ctx.n_src, ctx.ks, ctx.scatter_plans = act, n_src, ks, scatter_plans
        ctx.save_for_backward(weight, pre if act else None, *srcs, *[i for i in idxs if i is not None])
        ctx.idx_none = [i is None for i in idxs]
        if act == "silu":
            return F.silu(pre)
        return pre

    @staticmethod
    def backward(ctx, g):
        from dgraph_b200.ops.gemm_native import gemm_bf16_tn

        saved = ctx.saved_tensors
        weight, pre = saved[0], saved[1]
        srcs = list(saved[2:2 + ctx.n_src])
        it = iter(saved[2 + ctx.n_src:])
        idxs = [None if none else next(it) for none in ctx.idx_none]
        g = g.contiguous()
        if ctx.act == "silu":
            from dgraph_b200.ops.norm import silu_backward

            g = silu_backward(g, pre)
        elif ctx.act == "relu":
            g = torch.ops.aten.threshold_backward(g, pre, 0)  # `pre` holds the ReLU output here: g where out > 0 else 0
        wb = weight.to(torch.bfloat16)
        dcat = gemm_bf16_tn(g, wb.t().contiguous())  # [M, sum(ks)]
        dw = None
        if ctx.needs_input_grad[0]:
            # per-source split-K weight gradients; gathered sources are gathered inside the kernel (no [M, K] temporaries)
            from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w

            parts = []
            for s_, i_ in zip(srcs, idxs):
                k = s_.shape[1]
                if k <= 256:
                    parts.append(gemm_bf16_nn_grad_w(g, s_, i_))
                else:
                    parts.append(torch.cat([gemm_bf16_nn_grad_w(g, s_[:, o:o + 256], i_) for o in range(0, k, 256)], dim=1))
            dw = torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]
        db = None
        if ctx.needs_input_grad[1]:
            from dgraph_b200.ops.reduce import colsum

            db = colsum(g)
        grads, o = [], 0
        for s, i, k in zip(srcs, idxs, ctx.ks):
            part = dcat[:, o:o + k]
            o += k
            if i is None:
                grads.append(part)
            else:
                sp = ctx.scatter_plans[len(grads)] if ctx.scatter_plans is not None else None
                if sp is not None:  # atomic-free: CSR walk over (source row <- edge id) incidence
                    from dgraph_b200.ops.segment import _native_seg_sum

                    pc = part if (part.stride(1) == 1 and part.stride(0) % 8 == 0 and part.data_ptr() % 16 == 0) else part.contiguous()
                    grads.append(_native_seg_sum(sp.rowptr, sp.col, sp.num_local, pc, None, False))
                else:
                    grads.append(torch.zeros_like(s).index_add_(0, i.long(), part))
        return (dw, db, None, None, None, *grads, *([None] * len(idxs)))


def gather_linear(srcs: List[torch.Tensor], idxs: List[Optional[torch.Tensor]], weight: torch.Tensor,
                  bias: Optional[torch.Tensor], act: Optional[str] = None, scatter_plans=None) -> torch.Tensor:
    """``scatter_plans[k]``: optional :class:`GraphKernelPlan` of the (row of ``srcs[k]`` <- edge id) incidence; when given, the
    backward scatters that source's gradient with an atomic-free segment sum instead of ``index_add_``."""
    ok = (all(s.is_cuda and s.dtype == torch.bfloat16 and s.shape[1] % 8 == 0 for s in srcs) and 1 <= len(srcs) <= 3
          and N.available())
    if not ok:
        if srcs[0].is_cuda:
            N.require_for(srcs[0])
        return gather_linear_ref(srcs, idxs, weight, bias, act)
    idx32 = [None if i is None else (i if i.dtype == torch.int32 else i.to(torch.int32)) for i in idxs]
    return _GatherLinear.apply(weight, bias, act, len(srcs), scatter_plans, *srcs, *idx32)
