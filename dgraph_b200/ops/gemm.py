"""Dense / gathered GEMM ops.

``linear(x, w, b, act)`` computes ``act(x @ w.T + b)``.  CUDA bf16 inputs run the hand-written tcgen05 kernel in
``csrc/gemm_tcgen05.cu`` (TMA-fed, TMEM accumulators, fused bias/activation epilogue); other inputs use ATen.
Replaces the reference's ``nn.Linear`` -> cuBLAS sgemm calls on the hot paths (``experiments/OGB/GCN.py:51-52``,
``experiments/GraphCast/layers.py:52-63``).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _sparsity

_ACT = {None: 0, "none": 0, "relu": 1, "silu": 2}


def _act_ref(y, act):
    if act in (None, "none"):
        return y
    return {"relu": torch.relu, "silu": F.silu}[act](y)


def linear_ref(x, w, b=None, act=None):
    return _act_ref(F.linear(x, w, b), act)


def _tc_ok(x, w) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
            and os.environ.get("DGRAPH_B200_GEMM", "tcgen05") == "tcgen05" and N.available()
            and hasattr(N.lib(), "dgb_gemm_bf16_tn"))


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act=None,
            out_dtype=torch.bfloat16, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``C[M, N] = act(A[M, K] @ B[N, K]^T + bias)``; A, B bf16 row-major (K contiguous), fp32 accumulate."""
    from dgraph_b200.ops.gemm_native import gemm_bf16_tn

    return gemm_bf16_tn(a, b, bias, act, out_dtype, out)


class _LinearTC(Function):
    """tcgen05 forward + backward (dX via the same kernel; dW via a K-split transposed-operand kernel)."""

    @staticmethod
    def forward(ctx, x, w, b, act):
        y = gemm_tn(x, w, b, act)
        ctx.act = act
        ctx.save_for_backward(x, w, y if act not in (None, "none") else None)
        ctx.has_bias = b is not None
        return y

    @staticmethod
    def backward(ctx, g):
        from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w, act_backward

        x, w, y = ctx.saved_tensors
        hint = _sparsity.take(g) if ctx.act in (None, "none") else None
        if hint is not None:  # g is zero outside rows `idx`: work on the compact rows only (bit-identical results)
            idx, gcp = hint  # compact rows, possibly with zero padding columns up to a 16-byte row
            gc = gcp[:, :w.shape[0]]
            dx = dw = db = None
            if ctx.needs_input_grad[0]:
                dx_sel = gemm_tn(gcp, w.t().contiguous())
                dx = torch.zeros((x.shape[0], w.shape[1]), dtype=dx_sel.dtype, device=g.device)
                dx.index_copy_(0, idx, dx_sel)
                _sparsity.put(dx, idx, dx_sel)
            if ctx.needs_input_grad[1]:
                dw = gemm_bf16_nn_grad_w(gc, x, x_idx=idx)
            if ctx.has_bias and ctx.needs_input_grad[2]:
                from dgraph_b200.ops.reduce import colsum

                db = colsum(gcp)[:w.shape[0]]
            return dx, dw, db, None
        g = g.contiguous()
        if ctx.act not in (None, "none"):
            g = act_backward(g, y, ctx.act)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm_tn(g, w.t().contiguous())
        if ctx.needs_input_grad[1]:
            dw = gemm_bf16_nn_grad_w(g, x)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            from dgraph_b200.ops.reduce import colsum

            db = colsum(g)
        return dx, dw, db, None


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None, act=None) -> torch.Tensor:
    if _tc_ok(x, w):
        if act == "silu" and torch.is_grad_enabled() and (x.requires_grad or w.requires_grad):
            return F.silu(_LinearTC.apply(x, w, b, None))  # SiLU backward needs the pre-activation
        return _LinearTC.apply(x, w, b, act)
    if x.is_cuda:
        N.require_for(x)
    return linear_ref(x, w, b, act)
