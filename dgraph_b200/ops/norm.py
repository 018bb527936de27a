"""Fused LayerNorm (+ residual) and SiLU gradient (``csrc/norm_ops.cu``) with autograd; PyTorch formulation elsewhere."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd import Function

from dgraph_b200.ops import _native as N


def layer_norm_residual_ref(z, gamma, beta, eps, residual=None):
    y = F.layer_norm(z.float() if z.dtype in (torch.bfloat16, torch.float16) else z, (z.shape[-1],),
                     gamma.to(torch.float32 if z.dtype != torch.float64 else torch.float64),
                     beta.to(torch.float32 if z.dtype != torch.float64 else torch.float64), eps).to(z.dtype)
    return y if residual is None else y + residual


class _LNRes(Function):
    @staticmethod
    def forward(ctx, z, gamma, beta, eps, residual, c_valid):
        M, C = z.shape  # C is the padded width (multiple of 8)
        y = torch.empty((M, C), dtype=torch.bfloat16, device=z.device)
        mean = torch.empty(M, dtype=torch.float32, device=z.device)
        rstd = torch.empty(M, dtype=torch.float32, device=z.device)
        gf = torch.zeros(C, dtype=torch.float32, device=z.device)
        bf = torch.zeros(C, dtype=torch.float32, device=z.device)
        gf[:c_valid] = gamma
        bf[:c_valid] = beta
        N.check(N.lib().dgb_layernorm_fwd(N.ptr(z), N.i64(z.stride(0)), N.ptr(residual),
                                          N.i64(0 if residual is None else residual.stride(0)), N.ptr(gf), N.ptr(bf), N.ptr(y),
                                          N.i64(C), N.ptr(mean), N.ptr(rstd), N.i64(M), N.i32(C), N.i32(c_valid), N.f32(eps),
                                          N.stream()), "dgb_layernorm_fwd")
        N.count_launch()
        ctx.save_for_backward(z, gf, mean, rstd)
        ctx.has_res, ctx.c_valid = residual is not None, c_valid
        return y

    @staticmethod
    def backward(ctx, gy):
        z, gf, mean, rstd = ctx.saved_tensors
        M, C = z.shape
        cv = ctx.c_valid
        gy = gy.contiguous()
        dz = torch.empty_like(gy)
        lib = N.lib()
        lib.dgb_layernorm_bwd_scratch_bytes.restype = ctypes.c_int64
        nb = int(lib.dgb_layernorm_bwd_scratch_bytes(N.i64(M), N.i32(C)))
        scratch = torch.empty(max(nb // 4, 1), dtype=torch.float32, device=gy.device)
        dgb = torch.empty(2 * C, dtype=torch.float32, device=gy.device)
        N.check(lib.dgb_layernorm_bwd(N.ptr(gy), N.i64(C), N.ptr(z), N.i64(z.stride(0)), N.ptr(gf), N.ptr(mean), N.ptr(rstd),
                                      N.ptr(dz), N.i64(C), N.ptr(scratch), N.ptr(dgb), N.i64(M), N.i32(C), N.i32(cv), N.stream()),
                "dgb_layernorm_bwd")
        N.count_launch(2)
        return dz, dgb[:cv], dgb[C:C + cv], None, (gy if ctx.has_res else None), None


def layer_norm_residual(z: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
                        residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``LayerNorm(z) * gamma + beta (+ residual)`` over the last dimension of a 2-D tensor.  Widths that are not a multiple of
    8 (GraphCast's 73 output channels) run on a zero-padded copy with the statistics restricted to the valid columns."""
    C = z.shape[-1]
    Cp = (C + 7) // 8 * 8
    if z.is_cuda and z.dtype == torch.bfloat16 and z.dim() == 2 and Cp <= 512 and N.available():
        if Cp != C:
            zp = F.pad(z, (0, Cp - C))
            rp = None if residual is None else F.pad(residual.to(torch.bfloat16), (0, Cp - C))
            return _LNRes.apply(zp, gamma, beta, float(eps), rp, C)[:, :C]
        if z.stride(1) != 1 or z.stride(0) % 8:
            z = z.contiguous()
        if residual is not None and (residual.dtype != torch.bfloat16 or residual.stride(-1) != 1 or residual.stride(0) % 8):
            residual = residual.to(torch.bfloat16).contiguous()
        return _LNRes.apply(z, gamma, beta, float(eps), residual, C)
    if z.is_cuda:
        N.require_for(z)
    return layer_norm_residual_ref(z, gamma, beta, eps, residual)


class _SiLU(Function):
    """SiLU whose backward is one fused bf16 kernel (ATen: sigmoid + 4 elementwise kernels on fp32 temporaries)."""

    @staticmethod
    def forward(ctx, pre):
        ctx.save_for_backward(pre)
        return F.silu(pre)

    @staticmethod
    def backward(ctx, g):
        (pre,) = ctx.saved_tensors
        g = g.contiguous()
        out = torch.empty_like(g)
        N.check(N.lib().dgb_silu_bwd(N.ptr(g), N.ptr(pre), N.ptr(out), N.i64(g.numel()), N.stream()), "dgb_silu_bwd")
        N.count_launch()
        return out


def silu(pre: torch.Tensor) -> torch.Tensor:
    if pre.is_cuda and pre.dtype == torch.bfloat16 and pre.is_contiguous() and pre.numel() % 8 == 0 and N.available():
        return _SiLU.apply(pre)
    return F.silu(pre)


def silu_backward(g: torch.Tensor, pre: torch.Tensor) -> torch.Tensor:
    if g.is_cuda and g.dtype == torch.bfloat16 and pre.dtype == torch.bfloat16 and g.is_contiguous() and pre.is_contiguous() \
            and g.numel() % 8 == 0 and N.available():
        out = torch.empty_like(g)
        N.check(N.lib().dgb_silu_bwd(N.ptr(g), N.ptr(pre), N.ptr(out), N.i64(g.numel()), N.stream()), "dgb_silu_bwd")
        N.count_launch()
        return out
    sg = torch.sigmoid(pre.float())
    return (g.float() * (sg * (1 + pre.float() * (1 - sg)))).to(g.dtype)
