"""One-kernel ``MeshGraphMLP`` block (F4): ``LayerNorm(Linear(SiLU(Linear(cat([x_k[idx_k]])))))) (+ residual)``.

The default path of the GraphCast blocks whose shape it supports (``DGRAPH_B200_FUSED_MLP=0`` switches back to the 3-4 kernel
chain): ``csrc/mlp_fused_tcgen05.cu`` (``UTCHMMA`` / ``UTMALDG`` / ``UTMASTG`` / ``LDTM``), validated against the fp32 PyTorch
formulation by ``tests/test_mlp_fused_gpu.py``; measured on B200: GraphCast 0.25 degree step 54.9 -> 54.4 ms on 1 GPU, 25.2 -> 24.6
ms on 4.  Semantics: the reference's ``MeshGraphMLP`` with one hidden
layer inside ``MeshEdgeBlock`` / ``MeshNodeBlock`` (``experiments/GraphCast/layers.py:24-75,136-143,205-216``), restricted to
hidden = out = 128 (the GraphCast default).

Forward: ONE launch; it stores ``pre`` (first Linear's output), ``z`` (second Linear's output), ``mean`` and ``rstd`` for the
backward.  Backward: the existing kernels -- fused LayerNorm gradient, tcgen05 GEMMs for ``dh`` / ``d cat``, split-K tcgen05
weight gradients (gathered sources are gathered inside the kernel), column sums for the biases, fused SiLU gradient, atomic-free
segment sums for the gathered sources' gradients.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

import torch
import torch.nn.functional as F
from torch.autograd import Function

from dgraph_b200.ops import _native as N

ENABLED = os.environ.get("DGRAPH_B200_FUSED_MLP", "1") == "1"
WIDTH = 128
UNSUPPORTED = 801


def mlp2_ln_ref(srcs: Sequence[torch.Tensor], idxs: Sequence[Optional[torch.Tensor]], w1, b1, w2, b2, gamma, beta, eps=1e-5,
                residual=None):
    """Plain PyTorch statement of the block (the oracle of the GPU test)."""
    cat = torch.cat([s if i is None else s.index_select(0, i.long()) for s, i in zip(srcs, idxs)], dim=1)
    h = F.silu(F.linear(cat, w1.to(cat.dtype), None if b1 is None else b1.to(cat.dtype)))
    z = F.linear(h, w2.to(cat.dtype), None if b2 is None else b2.to(cat.dtype))
    y = F.layer_norm(z, (z.shape[-1],), gamma.to(cat.dtype), beta.to(cat.dtype), eps)
    return y if residual is None else y + residual


def _f32(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if t is None else t.detach().float().contiguous()


def launch(srcs: List[torch.Tensor], idx32: List[Optional[torch.Tensor]], w1p: torch.Tensor, b1, w2b: torch.Tensor, b2, gamma,
           beta, eps: float, residual: Optional[torch.Tensor], M: int):
    """Raw launch: returns ``(rc, y, pre, z, mean, rstd)``; ``rc == 801`` means "shape not handled", nothing was launched."""
    from dgraph_b200.ops.gemm_native import _alloc_out

    dev = w1p.device
    n = len(srcs)
    ks = [int(s.shape[1]) for s in srcs]
    pre = _alloc_out(M, WIDTH, torch.bfloat16, dev)
    z = _alloc_out(M, WIDTH, torch.bfloat16, dev)
    y = _alloc_out(M, WIDTH, torch.bfloat16, dev)
    mean = torch.empty(M, dtype=torch.float32, device=dev)
    rstd = torch.empty(M, dtype=torch.float32, device=dev)
    bases = (ctypes.c_void_p * 3)(*[s.data_ptr() for s in srcs] + [0] * (3 - n))
    lds = (ctypes.c_int64 * 3)(*[s.stride(0) for s in srcs] + [0] * (3 - n))
    ips = (ctypes.c_void_p * 3)(*[(0 if i is None else i.data_ptr()) for i in idx32] + [0] * (3 - n))
    kk = (ctypes.c_int * 3)(*ks + [0] * (3 - n))
    rc = N.lib().dgb_mlp2_ln_fused_bf16(
        N.i32(n), bases, lds, ips, kk, N.ptr(w1p), N.i64(w1p.stride(0)), N.ptr(b1), N.ptr(w2b), N.i64(w2b.stride(0)), N.ptr(b2),
        N.ptr(gamma), N.ptr(beta), N.f32(eps), N.ptr(residual), N.i64(0 if residual is None else residual.stride(0)),
        N.ptr(pre), N.i64(pre.stride(0)), N.ptr(z), N.i64(z.stride(0)), N.ptr(y), N.i64(y.stride(0)), N.ptr(mean), N.ptr(rstd),
        N.i32(M), N.stream())
    return rc, y, pre, z, mean, rstd


class _FusedMLP(Function):
    @staticmethod
    def forward(ctx, w1, b1, w2, b2, gamma, beta, eps, residual, n_src, scatter_plans, *args):
        from dgraph_b200.ops.gemm_native import pack_segment_weight

        srcs, idxs = list(args[:n_src]), list(args[n_src:])
        ks = [int(s.shape[1]) for s in srcs]
        M = int(srcs[0].shape[0])
        for s, i in zip(srcs, idxs):
            M = int(i.shape[0]) if i is not None else int(s.shape[0])
        w1p = pack_segment_weight(w1, ks)
        w2b = w2.detach().to(torch.bfloat16).contiguous()
        gf = _f32(gamma)
        rc, y, pre, z, mean, rstd = launch(srcs, idxs, w1p, _f32(b1), w2b, _f32(b2), gf, _f32(beta), float(eps), residual, M)
        N.check(rc, "dgb_mlp2_ln_fused_bf16")
        N.count_launch()
        ctx.n_src, ctx.ks, ctx.scatter_plans, ctx.has_res = n_src, ks, scatter_plans, residual is not None
        ctx.has_b1, ctx.has_b2 = b1 is not None, b2 is not None
        ctx.idx_none = [i is None for i in idxs]
        ctx.save_for_backward(w1, w2b, gf, pre, z, mean, rstd, *srcs, *[i for i in idxs if i is not None])
        return y

    @staticmethod
    def backward(ctx, gy):
        from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w, gemm_bf16_tn
        from dgraph_b200.ops.norm import silu_backward
        from dgraph_b200.ops.reduce import colsum
        from dgraph_b200.ops.segment import _native_seg_sum

        saved = ctx.saved_tensors
        w1, w2b, gf, pre, z, mean, rstd = saved[:7]
        srcs = list(saved[7:7 + ctx.n_src])
        it = iter(saved[7 + ctx.n_src:])
        idxs = [None if none else next(it) for none in ctx.idx_none]
        M, C = z.shape
        lib = N.lib()
        gy = gy.contiguous()
        # LayerNorm gradient (csrc/norm_ops.cu): dz, dgamma, dbeta in two launches
        dz = torch.empty((M, C), dtype=torch.bfloat16, device=gy.device)
        lib.dgb_layernorm_bwd_scratch_bytes.restype = ctypes.c_int64
        nb = int(lib.dgb_layernorm_bwd_scratch_bytes(N.i64(M), N.i32(C)))
        scratch = torch.empty(max(nb // 4, 1), dtype=torch.float32, device=gy.device)
        dgb = torch.empty(2 * C, dtype=torch.float32, device=gy.device)
        N.check(lib.dgb_layernorm_bwd(N.ptr(gy), N.i64(gy.stride(0)), N.ptr(z), N.i64(z.stride(0)), N.ptr(gf), N.ptr(mean),
                                      N.ptr(rstd), N.ptr(dz), N.i64(C), N.ptr(scratch), N.ptr(dgb), N.i64(M), N.i32(C), N.i32(C),
                                      N.stream()), "dgb_layernorm_bwd")
        N.count_launch(2)
        need = ctx.needs_input_grad
        h = F.silu(pre)  # recomputed: the forward keeps h on chip
        dw2 = gemm_bf16_nn_grad_w(dz, h) if need[2] else None
        db2 = colsum(dz) if (need[3] and ctx.has_b2) else None
        dh = gemm_bf16_tn(dz, w2b.t().contiguous())
        g = silu_backward(dh.contiguous(), pre.contiguous())
        dcat = gemm_bf16_tn(g, w1.detach().to(torch.bfloat16).t().contiguous())  # [M, sum(ks)]
        dw1 = None
        if need[0]:
            parts = []
            for s_, i_ in zip(srcs, idxs):
                k = s_.shape[1]
                if k <= 256:
                    parts.append(gemm_bf16_nn_grad_w(g, s_, i_))
                else:
                    parts.append(torch.cat([gemm_bf16_nn_grad_w(g, s_[:, o:o + 256], i_) for o in range(0, k, 256)], dim=1))
            dw1 = torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]
        db1 = colsum(g) if (need[1] and ctx.has_b1) else None
        grads, o = [], 0
        for pos, (s, i, k) in enumerate(zip(srcs, idxs, ctx.ks)):
            part = dcat[:, o:o + k]
            o += k
            if not need[10 + pos]:
                grads.append(None)
            elif i is None:
                grads.append(part)
            else:
                sp = ctx.scatter_plans[pos] if ctx.scatter_plans is not None else None
                if sp is not None:  # atomic-free: CSR walk over the (source row <- edge id) incidence
                    pc = part if (part.stride(1) == 1 and part.stride(0) % 8 == 0 and part.data_ptr() % 16 == 0) else part.contiguous()
                    grads.append(_native_seg_sum(sp.rowptr, sp.col, sp.num_local, pc, None, False))
                else:
                    grads.append(torch.zeros_like(s).index_add_(0, i.long(), part))
        return (dw1, db1, dw2, db2, dgb[:C], dgb[C:], None, (gy if ctx.has_res else None), None, None, *grads,
                *([None] * len(idxs)))


def supported(srcs: Sequence[torch.Tensor], w1: torch.Tensor, w2: torch.Tensor, residual: Optional[torch.Tensor]) -> bool:
    """Shapes / dtypes the fused kernel handles: CUDA bf16 sources with 16-byte rows, hidden = out = 128, at most three sources."""
    if not (ENABLED and N.available() and 1 <= len(srcs) <= 3):
        return False
    if tuple(w2.shape) != (WIDTH, WIDTH) or w1.shape[0] != WIDTH or w1.shape[1] != sum(int(s.shape[1]) for s in srcs):
        return False
    if not all(s.is_cuda and s.dtype == torch.bfloat16 and s.shape[1] % 8 == 0 and s.stride(1) == 1 and s.stride(0) % 8 == 0
               and s.data_ptr() % 16 == 0 for s in srcs):
        return False
    if residual is not None and not (residual.dtype == torch.bfloat16 and residual.shape[1] == WIDTH and residual.stride(1) == 1
                                     and residual.stride(0) % 8 == 0 and residual.data_ptr() % 16 == 0):
        return False
    return hasattr(N.lib(), "dgb_mlp2_ln_fused_bf16")


def mlp2_ln_fused(srcs: List[torch.Tensor], idxs: List[Optional[torch.Tensor]], w1, b1, w2, b2, gamma, beta, eps: float = 1e-5,
                  residual: Optional[torch.Tensor] = None, scatter_plans=None) -> torch.Tensor:
    """Caller checks :func:`supported` first.  ``scatter_plans`` as in :func:`dgraph_b200.ops.gather_linear.gather_linear`."""
    idx32 = [None if i is None else (i if i.dtype == torch.int32 else i.to(torch.int32)).contiguous() for i in idxs]
    return _FusedMLP.apply(w1, b1, w2, b2, gamma, beta, float(eps), residual, len(srcs), scatter_plans, *srcs, *idx32)
