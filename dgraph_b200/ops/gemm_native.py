"""Launch wrappers for ``csrc/gemm_tcgen05.cu`` (tcgen05 / TMEM / TMA GEMM and gather-GEMM)."""
from __future__ import annotations

import os

import ctypes
from typing import List, Optional, Sequence

import torch

from dgraph_b200.ops import _native as N

_ACT = {None: 0, "none": 0, "relu": 1, "silu": 2}


def _as_bf16_rowmajor(t: torch.Tensor) -> torch.Tensor:
    """bf16, unit inner stride, 16-byte aligned base and row stride (TMA requirements)."""
    if t.dtype != torch.bfloat16:
        t = t.to(torch.bfloat16)
    if t.stride(1) != 1 or t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
        K = t.shape[1]
        Kp = (K + 7) // 8 * 8
        buf = torch.zeros((t.shape[0], Kp), dtype=torch.bfloat16, device=t.device)
        buf[:, :K] = t
        t = buf[:, :K] if Kp == K else buf
    return t


# B-resident tall-skinny variant (csrc/gemm_bres_tcgen05.cu): its N tile of B stays in shared memory and 5-8 A stages stream past
# it.  Measured on B200 (profiles/bench_gemm.jsonl): 2.45 M x 256 x 104 0.473 -> 0.322 ms, x 512 x 104 0.887 -> 0.599 ms; no gain for
# K >= 256, so it is taken for K <= 128 only.  DGRAPH_B200_GEMM_BRES=0 switches it off, =1 forces it for every shape it supports.
_BRES = os.environ.get("DGRAPH_B200_GEMM_BRES", "auto")
_BRES_MAX_K = 128


def _alloc_out(M: int, Nn: int, dtype, device) -> torch.Tensor:
    """Output with a 16-byte aligned row stride (lets the epilogue use TMA stores); returned as an ``[M, N]`` view."""
    per16 = 16 // torch.empty((), dtype=dtype).element_size()
    ld = (Nn + per16 - 1) // per16 * per16
    buf = torch.empty((M, ld), dtype=dtype, device=device)
    return buf if ld == Nn else buf[:, :Nn]


def gemm_bf16_tn(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act=None,
                 out_dtype=torch.bfloat16, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``act(a @ b.T + bias)``; ``a [M, K]``, ``b [N, K]`` bf16.  Zero-padded extra K columns in either operand are fine."""
    a, b = _as_bf16_rowmajor(a), _as_bf16_rowmajor(b)
    M, K = a.shape[0], min(a.shape[1], b.shape[1]) if a.shape[1] != b.shape[1] else a.shape[1]
    Nn = b.shape[0]
    if out is None:
        out = _alloc_out(M, Nn, out_dtype, a.device)
    assert out.stride(1) == 1 and out.dtype in (torch.bfloat16, torch.float32)
    bias_f = None if bias is None else bias.float().contiguous()
    if out.dtype == torch.bfloat16 and (_BRES == "1" or (_BRES == "auto" and K <= _BRES_MAX_K)):
        # 801 = shape not supported by this variant -> default kernel
        rc = N.lib().dgb_gemm_bf16_tn_bres(N.ptr(a), N.i64(a.stride(0)), N.ptr(b), N.i64(b.stride(0)), N.ptr(out),
                                           N.i64(out.stride(0)), N.i32(M), N.i32(Nn), N.i32(K), N.ptr(bias_f),
                                           N.i32(_ACT[act]), N.stream())
        if rc != 801:
            N.check(rc, "dgb_gemm_bf16_tn_bres")
            N.count_launch()
            return out
    N.check(N.lib().dgb_gemm_bf16_tn(N.ptr(a), N.i64(a.stride(0)), N.ptr(b), N.i64(b.stride(0)), N.ptr(out),
                                     N.i64(out.stride(0)), N.i32(M), N.i32(Nn), N.i32(K), N.ptr(bias_f),
                                     N.i32(_ACT[act]), N.i32(1 if out.dtype == torch.float32 else 0), N.stream()),
            "dgb_gemm_bf16_tn")
    N.count_launch()
    return out


def pack_segment_weight(w: torch.Tensor, seg_k: Sequence[int]) -> torch.Tensor:
    """``w [N, sum(seg_k)]`` -> bf16 ``[N, sum(pad64(k))]`` with every segment's columns padded to a multiple of 64."""
    parts, o = [], 0
    for k in seg_k:
        kp = (k + 63) // 64 * 64
        seg = w[:, o:o + k].to(torch.bfloat16)
        parts.append(torch.nn.functional.pad(seg, (0, kp - k)))
        o += k
    return torch.cat(parts, dim=1).contiguous()


def gather_gemm_bf16(srcs: List[torch.Tensor], idxs: List[Optional[torch.Tensor]], w_packed: torch.Tensor,
                     bias: Optional[torch.Tensor], M: int, seg_k: Sequence[int], act=None,
                     out_dtype=torch.bfloat16, peers: Optional[List[Optional[tuple]]] = None) -> torch.Tensor:
    """``act(cat([srcs[s][idxs[s]] for s], 1) @ W.T + bias)`` with the gather and the concat done inside the GEMM's
    producer warps (F1b).  ``idxs[s] is None`` means row ``m`` of ``srcs[s]`` (per-edge features)."""
    n = len(srcs)
    assert 1 <= n <= 3
    srcs = [_as_bf16_rowmajor(s) for s in srcs]
    idx32 = [None if i is None else (i if i.dtype == torch.int32 else i.to(torch.int32)).contiguous() for i in idxs]
    Nn = w_packed.shape[0]
    out = _alloc_out(M, Nn, out_dtype, w_packed.device)
    bases = (ctypes.c_void_p * 3)(*[s.data_ptr() for s in srcs] + [0] * (3 - n))
    lds = (ctypes.c_int64 * 3)(*[s.stride(0) for s in srcs] + [0] * (3 - n))
    ips = (ctypes.c_void_p * 3)(*[(0 if i is None else i.data_ptr()) for i in idx32] + [0] * (3 - n))
    ks = (ctypes.c_int * 3)(*[int(k) for k in seg_k] + [0] * (3 - n))
    bias_f = None if bias is None else bias.float().contiguous()
    if peers is not None and any(p is not None for p in peers):
        # peers[s] = (ptr_table int64 [W], halo_owner int32 [n_halo], halo_row int32 [n_halo], n_local): gathered rows
        # >= n_local of segment s are pulled from the owner GPU inside the producer warps (NVLink loads)
        pt = (ctypes.c_void_p * 3)(*[(0 if p is None else p[0].data_ptr()) for p in peers] + [0] * (3 - n))
        ho = (ctypes.c_void_p * 3)(*[(0 if p is None else p[1].data_ptr()) for p in peers] + [0] * (3 - n))
        hr = (ctypes.c_void_p * 3)(*[(0 if p is None else p[2].data_ptr()) for p in peers] + [0] * (3 - n))
        nl = (ctypes.c_int64 * 3)(*[(0 if p is None else int(p[3])) for p in peers] + [0] * (3 - n))
        N.check(N.lib().dgb_gather_gemm_peer_bf16(N.i32(n), bases, lds, ips, ks, pt, ho, hr, nl, N.ptr(w_packed),
                                                  N.i64(w_packed.stride(0)), N.ptr(out), N.i64(out.stride(0)), N.i32(M),
                                                  N.i32(Nn), N.ptr(bias_f), N.i32(_ACT[act]),
                                                  N.i32(1 if out_dtype == torch.float32 else 0), N.stream()),
                "dgb_gather_gemm_peer_bf16")
        N.count_launch()
        return out
    N.check(N.lib().dgb_gather_gemm_bf16(N.i32(n), bases, lds, ips, ks, N.ptr(w_packed), N.i64(w_packed.stride(0)),
                                         N.ptr(out), N.i64(out.stride(0)), N.i32(M), N.i32(Nn), N.ptr(bias_f),
                                         N.i32(_ACT[act]), N.i32(1 if out_dtype == torch.float32 else 0), N.stream()),
            "dgb_gather_gemm_bf16")
    N.count_launch()
    return out


def gemm_bf16_nn_grad_w(g: torch.Tensor, x: torch.Tensor, x_idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Weight gradient ``g^T @ x`` (``x`` optionally row-gathered by ``x_idx`` inside the kernel) -> fp32 ``[N1, N2]``
    (reduction over the huge row dimension) with the split-K tcgen05
    kernel (``csrc/gemm_wgrad_tcgen05.cu``: MN-major operands via TMA, TMEM accumulators, vector reductions into ``dW``).
    Falls back to a library GEMM only for ``N2 > 256``."""
    N1, N2 = g.shape[1], x.shape[1]
    rows_ok = g.shape[0] == (x.shape[0] if x_idx is None else x_idx.shape[0])
    ok = (g.is_cuda and g.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and N2 <= 256 and rows_ok
          and hasattr(N.lib(), "dgb_gemm_wgrad_bf16"))
    if not ok:
        return torch.matmul(g.t(), x if x_idx is None else x.index_select(0, x_idx.long())).float()
    g, x = _as_bf16_rowmajor(g), _as_bf16_rowmajor(x)
    idx32 = None if x_idx is None else (x_idx if x_idx.dtype == torch.int32 else x_idx.to(torch.int32)).contiguous()
    out = torch.zeros((N1, (N2 + 3) // 4 * 4), dtype=torch.float32, device=g.device)
    N.check(N.lib().dgb_gemm_wgrad_bf16(N.ptr(g), N.i64(g.stride(0)), N.ptr(x), N.i64(x.stride(0)), N.ptr(out),
                                        N.i64(out.stride(0)), N.i64(g.shape[0]), N.i32(N1), N.i32(N2), N.ptr(idx32),
                                        N.stream()),
            "dgb_gemm_wgrad_bf16")
    N.count_launch()
    return out if out.shape[1] == N2 else out[:, :N2]


def act_backward(g: torch.Tensor, y: torch.Tensor, act: str) -> torch.Tensor:
    if act == "relu":
        return torch.ops.aten.threshold_backward(g, y, 0)  # g where y > 0 else 0, one kernel (was compare + cast + multiply)
    raise NotImplementedError(f"fused backward for activation {act}")
