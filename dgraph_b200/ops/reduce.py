"""Column reductions (bias gradients, BatchNorm statistics) -- ``csrc/reduce_ops.cu``."""
from __future__ import annotations

import torch

from dgraph_b200.ops import _native as N


def colsum(x: torch.Tensor, with_sq: bool = False) -> torch.Tensor:
    """``x.float().sum(0)`` (and ``(x.float() ** 2).sum(0)`` appended when ``with_sq``) for a bf16 ``[M, N]`` matrix."""
    Nc = x.shape[1]
    if not (x.is_cuda and x.dtype == torch.bfloat16 and Nc % 8 == 0 and 8 <= Nc <= 2048 and x.stride(1) == 1
            and x.stride(0) % 8 == 0):
        if x.is_cuda:
            N.require_for(x)
        xf = x if x.dtype == torch.float64 else x.float()
        return torch.cat([xf.sum(0), (xf * xf).sum(0)]) if with_sq else xf.sum(0)
    lib = N.lib()
    lib.dgb_colsum_scratch_bytes.restype = __import__("ctypes").c_int64
    nbytes = int(lib.dgb_colsum_scratch_bytes(N.i64(x.shape[0]), N.i32(Nc), N.i32(1 if with_sq else 0)))
    scratch = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=x.device)
    out = torch.empty(Nc * (2 if with_sq else 1), dtype=torch.float32, device=x.device)
    N.check(lib.dgb_colsum_bf16(N.ptr(x), N.i64(x.stride(0)), N.i64(x.shape[0]), N.i32(Nc), N.i32(1 if with_sq else 0),
                                N.ptr(scratch), N.ptr(out), N.stream()), "dgb_colsum_bf16")
    N.count_launch(2)
    return out
