"""Tiny driver for ncu: a few launches of the tcgen05 GEMM on the GCN layer-1 shape."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.ops.gemm_native import gemm_bf16_tn  # noqa: E402

M, N, K = int(os.environ.get("M", 2449029)), int(os.environ.get("N", 256)), int(os.environ.get("K", 104))
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(N, K, device="cuda").bfloat16()
bias = torch.randn(N, device="cuda")
for _ in range(4):
    y = gemm_bf16_tn(a, b, bias, "relu")
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
