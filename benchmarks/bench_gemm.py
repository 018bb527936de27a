"""tcgen05 GEMM vs cuBLAS (torch.matmul) on the framework's shapes; prints TFLOP/s, GB/s and roofline fractions."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.ops.gemm_native import gemm_bf16_tn  # noqa: E402


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    peaks = {"hbm_gbs": 6454.0, "bf16_tflops": 1682.9}
    try:
        peaks.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
    except Exception:
        pass
    shapes = [(2449029, 256, 104), (2449029, 512, 104), (2449029, 256, 256), (2449029, 512, 256), (2449029, 47, 256),
              (655320, 128, 384), (655320, 512, 512), (3114720, 128, 384), (16384, 2048, 8192)]
    rows = []
    for M, N, K in shapes:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        t_tc = timeit(lambda: gemm_bf16_tn(a, b))
        t_cb = timeit(lambda: torch.matmul(a, b.t()))
        fl = 2.0 * M * N * K
        by = 2.0 * (M * K + N * K + M * N)
        t_roof = max(fl / (peaks["bf16_tflops"] * 1e12), by / (peaks["hbm_gbs"] * 1e9)) * 1e3
        rows.append({"M": M, "N": N, "K": K, "tcgen05_ms": round(t_tc, 4), "cublas_ms": round(t_cb, 4),
                     "tcgen05_tflops": round(fl / t_tc / 1e9, 1), "tcgen05_gbs": round(by / t_tc / 1e6, 1),
                     "roofline_ms": round(t_roof, 4), "frac_of_measured_roofline": round(t_roof / t_tc, 3),
                     "speedup_vs_cublas": round(t_cb / t_tc, 3)})
        print(json.dumps(rows[-1]), flush=True)
        del a, b
    return rows


if __name__ == "__main__":
    main()
