"""Time the GraphConv aggregation kernels (forward add-ReLU-segment-sum, masked backward) on the products-shaped graph."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.data import make_graph  # noqa: E402
from dgraph_b200.ops.segment import graphconv_aggregate, segment_sum  # noqa: E402
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan  # noqa: E402


def timeit(fn, iters=10, warm=3):
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
    scale = float(os.environ.get("SCALE", "1.0"))
    H = int(os.environ.get("H", "256"))
    g = make_graph("ogbn-products", device="cuda", scale=scale)
    V, E = g.num_nodes, g.edge_list.shape[0]
    kp = GraphKernelPlan.from_edges(g.edge_list, V, V, device="cuda")
    del g
    P = torch.randn(V, H, device="cuda").bfloat16().requires_grad_(True)
    Q = torch.randn(V, H, device="cuda").bfloat16().requires_grad_(True)
    go = torch.randn(V, H, device="cuda").bfloat16()
    out = graphconv_aggregate(P, Q, kp)
    t_f = timeit(lambda: graphconv_aggregate(P, Q, kp))
    t_fb = timeit(lambda: torch.autograd.grad(graphconv_aggregate(P, Q, kp), [P, Q], go))
    x = Q.detach()
    t_s = timeit(lambda: segment_sum(x, kp, mean=True))
    gather_gb = E * H * 2 / 1e9
    print(json.dumps({"V": V, "E": E, "H": H, "graphconv_fwd_ms": round(t_f, 3), "graphconv_fwd_bwd_ms": round(t_fb, 3),
                      "sage_mean_ms": round(t_s, 3), "gather_GB": round(gather_gb, 2),
                      "fwd_gather_GBps": round(gather_gb / t_f * 1e3, 1), "sage_gather_GBps": round(gather_gb / t_s * 1e3, 1)}))


if __name__ == "__main__":
    main()
