"""One pass over every kernel family that sits on a default code path, at sizes large enough to be meaningful and small enough for
`ncu --set full` (about 40 replays per launch).  Not a benchmark: it exists so that ONE `ncu` capture yields a row per kernel for
`profiles/ncu_kernel_families.md` (`benchmarks/ncu_summary.py`).

    python benchmarks/kernel_zoo.py                                  # plain run first (must exit 0)
    ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:dgb:: -o gpurun_out/zoo \
        python benchmarks/kernel_zoo.py

Families: tcgen05 GEMM / B-resident GEMM / gather-GEMM / split-K weight gradient (plain and gathered), CSR / CSC segment kernels
(local and pull variants), halo pull, halo-gradient pull-reduce, peer row gather, fused gradient-sum + Adam, segment softmax
forward / backward, LayerNorm forward / backward, fused GraphCast MLP block, masked cross-entropy, BatchNorm apply / backward,
column sums, weight packing.  Peer kernels run with W ranks emulated as W buffers of this process (pointers are pointers).
"""
import os
import sys

import torch

STEPS = int(os.environ.get("ZOO_STEPS", "1"))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "experiments", "OGB-LSC"))


def gcn_one_rank(scale):
    from dgraph_b200 import Communicator
    from dgraph_b200.data import contiguous_partition, make_graph, node_features_and_labels
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.ops.loss import masked_cross_entropy
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from dgraph_b200.parallel.local_engine import LocalEngine

    dev = torch.device("cuda")
    g = make_graph("ogbn-products", device=dev, scale=scale)
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    pat = build_communication_pattern(g.edge_list, contiguous_partition(g.num_nodes, 1, device=dev), 0, 1)
    x, y = node_features_and_labels(g.num_nodes, g.num_features, g.num_classes, device=dev)
    xd = torch.zeros(g.num_nodes, 104, dtype=torch.bfloat16, device=dev)
    xd[:, :100] = x.to(torch.bfloat16)
    idx = torch.arange(0, g.num_nodes, 10, device=dev)
    model = CommAwareGCN(100, 256, g.num_classes, HaloExchange(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-3)
    for _ in range(STEPS):
        opt.zero_grad()
        masked_cross_entropy(model(xd, pat), y, idx).backward()
        opt.step()


def gcn_pull_emulated(W=2, V=200_000):
    from dgraph_b200.data.synthetic import contiguous_partition, symmetric_locality_graph
    from dgraph_b200.ops.graphconv import graphconv_pull_q
    from dgraph_b200.parallel import build_all_patterns
    from dgraph_b200.parallel.peer_engine import emulated_engines

    e = symmetric_locality_graph(V, V * 24, cut_fraction=0.1, device="cuda").cpu()
    pats = build_all_patterns(e, contiguous_partition(V, W), W)
    engines = emulated_engines(pats, arena_bytes=1 << 30)
    F, H = 256, 256
    w = (torch.randn(H, 2 * F, device="cuda") * 0.05).requires_grad_(True)
    b = torch.zeros(H, device="cuda", requires_grad=True)
    for mode in ("kernel", "buffer"):
        os.environ["DGRAPH_B200_PULL"] = mode
        xs = [torch.randn(p.num_local_vertices, F, device="cuda").bfloat16().requires_grad_(True) for p in pats]
        for _ in range(2):
            outs = [graphconv_pull_q(xs[r], w, b, pats[r].kernel_plan("cuda"), F, engines[r], pats[r], tag=mode) for r in range(W)]
        for r in range(W):
            outs[r].backward(torch.ones_like(outs[r]), retain_graph=False)
    os.environ.pop("DGRAPH_B200_PULL", None)
    # fused gradient sum + Adam over the "ranks"
    from dgraph_b200.ops import _native as N

    n = 1 << 20
    g = [torch.randn(n, device="cuda") for _ in range(W)]
    table = torch.tensor([t.data_ptr() for t in g], dtype=torch.int64, device="cuda")
    p, m, v = torch.randn(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    N.check(N.lib().dgb_adam_allreduce(N.ptr(p), N.ptr(table), N.i32(W), N.ptr(m), N.ptr(v), N.ptr(None), N.i64(n), N.f32(1e-3),
                                       N.f32(0.9), N.f32(0.999), N.f32(1e-8), N.f32(0.0), N.f32(1.0 / W), N.i32(1), N.stream()), "adam")


def rgat_one_rank():
    from config import ModelConfig
    from dgraph_b200 import Communicator
    from dgraph_b200.data.hetero import RELATIONS, SyntheticHeterogeneousDataset
    from dgraph_b200.models.rgat import CommAwareRGAT
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel.local_engine import LocalEngine

    dev = torch.device("cuda")
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    ds = SyntheticHeterogeneousDataset(0, 1, 32768 * 4, 16384 * 4, 256 * 4, 16, 153, comm=comm)
    ds.prepare_plans(fused=True)
    ds.to(dev)
    mc = ModelConfig()
    model = CommAwareRGAT(ds.num_features, ds.num_classes, 64, ds.num_relations, mc.num_layers, mc.heads, comm).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-4)
    xs = [f.to(torch.bfloat16) for f in ds.features]
    idx, tgt = ds.get_mask("train").to(dev), ds.get_target("train").to(dev)
    for _ in range(STEPS):
        opt.zero_grad()
        out = model(xs, RELATIONS, ds.patterns)
        torch.nn.functional.cross_entropy(out.index_select(0, idx).float(), tgt).backward()
        opt.step()


def graphcast_one_rank(fused):
    from dgraph_b200.communicator import Communicator
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset
    from dgraph_b200.models.graphcast import Config, DGraphCast
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel.local_engine import LocalEngine

    from dgraph_b200.ops import mlp_fused

    mlp_fused.ENABLED = bool(fused)
    cfg = Config()
    cfg.data.latlon_res, cfg.model.mesh_level = (181, 360), 5
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    sg = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level).shard(0, 1).to("cuda", torch.bfloat16)
    d = SyntheticWeatherDataset(73, cfg.data.latlon_res)[0]
    x, y = d["invar"].cuda().bfloat16(), d["outvar"].cuda().bfloat16()
    model = DGraphCast(cfg, comm).cuda()
    opt = FusedAdam(model.parameters(), lr=1e-3)
    for _ in range(STEPS):
        opt.zero_grad()
        torch.nn.functional.mse_loss(model(x, sg).float(), y.float()).backward()
        opt.step()


def lean():
    """One launch (or a few) of every family through the op-level APIs: ~60 launches, for a short `ncu -c 70` capture."""
    from dgraph_b200.nn.dist_bn import DistributedBatchNorm1D
    from dgraph_b200.ops import _native as N
    from dgraph_b200.ops import mlp_fused
    from dgraph_b200.ops.gather_linear import gather_linear
    from dgraph_b200.ops.loss import masked_cross_entropy
    from dgraph_b200.ops.norm import layer_norm_residual
    from dgraph_b200.ops.softmax import seg_softmax_aggregate
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

    gcn_pull_emulated(W=2, V=150_000)
    dev = "cuda"
    n, E, C = 300_000, 4_000_000, 128
    e = torch.stack([torch.randint(0, n, (E,), device=dev), torch.randint(0, n, (E,), device=dev)], 1)
    kp = GraphKernelPlan.from_edges(e, n, n, device=dev)
    # segment softmax + weighted aggregation (RGAT attention), forward and backward
    h = torch.randn(n, 64, device=dev).bfloat16().requires_grad_(True)
    s_i = torch.randn(n, device=dev, requires_grad=True)
    s_j = torch.randn(n, device=dev, requires_grad=True)
    seg_softmax_aggregate(s_i, s_j, h, kp, 0.2).sum().backward()
    # distributed BatchNorm (statistics, apply + ReLU, backward)
    bn = DistributedBatchNorm1D(64, activation="relu").cuda()
    xb = torch.randn(n, 64, device=dev).bfloat16().requires_grad_(True)
    bn(xb)[0].float().sum().backward()
    # GraphCast edge block: gather-GEMM (+SiLU) -> GEMM -> LayerNorm + residual, unfused chain and the one-kernel block
    src, dst = e[:, 0].int().contiguous(), e[:, 1].int().contiguous()
    xn = torch.randn(n, C, device=dev).bfloat16().requires_grad_(True)
    ef = torch.randn(E, C, device=dev).bfloat16().requires_grad_(True)
    w1 = (torch.randn(C, 3 * C, device=dev) * 0.05).requires_grad_(True)
    b1 = torch.zeros(C, device=dev, requires_grad=True)
    w2 = (torch.randn(C, C, device=dev) * 0.05).requires_grad_(True)
    b2 = torch.zeros(C, device=dev, requires_grad=True)
    gam, bet = torch.ones(C, device=dev, requires_grad=True), torch.zeros(C, device=dev, requires_grad=True)
    hcat = gather_linear([xn, xn, ef], [src, dst, None], w1, b1, act="silu")
    from dgraph_b200.ops.gemm import linear

    layer_norm_residual(linear(hcat, w2.bfloat16(), b2.bfloat16()), gam, bet, 1e-5, ef).float().sum().backward()
    if mlp_fused.supported([xn, xn, ef], w1, w2, ef):
        mlp_fused.mlp2_ln_fused([xn, xn, ef], [src, dst, None], w1, b1, w2, b2, gam, bet, 1e-5, ef).float().sum().backward()
    # masked cross-entropy, fused Adam
    logits = torch.randn(n, 48, device=dev).bfloat16().requires_grad_(True)
    masked_cross_entropy(logits, torch.randint(0, 47, (n,), device=dev), torch.arange(0, n, 10, device=dev)).backward()
    p, g, m, v = (torch.randn(1 << 20, device=dev) for _ in range(4))
    N.check(N.lib().dgb_adam_flat(N.ptr(p), N.ptr(g), N.ptr(m), N.ptr(v.abs_()), N.ptr(None), N.i64(p.numel()), N.f32(1e-3), N.f32(0.9),
                                  N.f32(0.999), N.f32(1e-8), N.f32(0.0), N.f32(1.0), N.i32(1), N.stream()), "adam")


def main():
    torch.manual_seed(0)
    if os.environ.get("ZOO_LEAN", "0") == "1":
        lean()
        torch.cuda.synchronize()
        print("kernel zoo ok (lean)")
        return
    gcn_one_rank(float(os.environ.get("ZOO_SCALE", "0.25")))
    gcn_pull_emulated()
    rgat_one_rank()
    graphcast_one_rank(False)
    graphcast_one_rank(True)
    torch.cuda.synchronize()
    print("kernel zoo ok")


if __name__ == "__main__":
    main()
