"""`bench.py --config <name>` workloads of this framework (BASELINE.json configs 3-5 and the RGAT experiment); the products GCN
(config 2, the headline) lives in `bench.py` itself.  Every builder returns a `runner.Workload`; the timing protocol is
`runner.run_workload`, shared with the reference arm (`baseline/ref_workloads.py`).

  papers100m  GraphSAGE, ogbn-papers100M-shaped (111 M nodes, ~1.6 B edges, 128-d, 172 classes) over 8 GPUs; with N < 8 GPUs the
              graph is scaled by N / 8 (fixed work per GPU: "weak").  Each rank generates only its own edge shard; shards are made
              symmetric and duplicate-free with one all-to-all of the reverse edges.
  graphcast   GraphCast 0.25 degree encoder-processor-decoder (1,038,240 grid nodes, level-6 multimesh), fwd + bwd + fused Adam.
  rgat        CommAwareRGAT on the synthetic MAG240M-shaped graph of the reference's SyntheticDatasetConfig, optionally x8.
  halo        halo gather / gather+scatter bandwidth sweep, 1 K - 10 M halo rows x 64 - 768-d.
"""
from __future__ import annotations

import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "experiments"), os.path.join(ROOT, "experiments", "Benchmarks"), os.path.join(ROOT, "experiments", "OGB-LSC")):
    if p not in sys.path:
        sys.path.insert(0, p)

from benchmarks.runner import Workload, median_ms, run_workload  # noqa: E402


def _launches():
    from dgraph_b200.ops import _native

    return _native.launches()


def _pin(t):
    return t.cpu().contiguous().pin_memory()


def _graphed(args, world, comm, forward_loss, opt, example_inputs):
    """`--cuda-graph`: zero_grad + forward + loss + backward + optimizer step captured once and replayed (`utils/cuda_graph.py`); with
    several ranks the peer engine's barrier epochs live in device memory (`DGRAPH_B200_DEVICE_EPOCH=1`, set in `run` before the
    engine exists).  Returns the replacement for the eager step, or None."""
    if not getattr(args, "cuda_graph", False) or not (world == 1 or getattr(comm.engine, "device_epoch", False)):
        return None
    from dgraph_b200.utils.cuda_graph import GraphedTrainStep, capture_or_eager

    snap = opt.snapshot()  # undo the three training steps the capture performs: losses then start from the initial weights
    g = capture_or_eager(lambda: GraphedTrainStep(forward_loss, opt, example_inputs, warmup=2), world)
    opt.restore(snap)
    return None if g is None else (lambda *xs: g(*xs).clone())


# ------------------------------------------------------------------------------------------------------------------------------
def symmetric_shard(V, E, rank, world, dev, cut, window=8192, seed=0):
    """In-neighbour lists of the vertices this rank owns, `[E_loc, 2]` (owned vertex, neighbour) in global ids, of an UNDIRECTED
    graph without duplicate edges: every rank draws E / (2 W) edges whose first endpoint it owns, ships the reverse of those whose
    other endpoint lives elsewhere to that endpoint's owner (one all-to-all), and removes duplicates.  No rank ever holds the
    global edge list (the reference needs it on every rank: 26 GB at this size)."""
    lo, hi = V * rank // world, V * (rank + 1) // world
    n = hi - lo
    e_half = E * n // V // 2
    g = torch.Generator(device=dev).manual_seed(seed * 1000 + rank)
    u = torch.randint(lo, hi, (e_half,), generator=g, device=dev)
    off = torch.randint(1, window + 1, (e_half,), generator=g, device=dev) * (torch.randint(0, 2, (e_half,), generator=g, device=dev) * 2 - 1)
    far = torch.rand(e_half, generator=g, device=dev) < cut
    v = torch.where(far, torch.randint(0, V, (e_half,), generator=g, device=dev), (u + off) % V)
    keep = u != v
    u, v = u[keep], v[keep]
    del off, far, keep
    bounds = torch.tensor([V * r // world for r in range(1, world)], device=dev)
    owner_v = torch.bucketize(v, bounds, right=True)
    mine = owner_v == rank
    parts = [torch.stack([u, v], 1), torch.stack([v[mine], u[mine]], 1)]
    if world > 1:
        away = ~mine
        ov, ru, rv = owner_v[away], v[away], u[away]  # reverse edge (v, u) belongs to owner(v)
        order = torch.argsort(ov)
        send = torch.stack([ru[order], rv[order]], 1).contiguous()
        counts = torch.bincount(ov, minlength=world)
        recv_counts = torch.empty_like(counts)
        dist.all_to_all_single(recv_counts, counts)
        recv = torch.empty((int(recv_counts.sum()), 2), dtype=torch.long, device=dev)
        dist.all_to_all_single(recv, send, output_split_sizes=recv_counts.tolist(), input_split_sizes=counts.tolist())
        parts.append(recv)
        del send, order, ov, ru, rv
    e = torch.cat(parts)
    del parts, u, v
    key = torch.unique((e[:, 0] - lo) * V + e[:, 1])
    return torch.stack([key // V + lo, key % V], 1)


def build_papers100m(args, comm, rank, world, dev):
    from dgraph_b200.data.synthetic import SHAPES
    from dgraph_b200.models.sage import CommAwareSAGE
    from dgraph_b200.ops.loss import masked_cross_entropy
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern_from_shard

    V0, E0, F, C = SHAPES["ogbn-papers100M"]
    scale = args.scale * world / 8.0
    V, E = int(V0 * scale), int(E0 * scale)
    edges = symmetric_shard(V, E, rank, world, dev, cut=args.cut)
    offsets = torch.tensor([V * r // world for r in range(world + 1)])
    pat = build_communication_pattern_from_shard(edges, offsets, rank, world)
    e_loc = torch.tensor([edges.shape[0]], dtype=torch.float64, device=dev)
    del edges
    if world > 1:
        dist.all_reduce(e_loc)
    E_global = float(e_loc)
    pat.kernel_plan(dev)
    torch.cuda.empty_cache()
    n = pat.num_local_vertices
    g = torch.Generator().manual_seed(rank)
    xh = _pin(torch.randn(n, F, generator=g).to(torch.bfloat16))
    yh = _pin(torch.randint(0, C, (n,), generator=g))
    ex = HaloExchange(comm)
    torch.manual_seed(0)
    model = CommAwareSAGE(F, args.hidden, C, ex, comm, compute_dtype=torch.bfloat16).to(dev)
    opt = FusedAdam(model.parameters(), lr=args.lr, comm=comm)
    idx = torch.arange(0, n, 100, device=dev)  # ~1 % labelled vertices (papers100M: 1.1 %)

    def step(x, y):
        opt.zero_grad()
        loss = masked_cross_entropy(model(x, pat), y, idx)
        loss.backward()
        opt.step()
        return loss

    use_graph = False  # a 25 ms step is not launch-bound

    def alloc(tag):
        return ex.symmetric_rows(pat, F, torch.bfloat16, device=dev, tag=tag), torch.empty(n, dtype=torch.long, device=dev)

    return Workload(
        metric="full-graph GraphSAGE training throughput (edges/s; epoch = zero_grad+fwd+loss+bwd+step)", unit="edges/s",
        work_per_step=E_global, scaling="weak", host_inputs=(xh, yh), alloc_inputs=alloc, step=step, launches=_launches,
        loss_weight=float(idx.numel()),
        config={"model": f"CommAwareSAGE 2xSAGEConv({args.hidden}) mean aggregator + Linear({C})",
                "graph": f"ogbn-papers100M-shaped V={V} E={int(E_global)} F={F} (x{scale:.3f} of 111,059,956 nodes: fixed work per GPU)",
                "global_batch": "full graph", "seq_len": None, "parallelism": f"graph-partition x{world} + DP weights",
                "partition": f"contiguous blocks, cut_fraction={args.cut}, undirected, duplicate-free, built from per-rank edge shards",
                "hidden": args.hidden, "l2": "inputs larger than L2", "engine": type(comm.engine).__name__,
                "labelled": "1 % of the vertices", "cuda_graph": use_graph})


# ------------------------------------------------------------------------------------------------------------------------------
def build_graphcast(args, comm, rank, world, dev):
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset
    from dgraph_b200.models.graphcast import Config, DGraphCast
    from dgraph_b200.optim import FusedAdam

    cfg = Config()
    if args.scale < 1.0:  # debug only
        cfg.data.latlon_res = (max(int(721 * args.scale), 19), max(int(1440 * args.scale), 36))
        cfg.model.mesh_level = max(2, int(round(6 + torch.log2(torch.tensor(args.scale)).item())))
    gen = DistributedGraphCastGraphGenerator(cfg.data.latlon_res, cfg.model.mesh_level, multi=cfg.model.multimesh)
    sg = gen.shard(rank, world).to(dev, torch.bfloat16)
    item = SyntheticWeatherDataset(cfg.model.input_grid_dim, cfg.data.latlon_res, num_samples=2, grid_rows=sg.grid_rows_global.cpu())[0]
    xh, yh = _pin(item["invar"].to(torch.bfloat16)), _pin(item["outvar"].to(torch.bfloat16))
    torch.manual_seed(0)
    model = DGraphCast(cfg, comm).to(dev)
    want_graph = bool(getattr(args, "cuda_graph", False)) and (world == 1 or getattr(comm.engine, "device_epoch", False))
    opt = FusedAdam(model.parameters(), lr=cfg.training.lr, comm=comm, capturable=want_graph)

    def forward_loss(x, y):
        return torch.nn.functional.mse_loss(model(x, sg).float(), y.float())

    def step(x, y):
        opt.zero_grad()
        loss = forward_loss(x, y)
        loss.backward()
        opt.step()
        return loss

    graphed = _graphed(args, world, comm, forward_loss, opt, (xh.to(dev), yh.to(dev)))
    step = graphed or step

    def alloc(tag):
        return torch.empty_like(xh, device=dev), torch.empty_like(yh, device=dev)

    n_params = sum(p.numel() for p in model.parameters())
    return Workload(
        metric="GraphCast 0.25 degree training step time (zero_grad+fwd+MSE+bwd+Adam)", unit="ms", work_per_step=0.0,
        host_inputs=(xh, yh), alloc_inputs=alloc, step=step, launches=_launches, loss_weight=float(xh.shape[0]),
        config={"model": f"DGraphCast encoder-processor({cfg.model.processor_layers})-decoder hidden {cfg.model.hidden_dim}, "
                         f"{n_params} parameters, random init",
                "graph": f"grid {cfg.data.latlon_res[0]}x{cfg.data.latlon_res[1]} = {cfg.data.latlon_res[0] * cfg.data.latlon_res[1]} nodes, "
                         f"icosahedral multimesh level {cfg.model.mesh_level}",
                "global_batch": 1, "seq_len": None, "parallelism": f"graph-partition x{world} + DP weights",
                "l2": "inputs larger than L2", "engine": type(comm.engine).__name__, "cuda_graph": graphed is not None,
                "fused_mlp_block": os.environ.get("DGRAPH_B200_FUSED_MLP", "0") == "1"})


# ------------------------------------------------------------------------------------------------------------------------------
def build_rgat(args, comm, rank, world, dev):
    from config import ModelConfig, SyntheticDatasetConfig, TrainingConfig
    from dgraph_b200.data.hetero import RELATIONS, SyntheticHeterogeneousDataset
    from dgraph_b200.models.rgat import CommAwareRGAT
    from dgraph_b200.nn.dist_bn import GetGlobalVal
    from dgraph_b200.optim import FusedAdam

    sc, mc, tc = SyntheticDatasetConfig(), ModelConfig(), TrainingConfig()
    m = max(int(round(args.rgat_mult)), 1)
    ds = SyntheticHeterogeneousDataset(rank, world, sc.num_papers * m, sc.num_authors * m, sc.num_institutions * m, sc.num_features,
                                       sc.num_classes, comm=comm)
    ds.prepare_plans(fused=True, exchange=getattr(comm.engine, "exchange_ids", None))
    ds.to(dev)
    hidden = args.rgat_hidden
    torch.manual_seed(0)
    model = CommAwareRGAT(ds.num_features, ds.num_classes, hidden, ds.num_relations, mc.num_layers, mc.heads, comm, dropout=mc.dropout).to(dev)
    model.first_relation_only = args.rgat_relations == "first"
    want_graph = bool(getattr(args, "cuda_graph", False)) and (world == 1 or getattr(comm.engine, "device_epoch", False))
    opt = FusedAdam(model.parameters(), lr=tc.lr, weight_decay=tc.weight_decay, comm=comm, grad_average=False, capturable=want_graph)
    train_idx = ds.get_mask("train").to(dev)
    target = ds.get_target("train").to(dev)
    n_global = GetGlobalVal(train_idx.numel(), comm, dev)
    feats_h = tuple(_pin(f.to(torch.bfloat16)) for f in ds.features)
    E_global = float(sum(int(e.shape[0]) for e in ds.edges)) if hasattr(ds, "edges") else 0.0

    def forward_loss(*xs):
        out = model(list(xs), RELATIONS, ds.patterns)
        out = out[0] if out.dim() == 3 else out
        return torch.nn.functional.cross_entropy(out.index_select(0, train_idx).float(), target, reduction="sum") / n_global

    def step(*xs):
        opt.zero_grad()
        loss = forward_loss(*xs)
        loss.backward()
        opt.step()
        return loss

    graphed = _graphed(args, world, comm, forward_loss, opt, tuple(f.to(dev) for f in feats_h))
    step = graphed or step

    def alloc(tag):
        return tuple(torch.empty_like(f, device=dev) for f in feats_h)

    return Workload(
        metric="RGAT full-graph epoch time (zero_grad+fwd+loss+bwd+step), synthetic MAG240M-shaped graph", unit="ms", work_per_step=0.0,
        host_inputs=feats_h, alloc_inputs=alloc, step=step, launches=_launches, loss_weight=1.0,
        config={"model": f"CommAwareRGAT {mc.num_layers} layers x {ds.num_relations} relations, hidden {hidden}, DistributedBatchNorm1D, "
                         f"{'relation 0 only (as the reference ships)' if model.first_relation_only else 'all relations'}, max-stabilised segment softmax",
                "graph": f"synthetic MAG240M-shaped: {sc.num_papers * m} papers / {sc.num_authors * m} authors / {sc.num_institutions * m} "
                         f"institutions, {sc.num_features}-d, {sc.num_classes} classes (reference SyntheticDatasetConfig x{m})",
                "edges": E_global, "global_batch": "full graph", "seq_len": None, "parallelism": f"graph-partition x{world} + DP weights",
                "l2": "flushed implicitly: working set >> L2 only at x8; at x1 it is L2-resident (stated)",
                "engine": type(comm.engine).__name__, "cuda_graph": graphed is not None})


# ------------------------------------------------------------------------------------------------------------------------------
HALO_ROWS = (1_000, 10_000, 100_000, 1_000_000, 10_000_000)
HALO_F = (64, 256, 768)
HALO_HEADLINE = (1_000_000, 256)


def halo_points(args):
    rows = [r for r in HALO_ROWS if getattr(args, "halo_min_rows", 0) <= r <= args.halo_max_rows]
    return [(r, F) for r in rows for F in HALO_F]


def run_halo(args, comm, rank, world, dev, out, clock_sampler):
    """Not a training step: for every (rows, F) point the time of one halo gather (`HaloExchange.augmented`) and of gather + its
    backward (the scatter-sum of halo gradients to the owners), median of `--steps` (>= 30) after 10 warm-up exchanges, max over ranks;
    GB/s = max over ranks of max(bytes in, bytes out) / time.  The headline value is the gather at 1 M rows x 256-d."""
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from graph_utils import halo_rows_graph

    ex = HaloExchange(comm)
    iters = max(args.steps, 30)
    sweep, headline = [], None
    if rank == 0:
        clock_sampler.start()
    l0 = _launches()
    for rows, F in halo_points(args):
        per_peer = max(rows // max(world - 1, 1), 1)
        gph = halo_rows_graph(world, max(per_peer, 1024), per_peer)
        pat = build_communication_pattern(gph.edge_list.to(dev), gph.partitioning.to(dev), rank, world)
        x = ex.symmetric_rows(pat, F, torch.bfloat16, device=dev, tag=("halo", rows))
        x.normal_()
        xg = x.detach().clone().requires_grad_(True)

        def gather():
            with torch.no_grad():
                return ex.augmented(x, pat)

        def both():
            a = ex.augmented(xg, pat)
            a.backward(a.detach())

        t = torch.tensor([median_ms(gather, iters), median_ms(both, iters)], dtype=torch.float64, device=dev)
        nb = torch.tensor([float(max(pat.num_halo_vertices, int(pat.send_offset[-1])) * F * 2)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(nb, op=dist.ReduceOp.MAX)
        pt = {"halo_rows": rows, "F": F, "gather_ms": float(t[0]), "gather_scatter_ms": float(t[1]),
              "gather_GBps": float(nb) / float(t[0]) / 1e6, "gather_scatter_GBps": 2 * float(nb) / float(t[1]) / 1e6,
              "gather_rows_per_s": rows / (float(t[0]) * 1e-3), "frac_of_770GBps": float(nb) / float(t[0]) / 1e6 / 770.0}
        sweep.append(pt)
        if (rows, F) == HALO_HEADLINE or headline is None:
            headline = dict(pt)
            if not args.no_e2e:  # end to end: the step's rows come from pinned host memory, a checksum goes back to the host
                xh = _pin(x)
                res = torch.zeros((), dtype=torch.float32).pin_memory()

                def e2e():
                    x.copy_(xh, non_blocking=True)
                    a = gather()
                    res.copy_(a[-1].float().sum(), non_blocking=True)
                    torch.cuda.current_stream().synchronize()

                te = torch.tensor([median_ms(e2e, iters)], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(te, op=dist.ReduceOp.MAX)
                headline["e2e"] = {"value": float(nb) / float(te) / 1e6, "unit": "GB/s", "ms_per_step": float(te),
                                   "h2d_bytes_per_step": int(xh.numel() * 2), "d2h_bytes_per_step": 4}
        del x, xg, pat, gph
        if hasattr(comm.engine, "reset_arena"):  # the arena never frees single slots: start over for the next point
            torch.cuda.synchronize()
            comm.engine.reset_arena()
    clocks = clock_sampler.stop() if rank == 0 else None
    if rank == 0:
        print(json.dumps({
            "metric": "halo gather bandwidth (GB/s, max over ranks of max(bytes in, bytes out) / time) at 1 M halo rows x 256-d bf16",
            "value": headline["gather_GBps"], "unit": "GB/s", "n_gpus": world, "steps": iters, "warmup": 10,
            "ms_per_step": headline["gather_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "impl": "ours",
            "config": {"model": "HaloExchange.augmented (forward = gather) and its backward (scatter-sum to the owners)",
                       "graph": f"every rank pulls rows / (W - 1) distinct rows from every other rank; headline point {headline['halo_rows']} rows x {headline['F']}",
                       "global_batch": None, "seq_len": None, "parallelism": f"{world} ranks, all pairs",
                       "l2": "rows live in remote HBM; local L2 is bypassed by peer loads", "engine": type(comm.engine).__name__,
                       "nvlink_roofline_GBps": 770.0},
            "e2e": headline.get("e2e"), "gpu_launches": _launches() - l0, "clocks": clocks, "sweep": sweep}), file=out, flush=True)


# ------------------------------------------------------------------------------------------------------------------------------
BUILDERS = {"papers100m": build_papers100m, "graphcast": build_graphcast, "rgat": build_rgat}


def run(args, out, clock_sampler):
    if args.config == "halo":
        os.environ.setdefault("DGRAPH_B200_ARENA_MB", "98304")
    elif args.config == "papers100m":
        os.environ.setdefault("DGRAPH_B200_ARENA_MB", "65536")
    if getattr(args, "cuda_graph", False) and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ["DGRAPH_B200_DEVICE_EPOCH"] = "1"  # barrier epochs in device memory: steps with exchanges become capturable
    from _common import cleanup, init_distributed
    from dgraph_b200.ops import _native

    _native.lib()
    comm, rank, world, dev = init_distributed("nccl")
    if args.config == "halo":
        run_halo(args, comm, rank, world, dev, out, clock_sampler)
    else:
        wl = BUILDERS[args.config](args, comm, rank, world, dev)
        wl.extra["max_mem_GB_rank0"] = round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)
        run_workload(wl, args, rank, world, dev, "ours", out, clock_sampler)
    cleanup(comm)
