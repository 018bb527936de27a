#!/usr/bin/env python
"""Headline benchmark: full-graph GCN training throughput (edges/s) on an ogbn-products-shaped graph.

Contract (driver): ``python bench.py --gpus N --steps K --warmup W [--impl reference]`` prints ONE JSON line on rank 0.
For N > 1 it is launched under ``torch.distributed.run`` with one rank per GPU.

Metric (BASELINE.json / reference ``experiments/OGB/main.py:129-157``): one "step" = one full-graph epoch
``zero_grad -> forward -> loss -> backward -> optimizer step`` on the 3-layer GCN (2 x GraphConv(256) + Linear(47)),
timed with CUDA events between barriers, max over ranks; ``value = E_global * K / t`` edges per second.
Scaling is *strong*: the same 2.45 M-node / 61.9 M-edge graph is split over N GPUs.

Two arms:
  * ours      -- dgraph_b200 (bf16 compute, fp32 master weights + fused Adam), public API: Communicator,
                 build_communication_pattern, HaloExchange (features in ``symmetric_rows``), CommAwareGCN, FusedAdam.  With several
                 GPUs the step is captured once and replayed as a CUDA graph (``--no-cuda-graph`` for eager).
  * reference -- the unmodified LBANN/DGraph NCCL path installed under baseline/_ref (fp32), see baseline/ref_bench.py.
Both arms run the same end-to-end loop (``dgraph_b200/utils/e2e.py``) for the ``e2e`` entry and report the GLOBAL loss of the first and
last step.  The line also carries ``config.powerlaw_variant``: the same model and step, device-timed, on the power-law variant of the
graph (hub vertices of degree > 17 k; ``--no-powerlaw`` skips it).

Other workloads, same JSON line (``benchmarks/workloads.py``, reference arms in ``baseline/ref_workloads.py``):
``--config papers100m`` (GraphSAGE, 111 M nodes over 8 GPUs; weak scaling for fewer), ``--config graphcast``, ``--config rgat``
(``--rgat-mult 8`` for the 8x graph, ``--rgat-relations first`` to mirror the reference), ``--config halo`` (bandwidth sweep, field ``sweep``).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_JSON_OUT = sys.stdout
BASELINE_VALUE = None  # BASELINE.md publishes no number for this metric (``published: {}``)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="products", choices=["products", "papers100m", "graphcast", "rgat", "halo"],
                    help="products = the headline (BASELINE.json config 2); the others: benchmarks/workloads.py")
    ap.add_argument("--rgat-mult", type=float, default=1.0, help="rgat: multiple of the reference's SyntheticDatasetConfig graph")
    ap.add_argument("--rgat-hidden", type=int, default=64)
    ap.add_argument("--rgat-relations", default="all", choices=["all", "first"],
                    help="rgat: 'first' mirrors the reference as shipped, which processes relation 0 only (RGAT.py:356-358)")
    ap.add_argument("--halo-max-rows", type=int, default=10_000_000)
    ap.add_argument("--halo-min-rows", type=int, default=1_000)
    ap.add_argument("--graph", default="ogbn-products")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the graph (debug only; <1 is NOT the named config)")
    ap.add_argument("--hidden", type=int, default=256)
    ap.add_argument("--cut", type=float, default=0.1, help="fraction of long-range edges in the synthetic graph")
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--engine", default=None, help="override DGRAPH_B200_ENGINE (torch | peer)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cuda-graph", dest="cuda_graph", action="store_true", default=None,
                    help="capture the whole step (zero_grad+fwd+loss+bwd+Adam) in a CUDA graph and replay it; with several GPUs the peer "
                         "engine then keeps its barrier epochs in device memory.  Default: on with several GPUs (the per-rank step is "
                         "short enough for the host to matter once every step's loss is read back), off with one")
    ap.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false")
    ap.add_argument("--cpu", action="store_true", help="CPU/gloo plumbing run (tiny graphs only)")
    ap.add_argument("--no-powerlaw", action="store_true",
                    help="skip the second, device-timed measurement on the power-law variant of the graph (config.powerlaw_variant)")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(self.idx)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def setup_dist(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.cpu:
        dev = torch.device("cpu")
    else:
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("gloo" if args.cpu else "nccl", rank=rank, world_size=world,
                                **({} if args.cpu else {"device_id": dev}))
    return rank, world, local, dev


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank, world, local, dev = setup_dist(args)
    if args.engine:
        os.environ["DGRAPH_B200_ENGINE"] = args.engine
    if args.cuda_graph is None:
        args.cuda_graph = world > 1 and not args.cpu
    if args.cuda_graph and world > 1:  # barrier epochs in device memory make the multi-GPU step capturable
        os.environ["DGRAPH_B200_DEVICE_EPOCH"] = "1"
    from dgraph_b200 import Communicator
    from dgraph_b200.data import make_graph, contiguous_partition, node_features_and_labels
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.ops import _native
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    if not args.cpu:
        _native.lib()  # fail loudly if the sm_100a library is missing on a GPU box
    comm = Communicator.init_process_group("single" if world == 1 else "nccl")
    cdt = torch.float32 if args.cpu else torch.bfloat16

    # ---- graph, partition, plan (identical on every rank: same seed) -------------------------------
    g = make_graph(args.graph, cut_fraction=args.cut, seed=0, device=dev, scale=args.scale)
    E_global = int(g.edge_list.shape[0])
    part = contiguous_partition(g.num_nodes, world, device=dev)
    pat = build_communication_pattern(g.edge_list, part, rank, world)
    del g.edge_list
    rows = pat.local_vertices_global
    F_in = g.num_features
    F_pad = (F_in + 7) // 8 * 8  # 16-byte rows for TMA / vector loads
    x_host, y_host = node_features_and_labels(g.num_nodes, F_in, g.num_classes, seed=0, device=dev, rows=rows)
    train_mask = (torch.arange(rows.numel(), device=dev) % 10) == 0  # ~10 % labelled (products: 8 %)
    train_idx = torch.nonzero(train_mask, as_tuple=True)[0]
    xh = torch.zeros(rows.numel(), F_pad, dtype=cdt)
    xh[:, :F_in] = x_host.to(cdt).cpu()
    yh = y_host.cpu()
    if not args.cpu:
        xh, yh = xh.pin_memory(), yh.pin_memory()
    del x_host, y_host
    kp = pat.kernel_plan(dev)  # CSR/CSC build is one-time setup, not part of the epoch

    torch.manual_seed(0)
    model = CommAwareGCN(F_in, args.hidden, g.num_classes, HaloExchange(comm), comm, compute_dtype=cdt).to(dev)
    use_graph = bool(args.cuda_graph) and not args.cpu and (world == 1 or getattr(comm.engine, "device_epoch", False))
    opt = FusedAdam(model.parameters(), lr=args.lr, comm=comm, capturable=use_graph)
    from dgraph_b200.ops import _sparsity
    from dgraph_b200.ops.loss import masked_cross_entropy

    def forward_loss(x_dev, y_dev):
        out = model(x_dev, pat)  # logits of ALL local vertices, as in the reference's epoch
        return masked_cross_entropy(out, y_dev, train_idx)  # CrossEntropyLoss(out[train], y[train]), one kernel each way

    def step(x_dev, y_dev):
        opt.zero_grad()
        loss = forward_loss(x_dev, y_dev)
        loss.backward()
        opt.step()
        return loss

    def sync():
        if world > 1:
            dist.barrier()
        if not args.cpu:
            torch.cuda.synchronize()

    def timed(fn, k):
        sync()
        if args.cpu:
            t0 = time.perf_counter()
            fn(k)
            ms = (time.perf_counter() - t0) * 1e3
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(k)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
        sync()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- device-resident arm (kernel-only number) ---------------------------------------------------
    # features live where the exchange can use them in place (symmetric memory on the peer engine, plain memory otherwise): the
    # role of the reference's comm.alloc_buffer for its one-sided back-ends
    ex = model.halo_exchanger

    def feature_buffer(tag):
        return ex.symmetric_rows(pat, F_pad, cdt, device=dev, tag=tag)

    x_dev, y_dev = feature_buffer(0), yh.to(dev, non_blocking=True)
    x_dev.copy_(xh, non_blocking=True)
    losses = []
    if use_graph:  # same work per step, replayed with one cudaGraphLaunch (utils/cuda_graph.py); inputs stay in symmetric memory
        from dgraph_b200.utils.cuda_graph import GraphedTrainStep, capture_or_eager

        snap = opt.snapshot()  # capture trains for three steps (two eager warm-ups + the captured one): undo them, so that the first
        graphed = capture_or_eager(lambda: GraphedTrainStep(forward_loss, opt, (x_dev, y_dev), warmup=2, static_inputs=(x_dev, y_dev)),
                                   world)
        opt.restore(snap)      # reported loss is the loss of the initial weights, as without a graph
        use_graph = graphed is not None  # a failed capture leaves every rank on the eager step (and the JSON line says so)
        if use_graph:
            def step(x_in, y_in):  # noqa: F811
                return graphed(x_in, y_in).clone()

    def loop_dev(k):
        for _ in range(k):
            losses.append(step(x_dev, y_dev))

    loop_dev(args.warmup)
    sampler = ClockSampler(local)
    if rank == 0 and not args.cpu:
        sampler.start()
    l0 = _native.launches()
    ms_total = timed(loop_dev, args.steps)
    n_launch = _native.launches() - l0
    clocks = sampler.stop() if (rank == 0 and not args.cpu) else None
    ms_step = ms_total / args.steps
    value = E_global / (ms_step * 1e-3)

    # ---- end-to-end arm: per-step H2D of the step's inputs from pinned memory + D2H of the loss -------
    e2e = None
    if not args.no_e2e:
        h2d = xh.numel() * xh.element_size() + yh.numel() * yh.element_size()
        e2e_losses = []
        if args.cpu:
            def loop_e2e(k):
                for _ in range(k):
                    e2e_losses.append(float(step(xh.clone(), yh.clone()).detach()))
        else:
            from dgraph_b200.utils.e2e import make_e2e_loop

            if use_graph:  # the graph owns x_dev / y_dev: host copies land in two separate buffers and are copied in before each replay
                bufs = [(torch.empty_like(xh, device=dev), torch.empty_like(yh, device=dev)) for _ in range(2)]
            else:
                bufs = [(x_dev, y_dev), (feature_buffer(1), torch.empty_like(yh, device=dev))]
            loop_e2e = make_e2e_loop(step, (xh, yh), bufs, e2e_losses)  # same loop as the reference arm (baseline/ref_bench.py)

        loop_e2e(max(2, args.warmup))
        ms_e2e = timed(loop_e2e, args.steps) / args.steps
        e2e = {"value": E_global / (ms_e2e * 1e-3), "unit": "edges/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
               "copy_scheme": "double-buffered H2D on a side stream; loss read on the host one step later (every step, inside the "
                              "timed region); identical in both arms"}

    # ---- same model and step on the power-law variant of the graph (hub vertices of degree > 17 k, ids scattered) ---------------
    powerlaw = None
    if not args.no_powerlaw and not args.cpu and args.graph == "ogbn-products":
        del model, opt, kp, pat
        torch.cuda.empty_cache()
        g2 = make_graph(args.graph, cut_fraction=args.cut, seed=0, device=dev, scale=args.scale, degree_model="powerlaw")
        E2 = int(g2.edge_list.shape[0])
        deg = torch.bincount(g2.edge_list[:, 0], minlength=g2.num_nodes)
        max_deg, n_hub = int(deg.max()), int((deg > 2048).sum())
        pat2 = build_communication_pattern(g2.edge_list, contiguous_partition(g2.num_nodes, world, device=dev), rank, world)
        del g2.edge_list, deg
        rows2 = pat2.local_vertices_global
        x2h, y2 = node_features_and_labels(g2.num_nodes, F_in, g2.num_classes, seed=0, device=dev, rows=rows2)
        x2 = ex.symmetric_rows(pat2, F_pad, cdt, device=dev, tag=2)
        x2.zero_()
        x2[:, :F_in] = x2h.to(cdt)
        idx2 = torch.nonzero((torch.arange(rows2.numel(), device=dev) % 10) == 0, as_tuple=True)[0]
        pat2.kernel_plan(dev)
        torch.manual_seed(0)
        model2 = CommAwareGCN(F_in, args.hidden, g2.num_classes, ex, comm, compute_dtype=cdt).to(dev)
        opt2 = FusedAdam(model2.parameters(), lr=args.lr, comm=comm)

        def loop2(k):
            for _ in range(k):
                opt2.zero_grad()
                masked_cross_entropy(model2(x2, pat2), y2, idx2).backward()
                opt2.step()

        loop2(args.warmup)
        k2 = max(5, args.steps // 2)
        ms2 = timed(loop2, k2) / k2
        powerlaw = {"graph": f"{g2.name} V={g2.num_nodes} E={E2}", "max_degree": max_deg, "vertices_with_degree_over_2048": n_hub,
                    "ms_per_step": ms2, "edges_per_s": E2 / (ms2 * 1e-3), "steps": k2, "timing": "device, eager, max over ranks",
                    "hub_rows": "rows / columns longer than 2048 slots are split into 512-slot chunks (GraphKernelPlan.hub_plan)"}

    # global loss (sum over all labelled vertices / their number) of the first and the last device-arm step: comparable across N
    n_lab = float(train_idx.numel())
    gl = torch.tensor([float(losses[0].detach()) * n_lab, float(losses[-1].detach()) * n_lab, n_lab], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(gl)
    loss_first_global, loss_last_global = float(gl[0] / gl[2]), float(gl[1] / gl[2])
    if hasattr(comm.engine, "check_timeouts"):
        comm.engine.check_timeouts()
    final_loss = float(losses[-1].detach())
    if rank == 0:
        out = {
            "metric": "full-graph GCN training throughput (edges/s; epoch = zero_grad+fwd+loss+bwd+step)",
            "value": value, "unit": "edges/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": (value / BASELINE_VALUE) if BASELINE_VALUE else None,
            "dtype": "fp32" if args.cpu else "bf16", "data": "synthetic", "impl": "ours",
            "config": {"model": "CommAwareGCN 2xGraphConv(256)+Linear(47) [3-layer GCN]",
                       "graph": f"{g.name} V={g.num_nodes} E={E_global} F={F_in}", "global_batch": "full graph",
                       "seq_len": None, "parallelism": f"graph-partition x{world} + DP weights",
                       "partition": f"contiguous blocks, cut_fraction={args.cut}", "hidden": args.hidden,
                       "l2": "inputs larger than L2 (edge structure + features >> 126 MB)",
                       "engine": type(comm.engine).__name__, "scale": args.scale,
                       "exchange": ("fused into the layer kernels: conv1 pulls feature rows in place (no publish copy), conv2 projects "
                                    "into the symmetric slot and the aggregation kernel / one pull kernel reads remote Q rows over NVLink"
                                    if world > 1 and not args.cpu else "none (1 rank)" if world == 1 else "gloo all-to-all"),
                       "gemm": os.environ.get("DGRAPH_B200_GEMM", "tcgen05"), "cuda_graph": use_graph,
                       "powerlaw_variant": powerlaw,
                       "loss": "fused masked cross-entropy over the labelled rows (10 %)",
                       "grad_row_sparsity": ("exact: classifier and last-conv backward skip rows whose loss gradient is "
                                             "structurally zero" if _sparsity.ENABLED and not args.cpu else "off")},
            "e2e": e2e, "gpu_launches": n_launch, "clocks": clocks, "final_loss": loss_last_global,
            "loss_first": loss_first_global, "loss_scope": "global (all-reduced over ranks)",
            "final_loss_rank0": final_loss,
        }
        print(json.dumps(out), file=_JSON_OUT, flush=True)
    comm.destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    # stdout carries exactly ONE line (the JSON): anything native libraries print there (NCCL's version banner under
    # NCCL_DEBUG=VERSION, for instance) is routed to stderr for the whole run.
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import builtins

    builtins._DGB_JSON_OUT = _JSON_OUT  # baseline/ref_bench.py prints its line through the same descriptor
    if args.config != "products":
        torch_local = int(os.environ.get("LOCAL_RANK", "0"))
        if args.impl == "reference":
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            try:
                import ref_workloads
            except Exception as e:  # noqa: BLE001
                if int(os.environ.get("RANK", "0")) == 0:
                    print(json.dumps({"impl": "reference", "unavailable": f"reference harness import failed: {e!r}"[:300]}),
                          file=_JSON_OUT, flush=True)
                return
            ref_workloads.run(args, _JSON_OUT, ClockSampler(torch_local))
            return
        from benchmarks import workloads

        workloads.run(args, _JSON_OUT, ClockSampler(torch_local))
        return
    if args.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        try:
            import ref_bench
        except Exception as e:  # noqa: BLE001
            if int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"reference harness import failed: {e!r}"[:300]}),
                      file=_JSON_OUT, flush=True)
            return
        ref_bench.run(args)
        return
    run_ours(args)


if __name__ == "__main__":
    main()
