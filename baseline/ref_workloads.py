"""Reference arms of `bench.py --config {halo, rgat, graphcast, papers100m}`: the UNMODIFIED LBANN/DGraph from `baseline/_ref`
(`tools/install_reference.sh`) driven through its own public API and stock code path.  Nothing of dgraph_b200 runs inside a timed
region: this file only generates the synthetic inputs (the same generators as our arm, outside the timed region) and loops.

  halo        reference `HaloExchange` -> `Communicator.put` -> `NCCLBackendEngine.put` (`all_to_all_single`, fp32) on the same
              all-pairs halo graphs as our arm (`DGraph/distributed/haloExchange.py:121-139`, `nccl/NCCLBackendEngine.py:281-306`).
  rgat        reference `CommAwareRGAT` (`experiments/OGB-LSC/RGAT.py:271-382`, CommPlan gather / scatter + `torch_local`) + its
              `SyntheticHeterogeneousDataset` + DDP + Adam, the step of `Trainer.train` (`Trainer.py:131-146`) without its prints,
              `empty_cache()` and per-epoch device synchronisation.
  graphcast   one GPU only: the reference's unmodified `model.py` / `layers.py` (`DGraphCast`, 1,025,627 parameters, fp32) + Adam + MSE.  Two
              pieces of the reference cannot be used as shipped and are replaced OUTSIDE `baseline/_ref`: its
              `SingleProcessDummyCommunicator` is abstract (a 20-line concrete equivalent is defined below, same `torch.gather` /
              `scatter_add` bodies), and its graph generator cannot construct its graph object, so the static graph comes from this
              repository's generator (same field names, same 0.25 degree sizes).  More than one rank: unavailable (distributed wiring
              incomplete upstream).
  papers100m  unavailable: the reference has no GraphSAGE model and needs the global edge list on every rank.
"""
from __future__ import annotations

import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def _unavailable(out, why):
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why[:300]}), file=out, flush=True)


def _setup():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    if not dist.is_initialized():  # the reference's NCCL engine needs a process group even for one rank
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    return rank, world, local, dev


def _halo(args, out, sampler):
    from DGraph import Communicator
    from DGraph.distributed import HaloExchange
    from DGraph.distributed.commInfo import build_communication_pattern

    sys.path.insert(0, os.path.join(ROOT, "experiments", "Benchmarks"))
    from benchmarks.runner import median_ms
    from benchmarks.workloads import HALO_HEADLINE, halo_points
    from graph_utils import halo_rows_graph  # synthetic graph generator only (rank-contiguous ids, as the reference requires: Q8)

    rank, world, local, dev = _setup()
    comm = Communicator.init_process_group("nccl")
    ex = HaloExchange(comm)
    iters = max(args.steps, 30)
    sweep, headline = [], None
    if rank == 0:
        sampler.start()
    for rows, F in halo_points(args):
        per_peer = max(rows // max(world - 1, 1), 1)
        g = halo_rows_graph(world, max(per_peer, 1024), per_peer)
        # the reference derives what q sends to r from q's own edges into r: give it the symmetric closure of the pull graph
        e = torch.cat([g.edge_list, g.edge_list.flip(1)]).to(dev)
        pat = build_communication_pattern(e, g.partitioning.to(dev), rank, world)
        x = torch.randn(pat.num_local_vertices, F, device=dev)  # fp32: the reference's only dtype
        xg = x.clone().requires_grad_(True)

        def gather():
            with torch.no_grad():
                return ex(x, pat)

        def both():
            h = ex(xg, pat)
            h.backward(h.detach())

        t = torch.tensor([median_ms(gather, iters), median_ms(both, iters)], dtype=torch.float64, device=dev)
        nb = torch.tensor([float(max(int(pat.recv_offset[-1]), int(pat.send_offset[-1])) * F * 4)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(nb, op=dist.ReduceOp.MAX)
        pt = {"halo_rows": int(pat.recv_offset[-1]), "F": F, "gather_ms": float(t[0]), "gather_scatter_ms": float(t[1]),
              "gather_GBps": float(nb) / float(t[0]) / 1e6, "gather_scatter_GBps": 2 * float(nb) / float(t[1]) / 1e6,
              "gather_rows_per_s": int(pat.recv_offset[-1]) / (float(t[0]) * 1e-3), "frac_of_770GBps": float(nb) / float(t[0]) / 1e6 / 770.0}
        sweep.append(pt)
        if (rows, F) == HALO_HEADLINE or headline is None:
            headline = dict(pt)
            if not args.no_e2e:
                xh = x.cpu().pin_memory()
                res = torch.zeros((), dtype=torch.float32).pin_memory()

                def e2e():
                    x.copy_(xh, non_blocking=True)
                    h = gather()
                    res.copy_(h[-1].sum(), non_blocking=True)
                    torch.cuda.current_stream().synchronize()

                te = torch.tensor([median_ms(e2e, iters)], dtype=torch.float64, device=dev)
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
                headline["e2e"] = {"value": float(nb) / float(te) / 1e6, "unit": "GB/s", "ms_per_step": float(te),
                                   "h2d_bytes_per_step": int(xh.numel() * 4), "d2h_bytes_per_step": 4}
        del x, xg, pat, e, g
        torch.cuda.empty_cache()
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        print(json.dumps({
            "metric": "halo gather bandwidth (GB/s, max over ranks of max(bytes in, bytes out) / time) at 1 M halo rows x 256-d",
            "value": headline["gather_GBps"], "unit": "GB/s", "n_gpus": world, "steps": iters, "warmup": 10,
            "ms_per_step": headline["gather_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic", "impl": "reference",
            "config": {"model": "reference HaloExchange (index -> NCCLBackendEngine.put = all_to_all_single; halo only, no cat) and its backward",
                       "graph": f"every rank receives rows / (W - 1) distinct rows from every other rank; headline point {headline['halo_rows']} rows x {headline['F']}",
                       "global_batch": None, "seq_len": None, "parallelism": f"{world} ranks, all pairs", "l2": "inputs larger than L2 at the headline point",
                       "note": "fp32 rows are twice as wide as our bf16 rows: compare gather_rows_per_s as well as GB/s"},
            "e2e": headline.get("e2e"), "gpu_launches": 0, "clocks": clocks, "sweep": sweep}), file=out, flush=True)
    dist.barrier()
    dist.destroy_process_group()


def _rgat(args, out, sampler):
    from benchmarks.runner import Workload, run_workload

    sys.path.insert(0, os.path.join(REF, "experiments", "OGB-LSC"))
    from DGraph import Communicator
    from config import ModelConfig, SyntheticDatasetConfig, TrainingConfig
    from distributed_layers import GetGlobalVal
    from lsc_datasets.synthetic_dataset import SyntheticHeterogeneousDataset
    from RGAT import CommAwareRGAT
    from torch.nn.parallel import DistributedDataParallel as DDP

    rank, world, local, dev = _setup()
    comm = Communicator.init_process_group("nccl")
    m = max(int(round(args.rgat_mult)), 1)
    sc = SyntheticDatasetConfig()
    sc.num_papers, sc.num_authors, sc.num_institutions = sc.num_papers * m, sc.num_authors * m, sc.num_institutions * m
    mc, tc = ModelConfig(), TrainingConfig()
    torch.manual_seed(0)  # the reference's generator draws an unseeded graph on every rank; all ranks must see the same one
    ds = SyntheticHeterogeneousDataset(sc, comm).add_batch_dimension()
    torch.manual_seed(0)
    model = CommAwareRGAT(in_channels=ds.num_features, out_channels=ds.num_classes, num_relations=ds.num_relations,
                          hidden_channels=args.rgat_hidden, num_layers=mc.num_layers, heads=mc.heads, comm=comm,
                          dropout=mc.dropout).to(dev)
    model = DDP(model, device_ids=[local], find_unused_parameters=True)
    opt = torch.optim.Adam(model.parameters(), lr=tc.lr, weight_decay=5e-4)
    xs, _, edge_type, _ = ds[0]
    plans = ds.get_NCCL_comm_plans()
    train_mask = ds.get_mask("train").to(dev)
    target = ds.get_target("train").flatten().to(dev)
    n_global = GetGlobalVal(target.size(0))
    feats_h = tuple(x.cpu().contiguous().pin_memory() for x in xs)
    model.train()

    def step(*x):
        opt.zero_grad(set_to_none=True)
        o = model(list(x), edge_type, plans)
        loss = torch.nn.functional.cross_entropy(o[:, train_mask, :].squeeze(0), target, reduction="sum") / n_global
        loss.backward()
        opt.step()
        return loss

    wl = Workload(
        metric="RGAT full-graph epoch time (zero_grad+fwd+loss+bwd+step), synthetic MAG240M-shaped graph", unit="ms", work_per_step=0.0,
        dtype="fp32", host_inputs=feats_h, alloc_inputs=lambda tag: tuple(torch.empty_like(f, device=dev) for f in feats_h), step=step,
        loss_weight=1.0,
        config={"model": f"reference CommAwareRGAT as shipped ({mc.num_layers} layers, hidden {args.rgat_hidden}; only relation 0 is "
                         "processed, RGAT.py:356-358), CommPlan gather/scatter + torch_local, DDP(find_unused_parameters), Adam",
                "graph": f"reference SyntheticHeterogeneousDataset: {sc.num_papers} papers / {sc.num_authors} authors / {sc.num_institutions} "
                         f"institutions, {sc.num_features}-d, {sc.num_classes} classes (SyntheticDatasetConfig x{m})",
                "global_batch": "full graph", "seq_len": None, "parallelism": f"graph-partition x{world} + DDP",
                "l2": "as our arm"})
    run_workload(wl, args, rank, world, dev, "reference", out, sampler)
    dist.barrier()
    dist.destroy_process_group()


def _graphcast(args, out, sampler):
    from benchmarks.runner import Workload, run_workload

    sys.path.insert(0, os.path.join(REF, "experiments", "GraphCast"))
    import graphcast_config
    from DGraph.CommunicatorBase import CommunicatorBase
    from model import DGraphCast

    rank, world, local, dev = _setup()
    if world != 1:
        dist.destroy_process_group()
        return _unavailable(out, "reference GraphCast has no working multi-rank path (SURVEY.md Q7); its reference arm is one GPU only")

    class _SingleProcessCommunicator(CommunicatorBase):
        """Concrete stand-in for the reference's abstract `SingleProcessDummyCommunicator` (`experiments/GraphCast/dist_utils.py:8-46`):
        the same two tensor expressions, plus the abstract methods it forgot."""

        def __init__(self):
            super().__init__()
            self.backend, self._ranks_per_sample = "single", 1

        def init_process_group(self, *a, **k):
            return self

        def get_rank(self):
            return 0

        def get_world_size(self):
            return 1

        def barrier(self):
            pass

        def destroy(self):
            pass

        def scatter(self, tensor, src, rank_mappings, num_local_nodes):
            src = src.unsqueeze(-1).expand(-1, tensor.shape[-1])
            return torch.zeros(num_local_nodes, tensor.shape[1]).to(tensor.device).scatter_add(0, src, tensor)

        def gather(self, tensor, dst, rank_mappings):
            return torch.gather(tensor, 0, dst.unsqueeze(-1).expand(-1, tensor.shape[-1]))

    # static graph and data: this repository's generators (data only, outside the timed region)
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator, SyntheticWeatherDataset

    cfg = graphcast_config.Config()
    res, level = (721, 1440), cfg.model.mesh_level
    if args.scale < 1.0:  # debug only
        res, level = (max(int(721 * args.scale), 19), max(int(1440 * args.scale), 36)), max(2, int(round(6 + torch.log2(torch.tensor(args.scale)).item())))
        cfg.model.mesh_level = level
    sg = DistributedGraphCastGraphGenerator(res, level, multi=cfg.model.multimesh).shard(0, 1).to(dev, torch.float32)
    for n in list(sg.__dataclass_fields__):
        v = getattr(sg, n)
        if isinstance(v, torch.Tensor) and not v.is_floating_point():
            setattr(sg, n, v.long())
    item = SyntheticWeatherDataset(cfg.model.input_grid_dim, res, num_samples=2, grid_rows=sg.grid_rows_global.cpu())[0]
    xh, yh = item["invar"].float().contiguous().pin_memory(), item["outvar"].float().contiguous().pin_memory()
    torch.manual_seed(0)
    model = DGraphCast(cfg, _SingleProcessCommunicator()).to(dev)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)

    def step(x, y):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(model(x.unsqueeze(0), sg), y)
        loss.backward()
        opt.step()
        return loss

    wl = Workload(
        metric="GraphCast 0.25 degree training step time (zero_grad+fwd+MSE+bwd+Adam)", unit="ms", work_per_step=0.0, dtype="fp32",
        host_inputs=(xh, yh), alloc_inputs=lambda tag: (torch.empty_like(xh, device=dev), torch.empty_like(yh, device=dev)), step=step,
        loss_weight=float(xh.shape[0]),
        config={"model": f"reference DGraphCast (model.py / layers.py unmodified, {sum(p.numel() for p in model.parameters())} parameters, fp32) + "
                         "torch.optim.Adam; single-process communicator and static graph from outside baseline/_ref (the reference's own "
                         "cannot be instantiated as shipped, SURVEY Q7)",
                "graph": f"grid {res[0]}x{res[1]} = {res[0] * res[1]} nodes, icosahedral multimesh level {level} (this repository's generator)",
                "global_batch": 1, "seq_len": None, "parallelism": "1 GPU", "l2": "inputs larger than L2"})
    run_workload(wl, args, rank, world, dev, "reference", out, sampler)
    dist.barrier()
    dist.destroy_process_group()


def run(args, out, sampler):
    if not os.path.isdir(os.path.join(REF, "DGraph")):
        return _unavailable(out, "baseline/_ref is missing: run tools/install_reference.sh (offline pip install of /root/reference)")
    if args.config == "papers100m":
        return _unavailable(out, "reference has no GraphSAGE model and its plan builder needs the global edge list (26 GB of int64 pairs) "
                                 "on every rank")
    if ROOT not in sys.path:
        sys.path.append(ROOT)
    sys.path.insert(0, REF)  # ahead of the repo root, whose DGraph/ directory is our import-compatible shim
    assert "DGraph" not in sys.modules, "our DGraph shim was imported before the reference"
    try:
        {"halo": _halo, "rgat": _rgat, "graphcast": _graphcast}[args.config](args, out, sampler)
    except torch.OutOfMemoryError as e:
        _unavailable(out, f"reference ran out of device memory: {str(e)[:200]}")
    except Exception as e:  # noqa: BLE001
        _unavailable(out, f"reference failed: {e!r}")
        try:
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass
