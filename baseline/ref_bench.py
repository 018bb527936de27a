"""Reference arm of ``bench.py``: the UNMODIFIED LBANN/DGraph NCCL path on the same metric / graph / config.

Nothing from ``dgraph_b200`` runs on the timed path: the graph plan comes from the reference's
``DGraph.distributed.commInfo.build_communication_pattern``, the model is the reference's
``experiments/OGB/GCN.py::CommAwareGCN`` (imported by path from ``baseline/_ref/experiments/OGB``), halo exchange is
``DGraph.distributed.HaloExchange`` over ``NCCLBackendEngine.put`` (``dist.all_to_all_single``), weights are synced by
PyTorch DDP and the optimizer is ``torch.optim.Adam`` -- i.e. the epoch of ``experiments/OGB/main.py:129-157`` "as
shipped" (fp32, including the ``TimingReport`` barriers inside the model's forward).

Only the *data* is ours (no network => no OGB download): the same synthetic products-shaped graph, same seed, same
rank-contiguous partition (required by the reference), same features / labels / train mask.
"""
from __future__ import annotations

import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def _json_out():
    import builtins

    return getattr(builtins, "_DGB_JSON_OUT", sys.stdout)  # bench.py keeps the real stdout here (native noise -> stderr)


def _unavailable(why: str):
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why[:400]}), file=_json_out(), flush=True)


def run(args):
    if not os.path.isdir(os.path.join(REF, "DGraph")):
        return _unavailable("baseline/_ref/DGraph missing: reference not installed (see DESIGN.md, 'Reference install')")
    ogb_dir = os.path.join(REF, "experiments", "OGB")
    if not os.path.isfile(os.path.join(ogb_dir, "GCN.py")):
        return _unavailable("baseline/_ref/experiments/OGB/GCN.py missing (copy of the reference's model file)")
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        return _unavailable("reference NCCL path needs CUDA (commInfo.py:151-152 forces .cuda())")

    sys.path.insert(0, REF)
    sys.path.insert(0, ogb_dir)
    root = os.path.dirname(HERE)
    if root not in sys.path:
        sys.path.append(root)
    from bench import ClockSampler, setup_dist  # timing utilities only

    rank, world, local, dev = setup_dist(args)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    try:
        from DGraph import Communicator
        from DGraph.distributed import HaloExchange
        from DGraph.distributed.commInfo import build_communication_pattern
        from DGraph.utils.TimingReport import TimingReport
        from GCN import CommAwareGCN
    except Exception as e:  # noqa: BLE001
        return _unavailable(f"reference import failed: {e!r}")
    # data generation (ours; outside the timed region)
    from dgraph_b200.data import make_graph, contiguous_partition, node_features_and_labels

    try:
        comm = Communicator.init_process_group("nccl")
        TimingReport.init(comm)
        g = make_graph(args.graph, cut_fraction=args.cut, seed=0, device=dev, scale=args.scale)
        E_global = int(g.edge_list.shape[0])
        part = contiguous_partition(g.num_nodes, world, device=dev)
        pat = build_communication_pattern(g.edge_list, part, rank, world)
        rows = torch.where(part == rank)[0]
        del g.edge_list
        x, y = node_features_and_labels(g.num_nodes, g.num_features, g.num_classes, seed=0, device=dev, rows=rows)
        xh, yh = x.cpu().pin_memory(), y.cpu().pin_memory()
        train_mask = (torch.arange(rows.numel(), device=dev) % 10) == 0
        torch.manual_seed(0)
        model = CommAwareGCN(g.num_features, args.hidden, g.num_classes, HaloExchange(comm), comm).to(dev)
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP

            model = DDP(model, device_ids=[local], output_device=local)
        opt = torch.optim.Adam(model.parameters(), lr=args.lr)
        crit = torch.nn.CrossEntropyLoss()

        def step(xd, yd):
            opt.zero_grad()
            out = model(xd, pat)
            loss = crit(out[train_mask], yd[train_mask].reshape(-1))
            loss.backward()
            opt.step()
            return loss

        def sync():
            comm.barrier()
            torch.cuda.synchronize()

        def timed(fn, k):
            sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(k)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            sync()
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0])

        losses = []

        def loop_dev(k):
            for _ in range(k):
                losses.append(step(x, y))

        loop_dev(args.warmup)
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        ms_step = timed(loop_dev, args.steps) / args.steps
        clocks = sampler.stop() if rank == 0 else None
        value = E_global / (ms_step * 1e-3)

        e2e = None
        if not args.no_e2e:
            # the same host <-> device loop as our arm (harness code only: double-buffered H2D, loss read one step later)
            from dgraph_b200.utils.e2e import make_e2e_loop

            bufs = [(torch.empty_like(xh, device=dev), torch.empty_like(yh, device=dev)) for _ in range(2)]
            loop_e2e = make_e2e_loop(step, (xh, yh), bufs, [])

            loop_e2e(2)
            ms_e2e = timed(loop_e2e, args.steps) / args.steps
            e2e = {"value": E_global / (ms_e2e * 1e-3), "unit": "edges/s", "ms_per_step": ms_e2e,
                   "h2d_bytes_per_step": int(xh.numel() * 4 + yh.numel() * 8), "d2h_bytes_per_step": 4}
        if rank == 0:
            print(json.dumps({
                "metric": "full-graph GCN training throughput (edges/s; epoch = zero_grad+fwd+loss+bwd+step)",
                "value": value, "unit": "edges/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "fp32", "data": "synthetic", "impl": "reference",
                "config": {"model": "reference experiments/OGB/GCN.py CommAwareGCN (as shipped, incl. TimingReport scopes)",
                           "graph": f"{g.name} V={g.num_nodes} E={E_global} F={g.num_features}",
                           "global_batch": "full graph", "seq_len": None,
                           "parallelism": f"graph-partition x{world} + DDP", "hidden": args.hidden, "scale": args.scale,
                           "partition": f"contiguous blocks, cut_fraction={args.cut}",
                           "l2": "inputs larger than L2"},
                "e2e": e2e, "gpu_launches": 0, "clocks": clocks, "final_loss": float(losses[-1].detach()),
            }), file=_json_out(), flush=True)
    except torch.OutOfMemoryError as e:
        _unavailable(f"reference ran out of device memory at {world} GPU(s) on this graph (fp32 [E,2F] edge intermediates): "
                     f"{str(e)[:160]}")
    except Exception as e:  # noqa: BLE001
        _unavailable(f"reference failed: {e!r}")
    finally:
        try:
            if dist.is_initialized():
                dist.barrier()
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass
