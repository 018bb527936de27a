"""One launch, several answers about the multi-GPU GCN step (diagnostic; `bench.py` alone produces bench values):

* ms/step of the same training step under each exchange route (default, in-kernel pull, pull into a buffer, the reference's
  order of operations), graph and plans built once;
* a per-kernel breakdown of one step of the default route on rank 0 (torch.profiler: launches, device time) together with the
  bytes every peer-memory kernel moves over NVLink, computed from the plan -- the input of `profiles/step_breakdown_<N>gpu.md`;
* device time of the whole step vs the sum of its kernels (what is launch / barrier wait).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/scale_modes.py [--modes a,b] [--out f.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200 import Communicator  # noqa: E402
from dgraph_b200.data import contiguous_partition, make_graph, node_features_and_labels  # noqa: E402
from dgraph_b200.models import gcn as gcn_mod  # noqa: E402
from dgraph_b200.models.gcn import CommAwareGCN  # noqa: E402
from dgraph_b200.ops.loss import masked_cross_entropy  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402
from dgraph_b200.parallel import HaloExchange, build_communication_pattern  # noqa: E402

MODES = {  # name -> (DGRAPH_B200_GCN_EXCHANGE, DGRAPH_B200_PULL, extra environment)
    "default": ("auto", None, {}),
    "pull_kernel": ("auto", "kernel", {}),
    "pull_buffer": ("auto", "buffer", {}),
    "exchange_x": ("x", None, {}),
    "overlap_gemm": ("auto", None, {"DGRAPH_B200_OVERLAP": "gemm"}),
    "pull_prefetch": ("auto", "kernel", {"DGRAPH_B200_PULL_PREFETCH": "1"}),
}
_EXTRA_KEYS = sorted({k for m in MODES.values() for k in m[2]})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--modes", default="default,pull_kernel,pull_buffer,exchange_x")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--graph", default="ogbn-products")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--out", default=None)
    ap.add_argument("--profile", default="default", help="mode whose step is profiled kernel by kernel ('' = none)")
    a = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    g = make_graph(a.graph, device=dev, scale=a.scale)
    E_global = int(g.edge_list.shape[0])
    comm = Communicator.init_process_group("single" if world == 1 else "nccl")
    pat = build_communication_pattern(g.edge_list, contiguous_partition(g.num_nodes, world, device=dev), rank, world)
    del g.edge_list
    rows = pat.local_vertices_global
    F_in = g.num_features
    F_pad = (F_in + 7) // 8 * 8
    x, y = node_features_and_labels(g.num_nodes, F_in, g.num_classes, seed=0, device=dev, rows=rows)
    ex = HaloExchange(comm)
    xd = ex.symmetric_rows(pat, F_pad, torch.bfloat16, device=dev)
    xd.zero_()
    xd[:, :F_in] = x.to(torch.bfloat16)
    idx = torch.arange(0, rows.numel(), 10, device=dev)
    kp = pat.kernel_plan(dev)
    facts = {"world": world, "local_rows": int(rows.numel()), "halo_rows": int(pat.num_halo_vertices),
             "local_edges": int(pat.local_edge_list.shape[0]), "remote_edges": int(kp.remote_edge_count()),
             "boundary_rows_sent": int(pat.send_offset[-1])}
    if world > 1:
        info = comm.engine.pattern_info(pat)
        facts.update(boundary_rows=int(info.n_bnd), halo_copies_of_my_rows=int(info.src_row.numel()),
                     pull_mode_auto=comm.engine.pull_mode(info, kp))
    out = {"facts_rank0": facts, "modes": {}}

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for name in [m for m in a.modes.split(",") if m]:
        exch, pull, extra = MODES[name]
        for k in _EXTRA_KEYS:
            os.environ.pop(k, None)
        os.environ.update(extra)
        gcn_mod._EXCHANGE = exch
        if pull is None:
            os.environ.pop("DGRAPH_B200_PULL", None)
        else:
            os.environ["DGRAPH_B200_PULL"] = pull
        if world > 1 and hasattr(comm.engine.pattern_info(pat), "_pull_mode"):
            del comm.engine.pattern_info(pat)._pull_mode
        torch.manual_seed(0)
        model = CommAwareGCN(F_in, 256, g.num_classes, ex, comm, compute_dtype=torch.bfloat16).to(dev)
        opt = FusedAdam(model.parameters(), lr=1e-3, comm=comm)

        def step():
            opt.zero_grad()
            loss = masked_cross_entropy(model(xd, pat), y, idx)
            loss.backward()
            opt.step()
            return loss

        for _ in range(a.warmup):
            step()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
        tmin = t.clone()
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        out["modes"][name] = {"ms_per_step_max": float(t), "ms_per_step_min": float(tmin), "loss": float(loss),
                              "edges_per_s": E_global / (float(t) * 1e-3)}
        if rank == 0:
            print(f"{name}: {float(t):.3f} ms/step (min over ranks {float(tmin):.3f})", flush=True)
        if name == a.profile:
            from torch.profiler import ProfilerActivity, profile

            sync()
            with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
                for _ in range(3):
                    step()
                torch.cuda.synchronize()
            rowsk = []
            for ev in prof.key_averages():
                dt = getattr(ev, "device_time_total", None)
                if dt is None:
                    dt = getattr(ev, "cuda_time_total", 0.0)
                if ev.device_type.name == "CUDA" and dt > 0:
                    rowsk.append({"kernel": ev.key[:110], "launches_per_step": ev.count / 3.0, "us_per_step": dt / 3.0})
            rowsk.sort(key=lambda r: -r["us_per_step"])
            out["profile_rank0"] = {"mode": name, "kernels": rowsk, "sum_us_per_step": sum(r["us_per_step"] for r in rowsk)}
            sync()
        del model, opt
    if rank == 0:
        s = json.dumps(out)
        if a.out:
            os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
            with open(a.out, "w") as f:
                f.write(s)
        print(s)
    comm.destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
