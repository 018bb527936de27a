"""Per-kernel device time of one GCN training step (all kernels, ATen included) via torch.profiler -- a diagnostic to find what is
left outside the hand-written kernels; numbers taken under the profiler are never benchmark values."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200 import Communicator  # noqa: E402
from dgraph_b200.data import contiguous_partition, make_graph, node_features_and_labels  # noqa: E402
from dgraph_b200.models.gcn import CommAwareGCN  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402
from dgraph_b200.parallel import HaloExchange, build_communication_pattern  # noqa: E402
from dgraph_b200.utils.tracing import profile_step  # noqa: E402


def main():
    import torch.distributed as dist

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:  # torchrun --nproc-per-node N --master-addr 127.0.0.1 benchmarks/prof_step.py (rank 0 prints)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    name = os.environ.get("GRAPH", "ogbn-products")
    g = make_graph(name, device=dev, scale=float(os.environ.get("SCALE", "1.0")))
    comm = Communicator.init_process_group("single" if world == 1 else "nccl")
    pat = build_communication_pattern(g.edge_list, contiguous_partition(g.num_nodes, world, device=dev), rank, world)
    del g.edge_list
    rows = pat.local_vertices_global
    F_in = g.num_features
    F_pad = (F_in + 7) // 8 * 8
    x, y = node_features_and_labels(g.num_nodes, F_in, g.num_classes, seed=0, device=dev, rows=rows)
    xd = torch.zeros(rows.numel(), F_pad, dtype=torch.bfloat16, device=dev)
    xd[:, :F_in] = x.to(torch.bfloat16)
    idx = torch.arange(0, rows.numel(), 10, device=dev)
    if rank == 0:
        print(f"world={world} local rows={rows.numel()} halo rows={pat.num_halo_vertices} local edges={pat.local_edge_list.shape[0]}")
    pat.kernel_plan(dev)
    model = CommAwareGCN(F_in, 256, g.num_classes, HaloExchange(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-3, comm=comm)
    from dgraph_b200.ops.loss import masked_cross_entropy

    def step():
        opt.zero_grad()
        loss = masked_cross_entropy(model(xd, pat), y, idx)
        loss.backward()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    table = profile_step(step, row_limit=70)
    if rank == 0:
        print(table)
    comm.destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
