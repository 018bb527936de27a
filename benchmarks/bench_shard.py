"""Compute-only cost of ONE rank's shard of a W-rank run, on one GPU: the halo exchange is replaced by a local stand-in
(halo rows = copies of local rows), everything else (GEMMs over [local | halo] rows, segment kernels, Adam) is the real
per-rank work.  Separates per-rank compute from communication / synchronisation when reading multi-GPU numbers."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.communicator import Communicator  # noqa: E402
from dgraph_b200.data import contiguous_partition, make_graph, node_features_and_labels  # noqa: E402
from dgraph_b200.models.gcn import CommAwareGCN  # noqa: E402
from dgraph_b200.optim import FusedAdam  # noqa: E402
from dgraph_b200.parallel import build_communication_pattern  # noqa: E402
from dgraph_b200.parallel.local_engine import LocalEngine  # noqa: E402


class FakeHalo:
    def __init__(self, comm):
        self.comm = comm

    def augmented(self, x, pat):
        idx = torch.arange(pat.num_halo_vertices, device=x.device) % x.shape[0]
        return torch.cat([x, x.index_select(0, idx)], 0)


def main():
    W = int(os.environ.get("W", "8"))
    rank = int(os.environ.get("R", "0"))
    steps = int(os.environ.get("STEPS", "10"))
    dev = torch.device("cuda")
    g = make_graph("ogbn-products", device=dev)
    E = g.edge_list.shape[0]
    pat = build_communication_pattern(g.edge_list, contiguous_partition(g.num_nodes, W, dev), rank, W)
    del g.edge_list
    x, y = node_features_and_labels(g.num_nodes, g.num_features, g.num_classes, device=dev, rows=pat.local_vertices_global)
    xb = torch.nn.functional.pad(x, (0, 4)).bfloat16()
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    model = CommAwareGCN(g.num_features, 256, g.num_classes, FakeHalo(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-3)
    pat.kernel_plan(dev)

    def step():
        opt.zero_grad()
        out = model(xb, pat)
        loss = torch.nn.functional.cross_entropy(out[::10].float(), y[::10])
        loss.backward()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({"world": W, "rank": rank, "local_rows": pat.num_local_vertices, "halo_rows": pat.num_halo_vertices,
                      "local_edges": pat.num_local_edges, "compute_ms_per_step": e0.elapsed_time(e1) / steps,
                      "halo_bytes_fwd": pat.num_halo_vertices * (208 + 512)}))


if __name__ == "__main__":
    main()
