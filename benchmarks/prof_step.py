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
    dev = torch.device("cuda")
    name = os.environ.get("GRAPH", "ogbn-products")
    g = make_graph(name, device=dev, scale=float(os.environ.get("SCALE", "1.0")))
    comm = Communicator.init_process_group("single")
    pat = build_communication_pattern(g.edge_list, contiguous_partition(g.num_nodes, 1, device=dev), 0, 1)
    del g.edge_list
    F_in = g.num_features
    F_pad = (F_in + 7) // 8 * 8
    x, y = node_features_and_labels(g.num_nodes, F_in, g.num_classes, seed=0, device=dev)
    xd = torch.zeros(g.num_nodes, F_pad, dtype=torch.bfloat16, device=dev)
    xd[:, :F_in] = x.to(torch.bfloat16)
    idx = torch.arange(0, g.num_nodes, 10, device=dev)
    pat.kernel_plan(dev)
    model = CommAwareGCN(F_in, 256, g.num_classes, HaloExchange(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    opt = FusedAdam(model.parameters(), lr=1e-3, comm=comm)
    crit = torch.nn.CrossEntropyLoss()

    def step():
        opt.zero_grad()
        loss = crit(model(xd, pat).index_select(0, idx).float(), y.index_select(0, idx))
        loss.backward()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    print(profile_step(step, row_limit=60))


if __name__ == "__main__":
    main()
