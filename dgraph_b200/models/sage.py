"""GraphSAGE (mean aggregator) for vertex-partitioned full-graph training -- north-star config 3 (ogbn-papers100M-shaped).

Not in the reference.  Layer: ``h_i = act(W_self x_i + W_nbr mean_{j in N(i)} x_j + b)``; per layer ONE bf16 CSR
mean-aggregation kernel (fp32 accumulate) that reads the rows of remote neighbours from their owners over NVLink while it reduces
(``PeerMemoryEngine.aggregate_pull``; or ONE halo pull first when the graph re-reads halo rows often, ``pull_mode``) and ONE tcgen05
gather-GEMM over the two un-concatenated sources ``[x_local | agg]``."""
from __future__ import annotations

import torch
import torch.nn as nn

from dgraph_b200.ops.gather_linear import gather_linear
from dgraph_b200.ops.segment import segment_mean
from dgraph_b200.parallel.comm_pattern import CommunicationPattern
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan


class SAGEConv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.lin = nn.Linear(2 * in_channels, out_channels)  # columns [W_self | W_nbr]

    def forward(self, x_aug: torch.Tensor, kp: GraphKernelPlan, num_local: int, act=None, agg: torch.Tensor = None):
        if agg is None:
            agg = segment_mean(x_aug, kp)
        x_loc = x_aug[:num_local]
        k = self.lin.in_features // 2
        w = self.lin.weight
        if x_loc.shape[1] > k:  # zero-padded feature columns
            pad = x_loc.shape[1] - k
            w = torch.cat([torch.nn.functional.pad(w[:, :k], (0, pad)), torch.nn.functional.pad(w[:, k:], (0, pad))], 1)
        return gather_linear([x_loc, agg], [None, None], w, self.lin.bias, act=act)


class CommAwareSAGE(nn.Module):
    def __init__(self, in_channels: int, hidden_dims: int, num_classes: int, halo_exchanger, comm, num_layers: int = 2,
                 compute_dtype=None):
        super().__init__()
        self.halo_exchanger, self.comm, self.compute_dtype = halo_exchanger, comm, compute_dtype
        dims = [in_channels] + [hidden_dims] * num_layers
        self.convs = nn.ModuleList([SAGEConv(dims[i], dims[i + 1]) for i in range(num_layers)])
        self.fc = nn.Linear(hidden_dims, num_classes)

    def forward(self, x: torch.Tensor, comm_pattern: CommunicationPattern):
        from dgraph_b200.ops.gemm import linear

        if self.compute_dtype is not None and x.dtype != self.compute_dtype:
            x = x.to(self.compute_dtype)
        n = x.shape[0]
        kp = comm_pattern.kernel_plan(x.device)
        engine = getattr(self.comm, "engine", None)
        # peer-memory engine, several ranks: remote neighbour rows are read inside the aggregation kernel (no halo buffer, no exchange
        # kernel) unless the graph re-reads every halo row so often that one pull into local memory is cheaper (engine.pull_mode)
        fused_pull = (hasattr(engine, "aggregate_pull") and x.is_cuda and x.dtype == torch.bfloat16 and comm_pattern.world_size > 1
                      and (x.shape[1] * 2) % 16 == 0
                      and engine.pull_mode(engine.pattern_info(comm_pattern), kp) == "kernel")
        for conv in self.convs:
            if fused_pull:
                x = conv(x, kp, n, act="relu", agg=engine.aggregate_pull(x, comm_pattern, mean=True))
            else:
                x = conv(self.halo_exchanger.augmented(x, comm_pattern), kp, n, act="relu")
        return linear(x, self.fc.weight.to(x.dtype), self.fc.bias.to(x.dtype))
