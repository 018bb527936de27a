"""OGB full-graph GCN (edge-MLP convolution).

Parity: reference ``experiments/OGB/GCN.py:28-118`` -- ``GraphConvLayer`` and ``CommAwareGCN`` with identical
``state_dict`` keys and shapes (``conv1.conv.{weight[H, 2*in], bias}``, ``conv2.conv.{weight[H, 2H], bias}``,
``fc.{weight[C, H], bias}``), same forward semantics:

    m_ij = ReLU(W [x_i ; x_j] + b),   out_i = sum_j m_ij      (edges owned by the rank of i; x_j may be halo)

B200 formulation (``fused=True``): ``W = [W_i | W_j]`` so ``W [x_i; x_j] + b = P[i] + Q[j]`` with
``P = x_local W_i^T + b`` and ``Q = x_aug W_j^T`` -- two node-level tcgen05 GEMMs (25x fewer FLOPs than the per-edge
GEMM at ogbn-products' average degree) followed by ONE fused gather-add-ReLU-segment-sum kernel; no ``[E, *]``
tensor exists in forward or backward.  ``fused=False`` is the literal reference formulation (oracle).

With several ranks on the peer-memory engine the exchange is part of the layer (``ops/graphconv.py::_GraphConvPullQ``): a layer whose
input needs a gradient (or is at least as wide as the layer) projects its owned rows straight into the symmetric slot and the
aggregation kernel pulls the remote ``Q`` rows over NVLink while it reduces; the first layer (static, narrow input features)
exchanges the features themselves -- without a copy when they live in ``HaloExchange.symmetric_rows``.
"""
from __future__ import annotations

import os
import warnings
from typing import Optional

import torch
import torch.nn as nn

from dgraph_b200.ops.gemm import linear
from dgraph_b200.ops.segment import graphconv_aggregate
from dgraph_b200.parallel.comm_pattern import CommunicationPattern
from dgraph_b200.parallel.kernel_plan import MAX_COUNTED_DEGREE, GraphKernelPlan
from dgraph_b200.utils.timing import TimingReport


# Debug / oracle switch: "auto" (default, rule in GraphConvLayer.exchanges_projected_rows), "x" = always exchange the input
# features (the reference's order of operations), "q" = always exchange the projected rows.
_EXCHANGE = os.environ.get("DGRAPH_B200_GCN_EXCHANGE", "auto")


class GraphConvLayer(nn.Module):
    _warned_hub = False

    def __init__(self, message_dim: int, out_channels: int, fused: bool = True):
        super().__init__()
        self.conv = nn.Linear(message_dim, out_channels)
        self.act = nn.ReLU(inplace=True)
        self.fused = fused

    @staticmethod
    def _wait_halo(x):
        """A buffer from ``HaloExchange.augmented(defer=True)`` may still be receiving its halo rows: every path that does not
        overlap with the transfer itself (``ops/graphconv.py`` does) waits here before the first read."""
        if x.is_cuda:
            from dgraph_b200.ops import _pending

            ev = _pending.take(x)
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)

    def forward(self, x, edge_index, num_local_nodes, edge_features=None, kernel_plan: Optional[GraphKernelPlan] = None):
        if self.fused and edge_features is None:
            return self._forward_fused(x, edge_index, num_local_nodes, kernel_plan)
        self._wait_halo(x)
        return self._forward_edges(x, edge_index, num_local_nodes, edge_features)

    def _forward_edges(self, x, edge_index, num_local_nodes, edge_features):
        """The reference's formulation, edge by edge (``GCN.py:43-65``)."""
        src, dst = edge_index[:, 0], edge_index[:, 1]
        F_in = (self.conv.weight.shape[1] - (0 if edge_features is None else edge_features.shape[1])) // 2
        x = x[:, :F_in]
        x_i, x_j = x[src], x[dst]
        parts = [x_i, x_j] if edge_features is None else [x_i, x_j, edge_features]
        w = self.conv.weight.to(x.dtype)
        b = self.conv.bias.to(x.dtype)
        m = torch.relu(torch.nn.functional.linear(torch.cat(parts, dim=1), w, b))
        out = torch.zeros(num_local_nodes, m.shape[1], dtype=m.dtype, device=m.device)
        return out.index_add_(0, src, m)

    def _forward_fused(self, x, edge_index, num_local_nodes, kp):
        if kp is None:
            kp = GraphKernelPlan.from_edges(edge_index, num_local_nodes, x.shape[0], device=x.device)
        F_in = self.conv.weight.shape[1] // 2
        if x.is_cuda and kp.max_row_degree() > MAX_COUNTED_DEGREE:
            # hub vertex beyond the kernels' 16-bit per-column counters: literal formulation (materialises [E, *] tensors)
            if not GraphConvLayer._warned_hub:
                GraphConvLayer._warned_hub = True
                warnings.warn(f"GraphConvLayer: vertex degree {kp.max_row_degree()} > {MAX_COUNTED_DEGREE}, "
                              "using the unfused edge formulation for this graph")
            self._wait_halo(x)
            return self._forward_edges(x, edge_index, num_local_nodes, None)
        if x.is_cuda and x.dtype == torch.bfloat16 and self.conv.weight.shape[0] % 8 == 0 and x.shape[1] % 8 == 0:
            from dgraph_b200.ops.graphconv import graphconv_fused

            return graphconv_fused(x, self.conv.weight, self.conv.bias, kp, num_local_nodes, F_in)
        self._wait_halo(x)
        W = self.conv.weight.to(x.dtype)
        w_i, w_j = W[:, :F_in], W[:, F_in:2 * F_in]
        pad = x.shape[1] - F_in  # features may arrive zero-padded to a 16-byte row (e.g. 100 -> 104 columns)
        assert pad >= 0, "feature matrix narrower than the layer's input width"
        if pad:
            w_i = torch.nn.functional.pad(w_i, (0, pad))
            w_j = torch.nn.functional.pad(w_j, (0, pad))
        P = linear(x[:num_local_nodes], w_i.contiguous(), self.conv.bias.to(x.dtype))
        Q = linear(x, w_j.contiguous(), None)
        return graphconv_aggregate(P, Q, kp)


    def exchanges_projected_rows(self, x_local) -> bool:
        """Which rows travel.  Exchanging ``Q = x W_j^T`` (``H`` wide) instead of ``x`` removes every halo-row GEMM (forward ``Q``,
        backward ``dX`` and half of ``dW``) but needs the transposed exchange of ``dQ`` in backward; exchanging ``x`` of a layer
        whose input needs no gradient (the first layer) has no backward exchange at all.  So: projected rows when the input needs
        a gradient or is at least as wide as the layer, features otherwise."""
        if _EXCHANGE in ("x", "q"):
            return _EXCHANGE == "q"
        return bool(x_local.requires_grad) or x_local.shape[1] >= self.conv.weight.shape[0]

    def forward_exchange_q(self, x_local, comm_pattern, exchanger, kp: Optional[GraphKernelPlan] = None, tag=0):
        """Same layer with the exchange moved behind the projection: every rank projects only the rows it owns, ``Q = x W_j^T``
        travels instead of ``x``.  On the peer-memory engine this is ONE autograd node whose kernels do the exchange themselves
        (``ops/graphconv.py::graphconv_pull_q``); on the other engines (CPU / gloo / NCCL all-to-all) it is the same algebra through
        ``exchanger.augmented``, which also serves as the oracle of the fused node."""
        n = x_local.shape[0]
        if kp is None:
            kp = comm_pattern.kernel_plan(x_local.device)
        F_in = self.conv.weight.shape[1] // 2
        engine = exchanger._fused_engine(x_local) if hasattr(exchanger, "_fused_engine") else None
        if (engine is not None and engine.get_world_size() > 1 and self.conv.weight.shape[0] % 8 == 0
                and kp.max_row_degree() <= MAX_COUNTED_DEGREE):
            from dgraph_b200.ops.graphconv import graphconv_pull_q

            return graphconv_pull_q(x_local, self.conv.weight, self.conv.bias, kp, F_in, engine, comm_pattern, tag)
        W = self.conv.weight.to(x_local.dtype)
        w_i, w_j = W[:, :F_in], W[:, F_in:2 * F_in]
        pad = x_local.shape[1] - F_in
        assert pad >= 0, "feature matrix narrower than the layer's input width"
        if pad:
            w_i = torch.nn.functional.pad(w_i, (0, pad))
            w_j = torch.nn.functional.pad(w_j, (0, pad))
        P = linear(x_local, w_i.contiguous(), self.conv.bias.to(x_local.dtype))
        Q = linear(x_local, w_j.contiguous(), None)
        Q_aug = exchanger.augmented(Q, comm_pattern)
        assert P.shape[0] == n
        return graphconv_aggregate(P, Q_aug, kp)


class CommAwareGCN(nn.Module):
    """2 x (halo exchange -> GraphConvLayer) -> Linear  (``GCN.py:70-118``)."""

    def __init__(self, in_channels: int, hidden_dims: int, num_classes: int, halo_exchanger, comm,
                 fused: bool = True, compute_dtype: Optional[torch.dtype] = None):
        super().__init__()
        self.halo_exchanger = halo_exchanger
        self.conv1 = GraphConvLayer(2 * in_channels, hidden_dims, fused=fused)
        self.conv2 = GraphConvLayer(2 * hidden_dims, hidden_dims, fused=fused)
        self.fc = nn.Linear(hidden_dims, num_classes)
        self.softmax = nn.Softmax(dim=1)
        self.comm = comm
        self.compute_dtype = compute_dtype

    def forward(self, local_node_features: torch.Tensor, comm_pattern: CommunicationPattern):
        n = local_node_features.shape[0]
        x = local_node_features
        if self.compute_dtype is not None and x.dtype != self.compute_dtype:
            x = x.to(self.compute_dtype)
        kp = comm_pattern.kernel_plan(x.device) if self.conv1.fused else None
        edges = comm_pattern.local_edge_list
        defer = self.conv1.fused  # the fused layer overlaps its local work with the in-flight halo rows (ops/_pending.py)
        multi = comm_pattern.world_size > 1
        for li, conv in enumerate((self.conv1, self.conv2), start=1):
            if conv.fused and multi and conv.exchanges_projected_rows(x):
                with TimingReport(f"process-{li}"):  # the exchange happens inside the layer's kernels
                    x = conv.forward_exchange_q(x, comm_pattern, self.halo_exchanger, kp, tag=li)
            else:
                with TimingReport(f"feature-exchange-{li}"):
                    aug = self.halo_exchanger.augmented(x, comm_pattern, defer=defer)
                with TimingReport(f"process-{li}"):
                    x = conv(aug, edges, n, kernel_plan=kp)
        with TimingReport("final-fc"):
            x = linear(x, self.fc.weight.to(x.dtype), self.fc.bias.to(x.dtype))
        return x
