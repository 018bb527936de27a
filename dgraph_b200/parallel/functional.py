"""Autograd functions for distributed gather / scatter-sum over a CommPlan.

Parity: reference ``DGraph/distributed/nccl/_torch_func_impl.py:27-352`` (``CommPlan_GatherFunction`` /
``CommPlan_ScatterFunction``): gather ships vertex rows outward, scatter ships locally pre-reduced edge sums
inward, and each one's backward is the other.  This module is the engine-independent formulation: the
engine only supplies ``all_to_all_rows``.  On B200 the same functions dispatch to the fused peer-memory
kernels through :mod:`dgraph_b200.ops` (one pull / push kernel instead of index + all-to-all + index).
"""
from __future__ import annotations

import torch
from torch.autograd import Function


def _rows(x):
    """Accept ``[N, F]`` or the reference's ``[1, N, F]`` layout; return 2-D view and a restore fn."""
    if x.dim() == 3:
        assert x.shape[0] == 1, "batch dimension must be 1 (NCCLBackendEngine.py:135-136)"
        return x[0], lambda y: y.unsqueeze(0)
    if x.dim() == 1:
        return x.unsqueeze(1), lambda y: y.squeeze(1)
    return x, lambda y: y


def plan_gather_rows(engine, x2d, plan):
    """``out[e] = X[idx[e]]`` with remote rows fetched from their owners."""
    send = x2d.index_select(0, plan.boundary_vertex_idx)
    recv = engine.all_to_all_rows(send, plan.boundary_vertex_splits, plan.boundary_edge_splits)
    aug = torch.cat([x2d, recv], dim=0) if recv.shape[0] else x2d
    return aug.index_select(0, plan.edge_to_row())


def plan_scatter_rows(engine, e2d, plan):
    """``Y[idx[e]] += E[e]``; contributions to remote rows are pre-reduced per unique vertex, then shipped."""
    n = plan.num_local_vertices
    r = int(sum(plan.boundary_edge_splits))
    acc = torch.zeros(n + r, e2d.shape[1], dtype=e2d.dtype, device=e2d.device)
    acc.index_add_(0, plan.edge_to_row(), e2d)
    out = acc[:n]
    recv = engine.all_to_all_rows(acc[n:].contiguous(), plan.boundary_edge_splits, plan.boundary_vertex_splits)
    if recv.shape[0]:
        out = out.index_add(0, plan.boundary_vertex_idx, recv)
    return out


class CommPlan_GatherFunction(Function):
    @staticmethod
    def forward(ctx, x, plan, engine):
        x2d, restore = _rows(x)
        ctx.plan, ctx.engine, ctx.restore_in = plan, engine, x.dim()
        return restore(plan_gather_rows(engine, x2d.contiguous(), plan))

    @staticmethod
    def backward(ctx, g):
        g2d, restore = _rows(g)
        return restore(plan_scatter_rows(ctx.engine, g2d.contiguous(), ctx.plan)), None, None


class CommPlan_ScatterFunction(Function):
    @staticmethod
    def forward(ctx, e, plan, engine):
        e2d, restore = _rows(e)
        ctx.plan, ctx.engine = plan, engine
        return restore(plan_scatter_rows(engine, e2d.contiguous(), plan))

    @staticmethod
    def backward(ctx, g):
        g2d, restore = _rows(g)
        return restore(plan_gather_rows(ctx.engine, g2d.contiguous(), ctx.plan)), None, None
