"""Distributed (sync) BatchNorm over the vertex dimension of a partitioned graph.

Parity: reference ``experiments/OGB-LSC/distributed_layers.py:22-214`` -- same module name, parameters
(``gamma``/``beta`` ``[1, F]``), buffers (``running_mean``/``running_var``/``num_batches_tracked``), ``[N, F]`` or
``[1, N, F]`` input, ``[1, N, F]`` output, ``recompute`` option and the ``GetGlobalVal`` helper.

Differences: statistics are the mathematically correct ones (the reference divides a mean by the global row count and
its backward uses ``(var+eps)**2`` where ``**-1.5`` belongs); the forward needs ONE all-reduce of a
``[2F+1]`` vector (sum, sum of squares, row count) and the backward ONE of ``[2F]`` instead of 3 + 4 collectives; local
sums come from the deterministic two-phase column-sum kernel (``csrc/reduce_ops.cu``) on CUDA bf16 inputs.
"""
from __future__ import annotations

import torch
from torch import nn
from torch.autograd import Function

from dgraph_b200.ops.reduce import colsum


class _NoComm:
    def all_reduce(self, t, op="sum"):
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(t)
        return t


def _cdt(x):
    return torch.float64 if x.dtype == torch.float64 else torch.float32


def _stats(x, comm):
    F = x.shape[1]
    buf = torch.empty(2 * F + 1, dtype=_cdt(x), device=x.device)
    buf[: 2 * F] = colsum(x, with_sq=True)
    buf[2 * F] = float(x.shape[0])
    comm.all_reduce(buf)
    n = buf[2 * F].clamp(min=1.0)
    mean = buf[:F] / n
    var = (buf[F:2 * F] / n - mean * mean).clamp(min=0.0)
    return mean, var, n


class DistributedBN_Impl(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, comm, eps, recompute):
        mean, var, n = _stats(x, comm)
        rstd = torch.rsqrt(var + eps)
        cd = _cdt(x)
        x_hat = (x.to(cd) - mean) * rstd
        y = x_hat * gamma.to(cd).view(1, -1) + beta.to(cd).view(1, -1) if gamma is not None else x_hat
        ctx.save_for_backward(x if recompute else x_hat.to(x.dtype), gamma, mean, rstd)
        ctx.comm, ctx.n, ctx.recompute = comm, n, recompute
        ctx.mark_non_differentiable(mean, var)
        return y.to(x.dtype), mean.view(1, -1), var.view(1, -1)

    @staticmethod
    def backward(ctx, gy, _gm, _gv):
        saved, gamma, mean, rstd = ctx.saved_tensors
        cd = _cdt(gy)
        x_hat = ((saved.to(cd) - mean) * rstd) if ctx.recompute else saved.to(cd)
        F = gy.shape[1]
        gyf = gy.to(cd)
        buf = torch.empty(2 * F, dtype=cd, device=gy.device)
        buf[:F] = gyf.sum(0)
        buf[F:] = (gyf * x_hat).sum(0)
        ctx.comm.all_reduce(buf)
        dbeta, dgamma = buf[:F], buf[F:]
        gscale = gamma.to(cd).view(1, -1) if gamma is not None else 1.0
        n = ctx.n
        dx = (gscale * rstd) * (gyf - dbeta / n - x_hat * (dgamma / n))
        dg = dgamma.view(1, -1).to(gamma.dtype) if gamma is not None else None
        db = dbeta.view(1, -1).to(gamma.dtype) if gamma is not None else None
        return dx.to(gy.dtype), dg, db, None, None, None


class DistributedBN_with_Recompute(DistributedBN_Impl):
    """Alias kept for API parity (``distributed_layers.py:77-107``); recompute is a flag of the same function here."""


class DistributedBatchNorm1D(nn.Module):
    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, recompute=False,
                 comm=None):
        super().__init__()
        if affine:
            self.gamma = nn.Parameter(torch.ones(1, num_features))
            self.beta = nn.Parameter(torch.zeros(1, num_features))
        else:
            self.register_parameter("gamma", None)
            self.register_parameter("beta", None)
        self.eps, self.momentum, self.track_running_stats, self.recompute = eps, momentum, track_running_stats, recompute
        if track_running_stats:
            self.register_buffer("running_mean", torch.zeros(1, num_features))
            self.register_buffer("running_var", torch.ones(1, num_features))
            self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        else:
            self.running_mean = self.running_var = self.num_batches_tracked = None
        self.comm = comm

    def forward(self, x):
        if x.dim() == 3:
            assert x.size(0) == 1, "only mini-batch size 1 is supported"
            x = x.squeeze(0)
        elif x.dim() != 2:
            raise ValueError(f"Expected 2D or 3D input (got {x.dim()}D input)")
        comm = self.comm if self.comm is not None else _NoComm()
        if self.training:
            y, mean, var = DistributedBN_Impl.apply(x, self.gamma, self.beta, comm, self.eps, self.recompute)
            if self.track_running_stats:
                with torch.no_grad():
                    self.num_batches_tracked += 1
                    self.running_mean.mul_(1 - self.momentum).add_(self.momentum * mean)
                    self.running_var.mul_(1 - self.momentum).add_(self.momentum * var)
        else:
            cd = _cdt(x)
            y = (x.to(cd) - self.running_mean.to(cd)) * torch.rsqrt(self.running_var.to(cd) + self.eps)
            if self.gamma is not None:
                y = y * self.gamma.to(cd) + self.beta.to(cd)
            y = y.to(x.dtype)
        return y.unsqueeze(0)


def GetGlobalVal(local_val, comm=None, device=None):
    """Global sum of a local scalar across all ranks (``distributed_layers.py:210-214``)."""
    dev = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
    t = torch.tensor([float(local_val)], dtype=torch.float64, device=dev)
    (comm if comm is not None else _NoComm()).all_reduce(t)
    return t.item()
