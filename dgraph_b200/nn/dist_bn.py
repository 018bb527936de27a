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

from dgraph_b200.ops import _native as N
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


def _native_ok(x) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.shape[1] % 8 == 0 and 8 <= x.shape[1] <= 1024
            and x.stride(1) == 1 and x.stride(0) % 8 == 0 and x.data_ptr() % 16 == 0 and N.use_native(x))


class _NativeBN(Function):
    """CUDA bf16 path (``csrc/bn_ops.cu``): per direction one statistics kernel pair, ONE all-reduce and one elementwise kernel that
    reads bf16 and writes bf16 -- no fp32 copy of ``[N, F]``, ``x_hat`` is recomputed from ``x`` instead of stored.  ``act=1`` fuses
    the ReLU that follows the layer in RGAT (its backward masks ``dy`` with the saved output)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, comm, eps, act):
        M, F = x.shape
        dev = x.device
        buf = torch.empty(2 * F + 1, dtype=torch.float32, device=dev)
        buf[: 2 * F] = colsum(x, with_sq=True)
        buf[2 * F:].fill_(float(M))
        comm.all_reduce(buf)
        stats = torch.empty((3, F), dtype=torch.float32, device=dev)  # mean, var, rstd
        lib = N.lib()
        N.check(lib.dgb_bn_finalize_stats(N.ptr(buf), N.i32(F), N.f32(eps), N.ptr(stats[0]), N.ptr(stats[1]), N.ptr(stats[2]),
                                          N.stream()), "dgb_bn_finalize_stats")
        g = None if gamma is None else gamma.detach().reshape(-1).float().contiguous()
        b = None if beta is None else beta.detach().reshape(-1).float().contiguous()
        y = torch.empty((M, F), dtype=torch.bfloat16, device=dev)
        N.check(lib.dgb_bn_apply_bf16(N.ptr(x), N.i64(x.stride(0)), N.i64(M), N.i32(F), N.ptr(stats[0]), N.ptr(stats[2]), N.ptr(g),
                                      N.ptr(b), N.i32(act), N.ptr(y), N.i64(F), N.stream()), "dgb_bn_apply_bf16")
        N.count_launch(2)
        ctx.save_for_backward(x, y if act else None, g, stats, buf)
        ctx.comm, ctx.act, ctx.has_affine, ctx.gshape = comm, act, gamma is not None, (None if gamma is None else gamma.shape)
        mean, var = stats[0].view(1, -1), stats[1].view(1, -1)
        ctx.mark_non_differentiable(mean, var)
        return y, mean, var

    @staticmethod
    def backward(ctx, gy, _gm, _gv):
        x, y, g, stats, buf = ctx.saved_tensors
        M, F = x.shape
        dev = x.device
        gy = gy.contiguous()
        if gy.dtype != torch.bfloat16:
            gy = gy.to(torch.bfloat16)
        lib = N.lib()
        lib.dgb_bn_bwd_scratch_bytes.restype = __import__("ctypes").c_int64
        scratch = torch.empty(int(lib.dgb_bn_bwd_scratch_bytes(N.i32(F))) // 4, dtype=torch.float32, device=dev)
        sums = torch.empty(2 * F, dtype=torch.float32, device=dev)
        ld_y = 0 if y is None else y.stride(0)
        N.check(lib.dgb_bn_bwd_stats_bf16(N.ptr(gy), N.i64(F), N.ptr(x), N.i64(x.stride(0)), N.ptr(y), N.i64(ld_y), N.i64(M), N.i32(F),
                                          N.ptr(stats[0]), N.ptr(stats[2]), N.ptr(scratch), N.ptr(sums), N.stream()),
                "dgb_bn_bwd_stats_bf16")
        local = sums.clone() if ctx.has_affine else None  # dgamma / dbeta are summed over the ranks by the optimizer's all-reduce
        ctx.comm.all_reduce(sums)
        dx = torch.empty((M, F), dtype=torch.bfloat16, device=dev)
        N.check(lib.dgb_bn_bwd_dx_bf16(N.ptr(gy), N.i64(F), N.ptr(x), N.i64(x.stride(0)), N.ptr(y), N.i64(ld_y), N.i64(M), N.i32(F),
                                       N.ptr(stats[0]), N.ptr(stats[2]), N.ptr(g), N.ptr(sums), N.ptr(buf[2 * F:]), N.ptr(dx),
                                       N.i64(F), N.stream()), "dgb_bn_bwd_dx_bf16")
        N.count_launch(3)
        dg = db = None
        if ctx.has_affine:
            db, dg = local[:F].view(ctx.gshape), local[F:].view(ctx.gshape)
        return dx, dg, db, None, None, None


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
        # gamma / beta are replicated parameters: every rank returns the partial sum over ITS rows and the optimizer's gradient
        # all-reduce adds the ranks up, like every other weight gradient (the reference hands every rank the global sum and lets
        # DDP average it, distributed_layers.py:62-68); dx needs the global sums
        local = buf.clone()
        ctx.comm.all_reduce(buf)
        dbeta, dgamma = buf[:F], buf[F:]
        gscale = gamma.to(cd).view(1, -1) if gamma is not None else 1.0
        n = ctx.n
        dx = (gscale * rstd) * (gyf - dbeta / n - x_hat * (dgamma / n))
        dg = local[F:].view(1, -1).to(gamma.dtype) if gamma is not None else None
        db = local[:F].view(1, -1).to(gamma.dtype) if gamma is not None else None
        return dx.to(gy.dtype), dg, db, None, None, None


class DistributedBN_with_Recompute(DistributedBN_Impl):
    """Alias kept for API parity (``distributed_layers.py:77-107``); recompute is a flag of the same function here."""


class DistributedBatchNorm1D(nn.Module):
    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, recompute=False,
                 comm=None, activation=None):
        """``activation="relu"``: the ReLU that follows the layer is applied by the same kernel (CUDA bf16 path) / right after the
        normalisation (elsewhere); the module then returns the activated tensor."""
        super().__init__()
        assert activation in (None, "relu")
        self.activation = activation
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
        act = getattr(self, "activation", None)
        if self.training:
            if _native_ok(x):
                y, mean, var = _NativeBN.apply(x, self.gamma, self.beta, comm, self.eps, 1 if act == "relu" else 0)
                act = None
            else:
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
        if act == "relu":
            y = torch.relu(y)
        return y.unsqueeze(0)


def GetGlobalVal(local_val, comm=None, device=None):
    """Global sum of a local scalar across all ranks (``distributed_layers.py:210-214``)."""
    dev = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
    t = torch.tensor([float(local_val)], dtype=torch.float64, device=dev)
    (comm if comm is not None else _NoComm()).all_reduce(t)
    return t.item()
