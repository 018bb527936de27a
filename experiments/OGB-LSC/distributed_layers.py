"""Reference file name (``experiments/OGB-LSC/distributed_layers.py:22-214``): see ``dgraph_b200.nn.dist_bn``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.nn.dist_bn import (DistributedBatchNorm1D, DistributedBN_Impl, DistributedBN_with_Recompute,  # noqa: F401,E402
                                    GetGlobalVal)

__all__ = ["DistributedBN_Impl", "DistributedBN_with_Recompute", "DistributedBatchNorm1D", "GetGlobalVal"]


def _compute_bn_forward(input, learned_gamma=None, learned_beta=None, comm=None, eps: float = 1e-5):
    """Functional form with the reference's return tuple ``(output, x_hat, global_mean, global_var, global_num_rows)``
    (``distributed_layers.py:22-41``), on the corrected statistics of ``dgraph_b200.nn.dist_bn`` (one all-reduce of
    ``[sum | sum of squares | rows]``)."""
    import torch

    from dgraph_b200.nn.dist_bn import _NoComm, _stats

    mean, var, n = _stats(input, comm if comm is not None else _NoComm())
    x_hat = (input.to(mean.dtype) - mean) * torch.rsqrt(var + eps)
    output = x_hat
    if learned_gamma is not None and learned_beta is not None:
        output = x_hat * learned_gamma.to(mean.dtype).view(1, -1) + learned_beta.to(mean.dtype).view(1, -1)
    return output.to(input.dtype), x_hat.to(input.dtype), mean, var, n.view(1)


def _compute_bn_backward(grad_output, x, x_hat, mean, var, num_rows, learned_gamma=None, learned_beta=None, comm=None,
                         eps: float = 1e-5):
    """``(dx, global_dgamma, global_dbeta)`` (``distributed_layers.py:44-74``): the exact gradient of the normalisation over
    ALL ranks' rows, one all-reduce of ``[sum g | sum g * x_hat]``."""
    import torch

    from dgraph_b200.nn.dist_bn import _NoComm

    cd = mean.dtype
    g, xh = grad_output.to(cd), x_hat.to(cd)
    F = g.shape[1]
    buf = torch.cat([g.sum(0), (g * xh).sum(0)])
    (comm if comm is not None else _NoComm()).all_reduce(buf)
    dbeta, dgamma = buf[:F], buf[F:]
    scale = learned_gamma.to(cd).view(1, -1) if learned_gamma is not None else 1.0
    n = num_rows.to(cd).reshape(-1)[0]
    dx = scale * torch.rsqrt(var.to(cd).view(1, -1) + eps) * (g - dbeta / n - xh * (dgamma / n))
    if learned_gamma is None or learned_beta is None:
        return dx.to(grad_output.dtype), None, None
    return dx.to(grad_output.dtype), dgamma.view(1, -1), dbeta.view(1, -1)
