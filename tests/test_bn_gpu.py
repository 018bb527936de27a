"""Distributed BatchNorm kernels (``csrc/bn_ops.cu``) against the plain fp32 PyTorch formulation of the same op, one rank."""
import pytest
import torch

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("M,F", [(1000, 64), (4097, 256), (333, 8), (20000, 16)])
@pytest.mark.parametrize("relu", [False, True])
def test_native_batchnorm_matches_fp32(M, F, relu):
    from dgraph_b200.nn.dist_bn import DistributedBatchNorm1D, _native_ok

    torch.manual_seed(0)
    x = (torch.randn(M, F, device="cuda") * 2 + 0.5).bfloat16()
    assert _native_ok(x)
    m = DistributedBatchNorm1D(F, activation="relu" if relu else None).cuda()
    m.gamma.data.uniform_(0.5, 1.5)
    m.beta.data.normal_()
    xg = x.clone().requires_grad_(True)
    y = m(xg)[0]
    assert y.dtype == torch.bfloat16
    go = torch.randn(M, F, device="cuda").bfloat16()
    y.backward(go)
    # fp32 oracle
    xr = x.float().requires_grad_(True)
    g, b = m.gamma.detach().float().clone().requires_grad_(True), m.beta.detach().float().clone().requires_grad_(True)
    mean, var = xr.mean(0), xr.var(0, unbiased=False)
    yr = (xr - mean) * torch.rsqrt(var + m.eps) * g + b
    if relu:
        yr = torch.relu(yr)
    yr.backward(go.float())
    assert torch.allclose(y.float(), yr.detach(), rtol=2e-2, atol=3e-2)
    assert torch.allclose(xg.grad.float(), xr.grad, rtol=3e-2, atol=3e-2 * float(xr.grad.abs().max()))
    assert torch.allclose(m.gamma.grad.float(), g.grad, rtol=2e-2, atol=2e-2 * float(g.grad.abs().max()))
    assert torch.allclose(m.beta.grad.float(), b.grad, rtol=2e-2, atol=2e-2 * float(b.grad.abs().max()))
    assert torch.allclose(m.running_mean.view(-1), 0.1 * x.float().mean(0), atol=2e-3)
