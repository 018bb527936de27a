"""sm_100a segment kernels vs. plain PyTorch fp32 formulations of the same ops."""
import pytest
import torch

from tests._graphs import random_graph

pytestmark = pytest.mark.gpu


def _kp(V, deg, seed, n_aug=None, symmetric=False):
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

    e = random_graph(V, deg, seed=seed, symmetric=symmetric)
    if n_aug is not None:  # send some neighbour ids into a fake halo range
        g = torch.Generator().manual_seed(seed)
        halo = torch.randint(V, n_aug, (e.shape[0],), generator=g)
        pick = torch.rand(e.shape[0], generator=g) < 0.3
        e = torch.stack([e[:, 0], torch.where(pick, halo, e[:, 1])], 1)
    return GraphKernelPlan.from_edges(e, V, n_aug or V, device="cuda")


@pytest.mark.parametrize("H", [8, 32, 64, 104, 128, 256, 512])
@pytest.mark.parametrize("mean", [False, True])
def test_segment_sum_matches_torch(H, mean):
    from dgraph_b200.ops.segment import segment_sum, segment_sum_ref

    torch.manual_seed(0)
    V, n_aug = 777, 1000
    kp = _kp(V, 9, 1, n_aug)
    x = torch.randn(n_aug, H, device="cuda").bfloat16().requires_grad_(True)
    y = segment_sum(x, kp, mean=mean)
    assert y.dtype == torch.bfloat16 and y.shape == (V, H)
    xf = x.detach().float().requires_grad_(True)
    yr = segment_sum_ref(xf, kp, mean=mean)
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=2e-2)
    g = torch.randn_like(yr)
    y.backward(g.bfloat16())
    yr.backward(g.bfloat16().float())
    assert torch.allclose(x.grad.float(), xf.grad, rtol=2e-2, atol=3e-2)


@pytest.mark.parametrize("H", [16, 64, 128, 256, 512])
def test_graphconv_aggregate_matches_torch(H):
    from dgraph_b200.ops.segment import graphconv_aggregate, graphconv_aggregate_ref

    torch.manual_seed(1)
    V, n_aug = 513, 700
    kp = _kp(V, 12, 2, n_aug)
    P = torch.randn(V, H, device="cuda").bfloat16().requires_grad_(True)
    Q = torch.randn(n_aug, H, device="cuda").bfloat16().requires_grad_(True)
    y = graphconv_aggregate(P, Q, kp)
    Pf, Qf = P.detach().float().requires_grad_(True), Q.detach().float().requires_grad_(True)
    yr = graphconv_aggregate_ref(Pf, Qf, kp)
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=5e-2)
    g = torch.randn_like(yr).bfloat16()
    y.backward(g)
    yr.backward(g.float())
    assert torch.allclose(P.grad.float(), Pf.grad, rtol=2e-2, atol=5e-2)
    assert torch.allclose(Q.grad.float(), Qf.grad, rtol=2e-2, atol=5e-2)


def test_strided_operands_and_empty_rows():
    from dgraph_b200.ops.segment import graphconv_aggregate, graphconv_aggregate_ref
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

    e = torch.tensor([[0, 1], [0, 2], [3, 0]])  # rows 1, 2 have no edges
    kp = GraphKernelPlan.from_edges(e, 4, 4, device="cuda")
    PQ = torch.randn(4, 128, device="cuda").bfloat16()
    P, Q = PQ[:, :64], PQ[:, 64:]  # row stride 128, column offset
    y = graphconv_aggregate(P, Q, kp)
    yr = graphconv_aggregate_ref(P.float(), Q.float(), kp)
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=2e-2)
    assert (y[1] == 0).all() and (y[2] == 0).all()


def test_native_library_is_loaded():
    from dgraph_b200.ops import _native

    assert _native.available()
    assert _native.lib().dgb_version() >= 100
    assert _native.lib().dgb_sm_count() == torch.cuda.get_device_properties(0).multi_processor_count
