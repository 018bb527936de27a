"""Segment-softmax-aggregate kernel vs the PyTorch fp32 formulation of the same op (forward and backward)."""
import pytest
import torch

from tests._graphs import random_graph

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]


@pytest.mark.parametrize("C", [8, 64, 128, 256, 512])
def test_seg_softmax_aggregate_matches_torch(C):
    from dgraph_b200.ops.softmax import seg_softmax_aggregate, seg_softmax_aggregate_ref
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

    torch.manual_seed(C)
    n_dst, n_src = 333, 400
    g = torch.Generator().manual_seed(1)
    e = torch.unique(torch.stack([torch.randint(0, n_dst - 5, (4000,), generator=g),  # last rows have no edges
                                  torch.randint(0, n_src, (4000,), generator=g)], 1), dim=0)
    kp = GraphKernelPlan.from_edges(e, n_dst, n_src, device="cuda", num_nbr_local=n_src)
    s_i = (torch.randn(n_dst, device="cuda") * 3).requires_grad_(True)
    s_j = (torch.randn(n_src, device="cuda") * 3).requires_grad_(True)
    h = torch.randn(n_src, C, device="cuda").bfloat16().requires_grad_(True)
    out = seg_softmax_aggregate(s_i, s_j, h, kp, 0.2)
    si2, sj2, h2 = s_i.detach().clone().requires_grad_(True), s_j.detach().clone().requires_grad_(True), h.detach().float().requires_grad_(True)
    ref = seg_softmax_aggregate_ref(si2, sj2, h2, kp, 0.2)
    assert torch.allclose(out.float(), ref, rtol=2e-2, atol=2e-2)
    assert (out[-5:] == 0).all()
    go = torch.randn_like(ref).bfloat16()
    out.backward(go)
    ref.backward(go.float())
    assert torch.allclose(h.grad.float(), h2.grad, rtol=2e-2, atol=3e-2)
    assert torch.allclose(s_i.grad, si2.grad, rtol=3e-2, atol=5e-2 * (C ** 0.5) / 8 + 5e-2)
    assert torch.allclose(s_j.grad, sj2.grad, rtol=3e-2, atol=5e-2 * (C ** 0.5) / 8 + 5e-2)
