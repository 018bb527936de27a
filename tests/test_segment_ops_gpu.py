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


@pytest.mark.parametrize("M,N", [(1, 8), (777, 256), (5000, 104), (100000, 512), (33, 2048)])
@pytest.mark.parametrize("with_sq", [False, True])
def test_colsum_matches_torch(M, N, with_sq):
    from dgraph_b200.ops.reduce import colsum

    torch.manual_seed(0)
    big = torch.randn(M, N + 8, device="cuda").bfloat16()
    x = big[:, :N]  # strided view
    got = colsum(x, with_sq)
    xf = x.float()
    want = torch.cat([xf.sum(0), (xf * xf).sum(0)]) if with_sq else xf.sum(0)
    assert torch.allclose(got, want, rtol=1e-3, atol=1e-2 * max(1.0, M ** 0.5))


def test_fused_graphconv_layer_matches_reference_formulation():
    from dgraph_b200.models.gcn import GraphConvLayer

    torch.manual_seed(3)
    V, n_aug, F, H = 400, 520, 100, 64
    kp = _kp(V, 10, 5, n_aug)
    e = torch.stack([torch.repeat_interleave(torch.arange(V, device="cuda"), (kp.rowptr[1:] - kp.rowptr[:-1]).long()),
                     kp.col.long()], 1)
    fused = GraphConvLayer(2 * F, H, fused=True).cuda()
    ref = GraphConvLayer(2 * F, H, fused=False).cuda()
    ref.load_state_dict(fused.state_dict())
    x = torch.zeros(n_aug, 104, device="cuda")
    x[:, :F] = torch.randn(n_aug, F, device="cuda") * 0.5
    xb = x.bfloat16().requires_grad_(True)
    xf = x.bfloat16().float().requires_grad_(True)
    y = fused(xb, e, V, kernel_plan=kp)
    yr = ref(xf, e, V)
    assert torch.allclose(y.float(), yr, rtol=3e-2, atol=0.15), (y.float() - yr).abs().max()
    g = torch.randn_like(yr).bfloat16()
    y.backward(g)
    yr.backward(g.float())
    # P/Q are rounded to bf16 before the ReLU, so a few near-zero pre-activations flip: compare in the Frobenius norm
    def rel(a, b):
        return float((a - b).norm() / b.norm())

    assert rel(xb.grad.float()[:, :F], xf.grad[:, :F]) < 3e-2
    assert rel(fused.conv.weight.grad, ref.conv.weight.grad) < 3e-2
    assert rel(fused.conv.bias.grad, ref.conv.bias.grad) < 3e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_rank_local_ops_match_python_loops(dtype):  # model: reference tests/test_local_kernels.py:19-154
    from dgraph_b200.parallel import rank_local as R

    torch.manual_seed(0)
    src = torch.randn(1, 8, 8, device="cuda").to(dtype)
    si = torch.tensor([0, 3, 3, 7, 1], device="cuda")
    di = torch.tensor([2, 2, 5, 0, 2], device="cuda")
    out = torch.zeros(1, 6, 8, device="cuda", dtype=dtype)
    R.OptimizedLocalScatterSumGather(src, si, di, out)
    want = torch.zeros(6, 8, device="cuda")
    for s, d in zip(si.tolist(), di.tolist()):
        want[d] += src[0, s].float()
    assert torch.allclose(out[0].float(), want, rtol=2e-2, atol=2e-2)
    out2 = torch.zeros(1, 6, 8, device="cuda", dtype=dtype)
    R.OptimizedLocalScatterGather(src, torch.tensor([4, 6], device="cuda"), torch.tensor([1, 3], device="cuda"), out2)
    assert torch.equal(out2[0, 1], src[0, 4]) and torch.equal(out2[0, 3], src[0, 6]) and (out2[0, 0] == 0).all()
    rm = torch.tensor([0, 1, 0, 1, 0], device="cuda")
    out3 = torch.full((1, 5, 8), -1.0, device="cuda", dtype=dtype)
    R.OptimizedRankLocalMaskedGather(src, si, rm, out3, 0)
    for i in range(5):
        assert torch.equal(out3[0, i], src[0, si[i]]) if rm[i] == 0 else (out3[0, i] == -1).all()
    neg = -src.abs()  # masked scatter must not apply a ReLU (reference Q9)
    out4 = torch.zeros(1, 8, 8, device="cuda", dtype=dtype)
    R.RankLocalMaskedScatter(neg[:, :5], out4, di, rm, 0)
    want4 = torch.zeros(8, 8, device="cuda")
    for i in range(5):
        if rm[i] == 0:
            want4[di[i]] += neg[0, i].float()
    assert torch.allclose(out4[0].float(), want4, rtol=2e-2, atol=2e-2) and (out4.float() <= 0).all()


@pytest.mark.parametrize("C", [8, 128, 256, 512])
@pytest.mark.parametrize("with_res", [False, True])
def test_fused_layernorm_residual_matches_torch(C, with_res):
    from dgraph_b200.ops.norm import layer_norm_residual, silu

    torch.manual_seed(C)
    M = 3001
    z = (torch.randn(M, C, device="cuda") * 2 + 0.5).bfloat16().requires_grad_(True)
    res = torch.randn(M, C, device="cuda").bfloat16().requires_grad_(True) if with_res else None
    gamma = (torch.rand(C, device="cuda") + 0.5).requires_grad_(True)
    beta = torch.randn(C, device="cuda").requires_grad_(True)
    y = layer_norm_residual(z, gamma, beta, 1e-5, res)
    zf, gf, bf = z.detach().float().requires_grad_(True), gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    rf = res.detach().float().requires_grad_(True) if with_res else None
    yr = torch.nn.functional.layer_norm(zf, (C,), gf, bf, 1e-5) + (rf if with_res else 0)
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=3e-2)
    go = torch.randn_like(yr).bfloat16()
    y.backward(go)
    yr.backward(go.float())
    assert torch.allclose(z.grad.float(), zf.grad, rtol=3e-2, atol=3e-2)
    assert torch.allclose(gamma.grad, gf.grad, rtol=2e-2, atol=0.5)
    assert torch.allclose(beta.grad, bf.grad, rtol=2e-2, atol=0.5)
    if with_res:
        assert torch.equal(res.grad, go)
    p = torch.randn(M, C, device="cuda").bfloat16().requires_grad_(True)
    pf = p.detach().float().requires_grad_(True)
    silu(p).backward(go)
    torch.nn.functional.silu(pf).backward(go.float())
    assert torch.allclose(p.grad.float(), pf.grad, rtol=2e-2, atol=2e-2)
