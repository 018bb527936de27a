"""tcgen05 GEMM / gather-GEMM numerics vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]


def _ref(a, b, bias, act):
    y = a.float() @ b.float().t()
    if bias is not None:
        y = y + bias.float()
    if act == "relu":
        y = torch.relu(y)
    elif act == "silu":
        y = torch.nn.functional.silu(y)
    return y


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (300, 256, 256), (1000, 512, 104), (77, 47, 256), (4096, 128, 128),
                                   (20000, 256, 512), (129, 300, 72), (5, 16, 8)])
@pytest.mark.parametrize("act", [None, "relu"])
def test_gemm_tn_matches_fp32_reference(M, N, K, act):
    from dgraph_b200.ops.gemm_native import gemm_bf16_tn

    torch.manual_seed(M + N + K)
    a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(N, K, device="cuda") * 0.5).bfloat16()
    bias = torch.randn(N, device="cuda")
    y = gemm_bf16_tn(a, b, bias, act)
    yr = _ref(a, b, bias, act)
    assert y.shape == (M, N) and y.dtype == torch.bfloat16
    err = (y.float() - yr).abs().max().item()
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=2e-2 * max(1.0, yr.abs().max().item() / 16)), err


def test_gemm_fp32_out_silu_and_strided_output():
    from dgraph_b200.ops.gemm_native import gemm_bf16_tn

    torch.manual_seed(0)
    a = torch.randn(513, 200, device="cuda").bfloat16()
    b = torch.randn(96, 200, device="cuda").bfloat16()
    y = gemm_bf16_tn(a, b, None, "silu", out_dtype=torch.float32)
    assert torch.allclose(y, _ref(a, b, None, "silu"), rtol=1e-2, atol=5e-2)
    big = torch.zeros(513, 256, device="cuda", dtype=torch.bfloat16)
    gemm_bf16_tn(a, b, None, None, out=big[:, 128:224])
    assert torch.allclose(big[:, 128:224].float(), _ref(a, b, None, None), rtol=2e-2, atol=2e-1)
    assert (big[:, :128] == 0).all() and (big[:, 224:] == 0).all()


def test_linear_autograd_matches_torch():
    from dgraph_b200.ops.gemm import linear

    torch.manual_seed(1)
    x = torch.randn(700, 104, device="cuda").bfloat16().requires_grad_(True)
    w = (torch.randn(256, 104, device="cuda") * 0.1).bfloat16().requires_grad_(True)
    b = torch.randn(256, device="cuda").bfloat16().requires_grad_(True)
    y = linear(x, w, b, "relu")
    xf, wf, bf = [t.detach().float().requires_grad_(True) for t in (x, w, b)]
    yr = torch.relu(torch.nn.functional.linear(xf, wf, bf))
    g = torch.randn_like(yr).bfloat16()
    y.backward(g)
    yr.backward(g.float())
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=5e-2)
    assert torch.allclose(x.grad.float(), xf.grad, rtol=3e-2, atol=5e-2)
    assert torch.allclose(w.grad.float(), wf.grad, rtol=3e-2, atol=0.5)
    assert torch.allclose(b.grad.float(), bf.grad, rtol=3e-2, atol=0.5)


@pytest.mark.parametrize("ks", [[128], [128, 128], [100, 100, 4], [64, 256, 8]])
def test_gather_gemm_matches_cat_linear(ks):
    from dgraph_b200.ops.gemm_native import gather_gemm_bf16, pack_segment_weight

    torch.manual_seed(2)
    M, N, V = 3000, 128, 500
    srcs, idxs = [], []
    for i, k in enumerate(ks):
        kp = (k + 7) // 8 * 8
        if i < 2:
            s = torch.zeros(V, kp, device="cuda", dtype=torch.bfloat16)
            s[:, :k] = torch.randn(V, k, device="cuda")
            srcs.append(s)
            idxs.append(torch.randint(0, V, (M,), device="cuda", dtype=torch.int32))
        else:
            s = torch.zeros(M, kp, device="cuda", dtype=torch.bfloat16)
            s[:, :k] = torch.randn(M, k, device="cuda")
            srcs.append(s)
            idxs.append(None)
    w = (torch.randn(N, sum(ks), device="cuda") * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda")
    wp = pack_segment_weight(w, ks)
    y = gather_gemm_bf16(srcs, idxs, wp, bias, M, [s.shape[1] for s in srcs], act="silu")
    cat = torch.cat([(s[:, :k] if i is None else s[i.long(), :k]) for s, i, k in zip(srcs, idxs, ks)], 1)
    yr = _ref(cat, w, bias, "silu")
    assert torch.allclose(y.float(), yr, rtol=2e-2, atol=5e-2), (y.float() - yr).abs().max().item()


@pytest.mark.parametrize("M,N1,N2", [(1000, 128, 64), (5000, 512, 104), (70000, 256, 256), (333, 47, 8), (100000, 512, 256)])
def test_wgrad_splitk_matches_fp32_reference(M, N1, N2):
    from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w

    torch.manual_seed(M)
    big = torch.randn(M, N1 + 8, device="cuda").bfloat16()
    g = big[:, :N1] if N1 % 8 == 0 else torch.randn(M, N1, device="cuda").bfloat16()
    x = torch.randn(M, N2, device="cuda").bfloat16()
    dw = gemm_bf16_nn_grad_w(g, x)
    ref = g.float().t() @ x.float()
    assert dw.shape == (N1, N2) and dw.dtype == torch.float32
    assert torch.allclose(dw, ref, rtol=1e-2, atol=2e-2 * (M ** 0.5)), (dw - ref).abs().max().item()


@pytest.mark.parametrize("M,N1,N2,V", [(5000, 128, 128, 700), (70001, 128, 104, 3000), (999, 256, 256, 50)])
def test_wgrad_with_gathered_rows_matches_reference(M, N1, N2, V):
    from dgraph_b200.ops.gemm_native import gemm_bf16_nn_grad_w

    torch.manual_seed(M)
    g = torch.randn(M, N1, device="cuda").bfloat16()
    x = torch.randn(V, N2, device="cuda").bfloat16()
    idx = torch.randint(0, V, (M,), device="cuda", dtype=torch.int32)
    dw = gemm_bf16_nn_grad_w(g, x, idx)
    ref = g.float().t() @ x.float()[idx.long()]
    assert torch.allclose(dw, ref, rtol=1e-2, atol=2e-2 * (M ** 0.5)), (dw - ref).abs().max().item()


def test_gather_linear_relu_forward_backward_matches_fp32():
    """The GraphSAGE layer's op: ``relu(Linear(cat([x_local, agg])))`` with the ReLU in the GEMM epilogue and a gradient that needs
    only the output -- against autograd through the fp32 statement."""
    from dgraph_b200.ops.gather_linear import gather_linear, gather_linear_ref

    torch.manual_seed(5)
    M, K, N = 3000, 64, 256
    a = torch.randn(M, K, device="cuda").bfloat16().requires_grad_()
    b = torch.randn(M, K, device="cuda").bfloat16().requires_grad_()
    w = (torch.randn(N, 2 * K, device="cuda") * 0.1).requires_grad_()
    bias = (torch.randn(N, device="cuda") * 0.1).requires_grad_()
    y = gather_linear([a, b], [None, None], w, bias, act="relu")
    go = torch.randn_like(y)
    y.backward(go)
    a32, b32 = a.detach().float().requires_grad_(), b.detach().float().requires_grad_()
    w32, bias32 = w.detach().bfloat16().float().requires_grad_(), bias.detach().clone().requires_grad_()
    yr = gather_linear_ref([a32, b32], [None, None], w32, bias32, act="relu")
    yr.backward(go.float())
    assert (y.float() - yr).norm() <= 1e-2 * yr.norm()
    assert ((y > 0) == (yr > 0)).float().mean() > 0.99
    for got, want in [(a.grad, a32.grad), (b.grad, b32.grad), (w.grad, w32.grad), (bias.grad, bias32.grad)]:
        assert (got.float() - want).norm() <= 3e-2 * want.norm() + 1e-3
