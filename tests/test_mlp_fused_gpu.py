"""One-kernel GraphCast MLP block (csrc/mlp_fused_tcgen05.cu, the default for the block shapes it supports) vs the fp32 PyTorch
statement of the same block."""
import pytest
import torch

pytestmark = [pytest.mark.gpu]


def _case(M, widths, n_rows, with_idx, with_res, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    srcs = [torch.randn(n, k, device="cuda", generator=g).to(torch.bfloat16) for n, k in zip(n_rows, widths)]
    idxs = [torch.randint(0, n, (M,), device="cuda", generator=g, dtype=torch.int32) if wi else None
            for n, wi in zip(n_rows, with_idx)]
    K = sum(widths)
    w1 = (torch.randn(128, K, device="cuda", generator=g) / K ** 0.5).requires_grad_()
    b1 = (torch.randn(128, device="cuda", generator=g) * 0.1).requires_grad_()
    w2 = (torch.randn(128, 128, device="cuda", generator=g) / 128 ** 0.5).requires_grad_()
    b2 = (torch.randn(128, device="cuda", generator=g) * 0.1).requires_grad_()
    gamma = (1 + 0.1 * torch.randn(128, device="cuda", generator=g)).requires_grad_()
    beta = (0.1 * torch.randn(128, device="cuda", generator=g)).requires_grad_()
    res = torch.randn(M, 128, device="cuda", generator=g).to(torch.bfloat16) if with_res else None
    return srcs, idxs, (w1, b1, w2, b2, gamma, beta), res


@pytest.mark.parametrize("M,widths,n_rows,with_idx,with_res", [
    (1000, [128, 128, 128], [300, 300, 1000], [True, True, False], True),    # edge block: cat[src, dst, e] + e
    (128 * 5 + 37, [128, 128], [128 * 5 + 37] * 2, [False, False], True),    # node block: cat[n, agg] + n
    (4096, [80], [4096], [False], False),                                    # embedder: 73 -> 128 (padded to 80)
    (777, [8], [777], [False], False),                                       # embedder: 3 / 4 -> 128 (padded to 8)
])
def test_fused_mlp_forward_matches_fp32(M, widths, n_rows, with_idx, with_res):
    from dgraph_b200.ops import mlp_fused as mf

    srcs, idxs, (w1, b1, w2, b2, gamma, beta), res = _case(M, widths, n_rows, with_idx, with_res)
    assert mf.supported(srcs, w1, w2, res)
    y = mf.mlp2_ln_fused(srcs, idxs, w1, b1, w2, b2, gamma, beta, 1e-5, res)
    # oracle on the bf16-rounded weights (the kernel's operands), fp32 arithmetic
    ref = mf.mlp2_ln_ref([s.float() for s in srcs], idxs, w1.detach().to(torch.bfloat16).float(), b1.detach(),
                         w2.detach().to(torch.bfloat16).float(), b2.detach(), gamma.detach(), beta.detach(), 1e-5,
                         None if res is None else res.float())
    assert y.shape == ref.shape and torch.isfinite(y.float()).all()
    assert (y.float() - ref).norm() <= 2e-2 * ref.norm()


def test_fused_mlp_backward_matches_unfused_chain():
    """Gradients of the fused block against autograd through the fp32 statement (bf16 tolerances)."""
    from dgraph_b200.ops import mlp_fused as mf

    M = 3000
    srcs, idxs, params, res = _case(M, [128, 128, 128], [500, 500, M], [True, True, False], True, seed=3)
    srcs = [s.requires_grad_() for s in srcs]
    y = mf.mlp2_ln_fused(srcs, idxs, *params, 1e-5, res)
    go = torch.randn_like(y)
    y.backward(go)
    got = [p.grad.float().clone() for p in params] + [s.grad.float().clone() for s in srcs]
    p32 = [p.detach().clone().requires_grad_() for p in params]
    s32 = [s.detach().float().requires_grad_() for s in srcs]
    ref = mf.mlp2_ln_ref(s32, idxs, *p32, 1e-5, res.float())
    ref.backward(go.float())
    want = [p.grad for p in p32] + [s.grad for s in s32]
    for a, b in zip(got, want):
        assert (a - b).norm() <= 5e-2 * b.norm() + 1e-3, ((a - b).norm().item(), b.norm().item())


def test_fused_mlp_rejects_unaligned_outputs():
    from dgraph_b200.ops import _native as Nn
    import ctypes

    x = torch.zeros(256, 128, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(128, 128, dtype=torch.bfloat16, device="cuda")
    f = torch.ones(128, device="cuda")
    out = torch.zeros(256 * 132 + 1, dtype=torch.bfloat16, device="cuda")[1:].view(256, 132)  # 2-byte aligned base
    st = torch.zeros(256, device="cuda")
    bases = (ctypes.c_void_p * 3)(x.data_ptr(), 0, 0)
    lds = (ctypes.c_int64 * 3)(128, 0, 0)
    ips = (ctypes.c_void_p * 3)(0, 0, 0)
    ks = (ctypes.c_int * 3)(128, 0, 0)
    rc = Nn.lib().dgb_mlp2_ln_fused_bf16(Nn.i32(1), bases, lds, ips, ks, Nn.ptr(w), Nn.i64(128), Nn.ptr(f), Nn.ptr(w), Nn.i64(128),
                                         Nn.ptr(f), Nn.ptr(f), Nn.ptr(f), Nn.f32(1e-5), Nn.ptr(None), Nn.i64(0), Nn.ptr(out),
                                         Nn.i64(132), Nn.ptr(out), Nn.i64(132), Nn.ptr(out), Nn.i64(132), Nn.ptr(st), Nn.ptr(st),
                                         Nn.i32(256), Nn.stream())
    assert rc == 801
