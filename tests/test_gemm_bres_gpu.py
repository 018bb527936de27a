"""B-resident tcgen05 GEMM (csrc/gemm_bres_tcgen05.cu; the default for K <= 128, see ops/gemm_native.py) vs the fp32 PyTorch
reference, called through the C ABI so that no dispatch rule can route the test around the kernel."""
import pytest
import torch

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("M,N,K", [(1000, 256, 104), (4096 + 37, 512, 104), (3000, 128, 128), (777, 256, 64), (5000, 200, 72)])
@pytest.mark.parametrize("act", [None, "relu", "silu"])
def test_bres_gemm_matches_fp32(M, N, K, act):
    from dgraph_b200.ops import _native as Nn
    from dgraph_b200.ops.gemm import linear_ref
    from dgraph_b200.ops.gemm_native import _ACT, _alloc_out

    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    Kp = (K + 7) // 8 * 8
    ap = torch.zeros(M, Kp, dtype=torch.bfloat16, device="cuda")
    ap[:, :K] = a
    wp = torch.zeros(N, Kp, dtype=torch.bfloat16, device="cuda")
    wp[:, :K] = w
    out = _alloc_out(M, N, torch.bfloat16, a.device)
    rc = Nn.lib().dgb_gemm_bf16_tn_bres(Nn.ptr(ap), Nn.i64(Kp), Nn.ptr(wp), Nn.i64(Kp), Nn.ptr(out), Nn.i64(out.stride(0)),
                                        Nn.i32(M), Nn.i32(N), Nn.i32(Kp), Nn.ptr(bias), Nn.i32(_ACT[act]), Nn.stream())
    assert rc == 0, rc
    ref = linear_ref(a.float(), w.float(), bias, act)
    assert (out.float() - ref).norm() <= 1e-2 * ref.norm()


def test_bres_reports_unsupported_shapes():
    from dgraph_b200.ops import _native as Nn

    a = torch.zeros(256, 1024, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(256, 1024, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(256, 256, dtype=torch.bfloat16, device="cuda")
    rc = Nn.lib().dgb_gemm_bf16_tn_bres(Nn.ptr(a), Nn.i64(1024), Nn.ptr(w), Nn.i64(1024), Nn.ptr(out), Nn.i64(256), Nn.i32(256),
                                        Nn.i32(256), Nn.i32(1024), Nn.ptr(None), Nn.i32(0), Nn.stream())
    assert rc == 801  # 16 k-blocks of B do not fit next to four A stages
