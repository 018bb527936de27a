"""The distributed GraphConv layer with the exchange inside its kernels (``ops/graphconv.py::_GraphConvPullQ``) on ONE GPU.

W ranks are W engines of one process whose symmetric arenas are plain device buffers; barriers are no-ops and every phase is run
rank after rank on one stream, so no kernel ever waits for another one (B200_PROFILING.md rule).  A rank's forward reads what its
peers' GEMMs wrote, therefore every phase is run twice: the first pass fills the slots, the second one is the result.  Oracle: the
layer on the whole graph in one piece (fp32 PyTorch formulation of the same op)."""
import pytest
import torch

from tests._graphs import random_graph, partition

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(180)]


def _make_engines(W, pats, arena_bytes=64 << 20):
    from dgraph_b200.parallel.peer_engine import emulated_engines

    assert len(pats) == W
    return emulated_engines(pats, arena_bytes)


@pytest.mark.parametrize("W,mode,how", [(2, "kernel", "random"), (3, "kernel", "random"), (3, "buffer", "random"),
                                        (4, "buffer", "random"), (8, "kernel", "contiguous")])
@pytest.mark.parametrize("F,f_in,H", [(24, 20, 32), (104, 100, 256)])  # F = padded feature width, f_in = the layer's input width
def test_pull_q_layer_matches_whole_graph(W, mode, how, F, f_in, H, monkeypatch):
    from dgraph_b200.ops.graphconv import graphconv_pull_q
    from dgraph_b200.parallel import build_all_patterns
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan
    from dgraph_b200.ops.segment import graphconv_aggregate_ref

    monkeypatch.setenv("DGRAPH_B200_PULL", mode)
    torch.manual_seed(0)
    V = 500
    e = random_graph(V, 8, seed=3 * W + 1, symmetric=True)
    part = partition(V, W, how)
    pats = build_all_patterns(e, part, W)
    engines = _make_engines(W, pats)
    xg = torch.zeros(V, F, device="cuda")
    xg[:, :f_in] = torch.randn(V, f_in, device="cuda")
    weight = (torch.randn(H, 2 * f_in, device="cuda") * 0.2).requires_grad_(True)
    bias = (torch.randn(H, device="cuda") * 0.1).requires_grad_(True)
    gout = torch.randn(V, H, device="cuda")

    # ---- oracle: the whole graph as one rank, fp32 -------------------------------------------------------------------------------
    xr = xg.clone().bfloat16().float().requires_grad_(True)
    wr = weight.detach().bfloat16().float().requires_grad_(True)
    br = bias.detach().clone().requires_grad_(True)
    kpg = GraphKernelPlan.from_edges(e.cuda(), V, V, device="cuda")
    P = xr[:, :f_in] @ wr[:, :f_in].t() + br
    Q = xr[:, :f_in] @ wr[:, f_in:].t()
    want = graphconv_aggregate_ref(P, Q, kpg)
    want.backward(gout.bfloat16().float())

    xs = [xg[p.local_vertices_global.cuda()].bfloat16().requires_grad_(True) for p in pats]
    kps = [p.kernel_plan("cuda") for p in pats]
    outs = [None] * W
    for _ in range(2):  # pass 1 fills every rank's slot, pass 2 reads complete peers
        for r in range(W):
            outs[r] = graphconv_pull_q(xs[r], weight, bias, kps[r], f_in, engines[r], pats[r], tag=1)
    for r in range(W):
        gids = pats[r].local_vertices_global.cuda()
        assert torch.allclose(outs[r].float(), want[gids].detach(), rtol=3e-2, atol=0.25), f"forward rank {r}"

    dws, dbs = [], []
    for it in range(2):
        for r in range(W):
            gids = pats[r].local_vertices_global.cuda()
            xs[r].grad = None
            dx_, dw_, db_ = torch.autograd.grad(outs[r], [xs[r], weight, bias], gout[gids].bfloat16(), retain_graph=True)
            if it == 1:
                dws.append(dw_)
                dbs.append(db_)
                ref = xr.grad[gids][:, :f_in]
                assert torch.allclose(dx_.float()[:, :f_in], ref, rtol=5e-2, atol=0.05 * float(ref.abs().max()) + 1e-3), f"dx rank {r}"
    dw, db = sum(dws), sum(dbs)
    assert torch.allclose(dw, wr.grad, rtol=5e-2, atol=0.03 * float(wr.grad.abs().max()))
    assert torch.allclose(db, br.grad, rtol=5e-2, atol=0.03 * float(br.grad.abs().max()))


def test_pull_modes_agree_bitwise():
    """In-kernel pull and pull-into-buffer read the same rows in the same order: identical outputs."""
    import os

    from dgraph_b200.ops.graphconv import graphconv_pull_q
    from dgraph_b200.parallel import build_all_patterns

    torch.manual_seed(1)
    W, V, F, H = 3, 400, 64, 64
    e = random_graph(V, 9, seed=11, symmetric=True)
    pats = build_all_patterns(e, partition(V, W, "random"), W)
    xg = torch.randn(V, F, device="cuda")
    weight = torch.randn(H, 2 * F, device="cuda") * 0.2
    bias = torch.randn(H, device="cuda") * 0.1
    res = {}
    for mode in ("kernel", "buffer"):
        os.environ["DGRAPH_B200_PULL"] = mode
        try:
            engines = _make_engines(W, pats)
            xs = [xg[p.local_vertices_global.cuda()].bfloat16() for p in pats]
            for _ in range(2):
                outs = [graphconv_pull_q(xs[r], weight, bias, pats[r].kernel_plan("cuda"), F, engines[r], pats[r], tag=0)
                        for r in range(W)]
            res[mode] = outs
        finally:
            os.environ.pop("DGRAPH_B200_PULL", None)
    for a, b in zip(res["kernel"], res["buffer"]):
        assert torch.equal(a, b)


def test_restricted_reduce_equals_dense_reduce():
    """A gradient that is zero outside the labelled rows: the backward that carries this fact through the exchange (restricted CSC walk,
    halo rows without an edge not written, owners read only the halo copies that can be non-zero) gives bit-identical ``dx`` and the same
    ``dW`` as the dense backward on the same gradient."""
    from dgraph_b200.ops import _sparsity
    from dgraph_b200.ops.graphconv import graphconv_pull_q
    from dgraph_b200.parallel import build_all_patterns

    torch.manual_seed(3)
    W, V, F, H = 4, 600, 64, 64
    e = random_graph(V, 7, seed=5, symmetric=True)
    pats = build_all_patterns(e, partition(V, W, "random"), W)
    engines = _make_engines(W, pats)
    xg = torch.randn(V, F, device="cuda")
    weight = (torch.randn(H, 2 * F, device="cuda") * 0.2).requires_grad_(True)
    bias = (torch.randn(H, device="cuda") * 0.1).requires_grad_(True)
    xs = [xg[p.local_vertices_global.cuda()].bfloat16().requires_grad_(True) for p in pats]
    kps = [p.kernel_plan("cuda") for p in pats]
    idxs = [torch.arange(0, p.num_local_vertices, 7, device="cuda") for p in pats]
    gs = []
    for r in range(W):
        g = torch.zeros(pats[r].num_local_vertices, H, device="cuda", dtype=torch.bfloat16)
        g[idxs[r]] = torch.randn(idxs[r].numel(), H, device="cuda").bfloat16()
        gs.append(g)
    for _ in range(2):
        outs = [graphconv_pull_q(xs[r], weight, bias, kps[r], F, engines[r], pats[r], tag=0) for r in range(W)]
    # the restricted plans are built collectively: emulate the two phases (publish flags everywhere, then gather)
    for it in range(2):
        for r in range(W):
            info = engines[r].pattern_info(pats[r])
            info.__dict__.pop("_restricted_reduce", None)
            engines[r].restricted_reduce_plan(info, kps[r], idxs[r], kps[r].restricted_csc(idxs[r])[0])
    res = {}
    for sparse in (False, True):
        for it in range(2):
            got = []
            for r in range(W):
                if sparse:
                    _sparsity.put(gs[r], idxs[r], gs[r][idxs[r]])
                else:
                    _sparsity.clear()
                got.append(torch.autograd.grad(outs[r], [xs[r], weight, bias], gs[r], retain_graph=True))
        res[sparse] = got
    assert _sparsity.STATS["hit"] >= W
    for r in range(W):
        fewer = engines[r].pattern_info(pats[r])._restricted_reduce
        plan = next(iter(fewer.values()))[1]
        assert plan[2].numel() < engines[r].pattern_info(pats[r]).src_row.numel()  # really restricted
        assert torch.equal(res[True][r][0], res[False][r][0])
        for a, b in zip(res[True][r][1:], res[False][r][1:]):
            assert torch.allclose(a, b, rtol=1e-3, atol=1e-3 * float(b.abs().max()))


def test_packed_weights_kernel_and_cache():
    from dgraph_b200.ops import _wcache
    from dgraph_b200.ops.graphconv import packed_weights

    torch.manual_seed(2)
    H, f_in, Fp = 48, 100, 104
    w = torch.randn(H, 2 * f_in, device="cuda")
    b = torch.randn(H, device="cuda")
    wcat, wcat_t, bcat = packed_weights(w, b, f_in, Fp)
    ref = torch.zeros(2 * H, Fp, device="cuda")
    ref[:H, :f_in] = w[:, :f_in]
    ref[H:, :f_in] = w[:, f_in:]
    assert torch.equal(wcat, ref.bfloat16()) and torch.equal(wcat_t, ref.bfloat16().t().contiguous())
    assert torch.equal(bcat[:H], b) and float(bcat[H:].abs().sum()) == 0.0
    assert packed_weights(w, b, f_in, Fp)[0] is wcat  # cached
    w.mul_(2.0)  # in-place update (stock optimizers): version counter moves
    assert packed_weights(w, b, f_in, Fp)[0] is not wcat
    w2 = packed_weights(w, b, f_in, Fp)[0]
    _wcache.bump()  # native optimizer kernels: explicit invalidation
    assert packed_weights(w, b, f_in, Fp)[0] is not w2


def test_hub_rows_are_split_and_match_the_unsplit_math():
    """A vertex of degree 5,000 (> GraphKernelPlan HUB_DEGREE): its CSR row and CSC column run as 512-slot chunks in a second launch
    whose partial sums / counts are added up; forward and all gradients match the fp32 PyTorch formulation of the layer."""
    from dgraph_b200.ops.graphconv import graphconv_fused
    from dgraph_b200.ops.segment import graphconv_aggregate_ref
    from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

    torch.manual_seed(4)
    V, F, H = 6000, 32, 64
    hub = torch.arange(1, 5001)
    e = torch.cat([torch.stack([torch.zeros_like(hub), hub], 1), torch.stack([hub, torch.zeros_like(hub)], 1),
                   random_graph(V, 6, seed=2, symmetric=True)])
    e = torch.unique(e, dim=0).cuda()
    kp = GraphKernelPlan.from_edges(e, V, V, device="cuda")
    assert kp.hub_plan("csr") is not None and kp.hub_plan("csc") is not None and kp.max_row_degree() >= 5000
    x = torch.randn(V, F, device="cuda").bfloat16().requires_grad_(True)
    w = (torch.randn(H, 2 * F, device="cuda") * 0.1).requires_grad_(True)
    b = (torch.randn(H, device="cuda") * 0.1).requires_grad_(True)
    go = torch.randn(V, H, device="cuda").bfloat16()
    out = graphconv_fused(x, w, b, kp, V, F)
    dx, dw, db = torch.autograd.grad(out, [x, w, b], go)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().bfloat16().float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True)
    ref = graphconv_aggregate_ref(xr @ wr[:, :F].t() + br, xr @ wr[:, F:].t(), kp)
    rx, rw, rb = torch.autograd.grad(ref, [xr, wr, br], go.float())
    assert torch.allclose(out.float(), ref.detach(), rtol=3e-2, atol=0.02 * float(ref.abs().max()))
    assert torch.allclose(dx.float(), rx, rtol=5e-2, atol=0.03 * float(rx.abs().max()))
    assert torch.allclose(dw, rw, rtol=5e-2, atol=0.03 * float(rw.abs().max()))
    assert torch.allclose(db, rb, rtol=5e-2, atol=0.03 * float(rb.abs().max()))
