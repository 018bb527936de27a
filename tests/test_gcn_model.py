"""GCN model: fused (split-weight) formulation == literal reference formulation, single process and distributed."""
import pytest
import torch

from dgraph_b200.models.gcn import CommAwareGCN, GraphConvLayer
from dgraph_b200.parallel import HaloExchange, build_communication_pattern
from dgraph_b200.parallel.emulator import run_world
from dgraph_b200.parallel.kernel_plan import GraphKernelPlan
from tests._graphs import random_graph, partition


def test_state_dict_keys_match_reference():
    m = CommAwareGCN(100, 256, 47, None, None)
    sd = m.state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {
        "conv1.conv.weight": (256, 200), "conv1.conv.bias": (256,),
        "conv2.conv.weight": (256, 512), "conv2.conv.bias": (256,),
        "fc.weight": (47, 256), "fc.bias": (47,),
    }
    assert sum(p.numel() for p in m.parameters()) == 194863


def test_fused_layer_equals_reference_layer():
    torch.manual_seed(0)
    e = random_graph(30, 5, seed=1)
    x = torch.randn(30, 6, dtype=torch.float64, requires_grad=True)
    a = GraphConvLayer(12, 8, fused=True).double()
    b = GraphConvLayer(12, 8, fused=False).double()
    b.load_state_dict(a.state_dict())
    ya = a(x, e, 30)
    yb = b(x, e, 30)
    assert torch.allclose(ya, yb)
    ga = torch.autograd.grad(ya.square().sum(), [x] + list(a.parameters()))
    gb = torch.autograd.grad(yb.square().sum(), [x] + list(b.parameters()))
    for u, v in zip(ga, gb):
        assert torch.allclose(u, v)


def test_kernel_plan_csr_csc():
    e = random_graph(20, 4, seed=2, symmetric=False)
    kp = GraphKernelPlan.from_edges(e, 20, 20)
    for r in range(20):
        cols = kp.col[kp.rowptr[r]:kp.rowptr[r + 1]].tolist()
        assert sorted(cols) == sorted(e[e[:, 0] == r, 1].tolist())
        rows = kp.row[kp.colptr[r]:kp.colptr[r + 1]].tolist()
        assert sorted(rows) == sorted(e[e[:, 1] == r, 0].tolist())
    assert torch.equal(e[kp.csr_eid.long(), 1].int(), kp.col)
    assert torch.equal(kp.csr_eid[kp.csc_to_csr.long()], kp.csc_eid)


def _train_job(comm, e, part, xg, yg, W, state, steps):
    r = comm.get_rank()
    pat = build_communication_pattern(e, part, r, W)
    torch.manual_seed(0)
    model = CommAwareGCN(xg.shape[1], 16, 5, HaloExchange(comm), comm).double()
    model.load_state_dict(state)
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    x, y = xg[pat.local_vertices_global], yg[pat.local_vertices_global]
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        out = model(x, pat)
        loss = torch.nn.functional.cross_entropy(out, y, reduction="sum") / xg.shape[0]
        loss.backward()
        for p in model.parameters():  # data-parallel weight sync (reference: DDP all-reduce; loss already global-normalised)
            comm.all_reduce(p.grad)
        opt.step()
        l = loss.detach().clone()
        comm.all_reduce(l)
        losses.append(float(l))
    return losses, {k: v.clone() for k, v in model.state_dict().items()}


@pytest.mark.parametrize("W", [2, 3])
def test_distributed_training_matches_single_process(W):
    torch.manual_seed(0)
    V = 60
    e = random_graph(V, 5, seed=7)
    xg = torch.randn(V, 7, dtype=torch.float64)
    yg = torch.randint(0, 5, (V,))
    ref = CommAwareGCN(7, 16, 5, None, None).double()
    state = {k: v.clone() for k, v in ref.state_dict().items()}
    single = run_world(1, _train_job, e, torch.zeros(V, dtype=torch.long), xg, yg, 1, state, 4)[0]
    multi = run_world(W, _train_job, e, partition(V, W, "random"), xg, yg, W, state, 4)
    assert single[0][-1] < single[0][0]  # loss decreases
    for losses, sd in multi:
        assert torch.allclose(torch.tensor(losses), torch.tensor(single[0]), atol=1e-9)
        for k in sd:
            assert torch.allclose(sd[k], single[1][k], atol=1e-9)


def test_kernel_plan_degree_guard():
    """The add-ReLU aggregation kernel counts active edges per (row, column) in 16 bits: plans report their largest segment and
    the launch wrappers refuse a hub vertex beyond 65535 neighbours instead of wrapping around silently."""
    import pytest
    import torch
    from dgraph_b200.ops.segment import _check_degree
    from dgraph_b200.parallel.kernel_plan import MAX_COUNTED_DEGREE, GraphKernelPlan

    e = torch.tensor([[0, 1], [0, 2], [0, 3], [1, 0], [2, 0]])
    kp = GraphKernelPlan.from_edges(e, 4, 4)
    assert kp.max_row_degree() == 3
    _check_degree(kp)
    hub = torch.stack([torch.zeros(MAX_COUNTED_DEGREE + 1, dtype=torch.long), torch.arange(1, MAX_COUNTED_DEGREE + 2)], dim=1)
    kp2 = GraphKernelPlan.from_edges(hub, 1, MAX_COUNTED_DEGREE + 2)
    assert kp2.max_row_degree() == MAX_COUNTED_DEGREE + 1
    with pytest.raises(RuntimeError, match="16-bit"):
        _check_degree(kp2)
    assert GraphKernelPlan.from_edges(torch.zeros(0, 2, dtype=torch.long), 0, 0).max_row_degree() == 0


@pytest.mark.parametrize("mode", ["auto", "q"])
def test_exchanging_projected_rows_gives_the_same_training(mode, monkeypatch):
    """A convolution may project only its owned rows and exchange ``Q = x W_j^T`` instead of ``x`` (``GraphConvLayer.
    exchanges_projected_rows``: the default for every layer whose input needs a gradient).  Same losses and weights as the
    reference's order of operations, exchange first (fp64, 3 emulated ranks, random partition)."""
    from dgraph_b200.models import gcn as gcn_mod

    torch.manual_seed(0)
    V, W = 60, 3
    e = random_graph(V, 5, seed=11)
    xg = torch.randn(V, 7, dtype=torch.float64)
    yg = torch.randint(0, 5, (V,))
    state = {k: v.clone() for k, v in CommAwareGCN(7, 16, 5, None, None).double().state_dict().items()}
    part = partition(V, W, "random")
    monkeypatch.setattr(gcn_mod, "_EXCHANGE", "x")
    base = run_world(W, _train_job, e, part, xg, yg, W, state, 3)
    monkeypatch.setattr(gcn_mod, "_EXCHANGE", mode)
    alt = run_world(W, _train_job, e, part, xg, yg, W, state, 3)
    for (l0, sd0), (l1, sd1) in zip(base, alt):
        assert torch.allclose(torch.tensor(l0), torch.tensor(l1), atol=1e-9)
        for k in sd0:
            assert torch.allclose(sd0[k], sd1[k], atol=1e-9)
