"""RGAT / GAT layer: fused halo path == reference gather/scatter path == single-process dense oracle; dist BatchNorm."""
import pytest
import torch

from dgraph_b200.models.rgat import CommAwareGAT, CommAwareRGAT
from dgraph_b200.nn.dist_bn import DistributedBatchNorm1D
from dgraph_b200.parallel import build_communication_pattern
from dgraph_b200.parallel.comm_plan import COO_to_NCCLEdgeConditionedCommPlan
from dgraph_b200.parallel.emulator import run_world
from tests._graphs import random_graph


def _oracle_gat(layer, x_dst, x_src, e):
    """Dense single-process GAT on edges e[:,0]=dst, e[:,1]=src."""
    h, hj = layer.conv1(x_dst), layer.conv1(x_src)
    sc = torch.nn.functional.leaky_relu(layer.project_message(torch.cat([h[e[:, 0]], hj[e[:, 1]]], 1)), 0.2).squeeze(1)
    out = torch.zeros_like(h)
    for i in range(h.shape[0]):
        m = e[:, 0] == i
        if m.any():
            a = torch.softmax(sc[m], 0)
            out[i] = (a.unsqueeze(1) * hj[e[m, 1]]).sum(0)
    return out + layer.res_net(x_dst) + layer.bias


def test_state_dict_keys():
    m = CommAwareRGAT(16, 5, 8, 3, 2, 4, comm=None)
    keys = set(m.state_dict().keys())
    for k in ["layers.0.0.bias", "layers.0.0.conv1.weight", "layers.0.0.project_message.weight", "layers.0.0.project_message.bias",
              "layers.0.0.res_net.weight", "bn_layers.1.gamma", "bn_layers.1.running_var", "bn_layers.0.num_batches_tracked",
              "skip_layers.0.weight", "skip_layers.1.bias", "mlp.0.weight", "mlp.4.bias", "mlp.1.gamma"]:
        assert k in keys, k


@pytest.mark.parametrize("W", [1, 3])
def test_gat_paths_match_oracle(W):
    torch.manual_seed(0)
    Vd, Vs, Fin, C = 30, 24, 6, 8
    g = torch.Generator().manual_seed(1)
    e = torch.unique(torch.stack([torch.randint(0, Vd, (150,), generator=g), torch.randint(0, Vs, (150,), generator=g)], 1), dim=0)
    xd = torch.randn(Vd, Fin, dtype=torch.float64)
    xs = torch.randn(Vs, Fin, dtype=torch.float64)
    ref = CommAwareGAT(Fin, C, comm=None, residual=True, hetero=True).double()
    ref.bias.data.normal_()
    want = _oracle_gat(ref, xd, xs, e)
    state = ref.state_dict()
    doff = torch.tensor([Vd * r // W for r in range(W + 1)])
    soff = torch.tensor([Vs * r // W for r in range(W + 1)])
    dpart = torch.bucketize(torch.arange(Vd), doff, right=True) - 1
    spart = torch.bucketize(torch.arange(Vs), soff, right=True) - 1

    def job(comm):
        r = comm.get_rank()
        layer = CommAwareGAT(Fin, C, comm=comm, residual=True, hetero=True).double()
        layer.load_state_dict(state)
        # fused path: edges with their destination
        pat = build_communication_pattern(e, dpart, r, W, neighbor_partitioning=spart)
        a = layer(xd[doff[r]:doff[r + 1]], pat, x_j=xs[soff[r]:soff[r + 1]])
        # reference layout: edges with their source
        mine = torch.nonzero(spart[e[:, 1]] == r, as_tuple=True)[0]
        plan = COO_to_NCCLEdgeConditionedCommPlan(r, W, e[:, 1], e[:, 0], mine, soff, doff, exchange=comm.engine.exchange_ids)
        b = layer(xd[doff[r]:doff[r + 1]].unsqueeze(0), plan, x_j=xs[soff[r]:soff[r + 1]].unsqueeze(0))
        return a, b[0]

    for r, (a, b) in enumerate(run_world(W, job)):
        assert torch.allclose(a, want[doff[r]:doff[r + 1]], atol=1e-9)
        assert torch.allclose(b, want[doff[r]:doff[r + 1]], atol=1e-9)


@pytest.mark.parametrize("W", [1, 2, 4])
def test_distributed_batchnorm_matches_torch(W):
    torch.manual_seed(0)
    N, Fd = 64, 5
    x = torch.randn(N, Fd, dtype=torch.float64) * 3 + 1
    bn = torch.nn.BatchNorm1d(Fd).double()
    bn.weight.data.uniform_(0.5, 1.5)
    bn.bias.data.normal_()
    xr = x.clone().requires_grad_(True)
    yr = bn(xr)
    (yr ** 3).sum().backward()

    def job(comm):
        r = comm.get_rank()
        lo, hi = N * r // W, N * (r + 1) // W
        m = DistributedBatchNorm1D(Fd, comm=comm).double()
        m.gamma.data.copy_(bn.weight.data.view(1, -1))
        m.beta.data.copy_(bn.bias.data.view(1, -1))
        xl = x[lo:hi].clone().requires_grad_(True)
        y = m(xl)
        (y ** 3).sum().backward()
        return lo, hi, y.detach()[0], xl.grad, m.gamma.grad, m.running_mean

    res = run_world(W, job)
    for lo, hi, y, gx, gg, rm in res:
        assert torch.allclose(y, yr.detach()[lo:hi], atol=1e-8)
        assert torch.allclose(gx, xr.grad[lo:hi], atol=1e-8)
        assert torch.allclose(rm.view(-1), 0.1 * x.mean(0), atol=1e-8)
    # gamma is a replicated weight: each rank holds the partial sum over its rows, the optimizer's gradient all-reduce adds them up
    assert torch.allclose(sum(gg.view(-1) for _, _, _, _, gg, _ in res), bn.weight.grad, atol=1e-8)


def test_rgat_model_runs_and_trains_emulated():
    torch.manual_seed(0)
    W = 2
    sizes = [40, 20, 8]
    rels = [(0, 0), (0, 1), (1, 0), (1, 2), (2, 1)]  # (src type, dst type), reference relation list
    g = torch.Generator().manual_seed(3)
    edges = []
    for s, d in rels:
        ee = torch.stack([torch.randint(0, sizes[d], (80,), generator=g), torch.randint(0, sizes[s], (80,), generator=g)], 1)
        edges.append(torch.unique(ee, dim=0))
    feats = [torch.randn(n, 6) for n in sizes]
    y = torch.randint(0, 3, (sizes[0],))
    offs = [torch.tensor([n * r // W for r in range(W + 1)]) for n in sizes]
    parts = [torch.bucketize(torch.arange(n), o, right=True) - 1 for n, o in zip(sizes, offs)]

    def job(comm):
        r = comm.get_rank()
        torch.manual_seed(0)
        model = CommAwareRGAT(6, 3, 8, len(rels), 2, 1, comm, dropout=0.0)
        pats = [build_communication_pattern(e, parts[d], r, W, neighbor_partitioning=parts[s]) for (s, d), e in zip(rels, edges)]
        xs = [f[o[r]:o[r + 1]] for f, o in zip(feats, offs)]
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        losses = []
        for _ in range(4):
            opt.zero_grad()
            out = model(xs, rels, pats)
            loss = torch.nn.functional.cross_entropy(out, y[offs[0][r]:offs[0][r + 1]], reduction="sum") / sizes[0]
            loss.backward()
            for p in model.parameters():
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
                comm.all_reduce(p.grad)
            opt.step()
            l = loss.detach().clone()
            comm.all_reduce(l)
            losses.append(float(l))
        return losses

    res = run_world(W, job)
    single = run_world(1, lambda comm: None)  # smoke: emulator with W=1 works
    assert res[0] == pytest.approx(res[1])
    assert res[0][-1] < res[0][0]
