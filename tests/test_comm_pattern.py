"""Plan-layer tests (model: reference tests/test_comm_info.py, plus the property tests SURVEY §4 asks for)."""
import pytest
import torch

from dgraph_b200.parallel.comm_pattern import (
    build_communication_pattern, build_all_patterns, compute_local_vertices, compute_halo_vertices,
    compute_boundary_vertices, compute_recv_offsets,
)
from dgraph_b200.parallel.emulator import run_world
from dgraph_b200.parallel import HaloExchange, DGraphMessagePassing
from tests._graphs import random_graph, partition


def _four_vertex():
    # 0-1, 0-2, 1-3, 2-3 (both directions); partition [0,0,1,1]  (SURVEY A.3 / tests/test_comm_info.py:25-52)
    und = [(0, 1), (0, 2), (1, 3), (2, 3)]
    e = torch.tensor(und + [(b, a) for a, b in und])
    return e, torch.tensor([0, 0, 1, 1])


def test_four_vertex_example():
    e, part = _four_vertex()
    p0 = build_communication_pattern(e, part, 0, 2)
    p1 = build_communication_pattern(e, part, 1, 2)
    assert p0.num_local_vertices == 2 and p0.num_halo_vertices == 2
    assert p0.halo_vertices_global.tolist() == [2, 3]
    assert p0.send_local_idx.tolist() == [0, 1]
    assert p0.send_offset.tolist() == [0, 0, 2]
    assert p0.recv_offset.tolist() == [0, 0, 2]
    assert p1.send_offset.tolist() == [0, 2, 2]
    assert p0.comm_map.tolist() == [[0, 2], [2, 0]]
    got = set(map(tuple, p0.local_edge_list.tolist()))
    assert got == {(0, 1), (1, 0), (0, 2), (1, 3)}
    assert p0.put_forward_remote_offset.tolist() == [0, 0]
    assert p1.put_forward_remote_offset.tolist() == [0, 2]


def test_helpers_match_reference_contract():
    e, part = _four_vertex()
    assert compute_local_vertices(part, 1).tolist() == [2, 3]
    assert compute_halo_vertices(e, part, 1).tolist() == [0, 1]
    idx, off = compute_boundary_vertices(e, part, compute_local_vertices(part, 0), 0, 2)
    assert idx.tolist() == [0, 1] and off.tolist() == [0, 0, 2]
    roff, fwd = compute_recv_offsets(torch.tensor([[0, 2], [2, 0]]), 0)
    assert roff.tolist() == [0, 0, 2] and fwd.tolist() == [0, 0]


@pytest.mark.parametrize("W", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("kind", ["contiguous", "round_robin", "random"])
@pytest.mark.parametrize("symmetric", [True, False])
def test_pattern_invariants(W, kind, symmetric):
    V = 97
    e = random_graph(V, 6, seed=W, symmetric=symmetric)
    part = partition(V, W, kind)
    pats = build_all_patterns(e, part, W)
    total_edges = 0
    for r, p in enumerate(pats):
        total_edges += p.num_local_edges
        nl, nh = p.num_local_vertices, p.num_halo_vertices
        assert (p.local_edge_list[:, 0] < nl).all() and (p.local_edge_list[:, 0] >= 0).all()
        assert (p.local_edge_list[:, 1] < nl + nh).all() and (p.local_edge_list[:, 1] >= 0).all()
        assert p.send_offset[0] == 0 and (p.send_offset[1:] >= p.send_offset[:-1]).all()
        assert p.send_offset[r + 1] == p.send_offset[r]  # nothing sent to self
        assert (p.send_local_idx < nl).all()
        assert torch.equal(p.comm_map[r], p.send_offset[1:] - p.send_offset[:-1])
        assert torch.equal(p.comm_map[:, r], p.recv_offset[1:] - p.recv_offset[:-1])
        assert int(p.recv_offset[-1]) == nh
        for d in range(W):  # unique per destination segment
            seg = p.send_local_idx[p.send_offset[d]:p.send_offset[d + 1]]
            assert seg.unique().numel() == seg.numel()
        # the halo rows received from q, in order, are exactly q's send list for r (receive-order numbering)
        for q in range(W):
            want = p.halo_vertices_global[p.recv_offset[q]:p.recv_offset[q + 1]]
            pq = pats[q]
            sent = pq.local_vertices_global[pq.send_local_idx[pq.send_offset[r]:pq.send_offset[r + 1]]]
            assert torch.equal(want, sent)
            assert (p.halo_owner[p.recv_offset[q]:p.recv_offset[q + 1]] == q).all()
            rows = p.halo_remote_row[p.recv_offset[q]:p.recv_offset[q + 1]]
            assert torch.equal(pq.local_vertices_global[rows], want)
        # global edges are recovered
        gl = torch.cat([p.local_vertices_global, p.halo_vertices_global])
        rec = torch.stack([p.local_vertices_global[p.local_edge_list[:, 0]], gl[p.local_edge_list[:, 1]]], 1)
        mine = e[part[e[:, 0]] == r]
        assert torch.equal(torch.unique(rec, dim=0), torch.unique(mine, dim=0))
        assert torch.equal(p.put_forward_remote_offset, p.comm_map[:r, :].sum(0))
        assert torch.equal(p.put_backward_remote_offset, p.comm_map[:, :r].sum(1))
    assert total_edges == e.shape[0]


def _halo_job(comm, e, part, x_global, W):
    r = comm.get_rank()
    pat = build_communication_pattern(e, part, r, W)
    x = x_global[pat.local_vertices_global].clone().requires_grad_(True)
    halo = HaloExchange(comm)(x, pat)
    aug = torch.cat([x, halo], 0)
    src, dst = pat.local_edge_list[:, 0], pat.local_edge_list[:, 1]
    msg = aug[src] * aug[dst].sin()
    out = torch.zeros_like(x).index_add(0, src, msg)
    loss = (out ** 2).sum()
    loss.backward()
    return pat.local_vertices_global, out.detach(), x.grad.detach(), halo.detach(), pat.halo_vertices_global


@pytest.mark.parametrize("W", [1, 2, 3, 5])
@pytest.mark.parametrize("kind", ["round_robin", "random", "contiguous"])
def test_halo_exchange_forward_backward_emulated(W, kind):
    torch.manual_seed(0)
    V, F = 61, 5
    e = random_graph(V, 5, seed=3, symmetric=False)
    part = partition(V, W, kind)
    xg = torch.randn(V, F, dtype=torch.float64)
    res = run_world(W, _halo_job, e, part, xg, W)
    # single-process oracle
    x = xg.clone().requires_grad_(True)
    msg = x[e[:, 0]] * x[e[:, 1]].sin()
    out = torch.zeros_like(x).index_add(0, e[:, 0], msg)
    (out ** 2).sum().backward()
    for gids, o, g, halo, hg in res:
        assert torch.allclose(halo, xg[hg])
        assert torch.allclose(o, out.detach()[gids])
        assert torch.allclose(g, x.grad[gids])


def test_bipartite_pattern():
    # 3 central vertices x 4 neighbour vertices (model: tests/test_comm_info.py:89-109)
    e = torch.tensor([[0, 0], [0, 2], [1, 1], [1, 3], [2, 0], [2, 3]])
    cpart = torch.tensor([0, 1, 1])
    npart = torch.tensor([0, 0, 1, 1])
    p0 = build_communication_pattern(e, cpart, 0, 2, neighbor_partitioning=npart)
    p1 = build_communication_pattern(e, cpart, 1, 2, neighbor_partitioning=npart)
    assert p0.num_local_vertices == 1 and p0.num_neighbor_local_vertices == 2
    assert p0.halo_vertices_global.tolist() == [2]
    assert p1.halo_vertices_global.tolist() == [0, 1]
    assert p0.send_local_idx.tolist() == [0, 1] and p0.send_offset.tolist() == [0, 0, 2]
    assert p1.send_local_idx.tolist() == [0] and p1.send_offset.tolist() == [0, 1, 1]
    assert set(map(tuple, p1.local_edge_list.tolist())) == {(0, 3), (0, 1), (1, 2), (1, 1)}


def test_message_passing_wrapper_emulated():
    V, W = 40, 4
    e = random_graph(V, 4, seed=5)
    part = partition(V, W, "random")
    xg = torch.randn(V, 3, dtype=torch.float64)

    class Layer(torch.nn.Module):
        def forward(self, x, ei, ef):
            n = int(ei[:, 0].max()) + 1 if ei.numel() else 0
            return torch.zeros(self.n, x.shape[1], dtype=x.dtype).index_add(0, ei[:, 0], x[ei[:, 1]])

    def job(comm):
        pat = build_communication_pattern(e, part, comm.get_rank(), W)
        layer = Layer()
        layer.n = pat.num_local_vertices
        mp = DGraphMessagePassing(HaloExchange(comm), layer)
        return pat.local_vertices_global, mp(xg[pat.local_vertices_global], pat)

    want = torch.zeros_like(xg).index_add(0, e[:, 0], xg[e[:, 1]])
    for gids, got in run_world(W, job):
        assert torch.allclose(got, want[gids])
