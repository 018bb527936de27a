"""CommPlan tests (model: reference tests/test_NCCLCommPlan.py:127-359), run on the in-process world emulator."""
import pytest
import torch

from dgraph_b200.parallel.comm_plan import COO_to_NCCLCommPlan, COO_to_NCCLEdgeConditionedCommPlan
from dgraph_b200.parallel.emulator import run_world
from tests._graphs import random_graph


def _offsets(V, W):
    return torch.tensor([V * r // W for r in range(W + 1)])


def _job(comm, e, off, xg, eg, W):
    r = comm.get_rank()
    src, dst = e[:, 0], e[:, 1]
    mine = torch.nonzero((src >= off[r]) & (src < off[r + 1]), as_tuple=True)[0]
    plan = COO_to_NCCLEdgeConditionedCommPlan(r, W, src, dst, mine, off, None, exchange=comm.engine.exchange_ids)
    sp, dp = plan.source_graph_plan, plan.dest_graph_plan
    assert sp.boundary_edge_idx.numel() == 0  # edges live with their source (tests/test_NCCLCommPlan.py:235-239)
    assert sum(dp.boundary_edge_splits) == dp.boundary_edge_buffer_map.unique().numel()
    x = xg[off[r]:off[r + 1]].clone().unsqueeze(0).requires_grad_(True)
    got = comm.gather(x, comm_plan=dp)  # [1, E_loc, F]
    (got * got).sum().backward()
    ef = eg[mine].clone().unsqueeze(0).requires_grad_(True)
    sc = comm.scatter(ef, comm_plan=dp)  # [1, N_loc, F]
    (sc * sc).sum().backward()
    return mine, got.detach()[0], x.grad[0], sc.detach()[0], ef.grad[0], plan.source_graph_plan.memory_usage()


@pytest.mark.parametrize("W", [1, 2, 3, 4])
def test_comm_plan_gather_scatter(W):
    torch.manual_seed(1)
    V = 32 * W
    e = random_graph(V, 6, seed=W)
    off = _offsets(V, W)
    xg = torch.randn(V, 4, dtype=torch.float64)
    eg = torch.randn(e.shape[0], 4, dtype=torch.float64)
    res = run_world(W, _job, e, off, xg, eg, W)
    x = xg.clone().requires_grad_(True)
    g = x[e[:, 1]]
    (g * g).sum().backward()
    ef = eg.clone().requires_grad_(True)
    s = torch.zeros_like(xg).index_add(0, e[:, 1], ef)
    (s * s).sum().backward()
    for r, (mine, got, gx, sc, gef, mem) in enumerate(res):
        assert torch.allclose(got, g.detach()[mine])
        assert torch.allclose(gx, x.grad[off[r]:off[r + 1]])
        assert torch.allclose(sc, s.detach()[off[r]:off[r + 1]])
        assert torch.allclose(gef, ef.grad[mine])
        assert mem["unit"] == "MB" and mem["total"] > 0


def test_legacy_gather_scatter_api():
    """``gather(x, indices, rank_mappings)`` / ``scatter(x, indices, rank_mappings, output_size)`` (broken in the
    reference snapshot) -- equal contiguous partitions, explicit ``[2, E]`` rank map."""
    W, n = 3, 8
    V = W * n
    e = random_graph(V, 4, seed=9)
    xg = torch.randn(V, 3, dtype=torch.float64)
    edge_rank = e[:, 0] // n
    rm = torch.stack([edge_rank, e[:, 1] // n])

    def job(comm):
        r = comm.get_rank()
        x = xg[r * n:(r + 1) * n].unsqueeze(0)
        got = comm.gather(x, e[:, 1].unsqueeze(0), rm)
        mine = edge_rank == r
        back = comm.scatter(got, e[:, 1].unsqueeze(0), rm, n)
        return mine, got[0], back[0]

    want_s = torch.zeros_like(xg).index_add(0, e[:, 1], xg[e[:, 1]])
    for r, (mine, got, back) in enumerate(run_world(W, job)):
        assert torch.allclose(got, xg[e[mine, 1]])
        assert torch.allclose(back, want_s[r * n:(r + 1) * n])


def test_plan_to_and_reverse():
    e = random_graph(16, 4, seed=2)
    off = _offsets(16, 1)
    p = COO_to_NCCLEdgeConditionedCommPlan(0, 1, e[:, 0], e[:, 1], torch.arange(e.shape[0]), off, None)
    p = p.to("cpu")
    q = p.reverse()
    assert q.source_graph_plan is p.dest_graph_plan
    assert p.source_graph_plan.num_local_edges == e.shape[0]
    single = COO_to_NCCLCommPlan(0, 1, e[:, 1], torch.arange(e.shape[0]), off)
    assert single.local_edge_idx.numel() == e.shape[0]


def test_legacy_plan_is_cached_per_index_tensor():
    """The ``indices + rank_mappings`` API builds its plan (two collectives) once per index tensor object, not once per call."""
    W, n = 2, 6
    e = random_graph(W * n, 3, seed=5)
    xg = torch.randn(W * n, 2, dtype=torch.float64)
    idx = e[:, 1].unsqueeze(0)
    rm = torch.stack([e[:, 0] // n, e[:, 1] // n])

    def job(comm):
        x = xg[comm.get_rank() * n:(comm.get_rank() + 1) * n].unsqueeze(0)
        comm.gather(x, idx, rm)
        n1 = len(comm.engine._legacy_plans)
        p1 = next(iter(comm.engine._legacy_plans.values()))[2]
        comm.gather(x, idx, rm)
        comm.gather(x, idx, rm)
        same = next(iter(comm.engine._legacy_plans.values()))[2] is p1
        idx2 = idx.clone()
        comm.gather(x, idx2, rm)
        return n1, len(comm.engine._legacy_plans), same

    for n1, n2, same in run_world(W, job):
        assert n1 == 1 and n2 == 2 and same


def test_offline_plan_generation_equals_collective_build():
    """``experiments/OGB/GenerateCache.py::generate_comm_plan`` (no process group, global COO list) == the plans the ranks build
    together through the exchange collective."""
    import importlib.util
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("_gen_cache", os.path.join(root, "experiments", "OGB", "GenerateCache.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    W, V = 3, 30
    e = random_graph(V, 4, seed=11, symmetric=False)
    off = torch.tensor([0, 7, 19, 30])

    def job(comm):
        r = comm.get_rank()
        mine = torch.nonzero((e[:, 0] >= off[r]) & (e[:, 0] < off[r + 1]), as_tuple=True)[0]
        return COO_to_NCCLEdgeConditionedCommPlan(r, W, e[:, 0], e[:, 1], mine, off, None, exchange=comm.engine.exchange_ids)

    for r, live in enumerate(run_world(W, job)):
        offline = gen.generate_comm_plan(e.t(), off, r, W)
        for a, b in ((live.source_graph_plan, offline.source_graph_plan), (live.dest_graph_plan, offline.dest_graph_plan)):
            for f in ("local_edge_idx", "local_vertex_idx", "boundary_edge_idx", "boundary_edge_buffer_map", "boundary_vertex_idx"):
                assert torch.equal(getattr(a, f), getattr(b, f)), (r, f)
            assert list(a.boundary_edge_splits) == list(b.boundary_edge_splits)
            assert list(a.boundary_vertex_splits) == list(b.boundary_vertex_splits)
