"""Property tests (SURVEY §4 "what the new framework should add"): random graphs x random partitions (non-contiguous, empty
ranks, isolated ranks, W in {1,2,3,4,8}, feature widths that are not multiples of 4/8) on the in-process world emulator.

Oracles are global single-process tensor expressions; every distributed result, forward and backward, must match exactly
(fp64)."""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

from dgraph_b200.parallel import HaloExchange, build_communication_pattern
from dgraph_b200.parallel.comm_plan import COO_to_NCCLCommPlan
from dgraph_b200.parallel.emulator import run_world
from dgraph_b200.utils import debug

SETTINGS = dict(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])


@st.composite
def graph_and_partition(draw):
    W = draw(st.sampled_from([1, 2, 3, 4, 8]))
    V = draw(st.integers(min_value=1, max_value=40))
    E = draw(st.integers(min_value=0, max_value=160))
    seed = draw(st.integers(min_value=0, max_value=10_000))
    F = draw(st.sampled_from([1, 3, 5, 8, 13]))
    g = torch.Generator().manual_seed(seed)
    e = torch.stack([torch.randint(0, V, (E,), generator=g), torch.randint(0, V, (E,), generator=g)], 1) if E else \
        torch.zeros(0, 2, dtype=torch.long)
    kind = draw(st.sampled_from(["random", "contiguous", "round_robin", "skewed", "one_rank"]))
    if kind == "random":
        part = torch.randint(0, W, (V,), generator=g)
    elif kind == "contiguous":
        part = (torch.arange(V) * W // V).clamp(max=W - 1)
    elif kind == "round_robin":
        part = torch.arange(V) % W
    elif kind == "skewed":  # half of the ranks own nothing
        part = torch.randint(0, max(1, W // 2), (V,), generator=g)
    else:  # everything on the last rank: all other ranks are empty, nobody has a halo
        part = torch.full((V,), W - 1, dtype=torch.long)
    return W, V, F, e, part, seed


@settings(**SETTINGS)
@given(graph_and_partition())
def test_halo_exchange_and_message_passing_match_global(gp):
    W, V, F, e, part, seed = gp
    xg = torch.randn(V, F, dtype=torch.float64, generator=torch.Generator().manual_seed(seed + 1))
    wg = torch.randn(V, F, dtype=torch.float64, generator=torch.Generator().manual_seed(seed + 2))

    def job(comm):
        r = comm.get_rank()
        pat = build_communication_pattern(e, part, r, W)
        debug.validate_pattern(pat)
        debug.validate_kernel_plan(pat.kernel_plan("cpu"))
        own = pat.local_vertices_global
        x = xg[own].clone().requires_grad_(True)
        halo = HaloExchange(comm)(x, pat)
        assert torch.equal(halo.detach(), xg[pat.halo_vertices_global])
        aug = torch.cat([x, halo], 0)
        le = pat.local_edge_list
        out = torch.zeros(own.numel(), F, dtype=torch.float64).index_add(0, le[:, 0], aug[le[:, 1]])
        (out * wg[own]).sum().backward()
        return own, out.detach(), x.grad

    x = xg.clone().requires_grad_(True)
    want = torch.zeros(V, F, dtype=torch.float64).index_add(0, e[:, 0], x[e[:, 1]])
    (want * wg).sum().backward()
    seen = torch.zeros(V, dtype=torch.long)
    for own, out, gx in run_world(W, job):
        seen[own] += 1
        assert torch.allclose(out, want.detach()[own])
        assert torch.allclose(gx, x.grad[own])
    assert bool((seen == 1).all())  # every vertex is owned exactly once


@st.composite
def plan_case(draw):
    W = draw(st.sampled_from([1, 2, 3, 4, 8]))
    V = draw(st.integers(min_value=W, max_value=48))
    E = draw(st.integers(min_value=0, max_value=150))
    seed = draw(st.integers(min_value=0, max_value=10_000))
    F = draw(st.sampled_from([1, 2, 5, 8]))
    g = torch.Generator().manual_seed(seed)
    cuts = torch.sort(torch.randint(0, V + 1, (W - 1,), generator=g)).values if W > 1 else torch.zeros(0, dtype=torch.long)
    off = torch.cat([torch.zeros(1, dtype=torch.long), cuts, torch.tensor([V])])  # uneven, possibly empty ranks
    e = torch.stack([torch.randint(0, V, (E,), generator=g), torch.randint(0, V, (E,), generator=g)], 1) if E else \
        torch.zeros(0, 2, dtype=torch.long)
    return W, V, F, e, off, seed


@settings(**SETTINGS)
@given(plan_case())
def test_comm_plan_gather_scatter_match_global(pc):
    W, V, F, e, off, seed = pc
    xg = torch.randn(V, F, dtype=torch.float64, generator=torch.Generator().manual_seed(seed + 1))
    eg = torch.randn(e.shape[0], F, dtype=torch.float64, generator=torch.Generator().manual_seed(seed + 2))
    src, dst = e[:, 0], e[:, 1]

    def job(comm):
        r = comm.get_rank()
        mine = torch.nonzero((src >= off[r]) & (src < off[r + 1]), as_tuple=True)[0]
        plan = COO_to_NCCLCommPlan(r, W, dst, mine, off, exchange=comm.engine.exchange_ids)
        debug.validate_comm_plan(plan)
        x = xg[off[r]:off[r + 1]].clone().unsqueeze(0).requires_grad_(True)
        got = comm.gather(x, comm_plan=plan)
        (got * got).sum().backward()
        ef = eg[mine].clone().unsqueeze(0).requires_grad_(True)
        sc = comm.scatter(ef, comm_plan=plan)
        (sc * sc).sum().backward()
        return mine, got.detach()[0], x.grad[0], sc.detach()[0], ef.grad[0]

    x = xg.clone().requires_grad_(True)
    g = x[dst]
    (g * g).sum().backward()
    ef = eg.clone().requires_grad_(True)
    s = torch.zeros_like(xg).index_add(0, dst, ef)
    (s * s).sum().backward()
    for r, (mine, got, gx, sc, gef) in enumerate(run_world(W, job)):
        assert torch.allclose(got, g.detach()[mine])
        assert torch.allclose(gx, x.grad[off[r]:off[r + 1]])
        assert torch.allclose(sc, s.detach()[off[r]:off[r + 1]])
        assert torch.allclose(gef, ef.grad[mine])
