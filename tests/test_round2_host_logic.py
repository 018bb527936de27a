"""Host-side pieces added with the fused exchange path (all CPU): remote-neighbour index encoding, hub-row chunking, the
power-law graph generator, symmetric edge shards, the packed-weight cache and the bench workload modules."""
import ast
import os

import pytest

import torch

from dgraph_b200.parallel import build_all_patterns
from dgraph_b200.parallel.kernel_plan import HUB_CHUNK, HUB_DEGREE, GraphKernelPlan
from tests._graphs import partition, random_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_col_peer_encodes_owner_and_row_of_every_remote_neighbour():
    V, W = 300, 4
    e = random_graph(V, 6, seed=1, symmetric=True)
    pats = build_all_patterns(e, partition(V, W, "random"), W)
    for r, pat in enumerate(pats):
        kp = pat.kernel_plan("cpu")
        shift = max(1, (max(p.num_local_vertices for p in pats) - 1).bit_length())
        cp = kp.col_peer(shift)
        assert kp.col_peer(shift) is cp  # cached
        local = kp.col < kp.num_nbr_local
        assert torch.equal(cp[local], kp.col[local]) and bool((cp[~local] < 0).all())
        code = ~cp[~local].long()
        owner, row = code >> shift, code & ((1 << shift) - 1)
        h = (kp.col[~local] - kp.num_nbr_local).long()
        assert torch.equal(owner, pat.halo_owner.long()[h]) and torch.equal(row, pat.halo_remote_row.long()[h])
        # the row it names is that global vertex in the owner's numbering
        for o, rr, hh in list(zip(owner.tolist(), row.tolist(), h.tolist()))[:50]:
            assert int(pats[o].local_vertices_global[rr]) == int(pat.halo_vertices_global[hh])
        assert kp.remote_edge_count() == int((~local).sum())


def test_hub_plan_chunks_cover_every_long_segment_exactly_once():
    V = 200
    hub = torch.arange(1, HUB_DEGREE + 700)
    e = torch.cat([torch.stack([torch.zeros_like(hub), hub % V + V], 1), random_graph(V, 4, seed=2)])
    n_aug = int(e.max()) + 1
    kp = GraphKernelPlan.from_edges(e, V, n_aug, device="cpu", num_nbr_local=V)
    hp = kp.hub_plan("csr")
    assert hp is not None and hp.hubs.tolist() == [0] and kp.hub_plan("csr") is hp
    rp = kp.rowptr.long()
    cov = torch.zeros(kp.num_edges, dtype=torch.int32)
    for b, en in zip(hp.vbeg.tolist(), hp.vend.tolist()):
        assert 0 < en - b <= HUB_CHUNK
        cov[b:en] += 1
    assert bool((cov[rp[0]:rp[1]] == 1).all()) and int(cov.sum()) == int(rp[1] - rp[0])
    assert int(hp.main_end[0]) == int(rp[0]) and torch.equal(hp.main_end[1:].long(), rp[2:])  # hub row empty in the main launch
    assert torch.equal(hp.vrow, torch.zeros_like(hp.vrow)) and torch.equal(hp.seg, torch.zeros_like(hp.seg))
    small = GraphKernelPlan.from_edges(random_graph(50, 4, seed=3), 50, 50, device="cpu")
    assert small.hub_plan("csr") is None and small.hub_plan("csc") is None


def test_powerlaw_generator_has_hubs_and_stays_symmetric():
    from dgraph_b200.data.synthetic import symmetric_locality_graph

    V, E = 20000, 400000
    e = symmetric_locality_graph(V, E, degree_model="powerlaw", seed=0)
    u = symmetric_locality_graph(V, E, degree_model="uniform", seed=0)
    key = e[:, 0] * V + e[:, 1]
    assert key.unique().numel() == key.numel() and bool((e[:, 0] != e[:, 1]).all())
    assert set(key.tolist()) == set((e[:, 1] * V + e[:, 0]).tolist())  # symmetric
    deg_p, deg_u = torch.bincount(e[:, 0], minlength=V), torch.bincount(u[:, 0], minlength=V)
    assert int(deg_p.max()) > 8 * int(deg_u.max())  # heavy tail vs near-Poisson
    assert abs(e.shape[0] - E) <= 0.02 * E


def test_symmetric_shard_single_rank_is_undirected_and_duplicate_free():
    from benchmarks.workloads import symmetric_shard

    V, E = 5000, 60000
    e = symmetric_shard(V, E, 0, 1, torch.device("cpu"), cut=0.1)
    key = e[:, 0] * V + e[:, 1]
    assert key.unique().numel() == key.numel()
    assert set(key.tolist()) == set((e[:, 1] * V + e[:, 0]).tolist())
    assert 0.9 * E <= e.shape[0] <= 1.02 * E


def test_weight_cache_follows_parameter_versions():
    from dgraph_b200.ops import _wcache

    w = torch.nn.Parameter(torch.randn(4, 6))
    calls = []

    def build():
        calls.append(1)
        return w.detach().clone()

    a = _wcache.get(w, "k", build)
    assert _wcache.get(w, "k", build) is a and len(calls) == 1
    with torch.no_grad():
        w.add_(1.0)
    assert _wcache.get(w, "k", build) is not a and len(calls) == 2
    _wcache.bump()
    _wcache.get(w, "k", build)
    assert len(calls) == 3
    _wcache.get(w, "other", build)
    assert len(calls) == 4


def test_bench_workload_modules_parse_and_declare_every_config():
    for f in ("bench.py", "benchmarks/workloads.py", "benchmarks/runner.py", "benchmarks/scale_modes.py", "benchmarks/kernel_zoo.py",
              "benchmarks/p2p_sweep.py", "baseline/ref_workloads.py", "baseline/ref_bench.py"):
        ast.parse(open(os.path.join(ROOT, f)).read())
    from benchmarks import workloads

    assert set(workloads.BUILDERS) == {"papers100m", "graphcast", "rgat"} and workloads.HALO_HEADLINE in [(r, f) for r in workloads.HALO_ROWS for f in workloads.HALO_F]
    src = open(os.path.join(ROOT, "bench.py")).read()
    for cfg in ("products", "papers100m", "graphcast", "rgat", "halo"):
        assert f'"{cfg}"' in src


def test_duplicate_label_rows_merge_into_unique_support():
    """The loss's row-support hint for an index tensor with repeated rows: the compact rows of the unique rows equal what
    `index_add_` over the repeated list gives, and the unique-row tensor is the same object on every call."""
    from dgraph_b200.ops.loss import _merge_duplicate_rows, _row_support

    g = torch.Generator().manual_seed(0)
    V, C = 50, 8
    idx = torch.randint(0, V, (40,), generator=g)
    assert _row_support(torch.arange(0, V, 3)) is None and _row_support(torch.zeros(0, dtype=torch.long)) is None
    sup = _row_support(idx)
    assert sup is not None and _row_support(idx)[0] is sup[0]
    dsel = torch.randn(40, C, generator=g).bfloat16()
    merged = _merge_duplicate_rows(dsel, sup)
    dense = torch.zeros(V, C).index_add_(0, idx, dsel.float())
    got = torch.zeros(V, C).index_copy_(0, sup[0], merged.float())
    assert torch.allclose(got, dense, atol=2e-2, rtol=1e-2)
    assert torch.equal(torch.nonzero(dense.abs().sum(1)).flatten(), sup[0])


def test_multi_node_launch_is_detected(monkeypatch):
    """Peer memory (CUDA IPC) ends at the node boundary: a launch with ranks on several nodes selects the NCCL engine."""
    from dgraph_b200.communicator import _spans_nodes

    for k in ("WORLD_SIZE", "LOCAL_WORLD_SIZE", "SLURM_NNODES", "SLURM_JOB_NUM_NODES"):
        monkeypatch.delenv(k, raising=False)
    assert not _spans_nodes()
    monkeypatch.setenv("WORLD_SIZE", "8")
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert not _spans_nodes()
    monkeypatch.setenv("WORLD_SIZE", "16")
    assert _spans_nodes()
    monkeypatch.delenv("LOCAL_WORLD_SIZE")
    assert not _spans_nodes()
    monkeypatch.setenv("SLURM_NNODES", "2")
    assert _spans_nodes()


def test_installed_package_uses_the_prebuilt_libraries(tmp_path, monkeypatch):
    """A pip-installed copy ships `lib/*.so` but not `csrc/`: `build()` / `build_host()` hand back the libraries built at install time."""
    from dgraph_b200 import build as B

    if not (B.LIB.exists() and B.HOST_LIB.exists()):
        pytest.skip("libraries not built")
    monkeypatch.setattr(B, "CSRC", tmp_path)  # no sources anywhere below
    assert B.build() == B.LIB and B.build_host() == B.HOST_LIB
