"""Helper functions that the reference's experiment modules define next to their entry points, under the same module paths and
names here (``experiments/``): they are built on this package's vectorised / native code, so each is checked against a direct
statement of what it computes."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

from dgraph_b200.parallel.emulator import run_world

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXP = os.path.join(ROOT, "experiments")


def _load(rel, extra_path=()):
    path = os.path.join(EXP, rel)
    added = [os.path.dirname(path), *extra_path]
    for p in added:
        sys.path.insert(0, p)
    try:
        spec = importlib.util.spec_from_file_location("_exp_" + rel.replace("/", "_").replace("-", "_")[:-3], path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    finally:
        for p in added:
            sys.path.remove(p)


def test_nccl_benchmark_object_runs_the_legacy_api():
    tn = _load("Benchmarks/TestNCCL.py")
    gu = _load("Benchmarks/graph_utils.py")
    W, F = 3, 4
    gd = gu.get_nccl_gather_benchmark_data(F, W, "cpu")
    sd = gu.get_nccl_scatter_benchmark_data(F, W, "cpu")

    def job(comm):
        b = tn.NCCLBenchmark(comm, W, comm.get_rank())
        st = b.comm_stats_gather(gd.edge_rank_placement[0], gd.edge_src_rank[0])
        tg = tn.run_gather_benchmark(b, 3, gd)
        ts = tn.run_scatter_benchmark(b, 3, sd, cache=False)
        return int(st["local_comm_rev"]), int(st["local_comm_send"]), tg.shape, ts.shape

    for rev, send, sg, ss in run_world(W, job):
        assert rev == W - 1 and send == W - 1  # all-pairs: one message from and to every other rank
        assert sg == (3,) and ss == (3,)

    # one-sided calling convention (TestNVSHMEM.py): the gathered rows are the other ranks' vertices
    nv = _load("Benchmarks/TestNVSHMEM.py")

    def job2(comm):
        b = nv.NVSHMEMBenchmark(comm)
        r = comm.get_rank()
        pl = gd.edge_rank_placement
        got = comm.gather(gd.vertex_data[r:r + 1].unsqueeze(0), nv._mine(gd.edge_indices, pl, r), nv._mine(gd.edge_src_rank, pl, r))
        return got[0], nv.run_gather_benchmark(b, 2, gd).shape, nv.run_scatter_benchmark(b, 2, sd).shape

    for r, (got, sg, ss) in enumerate(run_world(W, job2)):
        assert torch.allclose(got, gd.vertex_data[[j for j in range(W) if j != r]]) and sg == (2,) and ss == (2,)

    # the one-sided generators hand every rank only its own share (reference graph_utils.py:121-230): same results
    def job3(comm):
        r = comm.get_rank()
        lg = gu.get_nvshmem_gather_benchmark_data(F, r, W, "cpu")
        ls = gu.get_nvshmem_scatter_benchmark_data(F, r, W, "cpu")
        assert lg.vertex_data.shape == (1, 1, F) and lg.edge_indices.shape == (1, W - 1) and ls.vertex_data.shape == (1, W - 1, F)
        b = nv.NVSHMEMBenchmark(comm)
        got = comm.gather(lg.vertex_data, lg.edge_indices, lg.edge_src_rank)
        return got[0], nv.run_gather_benchmark(b, 2, lg).shape, nv.run_scatter_benchmark(b, 2, ls).shape

    for r, (got, sg, ss) in enumerate(run_world(W, job3)):
        assert torch.allclose(got, gd.vertex_data[[j for j in range(W) if j != r]]) and sg == (2,) and ss == (2,)


def test_graphcast_utils_and_preprocess_helpers():
    ut = _load("GraphCast/data_utils/utils.py")
    pp = _load("GraphCast/data_utils/preprocess.py")
    gg = _load("GraphCast/data_utils/graphcast_graph.py")
    assert ut.padded_size(10, 4) == 12 and ut.padded_size(12, 4) == 12
    assert ut.pad_indices(torch.arange(5), 4).tolist() == [0, 1, 2, 3, 4, -1, -1, -1]
    # node features: (cos lat, sin lon, cos lon) in radians; the reference's degree-valued variant on request
    pos = torch.tensor([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]])
    nf = ut.generate_node_features(pos)
    assert torch.allclose(nf, torch.tensor([[1.0, 0.0, 1.0], [1.0, 1.0, 0.0], [0.0, 0.0, 1.0]]), atol=1e-6)
    nf_deg = ut.generate_node_features(pos, degrees_as_radians=True)
    assert abs(float(nf_deg[1, 1]) - np.sin(90.0)) < 1e-5
    # edge features: [dx, dy, dz, len] / max len, receiver rotated to (1, 0, 0)
    ef = ut.generate_edge_features(pos[[1, 2]], pos[[0, 0]])
    assert ef.shape == (2, 4) and torch.allclose(ef[:, 3], torch.ones(2), atol=1e-6)
    assert torch.allclose(ef[:, :3].norm(dim=1), ef[:, 3], atol=1e-6)
    n, e, s, d = ut.create_graph(torch.tensor([0, 1]), torch.tensor([1, 2]), pos)
    assert s.tolist() == [0, 1, 1, 2] and d.tolist() == [1, 2, 0, 1] and e.shape == (4, 4) and n.shape == (3, 3)
    gen = gg.DistributedGraphCastGraphGenerator((19, 36), mesh_level=2)
    g = gen.global_graph()
    lat = torch.linspace(90, -90, 19)
    lon = torch.arange(36) * 10.0
    ll = torch.stack(torch.meshgrid(lat, lon, indexing="ij"), -1).reshape(-1, 2)
    f, sm, dg = ut.create_mesh2grid_graph(ll, g.mesh_xyz, np.asarray(gen_faces(gen)))
    assert sm.numel() == 3 * ll.shape[0] and torch.equal(dg, torch.arange(ll.shape[0]).repeat_interleave(3))
    assert torch.equal(torch.stack([sm, dg], 1), g.mesh2grid_edges) and torch.allclose(f, g.mesh2grid_edge_features, atol=1e-6)
    topo = gg.GraphCastTopology.from_generator(gen, 0, 2)
    assert topo.mesh_rank_placement.numel() == g.num_mesh and topo.grid_rank_placement.numel() == g.num_grid
    assert topo.ranks_per_graph == 2 and torch.equal(topo.grid2mesh_graph_dst_indices, g.grid2mesh_edges[:, 1])
    pytest.importorskip("networkx")
    G = pp.graphcast_graph_to_nxgraph((None, None, g.mesh_edges[:, 0], g.mesh_edges[:, 1]))
    pl = pp.partition_graph(G, 2)
    assert pl.shape == (g.num_mesh,) and abs(int((pl == 0).sum()) - g.num_mesh // 2) <= 0.05 * g.num_mesh
    assert (pp.partition_graph(G, 1) == 0).all()
    with pytest.raises(ValueError):
        pp.partition_graph(G, 0)


def gen_faces(gen):
    from dgraph_b200.data.graphcast_graph import multimesh

    return multimesh(gen.mesh_level, gen.multi)[1]


def test_lsc_dataset_helpers():
    sys.path.insert(0, os.path.join(EXP, "OGB-LSC"))
    try:
        from lsc_datasets import MAG240M_dataset as m, distributed_graph_dataset as d, synthetic_dataset as s
    finally:
        sys.path.remove(os.path.join(EXP, "OGB-LSC"))
    assert d.get_vertex_offsets(10, 4).tolist() == [0, 3, 6, 9, 10]
    owner, mine = d.get_rank_mappings(10, 4, 3)
    assert owner.tolist() == [0, 0, 0, 1, 1, 1, 2, 2, 2, 3] and mine == 1
    ei = torch.tensor([[0, 9], [4, 2]])
    a, b = d.edge_mapping_from_vertex_mapping(ei, owner, owner)
    assert a.tolist() == [0, 3] and b.tolist() == [1, 0]
    assert m.get_edge_mappings(ei[0], ei[1], owner).tolist() == [1, 0]
    mask, k = m.load_or_generate_vertex_rank_mask(None, 10, 4, 1)
    assert mask.tolist() == (owner == 1).tolist() and k == 3
    assert s.generate_config_hash([1, 2, 3]) == s.generate_config_hash((1, 2, 3)) != s.generate_config_hash([1, 2, 4])
    e = s._generate_author_2_paper_edges(50, 100)
    assert e.shape[0] == 2 and int(e[0].max()) < 50 and int(e[1].max()) < 100 and torch.unique(e, dim=1).shape == e.shape
    rng = np.random.default_rng(0)
    pf = rng.standard_normal((20, 70)).astype(np.float32)
    idx = np.stack([rng.integers(0, 8, 50), rng.integers(0, 20, 50)])
    out = np.zeros((8, 70), dtype=np.float32)
    m._generate_features_from_paper_features(out, 8, 20, pf, idx, 70, dim_chunk_size=32)
    ref, cnt = np.zeros((8, 70)), np.zeros(8)
    for r, c in idx.T:
        ref[r] += pf[c]
        cnt[r] += 1
    assert np.allclose(out, ref / np.maximum(cnt, 1)[:, None], atol=1e-5)
    blk = m.get_col_slice(pf, 2, 9, 3, 11, chunk=3)
    assert np.array_equal(blk, pf[2:9, 3:11])
    dst = np.zeros_like(pf)
    m.save_col_slice(blk, dst, 2, 9, 3, 11, chunk=4)
    assert np.array_equal(dst[2:9, 3:11], blk) and np.count_nonzero(dst) == np.count_nonzero(blk)


def test_rank_mapping_files_become_contiguous_offsets():
    from dgraph_b200.data.hetero import node_offsets, offsets_from_rank_mapping

    assert torch.equal(offsets_from_rank_mapping(None, 10, 4), node_offsets(10, 4))
    assert offsets_from_rank_mapping(torch.tensor([0, 0, 0, 0, 1, 2, 2]), 7, 4).tolist() == [0, 4, 5, 7, 7]
    with pytest.raises(ValueError):
        offsets_from_rank_mapping(torch.tensor([0, 1, 0]), 3, 2)
    with pytest.raises(ValueError):
        offsets_from_rank_mapping(torch.tensor([0, 1]), 3, 2)


def test_distributed_bn_functional_forms_and_misc():
    dl = _load("OGB-LSC/distributed_layers.py", extra_path=(EXP,))
    x = torch.randn(40, 5, dtype=torch.float64, requires_grad=True)
    gam = torch.randn(1, 5, dtype=torch.float64, requires_grad=True)
    bet = torch.randn(1, 5, dtype=torch.float64, requires_grad=True)
    out, xh, mean, var, n = dl._compute_bn_forward(x.detach(), gam.detach(), bet.detach())
    ref = torch.nn.functional.batch_norm(x, None, None, gam.view(-1), bet.view(-1), training=True, eps=1e-5)
    assert torch.allclose(out, ref, atol=1e-12) and float(n) == 40
    go = torch.randn_like(ref)
    ref.backward(go)
    dx, dg, db = dl._compute_bn_backward(go, x.detach(), xh, mean, var, n, gam.detach(), bet.detach())
    assert torch.allclose(dx, x.grad, atol=1e-12) and torch.allclose(dg, gam.grad, atol=1e-12) and torch.allclose(db, bet.grad, atol=1e-12)
    oc = _load("OGB/ogb_comm_dataset.py", extra_path=(EXP,))
    pl = torch.tensor([0, 1, 0, 1, 1])
    masks = oc._build_local_split_masks(pl, {"train": torch.tensor([0, 3]), "valid": torch.tensor([4])}, 1)
    assert masks["train"].tolist() == [False, True, False] and masks["valid"].tolist() == [False, False, True]
    pat = oc.generate_communication_pattern(torch.tensor([[0, 1], [1, 0], [2, 3], [3, 2]]), pl, 0, 2)
    assert pat.num_local_vertices == 2 and pat.num_halo_vertices == 2
    tg = _load("GraphCast/train_graphcast.py", extra_path=(EXP,))

    class _C:
        _ranks_per_sample = 2

        def get_world_size(self):
            return 4

    a, b = torch.zeros(6, 3), torch.ones(6, 3)
    assert abs(float(tg.compute_loss(a, b, _C())) - 0.5) < 1e-7


def test_graphcast_generator_dictionary_stages_are_consistent():
    """``get_mesh_graph`` / ``get_grid2mesh_graph`` / ``get_mesh2grid_graph`` / ``get_graphcast_graph`` (the reference generator's
    stage names): renumbering keeps the edge sets, every rank owns one contiguous range, edges are ordered by holder rank."""
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator as G, SyntheticWeatherDataset

    gen = G((19, 36), mesh_level=2)
    g = gen.global_graph()
    pl = G.get_mesh_graph_partition(2, 2)
    assert pl.numel() == g.num_mesh and abs(int((pl == 0).sum()) - g.num_mesh // 2) <= 0.05 * g.num_mesh
    m = gen.get_mesh_graph(pl)
    nondecreasing = lambda t: bool((t[1:] >= t[:-1]).all())  # noqa: E731
    assert nondecreasing(m["node_rank_placement"]) and nondecreasing(m["edge_rank_placement"])
    old = torch.stack([m["mesh_vertex_renumbering"][m["src_indices"]], m["mesh_vertex_renumbering"][m["dst_indices"]]], 1)
    assert set(map(tuple, old.tolist())) == set(map(tuple, g.mesh_edges.tolist()))
    assert torch.equal(m["renumbered_vertices"][m["mesh_vertex_renumbering"]], torch.arange(g.num_mesh))
    assert torch.equal(m["src_rank_placement"], m["node_rank_placement"][m["src_indices"]])
    assert torch.equal(gen.get_grid_placement(pl, g.grid2mesh_edges[:, 1]), pl[g.grid2mesh_edges[:, 1]])
    g2m = gen.get_grid2mesh_graph(m)
    assert nondecreasing(g2m["grid2mesh_edge_rank_placement"]) and nondecreasing(g2m["grid_vertex_rank_placement"])
    assert g2m["src_indices"].numel() == g.grid2mesh_edges.shape[0]
    m2g = gen.get_mesh2grid_graph(g2m["grid_vertex_rank_placement"], m["renumbered_vertices"], g2m["renumbered_grid"])
    assert nondecreasing(m2g["mesh2grid_edge_rank_placement"]) and m2g["dst_indices"].numel() == 3 * g.num_grid
    shards = [gen.get_graphcast_graph(pl, rank=r) for r in range(2)]
    assert sum(s.num_local_mesh for s in shards) == g.num_mesh and sum(s.num_local_grid for s in shards) == g.num_grid
    d = SyntheticWeatherDataset.generate_data(3, 4, (5, 6), base_temp=15.0, amplitude=10.0, noise_level=0.0)
    assert d.shape == (3, 4, 5, 6) and np.allclose(d[0, 0, :, 0], [5, 10, 15, 20, 25]) and np.allclose(d[0, 2] - d[0, 0], -1.0)


def test_reference_constructor_orders_and_topology_patterns():
    """Positional constructor orders of the reference's experiment classes, and ``build_graphcast_comm_patterns`` on a topology
    record (the reference's calling form) giving the plans the generator builds itself."""
    from dgraph_b200.data.graphcast_graph import DistributedGraphCastGraphGenerator as G, build_graphcast_comm_patterns
    from dgraph_b200.utils.dist_utils import CommAwareDistributedSampler

    gg = _load("GraphCast/data_utils/graphcast_graph.py")
    grid = torch.zeros(19, 36, 2)
    gen = G(grid, 2, 2, 1, 2)  # lat_lon_grid tensor, mesh_level, ranks_per_graph, rank, world_size
    assert gen.lat_lon == (19, 36) and gen.rank == 1 and gen.world_size == 2 and gen.local_rank == 1
    topo = gg.GraphCastTopology.from_generator(gen, 1, 2)
    a = build_graphcast_comm_patterns(topo)
    b = build_graphcast_comm_patterns(gen)  # rank / world size from the constructor
    for name in ("grid2mesh", "mesh", "mesh2grid"):
        pa, pb = getattr(a, name), getattr(b, name)
        assert pa.num_local_vertices == pb.num_local_vertices and pa.num_halo_vertices == pb.num_halo_vertices
        assert torch.equal(pa.send_offset, pb.send_offset) and torch.equal(pa.recv_offset, pb.recv_offset)
        assert torch.equal(torch.sort(pa.halo_vertices_global).values, torch.sort(pb.halo_vertices_global).values)
    ds = list(range(10))
    assert list(CommAwareDistributedSampler(ds, None, False, 0, False, num_replicas=2, replica_id=1)) == [1, 3, 5, 7, 9]


def test_hetero_datasets_take_the_reference_constructor_form():
    from dgraph_b200.data.hetero import SyntheticHeterogeneousDataset
    from dgraph_b200.utils.dist_utils import SingleProcessDummyCommunicator

    class Cfg:
        num_papers, num_authors, num_institutions, num_features, num_classes = 64, 32, 4, 6, 9

    comm = SingleProcessDummyCommunicator()
    a = SyntheticHeterogeneousDataset(Cfg(), comm)
    b = SyntheticHeterogeneousDataset(0, 1, 64, 32, 4, 6, 9)
    assert a.num_nodes == b.num_nodes == [64, 32, 4] and a.num_classes == 9 and a.comm is comm
    assert all(torch.equal(x, y) for x, y in zip(a.features, b.features)) and torch.equal(a.labels, b.labels)
    assert torch.equal(a.get_mask("val"), a.get_mask("valid"))
