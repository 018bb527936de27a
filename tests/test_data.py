import torch

from dgraph_b200.data.graph import DistributedGraph, get_round_robin_node_rank_map
from dgraph_b200.data.partition import bfs_blocks, contiguous_blocks, edge_cut_fraction, hashed, round_robin
from dgraph_b200.data.preprocess import process_homogenous_data
from dgraph_b200.utils import largest_split, split_per_rank
from tests._graphs import random_graph


def test_round_robin_map_edge_cases():  # model: reference tests/test_distributed_graph.py:4-30
    assert get_round_robin_node_rank_map(5, 2).tolist() == [0, 1, 0, 1, 0]
    assert get_round_robin_node_rank_map(0, 4).numel() == 0
    assert get_round_robin_node_rank_map(3, 8).tolist() == [0, 1, 2]


def test_process_homogenous_data_uses_inverse_permutation():
    V, W = 11, 3
    e = random_graph(V, 4, seed=1)
    x = torch.arange(V, dtype=torch.float32).unsqueeze(1)  # feature = original id
    place = torch.tensor([2, 0, 1, 1, 0, 2, 2, 0, 1, 0, 2])  # not an involution
    g = process_homogenous_data({"node_feat": x, "edge_index": e.t().contiguous(), "num_nodes": V}, torch.arange(V), 0, W,
                                {"train": torch.tensor([0, 5, 9])}, place)  # the reference's argument order
    assert isinstance(g, DistributedGraph)
    assert torch.equal(g.node_loc, torch.sort(place).values)
    orig = g.node_features[:, 0].long()  # new id -> original id
    got = torch.stack([orig[g.edge_index[0]], orig[g.edge_index[1]]], 1)
    assert torch.equal(torch.unique(got, dim=0), torch.unique(e, dim=0))  # edges survive the renumbering
    assert (g.edge_loc[1:] >= g.edge_loc[:-1]).all() and torch.equal(g.edge_loc, g.node_loc[g.edge_index[0]])
    assert sorted(orig[g.train_mask].tolist()) == [0, 5, 9]
    for r in range(W):
        assert g.get_local_node_features(r).shape[0] == int((place == r).sum())
        assert torch.equal(orig[g.get_local_mask("train", r) + int(g.get_nodes_per_rank()[:r].sum())],
                           orig[g.train_mask[(g.node_loc[g.train_mask] == r)]])


def test_partitioners_balanced_and_bfs_cuts_less_than_hash():
    V, W = 600, 4
    # ring-like graph: locality exists, ids shuffled so that contiguous blocks are meaningless
    perm = torch.randperm(V, generator=torch.Generator().manual_seed(0))
    base = torch.arange(V)
    e = torch.cat([torch.stack([base, (base + k) % V], 1) for k in (1, 2, 3)])
    e = torch.cat([e, e.flip(1)])
    e = perm[e]
    for p in (contiguous_blocks(V, W), round_robin(V, W), hashed(V, W), bfs_blocks(e, V, W)):
        sizes = torch.bincount(p, minlength=W)
        assert sizes.sum() == V and sizes.max() <= V / W * 1.1 + 1
    assert edge_cut_fraction(e, bfs_blocks(e, V, W)) < 0.5 * edge_cut_fraction(e, hashed(V, W))


def test_split_helpers():
    assert largest_split(10, 4) == 3
    assert [split_per_rank(10, r, 4) for r in range(4)] == [3, 3, 3, 1]  # shard sizes, as in the reference
    from dgraph_b200.utils.data_splitting import split_range

    assert [split_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 9), (9, 10)]


def test_native_multilevel_partitioner_quality_balance_determinism():
    """csrc/host/partition.cpp (the METIS role): balanced parts, far fewer cut edges than hashing, better than the BFS blocks on a
    shuffled 2-D grid, reproducible for a seed, vertex weights respected."""
    from dgraph_b200.data.partition import bfs_blocks, edge_cut_fraction, hashed, multilevel

    n = 60
    V = n * n
    idx = torch.arange(V).view(n, n)
    e = torch.cat([torch.stack([idx[:, :-1].reshape(-1), idx[:, 1:].reshape(-1)], 1),
                   torch.stack([idx[:-1].reshape(-1), idx[1:].reshape(-1)], 1)])
    perm = torch.randperm(V, generator=torch.Generator().manual_seed(0))
    e = perm[e]  # vertex ids carry no locality
    for W in (2, 4, 8):
        p = multilevel(e, V, W, imbalance=1.03)
        sizes = torch.bincount(p, minlength=W)
        assert int(p.min()) >= 0 and int(p.max()) == W - 1
        assert float(sizes.max()) <= 1.03 * (V / W) + 2
        cut = edge_cut_fraction(e, p)
        assert cut < 0.25 * edge_cut_fraction(e, hashed(V, W))
        assert cut <= edge_cut_fraction(e, bfs_blocks(e, V, W)) + 1e-9
        assert torch.equal(p, multilevel(e, V, W, imbalance=1.03))
    assert multilevel(e, V, 1).sum() == 0
    w = torch.ones(V, dtype=torch.long)
    w[: V // 10] = 10  # a few heavy vertices: parts are balanced by weight, not by count
    p = multilevel(e, V, 4, vertex_weights=w)
    pw = torch.zeros(4, dtype=torch.long).index_add_(0, p, w)
    assert float(pw.max()) <= 1.05 * float(w.sum()) / 4 + 10
    # isolated vertices and an empty edge list are fine
    assert multilevel(torch.zeros(0, 2, dtype=torch.long), 10, 3).numel() == 10
