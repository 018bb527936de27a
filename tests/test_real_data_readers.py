"""On-disk OGB / MAG240M layouts are parsed without the ``ogb`` package (tiny fake datasets written in the real layout)."""
import gzip
import os

import numpy as np
import torch

from dgraph_b200.parallel.emulator import run_world


def _csv_gz(path, arr):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with gzip.open(path, "wt") as f:
        for row in np.atleast_2d(arr):
            f.write(",".join(str(v) for v in np.atleast_1d(row)) + "\n")


def _fake_ogbn(root, name="ogbn_arxiv", V=60, E=300, F=5, C=4, seed=0):
    rng = np.random.default_rng(seed)
    raw = os.path.join(root, name, "raw")
    e = np.unique(rng.integers(0, V, size=(E, 2)), axis=0)
    e = e[e[:, 0] != e[:, 1]]
    x = rng.standard_normal((V, F)).astype(np.float32)
    y = rng.integers(0, C, size=(V, 1))
    _csv_gz(os.path.join(raw, "edge.csv.gz"), e)
    _csv_gz(os.path.join(raw, "node-feat.csv.gz"), x)
    _csv_gz(os.path.join(raw, "node-label.csv.gz"), y)
    _csv_gz(os.path.join(raw, "num-node-list.csv.gz"), np.array([[V]]))
    _csv_gz(os.path.join(raw, "num-edge-list.csv.gz"), np.array([[e.shape[0]]]))
    perm = rng.permutation(V)
    for part, idx in (("train", perm[:30]), ("valid", perm[30:45]), ("test", perm[45:])):
        _csv_gz(os.path.join(root, name, "split", "time", f"{part}.csv.gz"), idx.reshape(-1, 1))
    return e, x, y.reshape(-1), perm


def test_ogbn_raw_files_feed_the_trainer_dataset(tmp_path):
    from dgraph_b200.data.ogb_raw import load_ogbn_raw
    from dgraph_b200.data.ogbn_datasets import DGraphOGBDataset

    e, x, y, perm = _fake_ogbn(str(tmp_path))
    g, labels, split = load_ogbn_raw("ogbn-arxiv", str(tmp_path))
    assert g["num_nodes"] == 60 and torch.equal(g["edge_index"], torch.from_numpy(e.T))
    assert torch.allclose(g["node_feat"], torch.from_numpy(x), atol=1e-6) and torch.equal(labels, torch.from_numpy(y))
    assert sorted(split["train"].tolist()) == sorted(perm[:30].tolist())

    def job(comm):
        ds = DGraphOGBDataset("arxiv", comm, root_dir=str(tmp_path))
        feats, lab, pat = ds[0]
        rows = pat.local_vertices_global
        return ds.synthetic, rows, feats, lab, ds.get_masks()["train_mask"]

    seen = torch.zeros(60, dtype=torch.long)
    for synthetic, rows, feats, lab, tm in run_world(2, job):
        assert synthetic is False
        seen[rows] += 1
        assert torch.allclose(feats, torch.from_numpy(x)[rows], atol=1e-6) and torch.equal(lab, torch.from_numpy(y)[rows])
        assert sorted(rows[tm].tolist()) == sorted(set(perm[:30].tolist()) & set(rows.tolist()))
    assert bool((seen == 1).all())


def _fake_mag(root, P=50, A=30, I=6, F=8, seed=0):
    rng = np.random.default_rng(seed)
    d = os.path.join(root, "mag240m_kddcup2021", "processed")

    def rel(ns, nd, count):
        e = np.unique(np.stack([rng.integers(0, ns, count), rng.integers(0, nd, count)]), axis=1)
        return e.astype(np.int64)

    edges = {"paper___cites___paper": rel(P, P, 200), "author___writes___paper": rel(A, P, 120),
             "author___affiliated_with___institution": rel(A, I, 25)}
    for k, v in edges.items():
        os.makedirs(os.path.join(d, k), exist_ok=True)
        np.save(os.path.join(d, k, "edge_index.npy"), v)
    os.makedirs(os.path.join(d, "paper"), exist_ok=True)
    feat = rng.standard_normal((P, F)).astype(np.float16)
    lab = rng.integers(0, 153, P).astype(np.float32)
    lab[::7] = np.nan  # unlabelled papers
    np.save(os.path.join(d, "paper", "node_feat.npy"), feat)
    np.save(os.path.join(d, "paper", "node_label.npy"), lab)
    np.save(os.path.join(d, "paper", "node_year.npy"), rng.integers(2000, 2020, P))
    perm = rng.permutation(P)
    torch.save({"train": perm[:30], "valid": perm[30:40], "test-dev": perm[40:]}, os.path.join(os.path.dirname(d), "split_dict.pt"))
    torch.save({"paper": P, "author": A, "institution": I}, os.path.join(os.path.dirname(d), "meta.pt"))
    return edges, feat, lab


def test_mag240m_files_give_the_same_plans_as_the_full_edge_lists(tmp_path):
    from dgraph_b200.data.hetero import RELATIONS, DGraph_MAG240M_Dataset, DistributedHeteroGraphDataset, node_offsets

    edges, feat, lab = _fake_mag(str(tmp_path))
    W = 3
    cites = torch.from_numpy(edges["paper___cites___paper"].T.copy())
    writes = torch.from_numpy(edges["author___writes___paper"].T.copy())
    affil = torch.from_numpy(edges["author___affiliated_with___institution"].T.copy())
    full_edges = [cites, writes.flip(1), writes, affil, affil.flip(1)]
    featf = torch.from_numpy(feat.astype(np.float32))
    for r in range(W):
        ds = DGraph_MAG240M_Dataset(r, W, data_dir=str(tmp_path), features="mean").prepare_plans(fused=True)
        assert ds.num_nodes == [50, 30, 6] and ds.num_classes == 153 and len(ds.edges) == len(RELATIONS)
        offs = [node_offsets(n, W) for n in ds.num_nodes]
        lo, hi = int(offs[0][r]), int(offs[0][r + 1])
        assert torch.allclose(ds.features[0], featf[lo:hi])
        want_lab = torch.from_numpy(np.nan_to_num(lab, nan=-1.0).astype(np.int64))[lo:hi]
        assert torch.equal(ds.labels, want_lab)
        # author features = mean of the written papers' features
        alo, ahi = int(offs[1][r]), int(offs[1][r + 1])
        for a in range(alo, ahi):
            ps = writes[writes[:, 0] == a][:, 1]
            want = featf[ps].mean(0) if ps.numel() else torch.zeros(featf.shape[1])
            assert torch.allclose(ds.features[1][a - alo], want, atol=1e-5)
        ilo, ihi = int(offs[2][r]), int(offs[2][r + 1])
        for i in range(ilo, ihi):
            au = affil[affil[:, 1] == i][:, 0]
            rows = []
            for a in au.tolist():
                ps = writes[writes[:, 0] == a][:, 1]
                rows.append(featf[ps].mean(0) if ps.numel() else torch.zeros(featf.shape[1]))
            want = torch.stack(rows).mean(0) if rows else torch.zeros(featf.shape[1])
            assert torch.allclose(ds.features[2][i - ilo], want, atol=1e-5)
        # plans built from the rank's edge subset == plans built from the full edge lists
        ref = DistributedHeteroGraphDataset(r, W, ds.num_nodes, ds.features, full_edges, ds.labels, 153).prepare_plans(fused=True)
        for p, q in zip(ds.patterns, ref.patterns):
            assert p.num_local_vertices == q.num_local_vertices and p.num_halo_vertices == q.num_halo_vertices
            assert torch.equal(p.send_offset, q.send_offset) and torch.equal(p.recv_offset, q.recv_offset)
            assert torch.equal(p.halo_vertices_global, q.halo_vertices_global)
            assert torch.equal(torch.unique(p.local_edge_list, dim=0), torch.unique(q.local_edge_list, dim=0))
        assert ds.get_mask("train").numel() + ds.get_mask("valid").numel() + ds.get_mask("test").numel() == hi - lo


def test_mag240m_feature_files_and_reference_accessors(tmp_path):
    """``generate_feature_data`` writes the author / institution feature files (means over neighbours) once; the dataset's
    reference-style accessors (``ds[0]``, ``_get_vertex_rank_mask``, ``get_NCCL_comm_plans``) describe the same graph."""
    from dgraph_b200.data.hetero import RELATIONS, DGraph_MAG240M_Dataset

    edges, feat, _ = _fake_mag(str(tmp_path), P=40, A=20, I=5, F=8)
    ds = DGraph_MAG240M_Dataset(0, 1, data_dir=str(tmp_path))
    paths = ds.generate_feature_data()
    A, I, F = 20, 5, 8
    af = np.memmap(paths["author"], mode="r", dtype=np.float16, shape=(A, F))
    w = edges["author___writes___paper"]
    ref = np.zeros((A, F))
    cnt = np.zeros(A)
    for a, p in w.T:
        ref[a] += feat[p].astype(np.float64)
        cnt[a] += 1
    ref /= np.maximum(cnt, 1)[:, None]
    assert np.allclose(np.asarray(af, dtype=np.float64), ref, atol=2e-2)
    inf = np.memmap(paths["institution"], mode="r", dtype=np.float16, shape=(I, F))
    aff = edges["author___affiliated_with___institution"]
    ref_i = np.zeros((I, F))
    cnt_i = np.zeros(I)
    for a, i in aff.T:
        ref_i[i] += np.asarray(af[a], dtype=np.float64)
        cnt_i[i] += 1
    ref_i /= np.maximum(cnt_i, 1)[:, None]
    assert np.allclose(np.asarray(inf, dtype=np.float64), ref_i, atol=2e-2)
    mtime = os.path.getmtime(paths["author"])
    assert ds.generate_feature_data() == paths and os.path.getmtime(paths["author"]) == mtime  # kept, not rewritten
    feats, edge_index, edge_type, rank_mappings = ds[0]
    assert len(ds) == 1 and edge_type == list(RELATIONS) and len(feats) == 3 and len(edge_index) == 5
    assert edge_index[0].shape[0] == 2 and all((rm[0] == 0).all() and (rm[1] == 0).all() for rm in rank_mappings)
    ids, pos = ds._get_vertex_rank_mask("val")
    assert torch.equal(ids[pos], ids) and ids.numel() == 10
    import pytest

    with pytest.raises(ValueError):
        ds._get_vertex_rank_mask("nope")
    assert len(ds.get_NCCL_comm_plans()) == 5
