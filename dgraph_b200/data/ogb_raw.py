"""Readers for the on-disk layouts of OGB / OGB-LSC datasets that need neither the ``ogb`` package nor a network.

The reference goes through ``ogb.nodeproppred.NodePropPredDataset`` (``DGraph/data/ogbn_datasets.py:96-123``,
``experiments/OGB/ogb_comm_dataset.py:54-90``) and ``ogb.lsc.MAG240MDataset`` (``experiments/OGB-LSC/lsc_datasets/
MAG240M_dataset.py:116-340``).  Those classes only download, unzip and parse a fixed directory layout; the parsing half is
re-stated here so that a machine that already holds the files (copied in, no egress) can train on the real graphs:

``<root>/<name with '-' -> '_'>/raw/``  ``edge.csv.gz``, ``node-feat.csv.gz``, ``node-label.csv.gz``, ``num-node-list.csv.gz``
(or, for papers100M, ``data.npz`` + ``node-label.npz``) and ``split/<scheme>/{train,valid,test}.csv.gz``;

``<root>/mag240m_kddcup2021/processed/``  ``paper/{node_feat,node_label,node_year}.npy``,
``{author___writes___paper,author___affiliated_with___institution,paper___cites___paper}/edge_index.npy``, ``split_dict.pt``.

Large arrays are opened with ``mmap_mode="r"`` so that a rank only touches the rows it keeps.
"""
from __future__ import annotations

import glob
import os
from typing import Dict, Optional, Tuple

import numpy as np
import torch


def _dir_name(dname: str) -> str:
    return dname.replace("-", "_")


def _read_csv(path: str, dtype) -> np.ndarray:
    import pandas as pd

    return pd.read_csv(path, compression="gzip" if path.endswith(".gz") else None, header=None).values.astype(dtype)


def find_ogbn(dname: str, root: str) -> Optional[str]:
    """Directory of the raw files of ``dname`` under ``root`` or None."""
    d = os.path.join(root, _dir_name(dname), "raw")
    if os.path.exists(os.path.join(d, "edge.csv.gz")) or os.path.exists(os.path.join(d, "data.npz")):
        return d
    return None


def load_ogbn_raw(dname: str, root: str) -> Tuple[Dict, torch.Tensor, Dict[str, torch.Tensor]]:
    """``(graph, labels, split)`` with ``graph = {"edge_index" [2, E], "node_feat" [V, F], "num_nodes"}`` -- the objects the
    ``ogb`` package would return for the same files (``NodePropPredDataset.__getitem__`` / ``get_idx_split``)."""
    raw = find_ogbn(dname, root)
    if raw is None:
        raise FileNotFoundError(f"no raw OGB files for {dname} under {root}")
    if os.path.exists(os.path.join(raw, "data.npz")):  # binary layout (ogbn-papers100M)
        z = np.load(os.path.join(raw, "data.npz"), mmap_mode="r")
        edge_index = torch.from_numpy(np.ascontiguousarray(z["edge_index"]).astype(np.int64))
        feat = torch.from_numpy(np.ascontiguousarray(z["node_feat"]))
        n = int(np.asarray(z["num_nodes_list"]).reshape(-1)[0])
        lab = np.load(os.path.join(raw, "node-label.npz"))["node_label"].reshape(-1)
        labels = torch.from_numpy(np.nan_to_num(lab, nan=-1.0).astype(np.int64))  # unlabelled papers are NaN in the file
    else:
        edge_index = torch.from_numpy(_read_csv(os.path.join(raw, "edge.csv.gz"), np.int64).T.copy())
        feat = torch.from_numpy(_read_csv(os.path.join(raw, "node-feat.csv.gz"), np.float32))
        labels = torch.from_numpy(_read_csv(os.path.join(raw, "node-label.csv.gz"), np.int64).reshape(-1))
        nn_file = os.path.join(raw, "num-node-list.csv.gz")
        n = int(_read_csv(nn_file, np.int64).reshape(-1)[0]) if os.path.exists(nn_file) else feat.shape[0]
    split_dirs = sorted(glob.glob(os.path.join(os.path.dirname(raw), "split", "*")))
    if not split_dirs:
        raise FileNotFoundError(f"no split directory next to {raw}")
    split = {}
    for part in ("train", "valid", "test"):
        f = os.path.join(split_dirs[0], f"{part}.csv.gz")
        split[part] = torch.from_numpy(_read_csv(f, np.int64).reshape(-1))
    return {"edge_index": edge_index, "node_feat": feat, "num_nodes": n}, labels, split


# ---------------------------------------------------------------------------------------------------------------------------
# MAG240M (OGB-LSC)
# ---------------------------------------------------------------------------------------------------------------------------
MAG_RELATIONS = {
    "cites": "paper___cites___paper",
    "writes": "author___writes___paper",
    "affiliated_with": "author___affiliated_with___institution",
}


def find_mag240m(root: str) -> Optional[str]:
    d = os.path.join(root, "mag240m_kddcup2021", "processed")
    return d if os.path.exists(os.path.join(d, "paper", "node_label.npy")) else None


class MAG240MFiles:
    """Memory-mapped view of the processed MAG240M files (the role of ``ogb.lsc.MAG240MDataset``).

    ``paper_feat`` is fp16 ``[num_papers, 768]`` on disk (175 GB): it is never read as a whole -- ``paper_feat_rows(lo, hi)``
    returns one rank's contiguous block.  Edge indices are ``int64 [2, E]``.
    """

    def __init__(self, root: str):
        d = find_mag240m(root)
        if d is None:
            raise FileNotFoundError(f"no processed MAG240M files under {root}")
        self.dir = d
        self.paper_label = np.load(os.path.join(d, "paper", "node_label.npy"), mmap_mode="r")
        self.paper_year = np.load(os.path.join(d, "paper", "node_year.npy"), mmap_mode="r")
        f = os.path.join(d, "paper", "node_feat.npy")
        self.paper_feat = np.load(f, mmap_mode="r") if os.path.exists(f) else None
        self.num_papers = int(self.paper_label.shape[0])
        self.num_classes = 153
        self._edges: Dict[str, np.ndarray] = {}
        meta = os.path.join(os.path.dirname(d), "meta.pt")
        m = torch.load(meta, weights_only=False) if os.path.exists(meta) else {}
        w, a = self.edge_index("writes"), self.edge_index("affiliated_with")
        self.num_authors = int(m.get("author", int(w[0].max()) + 1 if w.shape[1] else 0))
        self.num_institutions = int(m.get("institution", int(a[1].max()) + 1 if a.shape[1] else 0))

    def edge_index(self, rel: str) -> np.ndarray:
        if rel not in self._edges:
            self._edges[rel] = np.load(os.path.join(self.dir, MAG_RELATIONS[rel], "edge_index.npy"), mmap_mode="r")
        return self._edges[rel]

    def split(self) -> Dict[str, torch.Tensor]:
        sd = torch.load(os.path.join(os.path.dirname(self.dir), "split_dict.pt"), weights_only=False)
        return {k: torch.as_tensor(np.asarray(v), dtype=torch.long) for k, v in sd.items()}

    def paper_feat_rows(self, lo: int, hi: int, dtype=torch.bfloat16) -> torch.Tensor:
        if self.paper_feat is None:
            raise FileNotFoundError("paper/node_feat.npy is missing")
        return torch.from_numpy(np.ascontiguousarray(self.paper_feat[lo:hi]).astype(np.float32)).to(dtype)

    def labels(self) -> torch.Tensor:
        return torch.from_numpy(np.nan_to_num(np.asarray(self.paper_label), nan=-1.0).astype(np.int64))


def mean_neighbor_features(out, num_nodes: int, num_src: int, src_feat, edge_index, num_features: int, dim_chunk_size: int = 64,
                           row_chunk: int = 100_000) -> None:
    """``out[v] = mean over edges (v, p) of src_feat[p]`` written into ``out`` (a ``np.memmap`` or array ``[num_nodes, F]``), 64
    columns at a time: the author / institution feature generation of MAG240M (the reference does it with a ``torch_sparse``
    SpMM, ``MAG240M_dataset.py:65-102``).  A CSR segment mean over the rows sorted by ``v``; only one column block of the source
    features is resident at a time."""
    row = torch.as_tensor(np.asarray(edge_index[0]), dtype=torch.long)
    col = torch.as_tensor(np.asarray(edge_index[1]), dtype=torch.long)
    order = torch.argsort(row, stable=True)
    col = col[order]
    lengths = torch.bincount(row, minlength=num_nodes)
    for i in range(0, num_features, dim_chunk_size):
        j = min(i + dim_chunk_size, num_features)
        block = np.empty((num_src, j - i), dtype=np.float32)
        for r0 in range(0, num_src, row_chunk):
            r1 = min(r0 + row_chunk, num_src)
            block[r0:r1] = src_feat[r0:r1, i:j]
        red = torch.segment_reduce(torch.from_numpy(block)[col], "mean", lengths=lengths, unsafe=True)
        red = torch.nan_to_num(red, nan=0.0).numpy()  # vertices without neighbours
        for r0 in range(0, num_nodes, row_chunk):
            r1 = min(r0 + row_chunk, num_nodes)
            out[r0:r1, i:j] = red[r0:r1]
    if hasattr(out, "flush"):
        out.flush()


def generate_mag240m_feature_files(files: "MAG240MFiles", num_features: int = 0) -> Dict[str, str]:
    """Writes ``author_feat.npy`` (mean of the written papers' features) and ``institution_feat.npy`` (mean of the affiliated
    authors' features) next to the processed files, fp16 ``np.memmap``s as in the OGB example the reference follows
    (``MAG240M_dataset.py:262-320``).  Existing files are kept.  Returns the two paths."""
    if files.paper_feat is None:
        raise FileNotFoundError("paper/node_feat.npy is missing")
    F = int(num_features) or int(files.paper_feat.shape[1])
    a_path, i_path = os.path.join(files.dir, "author_feat.npy"), os.path.join(files.dir, "institution_feat.npy")
    if not os.path.exists(a_path):
        out = np.memmap(a_path, mode="w+", dtype=np.float16, shape=(files.num_authors, F))
        mean_neighbor_features(out, files.num_authors, files.num_papers, files.paper_feat, files.edge_index("writes"), F)
        del out
    if not os.path.exists(i_path):
        authors = np.memmap(a_path, mode="r", dtype=np.float16, shape=(files.num_authors, F))
        aff = np.asarray(files.edge_index("affiliated_with"))
        out = np.memmap(i_path, mode="w+", dtype=np.float16, shape=(files.num_institutions, F))
        mean_neighbor_features(out, files.num_institutions, files.num_authors, authors, aff[::-1], F)  # (institution, author)
        del out
    return {"author": a_path, "institution": i_path}
