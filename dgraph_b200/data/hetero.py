"""Heterogeneous (MAG240M-shaped) distributed graph datasets.

Parity: reference ``experiments/OGB-LSC/lsc_datasets/distributed_graph_dataset.py:24-489`` (``DistributedHeteroGraphDataset``:
3 node types paper / author / institution, 5 relations, contiguous ``*_vertex_offset[W+1]``, plan generation / save /
load) and ``synthetic_dataset.py:29-199`` (``SyntheticHeterogeneousDataset``: MAG240M-shaped random graph with average
degrees 11 / 3.5 / 0.35, 70/15/15 split, per-rank plan cache).  The real MAG240M loader needs ``ogb`` + ``torch_sparse``
(absent here); :class:`DGraph_MAG240M_Dataset` raises a clear error unless ``ogb`` is importable.

Plans: for every relation both forms are produced -- the reference's ``NCCLEdgeConditionedGraphCommPlan`` (edges stored with
their source) and a bipartite :class:`CommunicationPattern` (edges stored with their destination) that feeds the fused
segment-softmax path.
"""
from __future__ import annotations

import hashlib
import os
from typing import Dict, List, Optional, Tuple

import torch

from dgraph_b200.parallel.comm_pattern import CommunicationPattern, build_communication_pattern
from dgraph_b200.parallel.comm_plan import COO_to_NCCLEdgeConditionedCommPlan, NCCLEdgeConditionedGraphCommPlan

NODE_TYPES = ["paper", "author", "institution"]
# (source type, destination type) -- same order as the reference (`distributed_graph_dataset.py`)
RELATIONS: List[Tuple[int, int]] = [(0, 0), (0, 1), (1, 0), (1, 2), (2, 1)]


def node_offsets(num_nodes: int, world_size: int) -> torch.Tensor:
    """Contiguous, nearly even vertex ranges: rank r owns ``[off[r], off[r+1])``."""
    return torch.tensor([num_nodes * r // world_size for r in range(world_size + 1)], dtype=torch.long)


def offsets_from_rank_mapping(rank_mapping: Optional[torch.Tensor], num_nodes: int, world_size: int) -> torch.Tensor:
    """Vertex ranges ``[W + 1]`` described by a per-vertex rank tensor (the ``*_rank_mapping_file`` arguments of the reference's
    OGB-LSC entry point).  The plans address vertices by contiguous ranges (as ``_NCCLCommPlan.py:210`` does with ``bucketize``),
    so the mapping has to be non-decreasing; ``None`` gives the even split."""
    if rank_mapping is None:
        return node_offsets(num_nodes, world_size)
    rm = torch.as_tensor(rank_mapping).reshape(-1).long()
    if rm.numel() != num_nodes:
        raise ValueError(f"rank mapping has {rm.numel()} entries for {num_nodes} vertices")
    if rm.numel() and (bool((rm[1:] < rm[:-1]).any()) or int(rm.min()) < 0 or int(rm.max()) >= world_size):
        raise ValueError("rank mapping must be non-decreasing with values in [0, world_size): renumber the vertices so that "
                         "every rank owns one contiguous range")
    off = torch.zeros(world_size + 1, dtype=torch.long)
    off[1:] = torch.bincount(rm, minlength=world_size).cumsum(0)
    return off


class DistributedHeteroGraphDataset:
    def __init__(self, rank: int, world_size: int, num_nodes: List[int], features: List[torch.Tensor],
                 edges: List[torch.Tensor], labels: torch.Tensor, num_classes: int, comm=None,
                 split: Optional[Dict[str, torch.Tensor]] = None, cache_dir: Optional[str] = None,
                 offsets: Optional[List[torch.Tensor]] = None):
        """``edges[k]`` is ``[E_k, 2]`` with column 0 = source id, column 1 = destination id of relation ``RELATIONS[k]``;
        ``features[t]`` holds this rank's rows of node type ``t``; ``labels`` this rank's paper labels."""
        self.rank, self.world_size, self.comm = rank, world_size, comm
        self.num_nodes, self.num_classes = list(num_nodes), num_classes
        self.offsets = list(offsets) if offsets is not None else [node_offsets(n, world_size) for n in num_nodes]
        self.paper_vertex_offset, self.author_vertex_offset, self.institution_vertex_offset = self.offsets
        self.features, self.edges, self.labels = features, edges, labels
        self.split = split or {}
        self.cache_dir = cache_dir
        self.comm_plans: Optional[List[NCCLEdgeConditionedGraphCommPlan]] = None
        self.patterns: Optional[List[CommunicationPattern]] = None
        self.batched = False

    # ---- reference accessors -----------------------------------------------------------------------
    @property
    def num_features(self) -> int:
        return int(self.features[0].shape[-1])

    @property
    def num_relations(self) -> int:
        return len(RELATIONS)

    def partition(self, t: int) -> torch.Tensor:
        off = self.offsets[t]
        return torch.bucketize(torch.arange(self.num_nodes[t]), off, right=True) - 1

    def local_range(self, t: int) -> Tuple[int, int]:
        return int(self.offsets[t][self.rank]), int(self.offsets[t][self.rank + 1])

    def get_mask(self, mask_type: str) -> torch.Tensor:
        """Global paper indices of a split (``"train"``, ``"valid"`` / ``"val"``, ``"test"``) -> local row indices on this rank."""
        name = mask_type if mask_type in self.split else {"val": "valid", "valid": "val"}.get(mask_type, mask_type)
        idx = self.split[name]
        lo, hi = self.local_range(0)
        m = (idx >= lo) & (idx < hi)
        return idx[m] - lo

    def get_target(self, mask_type: str) -> torch.Tensor:
        return self.labels[self.get_mask(mask_type).to(self.labels.device)]

    def __len__(self) -> int:
        return 1  # one full graph

    def __getitem__(self, idx):
        """``(features, edge_index, edge_type, rank_mappings)`` as in the reference (``distributed_graph_dataset.py:240-282``):
        per node type this rank's features; per relation the ``[2, E_k]`` (source, destination) edges this rank holds, the
        ``(source type, destination type)`` pair, and the owner ranks ``[source owners, destination owners]`` of the endpoints."""
        edge_index = [e.t().contiguous() for e in self.edges]
        rank_mappings = []
        for (st, dt), e in zip(RELATIONS, self.edges):
            rank_mappings.append([torch.bucketize(e[:, 0].contiguous(), self.offsets[st], right=True) - 1,
                                  torch.bucketize(e[:, 1].contiguous(), self.offsets[dt], right=True) - 1])
        return self.features, edge_index, list(RELATIONS), rank_mappings

    def _get_vertex_rank_mask(self, mask_type: str) -> Tuple[torch.Tensor, torch.Tensor]:
        """``(global paper ids of the split, positions inside that list of the papers this rank owns)``
        (``distributed_graph_dataset.py:332-355``); ``"val"`` and ``"valid"`` name the same split."""
        name = {"val": "valid"}.get(mask_type, mask_type)
        if name not in self.split and mask_type in self.split:
            name = mask_type
        if name not in self.split:
            raise ValueError(f"Invalid mask type: {mask_type}")
        ids = self.split[name].reshape(-1)
        lo, hi = self.local_range(0)
        return ids, torch.nonzero((ids >= lo) & (ids < hi), as_tuple=True)[0]

    def get_NCCL_comm_plans(self) -> List[NCCLEdgeConditionedGraphCommPlan]:
        """The five edge-conditioned plans in the order of ``RELATIONS`` (built on first use)."""
        if self.comm_plans is None:
            self.prepare_plans(fused=False, reference_layout=True)
        return self.comm_plans

    def add_batch_dimension(self):
        """``[N, F] -> [1, N, F]`` (the layout the reference's gather/scatter API expects)."""
        if not self.batched:
            self.features = [f.unsqueeze(0) for f in self.features]
            self.batched = True
        return self

    def to(self, device):
        self.features = [f.to(device) for f in self.features]
        self.labels = self.labels.to(device)
        if self.comm_plans is not None:
            self.comm_plans = [p.to(device) for p in self.comm_plans]
        if self.patterns is not None:
            self.patterns = [p.to(device) for p in self.patterns]
        return self

    # ---- plans -----------------------------------------------------------------------------------
    def _cache_path(self, kind: str) -> Optional[str]:
        if not self.cache_dir:
            return None
        h = hashlib.md5(str((self.num_nodes, [int(e.shape[0]) for e in self.edges])).encode()).hexdigest()[:12]
        return os.path.join(self.cache_dir, f"hetero_{h}_rank_{self.rank}_of_{self.world_size}_{kind}.pt")

    def _generate_comm_plans(self, exchange=None) -> List[NCCLEdgeConditionedGraphCommPlan]:
        plans = []
        for (s, d), e in zip(RELATIONS, self.edges):
            lo, hi = self.local_range(s)
            mine = torch.nonzero((e[:, 0] >= lo) & (e[:, 0] < hi), as_tuple=True)[0]
            plans.append(COO_to_NCCLEdgeConditionedCommPlan(self.rank, self.world_size, e[:, 0], e[:, 1], mine,
                                                           self.offsets[s], self.offsets[d], exchange=exchange))
        return plans

    def _generate_patterns(self) -> List[CommunicationPattern]:
        pats = []
        for (s, d), e in zip(RELATIONS, self.edges):
            ed = torch.stack([e[:, 1], e[:, 0]], dim=1)  # central = destination, neighbour = source
            pats.append(build_communication_pattern(ed, self.partition(d), self.rank, self.world_size,
                                                    neighbor_partitioning=self.partition(s)))
        return pats

    def _save_comm_plans(self, obj, kind):
        p = self._cache_path(kind)
        if p:
            os.makedirs(os.path.dirname(p), exist_ok=True)
            torch.save(obj, p)

    def _load_comm_plans(self, kind):
        p = self._cache_path(kind)
        if p and os.path.exists(p):
            return torch.load(p, weights_only=False)
        return None

    def prepare_plans(self, fused: bool = True, reference_layout: bool = False, exchange=None):
        if fused and self.patterns is None:
            self.patterns = self._load_comm_plans("patterns") or self._generate_patterns()
            self._save_comm_plans(self.patterns, "patterns")
        if reference_layout and self.comm_plans is None:
            self.comm_plans = self._load_comm_plans("comm_plans") or self._generate_comm_plans(exchange)
            self._save_comm_plans(self.comm_plans, "comm_plans")
        return self


class SyntheticHeterogeneousDataset(DistributedHeteroGraphDataset):
    """MAG240M-shaped random heterogeneous graph (``synthetic_dataset.py:29-199``)."""

    def __init__(self, rank, world_size=None, num_papers: int = 32768, num_authors: int = 16384,
                 num_institutions: int = 256, num_features: int = 16, num_classes: int = 153, seed: int = 0, comm=None,
                 cache_dir: Optional[str] = None, dtype=torch.float32):
        """Two calling forms: ``(rank, world_size, num_papers, ...)`` and the reference's
        ``(synthetic_config, comm, cached_comm_plans=None)`` (``synthetic_dataset.py:80-85``), where ``synthetic_config`` has the
        attributes ``num_papers``, ``num_authors``, ``num_institutions``, ``num_features``, ``num_classes``."""
        if hasattr(rank, "num_papers"):
            cfg, comm = rank, (world_size if world_size is not None else comm)
            cached = num_papers if isinstance(num_papers, str) else None  # third positional argument: plan-cache location
            cache_dir = cache_dir or (os.path.dirname(cached) or "." if cached else None)
            num_papers, num_authors, num_institutions = cfg.num_papers, cfg.num_authors, cfg.num_institutions
            num_features, num_classes = cfg.num_features, cfg.num_classes
            rank, world_size = comm.get_rank(), comm.get_world_size()
        g = torch.Generator().manual_seed(seed)
        n = [num_papers, num_authors, num_institutions]

        def rel(ns, nd, count):
            count = max(int(count), 1)
            e = torch.stack([torch.randint(0, ns, (count,), generator=g), torch.randint(0, nd, (count,), generator=g)], 1)
            return torch.unique(e, dim=0)

        cites = rel(num_papers, num_papers, num_papers * 11)  # paper -> paper, avg degree 11
        writes = rel(num_authors, num_papers, num_papers * 3.5)  # author -> paper
        affil = rel(num_authors, num_institutions, num_authors * 0.35)  # author -> institution
        edges = [cites, writes.flip(1), writes, affil, affil.flip(1)]  # order of RELATIONS
        offs = [node_offsets(k, world_size) for k in n]
        feats = []
        for t, k in enumerate(n):
            full = torch.randn(k, num_features, generator=g)
            feats.append(full[offs[t][rank]:offs[t][rank + 1]].to(dtype).contiguous())
        labels_full = torch.randint(0, num_classes, (num_papers,), generator=g)
        perm = torch.randperm(num_papers, generator=g)
        n_tr, n_va = int(0.7 * num_papers), int(0.15 * num_papers)
        split = {"train": perm[:n_tr].sort().values, "valid": perm[n_tr:n_tr + n_va].sort().values,
                 "test": perm[n_tr + n_va:].sort().values}
        super().__init__(rank, world_size, n, feats, edges, labels_full[offs[0][rank]:offs[0][rank + 1]].contiguous(),
                         num_classes, comm=comm, split=split, cache_dir=cache_dir)


def _edges_touching(ei, src_range, dst_range, chunk: int = 1 << 26) -> torch.Tensor:
    """Rows ``[E', 2]`` of a (memory-mapped) ``[2, E]`` edge index whose source lies in ``src_range`` or whose destination lies
    in ``dst_range`` -- everything one rank's plans can depend on -- read in chunks so the 1.3 B-edge citation list is never
    resident as a whole."""
    import numpy as np

    (slo, shi), (dlo, dhi) = src_range, dst_range
    keep = []
    E = int(ei.shape[1])
    for o in range(0, E, chunk):
        blk = np.asarray(ei[:, o:o + chunk])
        m = ((blk[0] >= slo) & (blk[0] < shi)) | ((blk[1] >= dlo) & (blk[1] < dhi))
        if m.any():
            keep.append(torch.from_numpy(np.ascontiguousarray(blk[:, m].T).astype(np.int64)))
    return torch.cat(keep, 0) if keep else torch.zeros(0, 2, dtype=torch.long)


class DGraph_MAG240M_Dataset(DistributedHeteroGraphDataset):
    """Real OGB-LSC MAG240M from the processed files on disk (``MAG240M_dataset.py:116-340``), without the ``ogb`` package:
    ``data/ogb_raw.MAG240MFiles`` memory-maps ``<data_dir>/mag240m_kddcup2021/processed``.

    Every rank keeps its contiguous block of each node type and only the edges that touch it.  Paper features are this rank's
    rows of ``paper/node_feat.npy``; author / institution features follow ``features``: ``"mean"`` = average of the written
    papers' / affiliated authors' features (what the reference computes with ``torch_sparse``), ``"random"`` = unit normal of
    the same width (what the reference currently trains with: ``randn(..., 1)``, ``MAG240M_dataset.py:150-170``)."""

    def __init__(self, rank, world_size=None, data_dir="data/MAG240M", comm=None, features: str = "random", feature_dim: int = 0,
                 cache_dir: Optional[str] = None, dtype=torch.float32, seed: int = 0, paper_rank_mappings=None,
                 author_rank_mappings=None, institution_rank_mappings=None, **kw):
        """Two calling forms: ``(rank, world_size, data_dir, comm=...)`` and the reference's
        ``(comm, data_dir, comm_plan_only=True, paper_rank_mappings=None, ...)`` (``MAG240M_dataset.py:119-130``)."""
        from dgraph_b200.data.ogb_raw import MAG240MFiles, find_mag240m

        if hasattr(rank, "get_rank"):  # reference form: the communicator comes first, the data directory second
            comm = rank
            if isinstance(world_size, str):
                data_dir = world_size
            rank, world_size = comm.get_rank(), comm.get_world_size()

        if find_mag240m(data_dir) is None:
            raise RuntimeError(f"no processed MAG240M files under {data_dir} (expected mag240m_kddcup2021/processed/...); there is "
                               "no network access here -- use SyntheticHeterogeneousDataset for MAG240M-shaped runs")
        f = MAG240MFiles(data_dir)
        n = [f.num_papers, f.num_authors, f.num_institutions]
        offs = [offsets_from_rank_mapping(m, k, world_size)
                for m, k in zip((paper_rank_mappings, author_rank_mappings, institution_rank_mappings), n)]
        rng = [(int(o[rank]), int(o[rank + 1])) for o in offs]
        cites = _edges_touching(f.edge_index("cites"), rng[0], rng[0])
        writes = _edges_touching(f.edge_index("writes"), rng[1], rng[0])  # author -> paper
        affil = _edges_touching(f.edge_index("affiliated_with"), rng[1], rng[2])  # author -> institution
        edges = [cites, writes.flip(1), writes, affil, affil.flip(1)]  # order of RELATIONS
        g = torch.Generator().manual_seed(seed + rank)
        if f.paper_feat is not None and features == "mean":
            F = int(f.paper_feat.shape[1])
            paper = f.paper_feat_rows(*rng[0], dtype=dtype)

            def mean_of(dst_lo, dst_hi, e, fetch):  # e: [E, 2] (dst-type id, src-type id); rows of the source via fetch(ids)
                sel = e[(e[:, 0] >= dst_lo) & (e[:, 0] < dst_hi)]
                out = torch.zeros(dst_hi - dst_lo, F, dtype=torch.float32)
                cnt = torch.zeros(dst_hi - dst_lo, dtype=torch.float32)
                if sel.numel():
                    out.index_add_(0, sel[:, 0] - dst_lo, fetch(sel[:, 1]).float())
                    cnt.index_add_(0, sel[:, 0] - dst_lo, torch.ones(sel.shape[0]))
                return out / cnt.clamp_(min=1).unsqueeze(1)

            def fetch_paper(ids):
                import numpy as np

                u, inv = torch.unique(ids, return_inverse=True)  # sorted unique ids: monotone reads from the memory map
                rows = torch.from_numpy(np.asarray(f.paper_feat[u.numpy()]).astype("float32"))
                return rows[inv]

            author = mean_of(*rng[1], writes, fetch_paper)

            def fetch_author(ids):  # authors of other ranks: recompute their mean from the papers they wrote
                import numpy as np

                u, inv = torch.unique(ids, return_inverse=True)
                w = np.asarray(f.edge_index("writes"))
                m = np.isin(w[0], u.numpy())
                ee = torch.from_numpy(np.ascontiguousarray(w[:, m].T).astype("int64"))
                pos = torch.searchsorted(u, ee[:, 0].contiguous())
                out = torch.zeros(u.numel(), F)
                cnt = torch.zeros(u.numel())
                if ee.numel():
                    out.index_add_(0, pos, fetch_paper(ee[:, 1]))
                    cnt.index_add_(0, pos, torch.ones(ee.shape[0]))
                return (out / cnt.clamp_(min=1).unsqueeze(1))[inv]

            inst = mean_of(*rng[2], affil.flip(1), fetch_author)
            feats = [paper, author.to(dtype), inst.to(dtype)]
        else:
            F = feature_dim or (int(f.paper_feat.shape[1]) if f.paper_feat is not None else 1)
            feats = [f.paper_feat_rows(*rng[0], dtype=dtype) if (f.paper_feat is not None and not feature_dim)
                     else torch.randn(rng[0][1] - rng[0][0], F, generator=g).to(dtype)]
            feats += [torch.randn(hi - lo, F, generator=g).to(dtype) for lo, hi in rng[1:]]
        labels = f.labels()[rng[0][0]:rng[0][1]].contiguous()
        sp = f.split()
        split = {"train": sp["train"], "valid": sp.get("valid", sp.get("val")), "test": sp.get("test-dev", sp.get("test"))}
        super().__init__(rank, world_size, n, feats, edges, labels, f.num_classes, comm=comm, split=split, cache_dir=cache_dir,
                         offsets=offs)
        self.files = f

    def generate_feature_data(self):
        """Rank 0 writes the author / institution feature files (means of neighbouring papers' / authors' features) next to the
        processed dataset unless they exist, everybody waits (``MAG240M_dataset.py:262-320``).  Returns their paths."""
        from dgraph_b200.data.ogb_raw import generate_mag240m_feature_files

        paths = None
        if self.rank == 0:
            paths = generate_mag240m_feature_files(self.files)
        if self.comm is not None:
            self.comm.barrier()
        return paths or {"author": os.path.join(self.files.dir, "author_feat.npy"),
                         "institution": os.path.join(self.files.dir, "institution_feat.npy")}
