"""OGB node-property datasets, distributed (parity: reference ``DGraph/data/ogbn_datasets.py:25-148`` and
``experiments/OGB/ogb_comm_dataset.py:54-145``).

Real data: if the OGB raw files are on disk (``<root>/ogbn_products/raw/*.csv.gz`` + ``split/``), ``data/ogb_raw.py`` parses them
directly; else the ``ogb`` package is tried; ``ogb`` is not installed here and there is no network, so otherwise a synthetic graph of the same
shape (``dgraph_b200.data.synthetic.SHAPES``) is generated instead -- every other step (rank-0-first preprocessing with
barriers, ``torch.save`` cache at ``{dir}/{dname}_graph_data_{W}.pt``, ``__getitem__`` layout) follows the reference."""
from __future__ import annotations

import os
from typing import Optional

import torch

from dgraph_b200.data.graph import DistributedGraph
from dgraph_b200.data.preprocess import process_homogenous_data
from dgraph_b200.data.synthetic import SHAPES, make_graph, node_features_and_labels
from dgraph_b200.parallel.comm_pattern import build_communication_pattern

SUPPORTED_DATASETS = ["ogbn-arxiv", "ogbn-proteins", "ogbn-papers100M", "ogbn-products"]
num_classes = {"ogbn-arxiv": 40, "ogbn-proteins": 112, "ogbn-papers100M": 172, "ogbn-products": 47}


def _load_raw(dname: str, root: str, scale: float = 1.0, seed: int = 0):
    from dgraph_b200.data import ogb_raw

    if ogb_raw.find_ogbn(dname, root) is not None:  # the files are on disk: parse them ourselves, no `ogb` package needed
        g, y, split = ogb_raw.load_ogbn_raw(dname, root)
        g = {"node_feat": g["node_feat"], "edge_index": g["edge_index"], "num_nodes": g["num_nodes"]}
        return g, y, split, False
    try:
        from ogb.nodeproppred import NodePropPredDataset  # type: ignore

        ds = NodePropPredDataset(dname, root=root)
        g, y = ds[0]
        return g, torch.as_tensor(y).reshape(-1), ds.get_idx_split(), False
    except Exception:  # noqa: BLE001 - no ogb / no files: synthetic stand-in of the same shape
        sg = make_graph(dname if dname in SHAPES else "tiny", seed=seed, scale=scale)
        x, y = node_features_and_labels(sg.num_nodes, sg.num_features, sg.num_classes, seed=seed)
        perm = torch.randperm(sg.num_nodes, generator=torch.Generator().manual_seed(seed))
        a, b = int(0.6 * sg.num_nodes), int(0.8 * sg.num_nodes)
        split = {"train": perm[:a], "valid": perm[a:b], "test": perm[b:]}
        return {"node_feat": x, "edge_index": sg.edge_list.t().contiguous(), "num_nodes": sg.num_nodes}, y, split, True


class DistributedOGBWrapper:
    def __init__(self, dname: str, comm_object, dir_name: str = "data", node_rank_placement: Optional[torch.Tensor] = None,
                 force_reprocess: bool = False, scale: float = 1.0):
        assert dname in SUPPORTED_DATASETS or dname == "tiny", f"Dataset {dname} not supported"
        self.dname, self.comm, self.num_classes = dname, comm_object, num_classes.get(dname, SHAPES["tiny"][3])
        self._rank, self._world = comm_object.get_rank(), comm_object.get_world_size()
        os.makedirs(dir_name, exist_ok=True)
        cache = os.path.join(dir_name, f"{dname}_graph_data_{self._world}.pt")
        if self._rank == 0 and (force_reprocess or not os.path.exists(cache)):
            g, y, split, self.synthetic = _load_raw(dname, dir_name, scale)
            graph = process_homogenous_data(g, y, self._rank, self._world, split, node_rank_placement)
            torch.save(graph, cache)
        comm_object.barrier()
        self.graph_obj: DistributedGraph = torch.load(cache, weights_only=False)

    def __len__(self):
        return 1

    def __getitem__(self, idx: int):
        g, r = self.graph_obj, self._rank
        return (g.get_local_node_features(r), g.get_global_edge_indices(), g.get_global_rank_mappings(), g.get_local_labels(r))


class DGraphOGBDataset:
    """Halo-plan flavour used by the GCN trainer: ``ds[0] -> (local_feats, local_labels, comm_pattern)``."""

    def __init__(self, dname: str, comm, node_rank_placement: Optional[torch.Tensor] = None, root_dir: Optional[str] = "data",
                 *args, scale: float = 1.0, seed: int = 0, device="cpu", partition: str = "contiguous", **kwargs):
        """``partition`` (used when no placement tensor is given): ``"contiguous"`` blocks of vertex ids, ``"multilevel"`` = the
        native multilevel partitioner (deterministic, so every rank computes the same placement), ``"round_robin"`` = the
        reference's default (``ogb_comm_dataset.py:54-90``)."""
        self.comm, self.rank, self.world = comm, comm.get_rank(), comm.get_world_size()
        name = {"arxiv": "ogbn-arxiv", "products": "ogbn-products"}.get(dname, dname)
        g, y, split, self.synthetic = _load_raw(name, root_dir or "dataset", scale, seed)
        self.num_classes = num_classes.get(name, int(y.max()) + 1)
        V = int(g["num_nodes"])
        edges = torch.as_tensor(g["edge_index"]).t().contiguous().long().to(device)
        if node_rank_placement is None and partition == "multilevel" and self.world > 1:
            from dgraph_b200.data.partition import multilevel

            node_rank_placement = multilevel(edges.cpu(), V, self.world)
        elif node_rank_placement is None and partition == "round_robin":
            node_rank_placement = torch.arange(V) % self.world
        elif node_rank_placement is None:
            # rank-contiguous blocks: what a METIS + renumber pass yields (the reference defaults to round robin)
            node_rank_placement = (torch.arange(V) * self.world // V).clamp(max=self.world - 1)
        part = node_rank_placement.to(device)
        self.comm_pattern = build_communication_pattern(edges, part, self.rank, self.world)
        rows = self.comm_pattern.local_vertices_global.cpu()
        self.local_features = torch.as_tensor(g["node_feat"])[rows].float()
        self.local_labels = y[rows].long()
        self._masks = {}
        for k_in, k_out in (("train", "train_mask"), ("valid", "val_mask"), ("test", "test_mask")):
            m = torch.zeros(V, dtype=torch.bool)
            m[torch.as_tensor(split[k_in]).long()] = True
            self._masks[k_out] = m[rows]
        self.num_nodes, self.num_edges = V, int(edges.shape[0])

    def get_masks(self):
        return self._masks

    def __len__(self):
        return 1

    def __getitem__(self, index):
        return self.local_features, self.local_labels, self.comm_pattern
