"""Partition an OGB(-shaped) graph and store the vertex placement (parity: reference ``experiments/OGB/preprocess.py:15-131``,
which uses networkx + METIS; here: the native multilevel k-way partitioner ``csrc/host/partition.cpp`` (default) or the
PyTorch BFS / label-propagation block partitioner of ``dgraph_b200.data.partition``).

``python experiments/OGB/preprocess.py --dataset products --world_size 8 --out placement_products_8.pt``
The file is what ``main.py --node_rank_placement_file`` loads."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from dgraph_b200.data.ogbn_datasets import _load_raw  # noqa: E402
from dgraph_b200.data.partition import bfs_blocks, contiguous_blocks, edge_cut_fraction, hashed, multilevel  # noqa: E402
from dgraph_b200.utils.cli import run_cli  # noqa: E402


def add_opposite_edges(edge_indices: torch.Tensor) -> torch.Tensor:
    """Symmetrise a ``[2, E]`` edge index and drop duplicates."""
    e = torch.cat([edge_indices, edge_indices.flip(0)], dim=1)
    return torch.unique(e, dim=1)


def node_renumbering(node_rank_placement, num_parts):
    """Old vertex ids listed rank by rank: position = new id, value = old id (reference ``preprocess.py:50-59``)."""
    pl = torch.as_tensor(node_rank_placement).long()
    return torch.argsort(pl, stable=True)


def edge_renumbering(edge_indices, renumbered_nodes):
    """``[E, 2]`` edges in old ids -> ``[2, E]`` edges in new ids.  ``renumbered_nodes`` is the new -> old list from
    :func:`node_renumbering`; the edges need its inverse (the reference indexes with the list itself, which is only right when the
    permutation is its own inverse)."""
    new_to_old = torch.as_tensor(renumbered_nodes).long()
    old_to_new = torch.empty_like(new_to_old)
    old_to_new[new_to_old] = torch.arange(new_to_old.numel())
    e = torch.as_tensor(edge_indices).long()
    return torch.stack([old_to_new[e[:, 0]], old_to_new[e[:, 1]]])


def partition_graph(coo_list, num_ranks: int):
    """Partition an undirected graph given as ``[E, 2]`` pairs and renumber it rank-contiguously: ``(new -> old vertex list,
    [2, E] edges in new ids)``.  METIS in the reference (``preprocess.py:15-29``), the native multilevel partitioner here."""
    e = torch.as_tensor(coo_list).long()
    V = int(e.max()) + 1 if e.numel() else 0
    placement = multilevel(e, V, num_ranks)
    nodes = node_renumbering(placement, num_ranks)
    return nodes, edge_renumbering(e, nodes)


def partition_directed_graph(coo_list, num_nodes: int, num_parts: int):
    """Same for a directed edge list with isolated vertices allowed (``preprocess.py:32-47``); direction is ignored for the cut."""
    e = torch.as_tensor(coo_list).long()
    placement = multilevel(e, num_nodes, num_parts)
    nodes = node_renumbering(placement, num_parts)
    return nodes, edge_renumbering(e, nodes)


def save_networkx_graph(coo_list, num_nodes, dname, directed=False):
    """Pickle the graph as a networkx object ``{dname}_directed={directed}.pkl`` (``preprocess.py:83-99``)."""
    import pickle

    import networkx as nx

    G = nx.DiGraph() if directed else nx.Graph()
    G.add_nodes_from(range(int(num_nodes)))
    pairs = [tuple(map(int, p)) for p in torch.as_tensor(coo_list).tolist()]
    G.add_edges_from(pairs)
    if directed:
        G.add_edges_from([(d, s_) for s_, d in pairs])
    with open(f"{dname}_directed={directed}.pkl", "wb") as f:
        pickle.dump(G, f)
    return G


def load_networkx_graph(dname):
    import pickle

    with open(f"{dname}.pkl", "rb") as f:
        return pickle.load(f)


def main(dset_name: str = "", world_size: int = 8, out: str = "", method: str = "multilevel", scale: float = 1.0,
         root_dir: str = "data", dataset: str = "arxiv"):
    """``dset_name`` is the reference's (positional) argument, e.g. ``ogbn-arxiv`` (``preprocess.py:108``); ``--dataset arxiv`` is the
    short form used by the other entry points here."""
    dataset = dset_name or dataset
    name = {"arxiv": "ogbn-arxiv", "products": "ogbn-products"}.get(dataset, dataset)
    g, _, _, synthetic = _load_raw(name, root_dir, scale)
    V = int(g["num_nodes"])
    e = add_opposite_edges(torch.as_tensor(g["edge_index"]).long()).t().contiguous()
    part = {"multilevel": lambda: multilevel(e, V, world_size), "bfs": lambda: bfs_blocks(e, V, world_size), "contiguous": lambda: contiguous_blocks(V, world_size),
            "hash": lambda: hashed(V, world_size)}[method]()
    out = out or f"placement_{dataset}_{world_size}.pt"
    torch.save(part, out)
    print(f"{name}{' (synthetic)' if synthetic else ''}: V={V} E={e.shape[0]} parts={world_size} "
          f"edge cut {edge_cut_fraction(e, part):.3f} sizes {torch.bincount(part, minlength=world_size).tolist()} -> {out}")


if __name__ == "__main__":
    run_cli(main)
