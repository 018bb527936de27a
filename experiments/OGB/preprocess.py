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


def add_opposite_edges(edge_index: torch.Tensor) -> torch.Tensor:
    """Symmetrise a ``[2, E]`` edge index and drop duplicates."""
    e = torch.cat([edge_index, edge_index.flip(0)], dim=1)
    return torch.unique(e, dim=1)


def main(dataset: str = "arxiv", world_size: int = 8, out: str = "", method: str = "multilevel", scale: float = 1.0, root_dir: str = "data"):
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
