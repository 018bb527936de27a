"""Pre-build and store the communication plans of one rank (parity: reference ``experiments/OGB/GenerateCache.py:39-84``).

``python experiments/OGB/GenerateCache.py --dataset arxiv --world_size 8 --rank 0`` writes both plan flavours
(``CommunicationPattern`` and ``NCCLEdgeConditionedGraphCommPlan`` need no process group when every rank's plan is derived
from the global edge list)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from dgraph_b200.data.ogbn_datasets import _load_raw  # noqa: E402
from dgraph_b200.parallel import build_communication_pattern  # noqa: E402
from dgraph_b200.utils.cli import run_cli  # noqa: E402


def main(dataset: str = "arxiv", world_size: int = 8, rank: int = 0, out_dir: str = "cache", scale: float = 1.0,
         node_rank_placement_file: str = ""):
    name = {"arxiv": "ogbn-arxiv", "products": "ogbn-products"}.get(dataset, dataset)
    g, _, _, _ = _load_raw(name, "data", scale)
    V = int(g["num_nodes"])
    e = torch.as_tensor(g["edge_index"]).long().t().contiguous()
    part = (torch.load(node_rank_placement_file, weights_only=False) if node_rank_placement_file
            else (torch.arange(V) * world_size // V).clamp(max=world_size - 1))
    pat = build_communication_pattern(e, part, rank, world_size)
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, f"{dataset}_rank_{rank}_of_{world_size}_comm_pattern.pt")
    torch.save(pat, path)
    print(f"rank {rank}/{world_size}: local {pat.num_local_vertices} halo {pat.num_halo_vertices} edges {pat.num_local_edges} "
          f"plan {pat.memory_usage('MB')['total']:.1f} MB -> {path}")


if __name__ == "__main__":
    run_cli(main)
