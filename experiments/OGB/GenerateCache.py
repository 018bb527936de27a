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


def generate_comm_plan(coo_list, offsets, rank, world_size, dest_offsets=None):
    """Edge-conditioned plan of one rank from the global COO list (reference ``GenerateCache.py:39-63``): a rank holds the edges
    whose source lies in its vertex range.  No process group is needed -- the requests of the other ranks are derived from the
    global list (``comm_plan.offline_exchange``), so one process can write the plans of every rank."""
    from dgraph_b200.parallel.comm_plan import COO_to_NCCLCommPlan, NCCLEdgeConditionedGraphCommPlan, offline_exchange

    src, dst = coo_list[0].long(), coo_list[1].long()
    d_off = offsets if dest_offsets is None else dest_offsets
    local_edges = torch.nonzero((src >= offsets[rank]) & (src < offsets[rank + 1]), as_tuple=True)[0]
    sp = COO_to_NCCLCommPlan(rank, world_size, src, local_edges, offsets,
                             exchange=offline_exchange(src, src, offsets, offsets, rank, world_size))
    dp = COO_to_NCCLCommPlan(rank, world_size, dst, local_edges, d_off,
                             exchange=offline_exchange(src, dst, offsets, d_off, rank, world_size))
    return NCCLEdgeConditionedGraphCommPlan(rank=rank, world_size=world_size, source_graph_plan=sp, dest_graph_plan=dp)


def main(dset: str = "", world_size: int = 8, node_rank_placement_file: str = "", rank: int = 0, out_dir: str = "cache",
         scale: float = 1.0, dataset: str = "arxiv"):
    """Argument order of the reference (``GenerateCache.py:66``): ``dset`` (e.g. ``ogbn-arxiv``), ``world_size``,
    ``node_rank_placement_file``; ``--dataset arxiv`` is the short form used by the other entry points here."""
    dataset = dset or dataset
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
    if bool((part[1:] >= part[:-1]).all()):  # contiguous placement: also the vector-addressed (CommPlan) flavour
        offs = torch.zeros(world_size + 1, dtype=torch.long)
        offs[1:] = torch.bincount(part, minlength=world_size).cumsum(0)
        plan = generate_comm_plan(e.t(), offs, rank, world_size)
        torch.save(plan, os.path.join(out_dir, f"{dataset}_rank_{rank}_of_{world_size}_comm_plans.pt"))
    print(f"rank {rank}/{world_size}: local {pat.num_local_vertices} halo {pat.num_halo_vertices} edges {pat.num_local_edges} "
          f"plan {pat.memory_usage('MB')['total']:.1f} MB -> {path}")


if __name__ == "__main__":
    run_cli(main)
