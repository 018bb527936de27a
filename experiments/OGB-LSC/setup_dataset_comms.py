"""Pre-build (and print) the communication plans of the heterogeneous dataset (parity: reference
``experiments/OGB-LSC/setup_dataset_comms.py:25-100``).  ``torchrun --nproc-per-node N setup_dataset_comms.py --data_dir cache``"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

from dgraph_b200.utils.cli import run_cli  # noqa: E402


def main(comm_type: str = "nccl", dataset: str = "synthetic", num_papers: int = 2048, num_authors: int = 512,
         num_institutions: int = 16, data_dir: str = "cache", reference_layout: bool = True):
    """The reference's flags (``setup_dataset_comms.py:25-32``): ``comm_type``, ``dataset`` (``synthetic`` / ``mag240m``), graph sizes
    for the synthetic case, ``data_dir`` (MAG240M location, and where the plans are cached)."""
    from dgraph_b200.data.hetero import RELATIONS, DGraph_MAG240M_Dataset, SyntheticHeterogeneousDataset

    assert comm_type in ("nccl", "nvshmem", "mpi") and dataset in ("synthetic", "mag240m")
    comm, rank, world, dev = init_distributed(comm_type)
    if dataset == "synthetic":
        ds = SyntheticHeterogeneousDataset(rank, world, num_papers, num_authors, num_institutions, comm=comm, cache_dir=data_dir)
    else:
        ds = DGraph_MAG240M_Dataset(rank, world, data_dir=data_dir, comm=comm, cache_dir=data_dir)
    ds.prepare_plans(fused=True, reference_layout=reference_layout, exchange=getattr(comm.engine, "exchange_ids", None))
    for k, (rel, pat) in enumerate(zip(RELATIONS, ds.patterns)):
        line = f"rank {rank} relation {k} {rel}: local {pat.num_local_vertices} halo {pat.num_halo_vertices} edges {pat.num_local_edges}"
        if ds.comm_plans is not None:
            p = ds.comm_plans[k].dest_graph_plan
            line += f" | CommPlan boundary edges {int(p.boundary_edge_idx.numel())} sends {sum(p.boundary_vertex_splits)}"
        print(line, flush=True)
    cleanup(comm)


if __name__ == "__main__":
    run_cli(main)
