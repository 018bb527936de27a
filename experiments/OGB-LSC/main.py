"""RGAT on a MAG240M-shaped heterogeneous graph (parity: reference ``experiments/OGB-LSC/main.py:30-119``).

``torchrun --nnodes 1 --nproc-per-node N experiments/OGB-LSC/main.py --dataset synthetic --num_papers 2048 ...``
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

from dgraph_b200.utils.cli import run_cli  # noqa: E402


def _load_optional_file(file_path: str):
    """``torch.load`` of a placement tensor if the file exists, else ``None`` (``main.py:24-27``)."""
    import torch

    if file_path and os.path.exists(file_path):
        return torch.load(file_path, weights_only=False)
    return None


def main(comm_type: str = "nccl", dataset: str = "synthetic", num_papers: int = 2048, num_authors: int = 512,
         num_institutions: int = 16, optimized_graph_structure: bool = True, paper_rank_mapping_file: str = "",
         author_rank_mapping_file: str = "", institution_rank_mapping_file: str = "", data_dir: str = "data/MAG240M",
         epochs: int = 20, hidden_channels: int = 64):
    assert comm_type in ("nccl", "nvshmem", "mpi")
    assert dataset in ("synthetic", "mag240m")
    from config import ModelConfig, TrainingConfig
    from Trainer import Trainer
    from dgraph_b200.data.hetero import DGraph_MAG240M_Dataset, SyntheticHeterogeneousDataset

    comm, rank, world, dev = init_distributed(comm_type)
    if dataset == "synthetic":
        ds = SyntheticHeterogeneousDataset(rank, world, num_papers, num_authors, num_institutions, comm=comm)
    else:
        ds = DGraph_MAG240M_Dataset(rank, world, data_dir=data_dir, comm=comm,
                                    paper_rank_mappings=_load_optional_file(paper_rank_mapping_file),
                                    author_rank_mappings=_load_optional_file(author_rank_mapping_file),
                                    institution_rank_mappings=_load_optional_file(institution_rank_mapping_file))
    trainer = Trainer(ds, comm, dev, ModelConfig(hidden_channels=hidden_channels), TrainingConfig(epochs=epochs),
                      fused=optimized_graph_structure).prepare_data()
    times = trainer.train()
    acc = trainer.evaluate("valid")
    if rank == 0:
        print(f"validation accuracy {acc:.4f}; mean epoch {sum(times[1:]) / max(len(times) - 1, 1):.2f} ms", flush=True)
    cleanup(comm)


if __name__ == "__main__":
    run_cli(main)
