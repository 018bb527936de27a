"""``backend="mpi"`` without mpi4py: on a CPU-only box the one-sided shared-memory window engine (remote-memory-access epochs like
the reference's ``MPI.Win`` code, ``dgraph_b200/parallel/shm_engine.py``), otherwise the ``torch.distributed`` engine."""
import torch as _torch

from dgraph_b200.parallel import shm_engine as _shm
from dgraph_b200.parallel.collective_engine import TorchDistEngine as _TorchDistEngine

_Base = _shm.ShmWindowEngine if (not _torch.cuda.is_available() and _shm.available()) else _TorchDistEngine


class MPIBackendEngine(_Base):
    """The engine under the MPI engine's name and constructor order (``MPIBackendEngine.py:268``: ``ranks_per_graph`` first).
    ``scatter`` accepts the MPI argument order ``(tensor, indices, num_output_rows, rank_mapping)``."""

    def __init__(self, ranks_per_graph: int = -1, *args, **kwargs):
        super().__init__(ranks_per_graph=ranks_per_graph, **kwargs)
