"""``backend="mpi"``: served by the gloo / peer-memory engines (mpi4py is not required)."""
from dgraph_b200.parallel.collective_engine import TorchDistEngine as _TorchDistEngine


class MPIBackendEngine(_TorchDistEngine):
    """The ``torch.distributed`` engine under the MPI engine's name and constructor order (``MPIBackendEngine.py:268``:
    ``ranks_per_graph`` first).  ``scatter`` accepts the MPI argument order ``(tensor, indices, num_output_rows, rank_mapping)``."""

    def __init__(self, ranks_per_graph: int = -1, *args, **kwargs):
        super().__init__(ranks_per_graph=ranks_per_graph, **kwargs)
