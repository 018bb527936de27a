"""``backend="mpi"``: served by the gloo / peer-memory engines (mpi4py is not required)."""
from dgraph_b200.parallel.collective_engine import TorchDistEngine as MPIBackendEngine  # noqa: F401
