"""Reference module path ``DGraph/distributed/mpi/MPIBackendEngine.py``.  There is no mpi4py dependency: ``backend="mpi"`` is served
by the same engines as the other backends; the one-sided gather / scatter functions map to the plan-based autograd functions."""
from DGraph.distributed.mpi import MPIBackendEngine  # noqa: F401
from dgraph_b200.parallel.functional import CommPlan_GatherFunction as MPIGatherFunction  # noqa: F401
from dgraph_b200.parallel.functional import CommPlan_ScatterFunction as MPIScatterFunction  # noqa: F401
