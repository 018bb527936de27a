"""Reference module path ``DGraph/distributed/nccl/NCCLBackendEngine.py``."""
from DGraph.distributed.nccl import NCCLBackendEngine, TIMINGS  # noqa: F401
