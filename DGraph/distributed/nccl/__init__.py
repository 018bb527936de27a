from dgraph_b200.parallel.collective_engine import CollectiveEngine, TorchDistEngine, TIMINGS


class _AnyCollectiveEngine(type(TorchDistEngine)):
    """Reference models assert ``isinstance(comm._Communicator__backend_engine, NCCLBackendEngine)``
    (``experiments/OGB-LSC/RGAT.py:189``).  Every engine of this framework that offers the plan gather / scatter API (NCCL, gloo,
    peer memory, the in-process emulator, the single-process engine) passes that check."""

    def __instancecheck__(cls, obj):
        return isinstance(obj, CollectiveEngine)


class NCCLBackendEngine(TorchDistEngine, metaclass=_AnyCollectiveEngine):
    """Constructing it gives the ``torch.distributed`` engine (NCCL on GPUs, gloo on CPU).  First positional argument:
    ``ranks_per_graph``, as in the reference (``NCCLBackendEngine.py:43``)."""

    def __init__(self, ranks_per_graph: int = -1, *args, **kwargs):
        super().__init__(ranks_per_graph=ranks_per_graph, **kwargs)

from dgraph_b200.parallel.comm_plan import (
    NCCLGraphCommPlan, NCCLEdgeConditionedGraphCommPlan, COO_to_NCCLCommPlan, COO_to_NCCLEdgeConditionedCommPlan,
)
from dgraph_b200.parallel.legacy_cache import (  # noqa: F401
    NCCLGatherCache, NCCLScatterCache, NCCLGatherCacheGenerator, NCCLScatterCacheGenerator,
)

__all__ = ["NCCLBackendEngine", "TIMINGS", "NCCLGraphCommPlan", "NCCLEdgeConditionedGraphCommPlan", "COO_to_NCCLCommPlan",
           "COO_to_NCCLEdgeConditionedCommPlan"]
