from dgraph_b200.parallel.collective_engine import TorchDistEngine as NCCLBackendEngine, TIMINGS
from dgraph_b200.parallel.comm_plan import (
    NCCLGraphCommPlan, NCCLEdgeConditionedGraphCommPlan, COO_to_NCCLCommPlan, COO_to_NCCLEdgeConditionedCommPlan,
)
from dgraph_b200.parallel.legacy_cache import (  # noqa: F401
    NCCLGatherCache, NCCLScatterCache, NCCLGatherCacheGenerator, NCCLScatterCacheGenerator,
)

__all__ = ["NCCLBackendEngine", "TIMINGS", "NCCLGraphCommPlan", "NCCLEdgeConditionedGraphCommPlan", "COO_to_NCCLCommPlan",
           "COO_to_NCCLEdgeConditionedCommPlan"]
