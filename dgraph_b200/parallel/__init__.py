from dgraph_b200.parallel.comm_pattern import (
    CommunicationPattern,
    build_communication_pattern,
    build_communication_pattern_from_shard,
    build_all_patterns,
)
from dgraph_b200.parallel.halo import HaloExchange, HaloExchangeImpl, DGraphMessagePassing
from dgraph_b200.parallel.comm_plan import (
    NCCLGraphCommPlan,
    NCCLEdgeConditionedGraphCommPlan,
    COO_to_NCCLCommPlan,
    COO_to_NCCLEdgeConditionedCommPlan,
)

__all__ = [
    "CommunicationPattern",
    "build_communication_pattern",
    "build_communication_pattern_from_shard",
    "build_all_patterns",
    "HaloExchange",
    "HaloExchangeImpl",
    "DGraphMessagePassing",
    "NCCLGraphCommPlan",
    "NCCLEdgeConditionedGraphCommPlan",
    "COO_to_NCCLCommPlan",
    "COO_to_NCCLEdgeConditionedCommPlan",
]
