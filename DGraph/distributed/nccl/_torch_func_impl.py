from dgraph_b200.parallel.functional import CommPlan_GatherFunction, CommPlan_ScatterFunction  # noqa: F401
