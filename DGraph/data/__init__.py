from dgraph_b200.data.graph import DistributedGraph, get_round_robin_node_rank_map  # noqa: F401
from dgraph_b200.data.ogbn_datasets import DistributedOGBWrapper  # noqa: F401
