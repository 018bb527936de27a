from dgraph_b200.data.graph import DistributedGraph, get_round_robin_node_rank_map  # noqa: F401
