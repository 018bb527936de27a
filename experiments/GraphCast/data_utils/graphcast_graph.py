"""Reference file name (``data_utils/graphcast_graph.py:34-437``)."""
from dgraph_b200.data.graphcast_graph import (DistributedGraphCastGraph, DistributedGraphCastGraphGenerator,  # noqa: F401
                                              GraphCastCommPatterns, GraphCastGraph, build_graphcast_comm_patterns)
