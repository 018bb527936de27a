from dgraph_b200.data.synthetic import (
    SHAPES, SyntheticGraph, make_graph, symmetric_locality_graph, contiguous_partition, contiguous_offsets,
    node_features_and_labels,
)
from dgraph_b200.data.graph import DistributedGraph, get_round_robin_node_rank_map
from dgraph_b200.data.ogbn_datasets import DistributedOGBWrapper, DGraphOGBDataset
