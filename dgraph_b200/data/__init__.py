from dgraph_b200.data.synthetic import (
    SHAPES, SyntheticGraph, make_graph, symmetric_locality_graph, contiguous_partition, contiguous_offsets,
    node_features_and_labels,
)
