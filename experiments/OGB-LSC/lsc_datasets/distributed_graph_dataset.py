"""Reference file name (``lsc_datasets/distributed_graph_dataset.py:24-489``)."""
from . import DistributedHeteroGraphDataset  # noqa: F401
from dgraph_b200.data.hetero import NODE_TYPES, RELATIONS, node_offsets  # noqa: F401
