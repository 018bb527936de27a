"""Reference package name (``experiments/OGB-LSC/lsc_datasets``): the datasets live in ``dgraph_b200.data.hetero``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import _common  # noqa: F401,E402

from dgraph_b200.data.hetero import (DGraph_MAG240M_Dataset, DistributedHeteroGraphDataset,  # noqa: F401,E402
                                     SyntheticHeterogeneousDataset)

__all__ = ["DistributedHeteroGraphDataset", "SyntheticHeterogeneousDataset", "DGraph_MAG240M_Dataset"]
