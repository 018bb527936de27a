"""Reference file name (``experiments/OGB/ogb_comm_dataset.py:54-145``): ``ds[0] -> (local_feats, local_labels, comm_pattern)``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.data.ogbn_datasets import DGraphOGBDataset  # noqa: F401,E402

__all__ = ["DGraphOGBDataset"]
