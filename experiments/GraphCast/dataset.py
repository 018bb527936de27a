"""Reference file name (``experiments/GraphCast/dataset.py:24-226``)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.data.graphcast_graph import (DistributedGraphCastGraph, DistributedGraphCastGraphGenerator,  # noqa: F401,E402
                                              SyntheticWeatherDataset)

__all__ = ["SyntheticWeatherDataset", "DistributedGraphCastGraphGenerator", "DistributedGraphCastGraph"]
