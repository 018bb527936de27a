"""Reference file name (``experiments/OGB/GCN.py:28-118``): the models live in ``dgraph_b200.models.gcn``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.gcn import CommAwareGCN, GraphConvLayer  # noqa: F401,E402

__all__ = ["GraphConvLayer", "CommAwareGCN"]
