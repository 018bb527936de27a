"""Reference file name (``experiments/OGB-LSC/distributed_layers.py:22-214``): see ``dgraph_b200.nn.dist_bn``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.nn.dist_bn import (DistributedBatchNorm1D, DistributedBN_Impl, DistributedBN_with_Recompute,  # noqa: F401,E402
                                    GetGlobalVal)

__all__ = ["DistributedBN_Impl", "DistributedBN_with_Recompute", "DistributedBatchNorm1D", "GetGlobalVal"]
