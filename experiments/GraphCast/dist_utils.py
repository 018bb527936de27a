"""Reference file name (``experiments/GraphCast/dist_utils.py:8-129``)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.utils.dist_utils import CommAwareDistributedSampler, SingleProcessDummyCommunicator  # noqa: F401,E402

__all__ = ["SingleProcessDummyCommunicator", "CommAwareDistributedSampler"]
