"""Reference file name (``experiments/OGB-LSC/RGAT.py:58-382``): the models live in ``dgraph_b200.models.rgat``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.rgat import CommAwareGAT, CommAwareRGAT, ConvLayer  # noqa: F401,E402

__all__ = ["ConvLayer", "CommAwareGAT", "CommAwareRGAT"]
