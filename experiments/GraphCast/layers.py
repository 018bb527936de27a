"""Reference file name (``experiments/GraphCast/layers.py:24-216``)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.graphcast import MeshEdgeBlock, MeshGraphMLP, MeshNodeBlock  # noqa: F401,E402

__all__ = ["MeshGraphMLP", "MeshEdgeBlock", "MeshNodeBlock"]
