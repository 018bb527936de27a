"""Reference file name (``experiments/GraphCast/model.py:24-394``): the model lives in ``dgraph_b200.models.graphcast``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.graphcast import (DGraphCast, GraphCastDecoder, GraphCastEmbedder, GraphCastEncoder,  # noqa: F401,E402
                                          GraphCastProcessor)

__all__ = ["GraphCastEmbedder", "GraphCastEncoder", "GraphCastProcessor", "GraphCastDecoder", "DGraphCast"]
