"""Reference package name (``experiments/GraphCast/data_utils``): geometry and graph construction live in
``dgraph_b200.data.graphcast_graph``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import _common  # noqa: F401,E402

