"""Reference file name (``lsc_datasets/MAG240M_dataset.py:31-340``): reads the processed MAG240M files from disk, no ``ogb``."""
from . import DGraph_MAG240M_Dataset  # noqa: F401
from dgraph_b200.data.ogb_raw import MAG240MFiles, find_mag240m  # noqa: F401
