"""Reference file name (``data_utils/spatial_utils.py:21-272``): lat/lon <-> xyz and rotated-displacement edge features."""
from dgraph_b200.data.graphcast_graph import edge_features, latlon_grid, sph2xyz, xyz2sph  # noqa: F401
