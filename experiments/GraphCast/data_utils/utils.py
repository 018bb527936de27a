"""Reference file name (``data_utils/utils.py:30-180``): bipartite grid<->mesh edge construction."""
from dgraph_b200.data.graphcast_graph import grid2mesh_edges, mesh2grid_edges  # noqa: F401
