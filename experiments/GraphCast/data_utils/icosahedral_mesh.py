"""Reference file name (``data_utils/icosahedral_mesh.py:32-300``): icosahedron, 4-way face refinement, multimesh edges."""
from dgraph_b200.data.graphcast_graph import icosahedron, multimesh, refine  # noqa: F401
