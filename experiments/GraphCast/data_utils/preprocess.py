"""Reference file name (``data_utils/preprocess.py:5-30``: METIS partition of the mesh): the native multilevel partitioner."""
from dgraph_b200.data.partition import edge_cut_fraction, multilevel  # noqa: F401


def partition_mesh(mesh_edges, num_mesh_vertices: int, world_size: int, vertex_weights=None):
    """Vertex placement of the multimesh for ``world_size`` ranks."""
    return multilevel(mesh_edges, num_mesh_vertices, world_size, vertex_weights=vertex_weights)
