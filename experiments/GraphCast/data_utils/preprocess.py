"""Reference file name (``data_utils/preprocess.py:5-30``: networkx graph + METIS partition of the mesh).  The partition runs on
the native multilevel k-way partitioner (``csrc/host/partition.cpp``); networkx is only a container here and optional."""
import numpy as np
import torch

from dgraph_b200.data.partition import edge_cut_fraction, multilevel  # noqa: F401


def partition_mesh(mesh_edges, num_mesh_vertices: int, world_size: int, vertex_weights=None):
    """Vertex placement of the multimesh for ``world_size`` ranks."""
    return multilevel(mesh_edges, num_mesh_vertices, world_size, vertex_weights=vertex_weights)


def graphcast_graph_to_nxgraph(mesh_graph):
    """Undirected ``networkx.Graph`` of a mesh graph tuple ``(node_features, edge_features, src, dst)`` (what
    ``data_utils.utils.create_graph`` returns)."""
    import networkx as nx

    g = nx.Graph()
    g.add_edges_from(np.stack([np.asarray(mesh_graph[2]), np.asarray(mesh_graph[3])], axis=1))
    return g


def partition_graph(G, num_ranks: int) -> np.ndarray:
    """Rank of every node of ``G`` (a ``networkx`` graph whose nodes are ``0 .. N-1``, or an ``[E, 2]`` edge array) for
    ``num_ranks`` balanced parts with a small edge cut."""
    if num_ranks < 1:
        raise ValueError("Number of ranks must be greater than 0.")
    if hasattr(G, "edges"):
        n = G.number_of_nodes()
        edges = torch.as_tensor(np.asarray(list(G.edges()), dtype=np.int64).reshape(-1, 2))
    else:
        edges = torch.as_tensor(np.asarray(G, dtype=np.int64).reshape(-1, 2))
        n = int(edges.max()) + 1 if edges.numel() else 0
    if num_ranks == 1:
        return np.zeros(n, dtype=int)
    return multilevel(edges, n, num_ranks).cpu().numpy().astype(int)
