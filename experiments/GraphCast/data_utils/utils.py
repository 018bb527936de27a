"""Reference file name (``data_utils/utils.py:30-180``): node / edge features and the bipartite grid <-> mesh graphs, with the
reference's function names and return conventions, on the vectorised builders of ``dgraph_b200.data.graphcast_graph`` (KD-tree
queries and array expressions instead of per-point Python loops)."""
import numpy as np
import torch
from torch import Tensor

from dgraph_b200.data.graphcast_graph import edge_features, grid2mesh_edges, mesh2grid_edges, sph2xyz, xyz2sph  # noqa: F401


def padded_size(x: int, y: int) -> int:
    """Smallest multiple of ``y`` that is >= ``x``."""
    return (x + y - 1) // y * y


def pad_indices(_indices: Tensor, ranks_per_graph: int) -> Tensor:
    """int32 copy of ``_indices`` padded with -1 up to a multiple of ``ranks_per_graph`` (equal shares per rank)."""
    n = _indices.shape[0]
    out = torch.full((padded_size(n, ranks_per_graph),), -1, dtype=torch.int32)
    out[:n] = _indices
    return out


def _np(t) -> np.ndarray:
    return t.detach().cpu().double().numpy() if isinstance(t, Tensor) else np.asarray(t, dtype=np.float64)


def generate_node_features(node_positions: Tensor, degrees_as_radians: bool = False) -> Tensor:
    """``[N, 3]`` unit vectors -> ``(cos lat, sin lon, cos lon)`` per node, angles in radians.  The reference takes the cosine /
    sine of the angles *in degrees* (``utils.py:52-56`` calls ``xyz2latlon`` with its default unit); ``degrees_as_radians=True``
    reproduces those numbers for anyone who needs inputs identical to a checkpoint trained there."""
    lat, lon = xyz2sph(_np(node_positions))
    if degrees_as_radians:
        lat, lon = np.rad2deg(lat), np.rad2deg(lon)
    return torch.from_numpy(np.stack([np.cos(lat), np.sin(lon), np.cos(lon)], axis=-1)).to(
        node_positions.dtype if isinstance(node_positions, Tensor) else torch.float32)


def generate_edge_features(src_pos: Tensor, dst_pos: Tensor) -> Tensor:
    """``[E, 4]``: displacement of the sender seen from the receiver's local frame and its length, both divided by the longest
    displacement of this edge set."""
    out = torch.from_numpy(edge_features(_np(src_pos), _np(dst_pos)))
    return out.to(src_pos.dtype if isinstance(src_pos, Tensor) else torch.float32)


def create_graph(src_indices: Tensor, dst_indices: Tensor, pos: Tensor, to_bidirected: bool = True):
    """``(node_features, edge_features, src, dst)`` of a mesh graph; ``to_bidirected`` appends every edge's reverse."""
    if to_bidirected:
        src_indices, dst_indices = torch.cat([src_indices, dst_indices]), torch.cat([dst_indices, src_indices])
    src, dst = src_indices.long(), dst_indices.long()
    return generate_node_features(pos), generate_edge_features(pos[src], pos[dst]), src, dst


def _grid_xyz(lat_lon_grid_flat: Tensor) -> np.ndarray:
    ll = np.deg2rad(_np(lat_lon_grid_flat))
    return sph2xyz(ll[:, 0], ll[:, 1])


def create_mesh2grid_graph(lat_lon_grid_flat: Tensor, mesh_vertices: np.ndarray, mesh_faces: np.ndarray):
    """Every grid point receives from the three corners of the mesh face whose centroid is nearest:
    ``(edge_features, src_mesh_indices, dst_grid_indices)``, ``lat_lon_grid_flat`` in degrees."""
    gxyz = _grid_xyz(lat_lon_grid_flat)
    mv = np.asarray(mesh_vertices, dtype=np.float64)
    e = mesh2grid_edges(gxyz, mv, np.asarray(mesh_faces, dtype=np.int64))
    feats = torch.from_numpy(edge_features(mv[e[:, 0]], gxyz[e[:, 1]])).float()
    return feats, torch.from_numpy(e[:, 0]).long(), torch.from_numpy(e[:, 1]).long()


def create_grid2mesh_graph(max_edge_len: float, lat_lon_grid_flat: Tensor, mesh_vertices: np.ndarray):
    """Every grid point sends to those of its four nearest mesh vertices that lie within ``0.6 * max_edge_len``:
    ``(edge_features, src_grid_indices, dst_mesh_indices)``."""
    gxyz = _grid_xyz(lat_lon_grid_flat)
    mv = np.asarray(mesh_vertices, dtype=np.float64)
    e = grid2mesh_edges(gxyz, mv, float(max_edge_len))
    feats = torch.from_numpy(edge_features(gxyz[e[:, 0]], mv[e[:, 1]])).float()
    return feats, torch.from_numpy(e[:, 0]).long(), torch.from_numpy(e[:, 1]).long()
