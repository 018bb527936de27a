"""GraphCast graphs: icosahedral multimesh, lat/lon grid, grid->mesh and mesh->grid bipartite graphs, geometric features.

Covers the reference's ``experiments/GraphCast/data_utils/{icosahedral_mesh,spatial_utils,utils,graphcast_graph}.py`` and
``dataset.py`` (``SyntheticWeatherDataset``), re-implemented from the published construction rather than from that code:

* multimesh level L: icosahedron refined L times (every triangle split in 4, new vertices projected onto the unit sphere),
  edges of ALL refinement levels kept on the finest vertex set, both directions (level 6: 40,962 nodes / 655,320 edges);
* grid2mesh: every grid point is linked to its <= 4 nearest mesh vertices within 0.6 x the longest finest-level edge;
* mesh2grid: every grid point receives from the 3 vertices of the finest-level face whose centroid is nearest;
* edge features (4-d): sender position relative to the receiver, expressed in the receiver's local frame (receiver
  rotated to lon = 0, lat = 0), as [length, dx, dy, dz] / max length;  mesh node features (3-d): (cos lat, sin lon, cos lon).

Distribution: grid rows are split in contiguous latitude bands, mesh vertices follow the band of their latitude and are
renumbered rank-major; each of the three graphs gets a :class:`CommunicationPattern` whose central side is the side that
aggregates (mesh for grid2mesh, the ``src`` endpoint for the multimesh -- as the reference's node block does -- and grid for
mesh2grid), so every edge block needs exactly one halo exchange and every node block none.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from dgraph_b200.parallel.comm_pattern import CommunicationPattern, build_communication_pattern


# ------------------------------------------------------------------------------------------------ geometry
def icosahedron(face_up: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """12 unit vertices, 20 outward-oriented faces.  ``face_up`` tilts the solid about the y axis by half the supplement of its
    dihedral angle, so that a face (not an edge) is perpendicular to the z axis: the orientation of the reference's mesh
    (``data_utils/icosahedral_mesh.py:159-177``), which fixes where the lat/lon poles fall relative to the mesh."""
    phi = (1.0 + 5.0 ** 0.5) / 2.0
    v = []
    for a in (-1.0, 1.0):
        for b in (-phi, phi):
            v += [(0.0, a, b), (a, b, 0.0), (b, 0.0, a)]
    v = np.asarray(v, dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    if face_up:
        t = (np.pi - 2.0 * np.arcsin(phi / 3.0 ** 0.5)) / 2.0
        c, s = np.cos(t), np.sin(t)
        v = np.stack([c * v[:, 0] - s * v[:, 2], v[:, 1], s * v[:, 0] + c * v[:, 2]], axis=1)
    # faces = all triples of mutually adjacent vertices (edge length of the unit icosahedron ~ 1.0515)
    d = np.linalg.norm(v[:, None] - v[None], axis=-1)
    adj = (d > 1e-6) & (d < 1.2)
    faces = []
    for i in range(12):
        for j in range(i + 1, 12):
            if not adj[i, j]:
                continue
            for k in range(j + 1, 12):
                if adj[i, k] and adj[j, k]:
                    a, b, c = v[i], v[j], v[k]
                    faces.append((i, j, k) if np.dot(np.cross(b - a, c - a), a) > 0 else (i, k, j))
    return v, np.asarray(faces, dtype=np.int64)


def refine(vertices: np.ndarray, faces: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Split every triangle into four; midpoints are shared between neighbouring faces and pushed onto the sphere."""
    edges = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], axis=0)
    key = np.sort(edges, axis=1)
    uniq, inv = np.unique(key, axis=0, return_inverse=True)
    mid = vertices[uniq[:, 0]] + vertices[uniq[:, 1]]
    mid /= np.linalg.norm(mid, axis=1, keepdims=True)
    nv = vertices.shape[0]
    new_vertices = np.concatenate([vertices, mid], axis=0)
    F = faces.shape[0]
    m01, m12, m20 = nv + inv[:F], nv + inv[F:2 * F], nv + inv[2 * F:]
    a, b, c = faces[:, 0], faces[:, 1], faces[:, 2]
    new_faces = np.concatenate([np.stack([a, m01, m20], 1), np.stack([m01, b, m12], 1), np.stack([m20, m12, c], 1),
                                np.stack([m01, m12, m20], 1)], axis=0)
    return new_vertices, new_faces


def multimesh(level: int, multi: bool = True, reference_duplicates: bool = True):
    """Returns ``(vertices [V,3], finest_faces [F,3], edges [E,2] bidirected, max finest edge length)``.

    ``reference_duplicates``: the reference derives directed edges from faces (which already yields both directions) and
    then adds the reverse of every edge again, so each directed edge appears twice (level 6: 655,320 rows,
    ``experiments/GraphCast/tests/test_single_graph_data.py:21-33``).  Kept by default so graph sizes match; ``False`` gives
    the 327,660 distinct directed edges."""
    v, f = icosahedron()
    all_edges = []

    def und(fc):
        e = np.concatenate([fc[:, [0, 1]], fc[:, [1, 2]], fc[:, [2, 0]]], axis=0)
        return np.unique(np.sort(e, axis=1), axis=0)

    all_edges.append(und(f))
    for _ in range(level):
        v, f = refine(v, f)
        all_edges.append(und(f))
    finest = all_edges[-1]
    e = np.unique(np.concatenate(all_edges if multi else [finest], axis=0), axis=0)
    max_len = float(np.linalg.norm(v[finest[:, 0]] - v[finest[:, 1]], axis=1).max())
    bid = np.concatenate([e, e[:, ::-1]], axis=0)
    if reference_duplicates:
        bid = np.concatenate([bid, bid[:, ::-1]], axis=0)
    return v, f, bid, max_len


def latlon_grid(n_lat: int, n_lon: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Regular grid, latitude from +90 to -90 (inclusive), longitude from 0 (exclusive end): ``[n_lat * n_lon]`` points."""
    lat = np.linspace(90.0, -90.0, n_lat)
    lon = np.arange(n_lon) * (360.0 / n_lon)
    la, lo = np.meshgrid(lat, lon, indexing="ij")
    la, lo = np.deg2rad(la.reshape(-1)), np.deg2rad(lo.reshape(-1))
    return la, lo, sph2xyz(la, lo)


def sph2xyz(lat: np.ndarray, lon: np.ndarray) -> np.ndarray:
    return np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], axis=-1)


def xyz2sph(xyz: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    return np.arcsin(np.clip(xyz[:, 2], -1, 1)), np.arctan2(xyz[:, 1], xyz[:, 0])


def edge_features(sender_xyz: np.ndarray, receiver_xyz: np.ndarray) -> np.ndarray:
    """[E, 4]: sender relative to receiver in the receiver's local frame, [dx, dy, dz, len] / max len -- the column order of the
    reference (``data_utils/utils.py:60-82``), so that the first Linear of an edge embedder means the same in both code bases."""
    lat, lon = xyz2sph(receiver_xyz)
    cl, sl, cp, sp = np.cos(lon), np.sin(lon), np.cos(lat), np.sin(lat)
    # R = R_y(lat) @ R_z(-lon): brings the receiver to (1, 0, 0)
    def rot(p):
        x = cl * p[:, 0] + sl * p[:, 1]
        y = -sl * p[:, 0] + cl * p[:, 1]
        z = p[:, 2]
        return np.stack([cp * x + sp * z, y, -sp * x + cp * z], axis=-1)

    disp = rot(sender_xyz) - rot(receiver_xyz)
    n = np.linalg.norm(disp, axis=1, keepdims=True)
    mx = max(float(n.max()) if n.size else 1.0, 1e-12)
    return np.concatenate([disp, n], axis=1) / mx


def grid2mesh_edges(grid_xyz: np.ndarray, mesh_xyz: np.ndarray, max_edge_len: float, k: int = 4, frac: float = 0.6):
    from scipy.spatial import cKDTree

    d, idx = cKDTree(mesh_xyz).query(grid_xyz, k=k, distance_upper_bound=frac * max_edge_len)
    ok = np.isfinite(d)
    src = np.repeat(np.arange(grid_xyz.shape[0]), k)[ok.reshape(-1)]
    dst = idx.reshape(-1)[ok.reshape(-1)]
    return np.stack([src, dst], axis=1)  # grid -> mesh


def mesh2grid_edges(grid_xyz: np.ndarray, mesh_xyz: np.ndarray, faces: np.ndarray):
    from scipy.spatial import cKDTree

    cent = mesh_xyz[faces].mean(axis=1)  # flat-face centroids (not pushed onto the sphere), as in data_utils/utils.py:117-120
    _, fi = cKDTree(cent).query(grid_xyz, k=1)
    src = faces[fi].reshape(-1)
    dst = np.repeat(np.arange(grid_xyz.shape[0]), 3)
    return np.stack([src, dst], axis=1)  # mesh -> grid


# ------------------------------------------------------------------------------------------------ containers
@dataclass
class GraphCastGraph:
    """Global (single-process) GraphCast graph; indices are global ids."""
    num_grid: int
    num_mesh: int
    grid_latlon: Tuple[int, int]
    mesh_node_features: torch.Tensor  # [V_mesh, 3]
    mesh_edges: torch.Tensor  # [E_m, 2] (src, dst)
    mesh_edge_features: torch.Tensor  # [E_m, 4]
    grid2mesh_edges: torch.Tensor  # [E_g2m, 2] (grid src, mesh dst)
    grid2mesh_edge_features: torch.Tensor
    mesh2grid_edges: torch.Tensor  # [E_m2g, 2] (mesh src, grid dst)
    mesh2grid_edge_features: torch.Tensor
    mesh_xyz: Optional[np.ndarray] = None


@dataclass
class GraphCastCommPatterns:
    grid2mesh: CommunicationPattern
    mesh: CommunicationPattern
    mesh2grid: CommunicationPattern


@dataclass
class DistributedGraphCastGraph:
    """This rank's shard (field names follow the reference's ``DistributedGraphCastGraph``, ``graphcast_graph.py:58-82``).

    Edge arrays are ordered by the aggregating (central) vertex, so a node block's scatter-sum is a contiguous segment sum.
    ``*_src_indices`` / ``*_dst_indices`` index the local rows of the aggregating side and the ``[local | halo]`` rows of the
    other side (see the module docstring for which side aggregates)."""
    rank: int
    world_size: int
    num_local_grid: int
    num_local_mesh: int
    mesh_graph_node_features: torch.Tensor
    mesh_graph_edge_features: torch.Tensor
    mesh_graph_src_indices: torch.Tensor
    mesh_graph_dst_indices: torch.Tensor
    grid2mesh_graph_edge_features: torch.Tensor
    grid2mesh_graph_src_indices: torch.Tensor
    grid2mesh_graph_dst_indices: torch.Tensor
    mesh2grid_graph_edge_features: torch.Tensor
    mesh2grid_graph_src_indices: torch.Tensor
    mesh2grid_graph_dst_indices: torch.Tensor
    patterns: GraphCastCommPatterns
    grid_rows_global: torch.Tensor  # global grid ids of the local rows
    mesh_rows_global: torch.Tensor
    _plans: Dict[str, object] = field(default_factory=dict, repr=False)

    def to(self, device, dtype=None):
        for n in list(self.__dataclass_fields__):
            v = getattr(self, n)
            if isinstance(v, torch.Tensor):
                v = v.to(device)
                if dtype is not None and v.is_floating_point():
                    v = v.to(dtype)
                setattr(self, n, v)
        for p in (self.patterns.grid2mesh, self.patterns.mesh, self.patterns.mesh2grid):
            p.to(device)
        self._plans.clear()
        return self


class DistributedGraphCastGraphGenerator:
    """Builds the global graph once and shards it (``graphcast_graph.py:162-437``)."""

    def __init__(self, lat_lon_grid=(721, 1440), mesh_level: int = 6, ranks_per_graph: int = 1, rank: int = 0, world_size: int = 1,
                 multi: bool = True, placement: str = "bands", lat_lon=None):
        """Positional order of the reference (``graphcast_graph.py:169-176``): ``lat_lon_grid`` -- there a ``[n_lat, n_lon, 2]``
        tensor of the regular grid's coordinates, here also just ``(n_lat, n_lon)`` --, ``mesh_level``, ``ranks_per_graph``,
        ``rank``, ``world_size`` (defaults for ``get_graphcast_graph`` / ``build_graphcast_comm_patterns``).  ``placement``: ``"bands"`` = contiguous latitude bands (no graph partitioner needed); ``"multilevel"`` = the mesh is cut
        by the native multilevel partitioner (the reference's METIS step, ``data_utils/preprocess.py:14-30``) with every mesh
        vertex weighted by the grid vertices it feeds, and each grid vertex follows the majority of its mesh2grid sources
        (the reference derives the grid placement from the bipartite edges as well, ``graphcast_graph.py:85-159``)."""
        assert placement in ("bands", "multilevel")
        if lat_lon is not None:
            lat_lon_grid = lat_lon
        if torch.is_tensor(lat_lon_grid):  # the reference passes the coordinate grid itself: only its extent is needed here
            assert lat_lon_grid.dim() == 3 and lat_lon_grid.shape[-1] == 2, "expected a [n_lat, n_lon, 2] grid tensor"
            self.lat_lon_grid, lat_lon_grid = lat_lon_grid, (int(lat_lon_grid.shape[0]), int(lat_lon_grid.shape[1]))
        lat_lon = tuple(int(v) for v in lat_lon_grid)
        self.ranks_per_graph, self.rank, self.world_size = int(ranks_per_graph), int(rank), int(world_size)
        self.local_rank = self.rank % max(self.ranks_per_graph, 1)
        self.lat_lon, self.mesh_level, self.multi, self.placement = lat_lon, mesh_level, multi, placement
        self._global: Optional[GraphCastGraph] = None
        self._placements = {}

    def global_graph(self) -> GraphCastGraph:
        if self._global is not None:
            return self._global
        n_lat, n_lon = self.lat_lon
        mv, faces, medges, max_len = multimesh(self.mesh_level, self.multi)
        _, _, gxyz = latlon_grid(n_lat, n_lon)
        g2m = grid2mesh_edges(gxyz, mv, max_len)
        m2g = mesh2grid_edges(gxyz, mv, faces)
        mlat, mlon = xyz2sph(mv)
        t = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dt)  # noqa: E731
        self._global = GraphCastGraph(
            num_grid=gxyz.shape[0], num_mesh=mv.shape[0], grid_latlon=self.lat_lon,
            mesh_node_features=t(np.stack([np.cos(mlat), np.sin(mlon), np.cos(mlon)], 1)),
            mesh_edges=t(medges, torch.long), mesh_edge_features=t(edge_features(mv[medges[:, 0]], mv[medges[:, 1]])),
            grid2mesh_edges=t(g2m, torch.long), grid2mesh_edge_features=t(edge_features(gxyz[g2m[:, 0]], mv[g2m[:, 1]])),
            mesh2grid_edges=t(m2g, torch.long), mesh2grid_edge_features=t(edge_features(mv[m2g[:, 0]], gxyz[m2g[:, 1]])),
            mesh_xyz=mv)
        return self._global

    def placements(self, world_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """``(grid_placement [V_grid], mesh_placement [V_mesh])``."""
        if world_size in self._placements:
            return self._placements[world_size]
        g = self.global_graph()
        if self.placement == "multilevel" and world_size > 1:
            from dgraph_b200.data.partition import multilevel

            m2g = g.mesh2grid_edges  # [E, 2] = (mesh source, grid destination)
            w = 1 + torch.bincount(m2g[:, 0], minlength=g.num_mesh)
            mesh_pl = multilevel(g.mesh_edges, g.num_mesh, world_size, vertex_weights=w)
            votes = torch.zeros(g.num_grid, world_size, dtype=torch.long)
            votes.index_put_((m2g[:, 1], mesh_pl[m2g[:, 0]]), torch.ones(m2g.shape[0], dtype=torch.long), accumulate=True)
            grid_pl = votes.argmax(1)
            self._placements[world_size] = (grid_pl, mesh_pl)
            return grid_pl, mesh_pl
        n_lat, n_lon = self.lat_lon
        row_rank = (torch.arange(n_lat) * world_size // n_lat).clamp(max=world_size - 1)
        grid_pl = row_rank.repeat_interleave(n_lon)
        mlat, _ = xyz2sph(g.mesh_xyz)
        lat_row = np.clip(np.round((90.0 - np.rad2deg(mlat)) / 180.0 * (n_lat - 1)), 0, n_lat - 1).astype(np.int64)
        mesh_pl = row_rank[torch.from_numpy(lat_row)]
        self._placements[world_size] = (grid_pl, mesh_pl)
        return grid_pl, mesh_pl

    def shard(self, rank: int, world_size: int) -> DistributedGraphCastGraph:
        g = self.global_graph()
        grid_pl, mesh_pl = self.placements(world_size)

        def build(edges, central_col, central_pl, nbr_pl, feats):
            ed = torch.stack([edges[:, central_col], edges[:, 1 - central_col]], 1)
            same = central_pl is nbr_pl
            pat = build_communication_pattern(ed, central_pl, rank, world_size,
                                              neighbor_partitioning=None if same else nbr_pl)
            owned = torch.nonzero(central_pl[ed[:, 0]] == rank, as_tuple=True)[0]
            le = pat.local_edge_list
            order = torch.argsort(le[:, 0] * (pat.num_neighbor_local_vertices + pat.num_halo_vertices + 1) + le[:, 1], stable=True)
            pat.local_edge_list = le[order].contiguous()
            return pat, feats[owned][order].contiguous()

        p_g2m, f_g2m = build(g.grid2mesh_edges, 1, mesh_pl, grid_pl, g.grid2mesh_edge_features)  # central = mesh (dst)
        p_m, f_m = build(g.mesh_edges, 0, mesh_pl, mesh_pl, g.mesh_edge_features)  # central = src
        p_m2g, f_m2g = build(g.mesh2grid_edges, 1, grid_pl, mesh_pl, g.mesh2grid_edge_features)  # central = grid (dst)
        grid_rows = torch.nonzero(grid_pl == rank, as_tuple=True)[0]
        mesh_rows = torch.nonzero(mesh_pl == rank, as_tuple=True)[0]
        return DistributedGraphCastGraph(
            rank=rank, world_size=world_size, num_local_grid=int(grid_rows.numel()), num_local_mesh=int(mesh_rows.numel()),
            mesh_graph_node_features=g.mesh_node_features[mesh_rows].contiguous(),
            mesh_graph_edge_features=f_m, mesh_graph_src_indices=p_m.local_edge_list[:, 0].contiguous(),
            mesh_graph_dst_indices=p_m.local_edge_list[:, 1].contiguous(),
            grid2mesh_graph_edge_features=f_g2m, grid2mesh_graph_src_indices=p_g2m.local_edge_list[:, 1].contiguous(),
            grid2mesh_graph_dst_indices=p_g2m.local_edge_list[:, 0].contiguous(),
            mesh2grid_graph_edge_features=f_m2g, mesh2grid_graph_src_indices=p_m2g.local_edge_list[:, 1].contiguous(),
            mesh2grid_graph_dst_indices=p_m2g.local_edge_list[:, 0].contiguous(),
            patterns=GraphCastCommPatterns(grid2mesh=p_g2m, mesh=p_m, mesh2grid=p_m2g),
            grid_rows_global=grid_rows, mesh_rows_global=mesh_rows)


# ---- dictionary-style stages of the reference's generator (``data_utils/graphcast_graph.py:198-437``) ------------------------------
# Same method names and dictionary keys; the values are the consistent versions of what the (work-in-progress) reference code
# computes: vertices renumbered so that every rank owns one contiguous range, edges ordered by the rank that holds them.
def _rank_sort(placement: torch.Tensor):
    """``(sorted placement, new -> old permutation, old -> new map)`` of a stable sort by rank."""
    sorted_pl, new_to_old = torch.sort(placement.reshape(-1).long(), stable=True)
    old_to_new = torch.empty_like(new_to_old)
    old_to_new[new_to_old] = torch.arange(new_to_old.numel())
    return sorted_pl, new_to_old, old_to_new


def _get_mesh_graph_partition(mesh_level: int, world_size: int) -> torch.Tensor:
    """Placement of the multimesh vertices on ``world_size`` ranks by the native multilevel partitioner (METIS upstream)."""
    from dgraph_b200.data.partition import multilevel

    v, _, edges, _ = multimesh(mesh_level, True)
    return multilevel(torch.from_numpy(np.ascontiguousarray(edges)).long(), v.shape[0], world_size)


def _get_mesh_graph(self, mesh_vertex_rank_placement: torch.Tensor) -> dict:
    g = self.global_graph()
    assert mesh_vertex_rank_placement.numel() == g.num_mesh
    node_pl, new_to_old, old_to_new = _rank_sort(mesh_vertex_rank_placement)
    src, dst = old_to_new[g.mesh_edges[:, 0]], old_to_new[g.mesh_edges[:, 1]]
    edge_pl, order = torch.sort(node_pl[src], stable=True)  # an edge lives with its source vertex
    src, dst = src[order], dst[order]
    return {"node_features": g.mesh_node_features[new_to_old], "edge_features": g.mesh_edge_features[order],
            "src_indices": src, "dst_indices": dst, "node_rank_placement": node_pl, "edge_rank_placement": edge_pl,
            "src_rank_placement": edge_pl, "dst_rank_placement": node_pl[dst], "mesh_vertex_renumbering": new_to_old,
            "renumbered_vertices": old_to_new}


def _get_grid_placement(self, mesh_vertex_rank_placement: torch.Tensor, grid2mesh_mesh_dst_indices: torch.Tensor) -> torch.Tensor:
    """Rank of every grid-to-mesh edge: the rank of the mesh vertex it ends at."""
    return mesh_vertex_rank_placement.reshape(-1)[grid2mesh_mesh_dst_indices.long()]


def _get_grid2mesh_graph(self, mesh_graph_dict: dict) -> dict:
    g = self.global_graph()
    node_pl, old_to_new = mesh_graph_dict["node_rank_placement"], mesh_graph_dict["renumbered_vertices"]
    world = int(node_pl.max()) + 1 if node_pl.numel() else 1
    src_grid, dst_mesh = g.grid2mesh_edges[:, 0], old_to_new[g.grid2mesh_edges[:, 1]]
    edge_rank = node_pl[dst_mesh]
    # a grid vertex goes where most of its mesh neighbours live (grid vertices without an edge: rank 0)
    votes = torch.zeros(g.num_grid, world, dtype=torch.long)
    votes.index_put_((src_grid, edge_rank), torch.ones_like(src_grid), accumulate=True)
    grid_pl_sorted, grid_new_to_old, grid_old_to_new = _rank_sort(votes.argmax(1))
    edge_pl, order = torch.sort(edge_rank, stable=True)
    return {"node_features": torch.tensor([]), "edge_features": g.grid2mesh_edge_features[order],
            "src_indices": grid_old_to_new[src_grid][order], "dst_indices": dst_mesh[order],
            "grid2mesh_edge_rank_placement": edge_pl, "grid_vertex_rank_placement": grid_pl_sorted,
            "renumbered_grid": grid_old_to_new, "grid_vertex_renumbering": grid_new_to_old}


def _get_mesh2grid_graph(self, grid_vertex_rank_placement: torch.Tensor, renumbered_vertices: torch.Tensor,
                         renumbered_grid: torch.Tensor) -> dict:
    g = self.global_graph()
    src = renumbered_vertices[g.mesh2grid_edges[:, 0]]
    dst = renumbered_grid[g.mesh2grid_edges[:, 1]]
    edge_pl, order = torch.sort(grid_vertex_rank_placement[dst], stable=True)  # an edge lives with the grid vertex it ends at
    return {"node_features": torch.tensor([]), "edge_features": g.mesh2grid_edge_features[order], "src_indices": src[order],
            "dst_indices": dst[order], "mesh2grid_edge_rank_placement": edge_pl}


def _get_graphcast_graph(self, mesh_vertex_rank_placement: torch.Tensor, rank: Optional[int] = None,
                         world_size: Optional[int] = None):
    """This rank's shard (features, local edge lists, the three halo plans) for a given mesh placement; the grid follows its
    mesh neighbours.  (Upstream returns one record of global tensors here and builds the plans in a second step.)"""
    pl = mesh_vertex_rank_placement.reshape(-1).long()
    rank = self.rank if rank is None else int(rank)
    if world_size is None:
        world_size = self.world_size if self.world_size > 1 else None
    world = int(pl.max()) + 1 if world_size is None else int(world_size)
    assert 0 <= rank < world and int(pl.max()) < world, "mesh vertex placement does not match rank / world size"
    m = self.get_mesh_graph(pl)
    g2m = self.get_grid2mesh_graph(m)
    grid_pl = torch.empty(self.global_graph().num_grid, dtype=torch.long)
    grid_pl[g2m["grid_vertex_renumbering"]] = g2m["grid_vertex_rank_placement"]
    self._placements[world] = (grid_pl, pl)
    return self.shard(rank, world)


DistributedGraphCastGraphGenerator.get_mesh_graph_partition = staticmethod(_get_mesh_graph_partition)
DistributedGraphCastGraphGenerator.get_mesh_graph = _get_mesh_graph
DistributedGraphCastGraphGenerator.get_grid_placement = _get_grid_placement
DistributedGraphCastGraphGenerator.get_grid2mesh_graph = _get_grid2mesh_graph
DistributedGraphCastGraphGenerator.get_mesh2grid_graph = _get_mesh2grid_graph
DistributedGraphCastGraphGenerator.get_graphcast_graph = _get_graphcast_graph


def build_graphcast_comm_patterns(graph, rank: Optional[int] = None, world_size: Optional[int] = None):
    """Reference name (``graphcast_graph.py:85-159``): the three plans (grid2mesh, mesh, mesh2grid) of one rank.  ``graph`` is a
    generator (then ``rank`` / ``world_size`` say which shard) or -- the reference's form -- a topology record with the fields
    ``rank``, ``world_size``, ``mesh_rank_placement``, ``grid_rank_placement`` and the six global ``*_src_indices`` /
    ``*_dst_indices`` tensors (``GraphCastTopology`` in ``experiments/GraphCast/data_utils/graphcast_graph.py``)."""
    if isinstance(graph, DistributedGraphCastGraphGenerator):
        return graph.shard(graph.rank if rank is None else rank, graph.world_size if world_size is None else world_size).patterns
    r, W = int(graph.rank if rank is None else rank), int(graph.world_size if world_size is None else world_size)
    mesh_pl, grid_pl = graph.mesh_rank_placement.reshape(-1).long(), graph.grid_rank_placement.reshape(-1).long()

    def build(src, dst, central_is_dst, central_pl, nbr_pl):
        c, n = (dst, src) if central_is_dst else (src, dst)
        ed = torch.stack([c.reshape(-1).long(), n.reshape(-1).long()], 1)
        return build_communication_pattern(ed, central_pl, r, W, neighbor_partitioning=None if central_pl is nbr_pl else nbr_pl)

    return GraphCastCommPatterns(
        grid2mesh=build(graph.grid2mesh_graph_src_indices, graph.grid2mesh_graph_dst_indices, True, mesh_pl, grid_pl),
        mesh=build(graph.mesh_graph_src_indices, graph.mesh_graph_dst_indices, False, mesh_pl, mesh_pl),
        mesh2grid=build(graph.mesh2grid_graph_src_indices, graph.mesh2grid_graph_dst_indices, True, grid_pl, mesh_pl))


# ------------------------------------------------------------------------------------------------ synthetic data
class SyntheticWeatherDataset(torch.utils.data.Dataset):
    """Sinusoid + noise fields ``[channels, n_lat, n_lon]`` flattened to ``[n_lat * n_lon, channels]`` (reference
    ``experiments/GraphCast/dataset.py:24-226``).  With ``grid_rows`` set it returns only this rank's grid rows."""

    def __init__(self, channels: int = 73, lat_lon: Tuple[int, int] = (721, 1440), num_samples: int = 8, seed: int = 0,
                 grid_rows: Optional[torch.Tensor] = None, dtype=torch.float32):
        self.channels, self.lat_lon, self.num_samples, self.seed = channels, lat_lon, num_samples, seed
        self.grid_rows, self.dtype = grid_rows, dtype

    def __len__(self):
        return self.num_samples

    def _field(self, t: int) -> torch.Tensor:
        n_lat, n_lon = self.lat_lon
        rows = self.grid_rows if self.grid_rows is not None else torch.arange(n_lat * n_lon)
        la = (rows // n_lon).to(torch.float32) / max(n_lat - 1, 1) * torch.pi
        lo = (rows % n_lon).to(torch.float32) / n_lon * 2 * torch.pi
        c = torch.arange(self.channels, dtype=torch.float32).view(1, -1)
        base = torch.sin(la.view(-1, 1) * (1 + c % 3) + 0.1 * t) * torch.cos(lo.view(-1, 1) * (1 + c % 5) + 0.05 * t * (1 + c % 2))
        g = torch.Generator().manual_seed(self.seed * 1000003 + t)
        noise = torch.randn(rows.numel(), self.channels, generator=g) * 0.05
        return (base + noise).to(self.dtype)

    def __getitem__(self, i):
        return {"invar": self._field(i), "outvar": self._field(i + 1)}

    @staticmethod
    def generate_data(num_days: int, num_channels: int, grid_size: Tuple[int, int], base_temp: float = 15.0,
                      amplitude: float = 10.0, noise_level: float = 2.0, seed: Optional[int] = None) -> np.ndarray:
        """Dense ``[num_days, num_channels, n_lat, n_lon]`` temperature-like field of the reference's generator
        (``dataset.py:100-171``): base + 0.5 colder per channel ("altitude") + a linear south-north ramp of ``+-amplitude`` +
        an annual sine + Gaussian noise.  Written with broadcasting, so only the result is materialised.  The streaming
        ``__getitem__`` above never builds this array (it is 4 x 73 x 721 x 1440 x 8 bytes = 2.4 GB at the default size)."""
        n_lat, n_lon = grid_size
        rng = np.random.default_rng(seed)
        altitude = (-0.5 * np.arange(num_channels))[None, :, None, None]
        ramp = np.linspace(-amplitude, amplitude, n_lat)[None, None, :, None]
        season = np.sin(2 * np.pi * np.arange(num_days) / 365.0)[:, None, None, None]
        out = rng.normal(scale=noise_level, size=(num_days, num_channels, n_lat, n_lon))
        out += base_temp + altitude + ramp + season
        return out
