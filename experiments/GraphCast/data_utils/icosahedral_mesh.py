"""Reference file name (``data_utils/icosahedral_mesh.py:32-300``): the triangular-mesh API (``TriangularMesh``,
``get_icosahedron``, ``get_hierarchy_of_triangular_meshes_for_sphere``, ``merge_meshes``, ``faces_to_edges``) on top of the mesh
code of ``dgraph_b200.data.graphcast_graph`` (vectorised refinement: midpoints are shared through a unique-edge table instead of
a per-vertex dictionary)."""
from typing import List, NamedTuple, Sequence, Tuple

import numpy as np

from dgraph_b200.data.graphcast_graph import icosahedron, multimesh, refine  # noqa: F401


class TriangularMesh(NamedTuple):
    """``vertices [V, 3]`` on the unit sphere, ``faces [F, 3]`` integer corner indices (consistently oriented)."""
    vertices: np.ndarray
    faces: np.ndarray


def get_icosahedron() -> TriangularMesh:
    """Regular icosahedron inscribed in the unit sphere: 12 vertices, 20 outward-oriented faces."""
    v, f = icosahedron()
    return TriangularMesh(vertices=v.astype(np.float64), faces=f.astype(np.int32))


def _two_split_unit_sphere_triangle_faces(triangular_mesh: TriangularMesh) -> TriangularMesh:
    """Every face split into four, new vertices at the edge midpoints projected onto the sphere.  The parent's vertices keep
    their indices, so a coarser mesh's faces stay valid on the finer vertex array."""
    v, f = refine(np.asarray(triangular_mesh.vertices, dtype=np.float64), np.asarray(triangular_mesh.faces, dtype=np.int64))
    return TriangularMesh(vertices=v.astype(np.float64), faces=f.astype(np.int32))


class _ChildVerticesBuilder(object):
    """Incremental variant of the refinement: hands out one child vertex per parent edge, creating it on first request."""

    def __init__(self, parent_vertices):
        self._parent = np.asarray(parent_vertices, dtype=np.float64)
        self._index = {}
        self._children = []

    def _get_child_vertex_key(self, parent_vertex_indices):
        return tuple(sorted(int(i) for i in parent_vertex_indices))

    def _create_child_vertex(self, parent_vertex_indices):
        p = self._parent[list(parent_vertex_indices)].mean(axis=0)
        self._index[self._get_child_vertex_key(parent_vertex_indices)] = len(self._parent) + len(self._children)
        self._children.append(p / np.linalg.norm(p))

    def get_new_child_vertex_index(self, parent_vertex_indices):
        key = self._get_child_vertex_key(parent_vertex_indices)
        if key not in self._index:
            self._create_child_vertex(parent_vertex_indices)
        return self._index[key]

    def get_all_vertices(self):
        return np.concatenate([self._parent, np.asarray(self._children).reshape(-1, 3)], axis=0)


def get_hierarchy_of_triangular_meshes_for_sphere(splits: int) -> List[TriangularMesh]:
    """``splits + 1`` meshes, each the 4-way refinement of the previous one, starting from the icosahedron."""
    meshes = [get_icosahedron()]
    for _ in range(splits):
        meshes.append(_two_split_unit_sphere_triangle_faces(meshes[-1]))
    return meshes


def merge_meshes(mesh_list: Sequence[TriangularMesh]) -> TriangularMesh:
    """Multimesh: the finest mesh's vertices with the faces of every level (coarser vertex sets are prefixes of finer ones)."""
    for a, b in zip(mesh_list[:-1], mesh_list[1:]):
        n = a.vertices.shape[0]
        assert np.allclose(a.vertices, b.vertices[:n]), "coarser vertices must be a prefix of the finer mesh's vertices"
    return TriangularMesh(vertices=mesh_list[-1].vertices, faces=np.concatenate([m.faces for m in mesh_list], axis=0))


def faces_to_edges(faces: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """``(senders, receivers)``, three directed edges per face (0->1, 1->2, 2->0); on a closed, consistently oriented surface
    every edge then appears in both directions."""
    faces = np.asarray(faces)
    assert faces.ndim == 2 and faces.shape[-1] == 3
    return (np.concatenate([faces[:, 0], faces[:, 1], faces[:, 2]]), np.concatenate([faces[:, 1], faces[:, 2], faces[:, 0]]))
