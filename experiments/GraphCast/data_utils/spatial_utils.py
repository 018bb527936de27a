"""Reference file name (``data_utils/spatial_utils.py:21-272``): spherical geometry helpers with the reference's function names.
The graph builder itself uses the vectorised numpy versions in ``dgraph_b200.data.graphcast_graph`` (re-exported below)."""
import math

import numpy as np
import torch
from torch import Tensor

from dgraph_b200.data.graphcast_graph import edge_features, latlon_grid, sph2xyz, xyz2sph  # noqa: F401


def deg2rad(deg):
    return deg * math.pi / 180.0


def rad2deg(rad):
    return rad * 180.0 / math.pi


def max_edge_length(vertices, source_nodes, destination_nodes) -> float:
    """Longest Euclidean edge of a graph given vertex coordinates ``[V, 3]`` and the two endpoint index arrays."""
    v = np.asarray(vertices, dtype=np.float64)
    d = v[np.asarray(source_nodes)] - v[np.asarray(destination_nodes)]
    return float(np.sqrt((d * d).sum(1).max())) if d.size else 0.0


def polar_angle(lat: Tensor) -> Tensor:
    """Latitude (rad) mapped to ``[0, 2 pi)``: negative latitudes wrap around."""
    return torch.where(lat >= 0.0, lat, lat + 2.0 * math.pi)


def azimuthal_angle(lon: Tensor) -> Tensor:
    """Rotation angle (rad) that brings a point of longitude ``lon`` onto the zero meridian."""
    return torch.where(lon >= 0.0, 2.0 * math.pi - lon, -lon)


def geospatial_rotation(invar: Tensor, theta: Tensor, axis: str, unit: str = "rad") -> Tensor:
    """Right-hand-rule rotation of points ``[N, 3]`` by per-point angles ``theta [N]`` about a coordinate axis.  ``unit="deg"``
    converts the angle to radians (the reference rescales the points instead, ``spatial_utils.py:107-108``; its callers only
    use radians)."""
    if unit == "deg":
        theta = deg2rad(theta)
    elif unit != "rad":
        raise ValueError("Not a valid unit")
    c, s = torch.cos(theta), torch.sin(theta)
    x, y, z = invar[:, 0], invar[:, 1], invar[:, 2]
    if axis == "x":
        out = (x, c * y - s * z, s * y + c * z)
    elif axis == "y":
        out = (c * x + s * z, y, -s * x + c * z)
    elif axis == "z":
        out = (c * x - s * y, s * x + c * y, z)
    else:
        raise ValueError("Invalid axis")
    return torch.stack(out, dim=1)


def xyz2latlon(xyz: Tensor, radius: float = 1, unit: str = "deg") -> Tensor:
    """``[N, 3]`` points on a sphere -> ``[N, 2]`` (latitude, longitude); x through (lat 0, lon 0), z through the poles."""
    lat = torch.arcsin((xyz[:, 2] / radius).clamp(-1.0, 1.0))
    lon = torch.arctan2(xyz[:, 1], xyz[:, 0])
    if unit == "deg":
        return torch.stack((rad2deg(lat), rad2deg(lon)), dim=1)
    if unit == "rad":
        return torch.stack((lat, lon), dim=1)
    raise ValueError("Not a valid unit")


def latlon2xyz(latlon: Tensor, radius: float = 1, unit: str = "deg") -> Tensor:
    if unit == "deg":
        latlon = deg2rad(latlon)
    elif unit != "rad":
        raise ValueError("Not a valid unit")
    lat, lon = latlon[:, 0], latlon[:, 1]
    return torch.stack((radius * torch.cos(lat) * torch.cos(lon), radius * torch.cos(lat) * torch.sin(lon),
                        radius * torch.sin(lat)), dim=1)


def get_face_centroids(vertices, faces) -> np.ndarray:
    """Centroid of every triangular face: mean of its three corner coordinates."""
    v = np.asarray(vertices, dtype=np.float64)
    f = np.asarray(faces, dtype=np.int64)
    return v[f].mean(axis=1)
