"""Synthetic graphs of the shapes named in BASELINE.json (there is no network for OGB downloads).

Generators run on whatever device they are given (GPU generation of the 62 M-edge products-shaped graph takes well
under a second) and are deterministic in ``seed`` so every rank can regenerate the same global graph, as the
reference's tests do (``tests/test_nccl_backend.py:33-52``).  ``cut_fraction`` controls how many edges leave a
vertex's locality window, i.e. the halo volume a contiguous-block partition sees (stand-in for a METIS cut,
SURVEY §7.3 item 6).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

SHAPES = {
    # name: (nodes, directed edges, feature dim, classes)
    "ogbn-arxiv": (169_343, 1_166_243 * 2, 128, 40),
    "ogbn-products": (2_449_029, 61_859_140, 100, 47),
    "ogbn-papers100M": (111_059_956, 1_615_685_872, 128, 172),
    "tiny": (10_000, 200_000, 32, 8),
}


@dataclass
class SyntheticGraph:
    num_nodes: int
    edge_list: torch.Tensor  # [E, 2] int64, symmetric, no self loops, unique
    num_features: int
    num_classes: int
    name: str = "synthetic"


def symmetric_locality_graph(num_nodes: int, num_directed_edges: int, cut_fraction: float = 0.1,
                             window: Optional[int] = None, seed: int = 0, device="cpu", degree_model: str = "uniform",
                             zipf_exponent: float = 0.55) -> torch.Tensor:
    """Symmetric edge list with ~``num_directed_edges`` rows: a fraction ``1 - cut_fraction`` of the undirected pairs
    connects vertices within ``window`` ids of each other, the rest are uniform over the whole id range.

    ``degree_model="powerlaw"``: edge endpoints are drawn with probability proportional to ``rank^-zipf_exponent`` over a random
    permutation of the vertices (Chung-Lu style), which gives the heavy-tailed degrees of a real co-purchase graph -- at the
    ogbn-products size a few vertices of degree > 17,000 (the real graph's maximum is 17,481), scattered over the id range -- instead
    of the near-Poisson degrees of the uniform model.  The locality window widens to 65,536 ids so that a hub finds enough distinct
    neighbours."""
    dev = torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed)
    powerlaw = degree_model == "powerlaw"
    assert degree_model in ("uniform", "powerlaw")
    n_und = int(num_directed_edges // 2 * (1.08 if powerlaw else 1.02)) + 16  # slack for self loops / duplicates removed below
    if window is None:
        window = max(16, min(num_nodes // 64, 65536 if powerlaw else 8192))
    if powerlaw:
        w = torch.arange(1, num_nodes + 1, device=dev, dtype=torch.float64).pow_(-zipf_exponent)
        cdf = torch.cumsum(w, 0)
        cdf /= cdf[-1].clone()
        perm = torch.randperm(num_nodes, generator=g, device=dev)

        def draw(k):
            r = torch.rand(k, generator=g, device=dev, dtype=torch.float64)
            return perm[torch.searchsorted(cdf, r).clamp_(max=num_nodes - 1)]

        u = draw(n_und)
    else:
        u = torch.randint(0, num_nodes, (n_und,), generator=g, device=dev)
    off = torch.randint(1, window + 1, (n_und,), generator=g, device=dev)
    sign = torch.randint(0, 2, (n_und,), generator=g, device=dev) * 2 - 1
    local_v = (u + off * sign) % num_nodes
    far_v = draw(n_und) if powerlaw else torch.randint(0, num_nodes, (n_und,), generator=g, device=dev)
    is_far = torch.rand(n_und, generator=g, device=dev) < cut_fraction
    v = torch.where(is_far, far_v, local_v)
    keep = u != v
    u, v = u[keep], v[keep]
    lo, hi = torch.minimum(u, v), torch.maximum(u, v)
    key = torch.unique(lo * num_nodes + hi)
    key = key[: num_directed_edges // 2]
    lo, hi = key // num_nodes, key % num_nodes
    e = torch.stack([torch.cat([lo, hi]), torch.cat([hi, lo])], dim=1)
    return e


def make_graph(name: str = "ogbn-products", cut_fraction: float = 0.1, seed: int = 0, device="cpu",
               scale: float = 1.0, degree_model: str = "uniform") -> SyntheticGraph:
    V, E, F, C = SHAPES[name]
    V, E = max(int(V * scale), 64), max(int(E * scale), 256)
    e = symmetric_locality_graph(V, E, cut_fraction=cut_fraction, seed=seed, device=device, degree_model=degree_model)
    return SyntheticGraph(V, e, F, C, name=f"{name}-shaped" + ("-powerlaw" if degree_model == "powerlaw" else ""))


def contiguous_partition(num_nodes: int, world_size: int, device="cpu") -> torch.Tensor:
    """Rank-contiguous block partition (ids of rank r form one range) -- what a METIS + renumber pass yields."""
    return (torch.arange(num_nodes, device=device) * world_size // num_nodes).clamp_(max=world_size - 1)


def contiguous_offsets(num_nodes: int, world_size: int) -> torch.Tensor:
    part = contiguous_partition(num_nodes, world_size)
    off = torch.zeros(world_size + 1, dtype=torch.long)
    off[1:] = torch.bincount(part, minlength=world_size).cumsum(0)
    return off


def node_features_and_labels(num_nodes: int, num_features: int, num_classes: int, seed: int = 0, device="cpu",
                             dtype=torch.float32, rows: Optional[torch.Tensor] = None):
    """Deterministic per-vertex features/labels; ``rows`` selects a shard without generating the rest."""
    dev = torch.device(device)
    if rows is None:
        g = torch.Generator(device=dev).manual_seed(seed + 1)
        x = torch.randn(num_nodes, num_features, generator=g, device=dev, dtype=torch.float32).to(dtype)
        y = torch.randint(0, num_classes, (num_nodes,), generator=g, device=dev)
        return x, y
    # hash-based generation so a rank can build only its shard
    r = rows.to(dev).to(torch.long)
    cols = torch.arange(num_features, device=dev)
    h = (r.unsqueeze(1) * 1_000_003 + cols * 7_919 + seed * 104_729) % 2_147_483_647
    h = (h * 48_271) % 2_147_483_647
    x = ((h.to(torch.float32) / 2_147_483_647.0) * 2.0 - 1.0).to(dtype)
    y = ((r * 2_654_435_761 + seed) % 4_294_967_291 % num_classes).to(torch.long)
    return x, y
