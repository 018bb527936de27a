"""Shared helpers: random graphs / partitions and brute-force oracles."""
import torch


def random_graph(num_nodes, avg_deg, seed=0, symmetric=True):
    g = torch.Generator().manual_seed(seed)
    E = num_nodes * avg_deg // (2 if symmetric else 1)
    src = torch.randint(0, num_nodes, (E,), generator=g)
    dst = torch.randint(0, num_nodes, (E,), generator=g)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    if symmetric:
        src, dst = torch.cat([src, dst]), torch.cat([dst, src])
    e = torch.unique(torch.stack([src, dst], 1), dim=0)
    return e


def partition(num_nodes, W, kind, seed=0):
    if kind == "contiguous":
        return (torch.arange(num_nodes) * W // num_nodes).clamp(max=W - 1)
    if kind == "round_robin":
        return torch.arange(num_nodes) % W
    g = torch.Generator().manual_seed(seed + 17)
    return torch.randint(0, W, (num_nodes,), generator=g)
