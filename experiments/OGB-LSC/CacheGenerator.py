"""Legacy gather / scatter cache generation (parity: reference ``experiments/OGB-LSC/CacheGenerator.py:24-184``)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

import torch  # noqa: E402

from dgraph_b200.parallel.legacy_cache import NCCLGatherCacheGenerator, NCCLScatterCacheGenerator  # noqa: E402


def get_cache(indices, edge_placement, edge_dest_ranks, num_rows, rank, world_size, kind="gather", cache_dir="cache",
              name="rel", exchange=None):
    """Build (or load) the legacy cache object of one relation; stored with ``torch.save`` like the reference."""
    os.makedirs(cache_dir, exist_ok=True)
    path = os.path.join(cache_dir, f"{name}_{kind}_cache_rank_{rank}_of_{world_size}.pt")
    if os.path.exists(path):
        return torch.load(path, weights_only=False)
    gen = NCCLGatherCacheGenerator if kind == "gather" else NCCLScatterCacheGenerator
    cache = gen(indices, edge_placement, edge_dest_ranks, num_rows, rank, world_size, exchange=exchange)
    torch.save(cache, path)
    return cache
