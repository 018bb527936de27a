"""Reference file name (``experiments/OGB-LSC/RGAT.py:58-382``): the models live in ``dgraph_b200.models.rgat``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _common  # noqa: F401,E402

from dgraph_b200.models.rgat import CommAwareGAT, CommAwareRGAT, ConvLayer  # noqa: F401,E402

__all__ = ["ConvLayer", "CommAwareGAT", "CommAwareRGAT", "print_on_rank_zero", "print_on_all_ranks"]


def _dist():
    import torch.distributed as dist

    return dist if (dist.is_available() and dist.is_initialized()) else None


def print_on_rank_zero(*args, **kwargs):
    """Rank 0 prints, fenced by barriers so the line lands between the other ranks' output (``RGAT.py:31-35``); a plain print
    without a process group."""
    d = _dist()
    if d is None:
        print(*args, **kwargs)
        return
    d.barrier()
    if d.get_rank() == 0:
        print(*args, **kwargs)
    d.barrier()


def print_on_all_ranks(*args, **kwargs):
    """Every rank prints in rank order (``RGAT.py:38-43``)."""
    d = _dist()
    if d is None:
        print(*args, **kwargs)
        return
    d.barrier()
    for r in range(d.get_world_size()):
        if d.get_rank() == r:
            print(*args, **kwargs)
        d.barrier()
