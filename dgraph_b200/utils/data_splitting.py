"""Small splitting helpers (parity: ``DGraph/utils/data_splitting.py:14-17`` and the shadowed
``DGraph/utils.py:17-34``)."""
import torch


def largest_split(global_size: int, world_size: int) -> int:
    """Ceil-div: rows of the largest shard when ``global_size`` rows are split over ``world_size`` ranks."""
    return (global_size + world_size - 1) // world_size


def split_per_rank(global_size: int, rank: int, world_size: int):
    """``(start, end)`` of this rank's contiguous shard; the last shards may be short or empty."""
    sz = largest_split(global_size, world_size)
    start = min(rank * sz, global_size)
    return start, min(start + sz, global_size)


def try_barrier(comm=None) -> bool:
    """Barrier if a process group / communicator exists; returns whether one ran (never raises)."""
    try:
        if comm is not None:
            comm.barrier()
            return True
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            dist.barrier()
            return True
    except Exception:  # noqa: BLE001
        return False
    return False
