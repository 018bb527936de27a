"""Small splitting helpers (parity: ``DGraph/utils/data_splitting.py:14-17`` and the shadowed
``DGraph/utils.py:17-34``)."""
import torch


def largest_split(size: int, num_splits: int) -> int:
    """Ceil-div: rows of the largest shard when ``size`` rows are split into ``num_splits`` shards."""
    return (size + num_splits - 1) // num_splits


def split_range(global_size: int, current_rank: int, world_size: int):
    """``(start, end)`` of this rank's contiguous shard; the last shards may be short or empty."""
    sz = largest_split(global_size, world_size)
    start = min(current_rank * sz, global_size)
    return start, min(start + sz, global_size)


def split_per_rank(global_size: int, current_rank: int, world_size: int) -> int:
    """Number of rows of ``current_rank``'s shard (the reference's return value, ``DGraph/utils.py:21-26``; never negative
    here -- ranks beyond the data get 0)."""
    start, end = split_range(global_size, current_rank, world_size)
    return end - start


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
