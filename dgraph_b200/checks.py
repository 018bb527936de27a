"""Environment checks (parity: reference ``DGraph/checks.py:17-26``)."""
import torch
import torch.distributed as dist


def check_dist_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def check_nccl_availability() -> bool:
    return dist.is_available() and dist.is_nccl_available() and torch.cuda.is_available()


def check_native_available() -> bool:
    """True when the in-tree sm_100a kernel library has been built and can be loaded."""
    from dgraph_b200.ops import _native

    return _native.available()
