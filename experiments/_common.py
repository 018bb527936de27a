"""Shared bootstrap for the entry points: repo on ``sys.path``, process-group + device setup from torchrun's env."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def init_distributed(backend: str = "nccl", **comm_kwargs):
    """Returns ``(comm, rank, world, device)``.  Single process -> ``single`` engine; CPU-only box -> gloo."""
    import torch
    import torch.distributed as dist

    from dgraph_b200 import Communicator

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cuda = torch.cuda.is_available()
    dev = torch.device("cuda", local % max(torch.cuda.device_count(), 1)) if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl" if cuda else "gloo", rank=rank, world_size=world,
                                **({"device_id": dev} if cuda else {}))
    Communicator._is_initialized = False
    comm = Communicator.init_process_group("single" if world == 1 else backend, **(comm_kwargs if world > 1 else {}))
    return comm, rank, world, dev  # rank / world are GLOBAL; comm.get_rank() / get_world_size() are per graph replica


def cleanup(comm):
    import torch.distributed as dist

    try:
        comm.destroy()
        eng = comm.engine
        if hasattr(eng, "arena") and eng.arena is not None:
            eng.finalize()
    finally:
        if dist.is_available() and dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
