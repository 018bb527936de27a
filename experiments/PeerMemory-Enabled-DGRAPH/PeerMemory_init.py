"""Smoke test of the one-sided runtime (role of the reference's ``NVSHMEM_init.py:7-74``): symmetric allocation, fill, and an
all-to-all ``dist_get`` of one row per peer.   ``torchrun --nproc-per-node N PeerMemory_init.py``"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _common import cleanup, init_distributed  # noqa: E402

import torch  # noqa: E402


def main():
    from dgraph_b200.parallel.peer_p2p import PeerP2P

    comm, rank, world, dev = init_distributed("nvshmem")
    if world < 2 or dev.type != "cuda":
        print("needs >= 2 GPUs")
        return cleanup(comm)
    PeerP2P.init(comm.engine)
    buf = PeerP2P.allocate_symmetric_memory(world * 4, dtype=torch.float32, name="smoke").view(world, 4)
    buf.fill_(float(rank))
    PeerP2P.barrier()
    out = torch.empty(world, 4, device=dev)
    PeerP2P.dist_get(out, buf, torch.full((world,), rank, device=dev), torch.arange(world, device=dev))
    PeerP2P.barrier()
    assert torch.equal(out[:, 0].cpu(), torch.arange(world, dtype=torch.float32)), out
    print(f"rank {rank}: dist_get from every peer ok", flush=True)
    cleanup(comm)


if __name__ == "__main__":
    main()
