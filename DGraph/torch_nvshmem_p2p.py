"""``DGraph.torch_nvshmem_p2p`` -- import path of the reference's NVSHMEM extension (``csrc/torch_nvshmem_p2p_bindings.cpp:19-36``:
``NVSHMEMP2P.{init, finalize, dist_put, dist_get, allocate_symmetric_memory, register_memory, deregister_memory,
clone_tensor, padded_clone_tensor, barrier, barrier_stream}``).  Here the one-sided runtime is the CUDA-IPC symmetric arena of
``dgraph_b200.parallel.peer_engine`` (no NVSHMEM, no MPI): :class:`NVSHMEMP2P` is :class:`dgraph_b200.parallel.peer_p2p.PeerP2P`."""
from dgraph_b200.parallel.peer_p2p import PeerP2P as NVSHMEMP2P  # noqa: F401

__all__ = ["NVSHMEMP2P"]
