"""``backend="nvshmem"``: served by the peer-memory engine (CUDA-IPC symmetric arena instead of NVSHMEM + MPI)."""
try:
    from dgraph_b200.parallel.peer_engine import PeerMemoryEngine as NVSHMEMBackendEngine  # noqa: F401
    from dgraph_b200.parallel.peer_p2p import PeerP2P as NVSHMEMP2P  # noqa: F401
except Exception:  # noqa: BLE001
    NVSHMEMBackendEngine = None
