"""Reference path (``experiments/NVSHMEM-Enabled-DGRAPH/NVSHMEM_init.py``).  There is no NVSHMEM / MPI dependency in this framework: the
one-sided runtime is the CUDA-IPC symmetric arena, and this script runs ``experiments/PeerMemory-Enabled-DGRAPH/PeerMemory_init.py``."""
import os
import runpy

runpy.run_path(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "PeerMemory-Enabled-DGRAPH", "PeerMemory_init.py"),
               run_name="__main__")
