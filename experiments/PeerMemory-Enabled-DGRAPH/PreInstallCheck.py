"""Environment check for the peer-memory back-end (role of the reference's ``NVSHMEM-Enabled-DGRAPH/PreInstallCheck.py``):
CUDA toolkit with sm_100a support, built kernel library, >= 2 peer-capable GPUs.  No NVSHMEM / MPI is needed."""
import os
import shutil
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main():
    ok = True
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if os.path.exists(nvcc):
        v = subprocess.run([nvcc, "--list-gpu-arch"], capture_output=True, text=True).stdout
        print(f"[ok] nvcc at {nvcc}" if "compute_100" in v else f"[FAIL] {nvcc} does not know compute_100a")
        ok &= "compute_100" in v
    else:
        print("[FAIL] nvcc not found (set NVCC or CUDA_HOME)")
        ok = False
    from dgraph_b200.ops import _native

    print("[ok] kernel library loads" if _native.available() else "[FAIL] kernel library missing: python -m dgraph_b200.build")
    ok &= _native.available()
    import torch

    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    print(f"[info] {n} CUDA device(s)")
    for i in range(n):
        for j in range(n):
            if i != j and not torch.cuda.can_device_access_peer(i, j):
                print(f"[FAIL] GPU {i} cannot access GPU {j} (peer memory path unavailable; DGRAPH_B200_ENGINE=torch still works)")
                ok = False
    print("all checks passed" if ok else "some checks failed")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
