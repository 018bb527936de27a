"""Build the in-tree sm_100a kernel library: ``python -m dgraph_b200.build [--force] [--verbose]``.

Replaces the reference's CMake / scikit-build-core flow (``CMakeLists.txt:1-192``, ``pyproject.toml``) with a
dependency-free nvcc driver: the kernels expose a C ABI (``include/dgraph_b200.h``) and need neither torch nor
pybind11 headers, so a full rebuild takes seconds instead of minutes.  Output: ``dgraph_b200/lib/libdgraph_b200.so``
(travels to the GPU box with the source snapshot; git-ignored).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "dgraph_b200" / "lib"
OBJ_DIR = ROOT / "build" / "obj"
LIB = OUT_DIR / "libdgraph_b200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
              "-Xcompiler", "-fvisibility=hidden", "-Xptxas", "-v"]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest(paths) -> str:
    h = hashlib.sha1()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(ARCH_FLAGS + NVCC_FLAGS).encode())
    return h.hexdigest()


def sources():
    return sorted(CSRC.glob("*.cu"))


def build(force: bool = False, verbose: bool = False) -> Path:
    srcs = sources()
    if not srcs and LIB.exists():  # installed package: built at install time, sources not shipped
        return LIB
    hdrs = sorted(CSRC.glob("*.cuh")) + sorted((ROOT / "include").glob("*.h"))
    stamp = OUT_DIR / ".build_stamp"
    digest = _digest(srcs + hdrs)
    if not force and LIB.exists() and stamp.exists() and stamp.read_text().strip() == digest:
        return LIB
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: Path):
        obj = OBJ_DIR / (src.stem + ".o")
        cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        (OBJ_DIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stderr[-6000:]}")
        if verbose:
            print(r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    link = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(LIB), *map(str, objs), "-Xcompiler", "-fPIC"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stderr[-4000:]}")
    stamp.write_text(digest)
    return LIB


HOST_LIB = OUT_DIR / "libdgraph_b200_host.so"


def build_host(force: bool = False) -> Path:
    """Host-only C++ runtime pieces (``csrc/host/*.cpp``: the multilevel graph partitioner, the shared-memory one-sided
    windows of the CPU engine) -> ``libdgraph_b200_host.so``.
    Separate from the CUDA library so that a host-code change never triggers an nvcc rebuild."""
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    if not srcs:
        if HOST_LIB.exists():  # installed package: the library was built at install time and the sources stayed behind
            return HOST_LIB
        raise RuntimeError("no host sources")
    stamp = OUT_DIR / ".build_stamp_host"
    h = hashlib.sha1()
    for p in srcs:
        h.update(p.read_bytes())
    digest = h.hexdigest()
    if not force and HOST_LIB.exists() and stamp.exists() and stamp.read_text().strip() == digest:
        return HOST_LIB
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX") or shutil.which("g++") or shutil.which("c++")
    if not cxx:
        raise RuntimeError("no C++ compiler found")
    cmd = [cxx, "-O3", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-pthread", "-o", str(HOST_LIB), *map(str, srcs), "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"host build failed:\n{r.stderr[-4000:]}")
    stamp.write_text(digest)
    return HOST_LIB


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    try:
        print("BUILD OK", build(a.force, a.verbose), build_host(a.force))
    except Exception as e:  # noqa: BLE001
        print(str(e)[:3000])
        print("BUILD FAILED")
        sys.exit(1)


if __name__ == "__main__":
    main()
