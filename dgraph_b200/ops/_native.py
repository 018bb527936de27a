"""ctypes binding to ``dgraph_b200/lib/libdgraph_b200.so`` (C ABI declared in ``include/dgraph_b200.h``).

Plays the role of the reference's pybind11 modules ``torch_local`` / ``torch_nvshmem_p2p``
(``csrc/torch_local_bindings.cpp:20-26``, ``csrc/torch_nvshmem_p2p_bindings.cpp:19-36``), without torch headers:
launchers take raw device pointers plus the current CUDA stream, so kernels run on PyTorch's stream (the
reference launches on the legacy default stream, SURVEY §5 "latent ordering hazard").
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path
from typing import Optional

import torch

_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libdgraph_b200.so"
_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None


class NativeUnavailable(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def _try_load():
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return
    if not _LIB_PATH.exists():
        if os.environ.get("DGRAPH_B200_AUTOBUILD", "1") == "1":
            try:
                from dgraph_b200.build import build

                build()
            except Exception as e:  # noqa: BLE001
                _load_error = f"library missing and build failed: {e}"
                return
        else:
            _load_error = f"{_LIB_PATH} not built (run `python -m dgraph_b200.build`)"
            return
    try:
        _lib = ctypes.CDLL(str(_LIB_PATH))
    except OSError as e:
        _load_error = f"cannot load {_LIB_PATH}: {e}"


def available() -> bool:
    """Library is built and loadable (does not require a GPU: libcudart is linked statically)."""
    _try_load()
    return _lib is not None


def lib() -> ctypes.CDLL:
    _try_load()
    if _lib is None:
        raise NativeUnavailable(
            f"dgraph_b200 native kernels are required on CUDA tensors but unavailable: {_load_error}"
        )
    return _lib


def require_for(t: torch.Tensor):
    """Fail loudly when a CUDA tensor reaches an op whose sm_100a kernel is not loadable (no silent fallback)."""
    if t.is_cuda:
        lib()


def use_native(t: torch.Tensor) -> bool:
    if not t.is_cuda:
        return False
    lib()
    return True


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def i64(v) -> ctypes.c_int64:
    return ctypes.c_int64(int(v))


def i32(v) -> ctypes.c_int:
    return ctypes.c_int(int(v))


def f32(v) -> ctypes.c_float:
    return ctypes.c_float(float(v))


# Observers called as ``hook(what)`` after every successful native launch (tracing: ``utils/tracing.KernelTrace``;
# sync-checking: ``utils/debug.enable_sync_checks``).  Empty on the hot path.
_post_launch_hooks: list = []


def check(code: int, what: str):
    if code == 0 and _post_launch_hooks:
        for h in tuple(_post_launch_hooks):
            h(what)
    if code != 0:
        name = ""
        try:
            fn = lib().dgb_error_string
            fn.restype = ctypes.c_char_p
            name = fn(ctypes.c_int(code)).decode()
        except Exception:  # noqa: BLE001
            pass
        raise RuntimeError(f"{what} failed: CUDA error {code} {name}")


_launches = 0


def count_launch(n: int = 1):
    """Every wrapper bumps this once per kernel it launches; ``bench.py`` reports it as ``gpu_launches``."""
    global _launches
    _launches += n


def launches() -> int:
    return _launches
