"""Public communicator façade.

Parity: reference ``DGraph/Communicator.py:21-141`` -- same constructor, static ``init_process_group`` with the
"already initialized" ``RuntimeError``, class-level ``_is_initialized`` flag, pass-through ``gather`` / ``scatter`` /
``put`` and the name-mangled ``_Communicator__backend_engine`` attribute that reference models reach into
(``experiments/OGB-LSC/RGAT.py:189``).

Backend strings: ``"nccl"``, ``"mpi"``, ``"nvshmem"`` are kept for API compatibility.  On a CUDA
box every one of them maps to the single sm_100a peer-memory engine (``PeerMemoryEngine``: NVLink loads/stores
issued from inside the kernels, NCCL only for bootstrap and small metadata); ``DGRAPH_B200_ENGINE=torch``
selects the plain ``torch.distributed`` baseline engine instead.  On a CPU-only box they map to gloo.  Extra
names: ``"gloo"``, ``"torch"`` (baseline engine), ``"b200"`` (peer-memory engine, explicit).
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from dgraph_b200.communicator_base import CommunicatorBase

SUPPORTED_BACKENDS = ["nccl", "mpi", "nvshmem"]
_EXTRA_BACKENDS = ["gloo", "torch", "b200", "emulated", "single"]


def _make_engine(backend: str, **kwargs):
    from dgraph_b200.parallel.collective_engine import TorchDistEngine

    if backend == "single":
        from dgraph_b200.parallel.local_engine import LocalEngine

        return LocalEngine()
    use_cuda = torch.cuda.is_available()
    forced = os.environ.get("DGRAPH_B200_ENGINE", "").lower()
    if backend == "gloo" or not use_cuda:
        return TorchDistEngine(dist_backend="gloo", **kwargs)
    if backend == "torch" or forced == "torch":
        return TorchDistEngine(dist_backend="nccl", **kwargs)
    from dgraph_b200.parallel.peer_engine import PeerMemoryEngine

    return PeerMemoryEngine(**kwargs)


class Communicator(CommunicatorBase):
    """All inter-rank communication goes through this object."""

    def __init__(self, backend: str, **kwargs) -> None:
        assert backend in SUPPORTED_BACKENDS + _EXTRA_BACKENDS, (
            f"Backend {backend} not supported. Supported backends: {SUPPORTED_BACKENDS}"
        )
        self.backend = backend
        self.kwargs = kwargs
        engine = kwargs.pop("_engine", None)
        self.__backend_engine = engine if engine is not None else _make_engine(backend, **kwargs)
        if engine is None:
            Communicator._is_initialized = True
        self._private = engine is not None

    @classmethod
    def _from_engine(cls, engine, backend: str = "emulated") -> "Communicator":
        """Bind to an existing engine without touching the process-wide singleton flag (emulator, tests)."""
        return cls(backend, _engine=engine)

    @staticmethod
    def init_process_group(backend: str, **kwargs) -> "Communicator":
        if Communicator._is_initialized:
            raise RuntimeError("Communicator already initialized")
        return Communicator(backend, **kwargs)

    @property
    def engine(self):
        return self.__backend_engine

    def get_rank(self) -> int:
        self.__check_init()
        return self.__backend_engine.get_rank()

    def get_world_size(self) -> int:
        self.__check_init()
        return self.__backend_engine.get_world_size()

    def get_local_rank_slice(self, tensor: torch.Tensor, dim: int = -1) -> torch.Tensor:
        self.__check_init()
        return self.__backend_engine.get_local_rank_slice(tensor, dim)

    def get_local_tensor(self, tensor: torch.Tensor, placement_tensor: torch.Tensor, dim: int = -1) -> torch.Tensor:
        """Entries of ``tensor`` along ``dim`` whose placement equals this rank (``Communicator.py:73-97``)."""
        self.__check_init()
        keep = torch.nonzero(placement_tensor.to(tensor.device) == self.get_rank(), as_tuple=True)[0]
        return tensor.index_select(dim, keep)

    def alloc_buffer(self, size: Tuple[int, ...], dtype: torch.dtype, device) -> torch.Tensor:
        return self.__backend_engine.allocate_buffer(size, dtype, device)

    def scatter(self, *args, **kwargs) -> torch.Tensor:
        self.__check_init()
        return self.__backend_engine.scatter(*args, **kwargs)

    def gather(self, *args, **kwargs) -> torch.Tensor:
        self.__check_init()
        return self.__backend_engine.gather(*args, **kwargs)

    def put(self, send_buffer, recv_buffer, send_offsets, recv_offsets, remote_offsets: Optional[torch.Tensor] = None) -> None:
        return self.__backend_engine.put(send_buffer, recv_buffer, send_offsets, recv_offsets,
                                         remote_offsets=remote_offsets)

    def all_reduce(self, tensor: torch.Tensor, op: str = "sum") -> torch.Tensor:
        """In-place all-reduce across the communicator (used by distributed BatchNorm / loss normalisation)."""
        return self.__backend_engine.all_reduce(tensor, op)

    def barrier(self) -> None:
        self.__check_init()
        self.__backend_engine.barrier()

    def destroy(self) -> None:
        self.__check_init()
        if not self._private:
            Communicator._is_initialized = False

    def __check_init(self) -> None:
        assert self._private or Communicator._is_initialized, "Communicator not initialized"
