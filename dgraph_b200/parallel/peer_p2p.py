"""One-sided primitives on the symmetric arena -- the role of the reference's ``NVSHMEMP2P`` class
(``DGraph/distributed/csrc/torch_nvshmem_p2p.cu:28-376``, bound in ``torch_nvshmem_p2p_bindings.cpp:19-36``).

Same method names (``init``, ``finalize``, ``allocate_symmetric_memory``, ``clone_tensor``, ``padded_clone_tensor``,
``register_memory``, ``deregister_memory``, ``dist_get``, ``dist_put``, ``get_rank``, ``get_world_size``, ``barrier``,
``barrier_stream``); implementation: CUDA-IPC arena + plain peer loads / vector reductions instead of NVSHMEM + MPI.
Symmetric allocations are named slots that persist (no ``nvshmem_malloc`` per call)."""
from __future__ import annotations

from typing import Optional

import torch

from dgraph_b200.ops import _native as N


class PeerP2P:
    _engine = None
    _count = 0

    @staticmethod
    def init(engine=None):
        """Bind to a :class:`PeerMemoryEngine` (created from the default process group when omitted)."""
        if engine is None:
            from dgraph_b200.parallel.peer_engine import PeerMemoryEngine

            engine = PeerMemoryEngine()
        PeerP2P._engine = engine
        return engine

    @staticmethod
    def finalize():
        if PeerP2P._engine is not None:
            PeerP2P._engine.finalize()
            PeerP2P._engine = None

    @staticmethod
    def get_rank() -> int:
        return PeerP2P._engine.get_rank()

    @staticmethod
    def get_world_size() -> int:
        return PeerP2P._engine.get_world_size()

    @staticmethod
    def allocate_symmetric_memory(numel: int, device=None, dtype=torch.float32, name: Optional[str] = None) -> torch.Tensor:
        """Collective: every rank must call it with the same ``numel`` / ``name`` sequence.  Returns a zeroed tensor."""
        PeerP2P._count += 1
        t, off = PeerP2P._engine.slot_tensor(("p2p", name or PeerP2P._count), (int(numel),), dtype)
        t.zero_()
        return t

    @staticmethod
    def clone_tensor(t: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
        s = PeerP2P.allocate_symmetric_memory(t.numel(), dtype=t.dtype, name=name)
        s.copy_(t.reshape(-1))
        return s.view(t.shape)

    @staticmethod
    def padded_clone_tensor(t: torch.Tensor, padded_numel: int, name: Optional[str] = None) -> torch.Tensor:
        s = PeerP2P.allocate_symmetric_memory(max(padded_numel, t.numel()), dtype=t.dtype, name=name)
        s[: t.numel()].copy_(t.reshape(-1))
        return s

    @staticmethod
    def register_memory(t: torch.Tensor):  # arena memory is already mapped by every peer
        return None

    @staticmethod
    def deregister_memory(t: torch.Tensor):
        return None

    @staticmethod
    def barrier():
        PeerP2P._engine.barrier_on_stream()
        torch.cuda.current_stream().synchronize()

    @staticmethod
    def barrier_stream():
        PeerP2P._engine.barrier_on_stream()

    @staticmethod
    def dist_get(output: torch.Tensor, symm_src: torch.Tensor, rows: torch.Tensor, owners: torch.Tensor):
        """``output[i, :] = symm_src@owners[i][rows[i], :]`` (one pull kernel; the reference issues one
        ``nvshmemx_getmem_nbi_warp`` per row, ``nvshmem_comm_kernels.cuh:255-299``)."""
        eng = PeerP2P._engine
        rb = symm_src.shape[-1] * symm_src.element_size()
        own32, row32 = owners.to(torch.int32).contiguous(), rows.to(torch.int32).contiguous()  # keep alive across the launch
        N.check(N.lib().dgb_gather_rows_peer(N.ptr(output), N.i64(rb), N.i64(rows.numel()), N.i32(rb),
                                             N.ptr(eng.arena.ptr_table(symm_src.data_ptr() - eng.arena.base)), N.i64(rb),
                                             N.ptr(own32), N.ptr(row32), N.stream()),
                "dgb_gather_rows_peer")
        N.count_launch()
        return output

    @staticmethod
    def dist_put(values: torch.Tensor, symm_dst: torch.Tensor, rows: torch.Tensor, owners: torch.Tensor):
        """``symm_dst@owners[i][rows[i], :] += values[i, :]`` (fp32 vector reductions on peer memory)."""
        eng = PeerP2P._engine
        F = symm_dst.shape[-1]
        own32, row32 = owners.to(torch.int32).contiguous(), rows.to(torch.int32).contiguous()  # keep alive across the launch
        N.check(N.lib().dgb_scatter_add_rows_peer_f32(N.ptr(values), N.i64(values.stride(0)), N.i64(rows.numel()), N.i32(F),
                                                      N.ptr(eng.arena.ptr_table(symm_dst.data_ptr() - eng.arena.base)), N.i64(F),
                                                      N.ptr(own32), N.ptr(row32), N.stream()),
                "dgb_scatter_add_rows_peer_f32")
        N.count_launch()
        return symm_dst
