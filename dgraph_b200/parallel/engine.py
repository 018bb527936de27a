"""Backend-engine contract used by :class:`dgraph_b200.Communicator`.

Parity: reference ``DGraph/distributed/Engine.py:18-106`` (same method set; ``put`` is the halo-exchange
primitive, ``allocate_buffer`` lets one-sided engines hand out peer-mapped memory).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch


class BackendEngine(object):
    def init_process_group(self, *args, **kwargs):
        raise NotImplementedError

    def get_rank(self) -> int:
        raise NotImplementedError

    def get_world_size(self) -> int:
        raise NotImplementedError

    def get_local_rank_slice(self, tensor: torch.Tensor, dim: int = -1) -> torch.Tensor:
        """Even split of ``tensor`` along ``dim``; this rank's slice (``NCCLBackendEngine.py:86-94``; the
        reference ignores ``dim`` and always slices dim 1 -- we honour it)."""
        W, r = self.get_world_size(), self.get_rank()
        sz = tensor.shape[dim] // W
        return tensor.narrow(dim, r * sz, sz)

    def scatter(self, src_tensor, indices=None, output_size=None, rank_mappings=None, *args, **kwargs) -> torch.Tensor:
        """Abstract, in the reference's parameter order (``Engine.py:46-54``).  The concrete engines take
        ``(src_tensor, indices, rank_mappings, output_size)`` like the reference's NCCL / NVSHMEM engines and recognise the row
        count by its type, so both orders work."""
        raise NotImplementedError

    def gather(self, src_tensor, indices=None, rank_mappings=None, *args, **kwargs) -> torch.Tensor:
        raise NotImplementedError

    def put(
        self,
        send_buffer: torch.Tensor,
        recv_buffer: torch.Tensor,
        send_offsets: torch.Tensor,
        recv_offsets: torch.Tensor,
        remote_offsets: Optional[torch.Tensor] = None,
    ) -> None:
        """Chunk ``send_buffer`` by ``send_offsets`` and deliver chunk *i* to rank *i*'s ``recv_buffer``.

        Synchronous with respect to the current stream: when it returns, work queued afterwards on the same
        stream sees a fully populated ``recv_buffer``.  Two-sided engines ignore ``remote_offsets``; one-sided
        engines write chunk *i* at row ``remote_offsets[i]`` of rank *i*'s buffer (``Engine.py:67-86``).
        """
        raise NotImplementedError

    def allocate_buffer(self, size: Tuple[int, ...], dtype: torch.dtype, device) -> torch.Tensor:
        return torch.empty(size, dtype=dtype, device=device)

    def all_reduce(self, tensor: torch.Tensor, op: str = "sum") -> torch.Tensor:
        raise NotImplementedError

    def finalize(self) -> None:
        raise NotImplementedError

    def barrier(self) -> None:
        raise NotImplementedError
