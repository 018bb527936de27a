"""Halo exchange and the message-passing wrapper.

Parity: reference ``DGraph/distributed/haloExchange.py:9-223`` (``HaloExchangeImpl``, ``HaloExchange``,
``DGraphMessagePassing``; same call signatures, 1-D and 2-D buffers).

On the peer-memory engine a standalone exchange is ONE pull kernel (``dgb_halo_pull``): it reads the boundary rows straight out
of the owners' feature matrices over NVLink into the rows behind the local ones -- no send-side index kernel, no all-to-all, no
staging buffer, no ``cat`` -- and its backward is ONE deterministic pull-reduce kernel on the owners.  (The GCN / SAGE layers go one
step further and read remote rows inside their aggregation kernels, ``ops/graphconv.py``; this module is the general-purpose
exchange that GraphCast, user layers and the reference-style ``DGraphMessagePassing`` use.)  The torch engines take the
reference's route (index -> ``put`` -> autograd index_add); the CPU window engine does the ``put`` one-sidedly.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
from torch.autograd import Function

from dgraph_b200.parallel.comm_pattern import CommunicationPattern


class HaloExchangeImpl(Function):
    """Communication half of the exchange: forward ``put``; backward = the transposed ``put``."""

    @staticmethod
    def forward(ctx, send_buffer, comm, comm_pattern: CommunicationPattern):
        two_d = send_buffer.dim() == 2
        F = send_buffer.shape[1] if two_d else 1
        n_recv = int(comm_pattern.recv_offset[-1])
        ctx.comm, ctx.pat, ctx.F = comm, comm_pattern, F
        recv = comm.alloc_buffer((n_recv, F) if two_d else (n_recv,), dtype=send_buffer.dtype, device=send_buffer.device)
        comm.put(send_buffer.contiguous(), recv, comm_pattern.send_offset, comm_pattern.recv_offset,
                 remote_offsets=comm_pattern.put_forward_remote_offset)
        return recv

    @staticmethod
    def backward(ctx, grad_recv):
        pat, comm = ctx.pat, ctx.comm
        n_sent = int(pat.send_offset[-1])
        two_d = grad_recv.dim() == 2
        grad_send = comm.alloc_buffer((n_sent, ctx.F) if two_d else (n_sent,), dtype=grad_recv.dtype, device=grad_recv.device)
        comm.put(grad_recv.contiguous(), grad_send, pat.recv_offset, pat.send_offset,
                 remote_offsets=pat.put_backward_remote_offset)
        return grad_send, None, None


class HaloExchange:
    """``halo = HaloExchange(comm)(x_local, pattern)`` -> ``[num_halo, F]`` in ``pattern.recv_offset`` order."""

    def __init__(self, comm):
        self.comm = comm

    def _fused_engine(self, x_local):
        engine = getattr(self.comm, "engine", None)
        ok = (engine is not None and hasattr(engine, "halo_exchange_aug") and x_local.is_cuda
              and x_local.dtype == torch.bfloat16 and x_local.dim() == 2 and (x_local.shape[1] * 2) % 16 == 0)
        return engine if ok else None

    def symmetric_rows(self, comm_pattern: CommunicationPattern, num_features: int, dtype=torch.bfloat16, device=None,
                       tag=0) -> torch.Tensor:
        """``[num_local, num_features]`` buffer for features that will be exchanged.  On the peer-memory engine it sits at the head of
        a symmetric ``[local | halo]`` buffer, so ``augmented`` / the pull kernels use it in place (no copy into a slot); elsewhere
        it is ordinary memory.  Plays the role of the reference's ``comm.alloc_buffer`` for one-sided back-ends
        (``DGraph/Communicator.py:99-104``).  Fill it between training steps (see ``PeerMemoryEngine.symmetric_rows``)."""
        engine = getattr(self.comm, "engine", None)
        if engine is not None and hasattr(engine, "symmetric_rows") and dtype == torch.bfloat16 and (num_features * 2) % 16 == 0:
            return engine.symmetric_rows(comm_pattern, num_features, dtype, tag=tag)
        if device is None:
            device = comm_pattern.local_edge_list.device
        return torch.empty((comm_pattern.num_local_vertices, num_features), dtype=dtype, device=device)

    def augmented(self, x_local: torch.Tensor, comm_pattern: CommunicationPattern, defer: bool = False) -> torch.Tensor:
        """``cat([x_local, halo])`` -- on the peer-memory engine this is one symmetric buffer filled by ONE pull
        kernel (no ``cat`` copy); elsewhere it is the reference's exchange followed by ``torch.cat``."""
        engine = self._fused_engine(x_local)
        if engine is not None:
            return engine.halo_exchange_aug(x_local, comm_pattern, defer=defer)
        if comm_pattern.world_size == 1 and comm_pattern.num_halo_vertices == 0:
            return x_local
        return torch.cat([x_local, self(x_local, comm_pattern)], dim=0)

    def __call__(self, x_local: torch.Tensor, comm_pattern: CommunicationPattern) -> torch.Tensor:
        engine = self._fused_engine(x_local)
        if engine is not None:
            return engine.halo_exchange(x_local, comm_pattern)
        # plain indexing on purpose: autograd's index backward does the scatter-add of rows sent to several ranks
        send = x_local[comm_pattern.send_local_idx]
        return HaloExchangeImpl.apply(send, self.comm, comm_pattern)


class DGraphMessagePassing(nn.Module):
    """halo exchange -> ``cat([local, halo])`` -> user layer ``(x_aug, local_edge_list, edge_feats)``.

    The wrapped layer must return features for the local vertices only (``[num_local, F_out]``).
    """

    def __init__(self, exchanger: HaloExchange, message_passing_layer: nn.Module):
        super().__init__()
        self.message_passing_layer = message_passing_layer
        self.exchanger = exchanger

    def forward(self, local_node_features: torch.Tensor, comm_pattern: CommunicationPattern,
                local_edge_features: Optional[torch.Tensor] = None) -> torch.Tensor:
        # one buffer holding [local | halo]: on the peer-memory engine the halo rows are pulled in right behind the local ones, so the
        # reference's ``cat`` (``haloExchange.py:214``) costs nothing; elsewhere ``augmented`` is exchange + ``torch.cat``
        aug = self.exchanger.augmented(local_node_features, comm_pattern)
        return self.message_passing_layer(aug, comm_pattern.local_edge_list, local_edge_features)
