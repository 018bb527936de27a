"""Engines built on a two-sided row all-to-all.

* :class:`TorchDistEngine` -- ``torch.distributed`` (NCCL on GPUs, gloo on CPU).  This is the portable
  baseline path and the correctness oracle of the fused kernels; parity with
  ``DGraph/distributed/nccl/NCCLBackendEngine.py:35-318``.
* :class:`ThreadEngine` -- W ranks as W Python threads of one process exchanging through shared lists.  It
  is the *world emulator* the reference lacks: plans, autograd transposition and model code can
  be tested for W in {1..8} on a CPU-only box without spawning processes.
"""
from __future__ import annotations

import threading
from typing import Dict, List, Optional, Sequence, Tuple

import weakref

import torch
import torch.distributed as dist

from dgraph_b200.parallel.engine import BackendEngine
from dgraph_b200.parallel.functional import CommPlan_GatherFunction, CommPlan_ScatterFunction
from dgraph_b200.parallel.comm_plan import NCCLGraphCommPlan, plan_from_requests

TIMINGS: Dict[str, List[float]] = {}


def _splits(offsets) -> List[int]:
    if isinstance(offsets, torch.Tensor):
        offsets = offsets.tolist()
    return [int(offsets[i + 1] - offsets[i]) for i in range(len(offsets) - 1)]


class CollectiveEngine(BackendEngine):
    """Everything that only needs ``all_to_all_rows`` / ``all_reduce`` / ``barrier``."""

    def __init__(self, ranks_per_graph: int = -1, **kwargs):
        self._ranks_per_graph = ranks_per_graph
        self._legacy_plans: Dict[Tuple, NCCLGraphCommPlan] = {}

    # -- subclasses implement -------------------------------------------------------------------
    def all_to_all_rows(self, send: torch.Tensor, in_splits: Sequence[int], out_splits: Sequence[int]) -> torch.Tensor:
        raise NotImplementedError

    def exchange_ids(self, ids: torch.Tensor, send_splits: List[int]):
        """Metadata exchange used while building plans: returns ``(recv_ids, recv_splits)``."""
        W = self.get_world_size()
        counts = self.all_to_all_rows(torch.tensor(send_splits, dtype=torch.long, device=ids.device).view(W, 1),
                                      [1] * W, [1] * W).view(-1).tolist()
        recv = self.all_to_all_rows(ids.view(-1, 1), send_splits, counts).view(-1)
        return recv, [int(c) for c in counts]

    # -- partition arithmetic (NCCLBackendEngine.py:56-64) -----------------------------------------
    @property
    def ranks_per_graph(self) -> int:
        W = self.get_world_size()
        return W if self._ranks_per_graph in (-1, None) else int(self._ranks_per_graph)

    def get_partition_id(self) -> int:
        return self.get_rank() // self.ranks_per_graph

    def get_partition_rank(self) -> int:
        return self.get_rank() % self.ranks_per_graph

    # Names the three reference engines use for the same quantities (``NCCLBackendEngine.py:70-81``,
    # ``NVSHMEMBackendEngine.py:212-222``, ``MPIBackendEngine.py:374-402``).  With one process group per graph replica
    # (``Communicator.init_process_group(..., ranks_per_graph=k)``) ``get_rank()`` already is the rank inside the replica.
    def get_local_rank(self) -> int:
        return self.get_partition_rank() if not hasattr(self, "replica_id") else self.get_rank()

    def get_partition_size(self) -> int:
        return self.ranks_per_graph

    def get_partition_num(self) -> int:
        return getattr(self, "replica_id", self.get_partition_id())

    def to_global_rank(self, local_rank: int) -> int:
        """Rank of ``local_rank`` of *this* replica in the whole job."""
        return self.get_partition_num() * self.ranks_per_graph + int(local_rank)

    def get_max(self, val: torch.Tensor) -> torch.Tensor:
        """In-place all-reduce(MAX) across the engine's ranks (used to agree on symmetric buffer sizes)."""
        return self.all_reduce(val, "max")

    def destroy(self) -> None:
        """The reference only clears its initialised flag here; releasing resources is ``finalize``."""

    def _get_splits(self, send_offset, recv_offset):
        """``(send_splits, recv_splits)`` in rows, from the CSR-style offsets of a halo plan."""
        return _splits(send_offset), _splits(recv_offset)

    @staticmethod
    def _scale_splits(splits, factor: int):
        return [int(s) * int(factor) for s in splits]

    # MPI window management of the reference's MPI engine (``MPIBackendEngine.py:343-365``): there are no windows here; the
    # calls hand back ordinary buffers so that code written against them keeps running.
    @staticmethod
    def Malloc(size, device):
        return None, torch.zeros(int(size), dtype=torch.float32, device=device)

    @staticmethod
    def Attach(tensor: torch.Tensor):
        return None, tensor

    @staticmethod
    def Detach(win) -> None:
        return None

    def get_comm(self):
        """The process group the engine communicates over (``None`` = default group / no group)."""
        return getattr(self, "_group", None)

    # -- halo primitive --------------------------------------------------------------------------
    def put(self, send_buffer, recv_buffer, send_offsets, recv_offsets, remote_offsets=None) -> None:
        ins, outs = _splits(send_offsets), _splits(recv_offsets)
        F = 1
        for d in send_buffer.shape[1:]:
            F *= int(d)
        s2 = send_buffer.reshape(send_buffer.shape[0], F)
        r2 = recv_buffer.view(recv_buffer.shape[0], F)
        got = self.all_to_all_rows(s2.contiguous(), ins, outs, out=r2)
        if got is not r2:
            r2.copy_(got)

    # -- gather / scatter -------------------------------------------------------------------------
    @staticmethod
    def _reference_keywords(src_tensor, rank_mappings, output_size, kwargs):
        """Keyword names the three reference engines use for the same arguments (``NCCLBackendEngine.py:113,183``:
        ``local_send_tensor``; ``MPIBackendEngine.py:417,478``: ``send_tensor``, ``rank_mapping``, ``num_output_rows``;
        ``NVSHMEMBackendEngine.py``: ``input_tensor``, ``num_output_rows``)."""
        for k in ("local_send_tensor", "send_tensor", "input_tensor"):
            if src_tensor is None and k in kwargs:
                src_tensor = kwargs.pop(k)
        if rank_mappings is None and "rank_mapping" in kwargs:
            rank_mappings = kwargs.pop("rank_mapping")
        if output_size is None and "num_output_rows" in kwargs:
            output_size = kwargs.pop("num_output_rows")
        return src_tensor, rank_mappings, output_size

    def gather(self, src_tensor=None, indices=None, rank_mappings=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, _ = self._reference_keywords(src_tensor, rank_mappings, None, kwargs)
        plan = comm_plan
        if plan is None and isinstance(indices, NCCLGraphCommPlan):
            plan = indices
        if plan is None:
            plan = self._legacy_plan("g", src_tensor, indices, rank_mappings)
        return CommPlan_GatherFunction.apply(src_tensor, plan, self)

    def scatter(self, src_tensor=None, indices=None, rank_mappings=None, output_size=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, output_size = self._reference_keywords(src_tensor, rank_mappings, output_size, kwargs)
        plan = comm_plan
        if plan is None and isinstance(indices, NCCLGraphCommPlan):
            plan = indices
        if plan is None:
            # the MPI engine of the reference takes (tensor, indices, num_output_rows, rank_mapping=None): a row count in the
            # third position is recognised by its type
            if isinstance(rank_mappings, int) and (output_size is None or isinstance(output_size, torch.Tensor)):
                rank_mappings, output_size = output_size, rank_mappings
            plan = self._legacy_plan("s", src_tensor, indices, rank_mappings, output_size)
        return CommPlan_ScatterFunction.apply(src_tensor, plan, self)

    def _legacy_plan(self, kind, src_tensor, indices, rank_mappings, output_size=None):
        """Plan for the ``indices + rank_mappings`` API (``NCCLBackendEngine.py:96-250``; broken in the reference
        snapshot).  ``indices`` holds the global row each of *my* edges addresses; ownership comes from
        ``rank_mappings`` (``[2, E]``: row 1 = owner of the addressed vertex; or ``[E]`` / ``[1, E]``: owner), or from
        an even contiguous split when absent.  Plans are cached per (indices, mapping) storage.
        """
        idx = indices.reshape(-1).to(torch.long)
        W, r = self.get_world_size(), self.get_rank()
        if kind == "g":
            n_rows_local = src_tensor.shape[-2]
        else:
            assert output_size is not None, "scatter needs output_size (rows of the local output)"
            n_rows_local = int(output_size)
        # Keyed on tensor identity + in-place version (not on the data pointer, which the allocator recycles): a hit needs
        # the very same index tensor object, unmodified, so every rank of an SPMD program hits or misses together.
        # (The caller's tensor is the key, not the reshaped / cast view derived from it above -- that is a new object per call.)
        key = (kind, id(indices), indices._version,
               None if rank_mappings is None else (id(rank_mappings), rank_mappings._version), n_rows_local)
        hit = self._legacy_plans.get(key)
        if hit is not None and hit[0]() is indices and (rank_mappings is None or hit[1]() is rank_mappings):
            return hit[2]
        idx_in, rm_in = indices, rank_mappings
        if rank_mappings is None:
            owner = torch.div(idx, n_rows_local, rounding_mode="floor").clamp_(max=W - 1)
            mine = torch.ones_like(idx, dtype=torch.bool)
        else:
            rm = rank_mappings.to(idx.device)
            if rm.dim() == 2 and rm.shape[0] == 2:
                owner, mine = rm[1].to(torch.long), rm[0] == r
            else:
                owner, mine = rm.reshape(-1).to(torch.long), torch.ones_like(idx, dtype=torch.bool)
        if not bool(mine.all()):
            idx, owner = idx[mine], owner[mine]
        plan = plan_from_requests(r, W, owner, idx % n_rows_local, n_rows_local, self.exchange_ids)
        if len(self._legacy_plans) >= 256:  # drop entries whose tensors died; bounded either way
            self._legacy_plans = {k: v for k, v in self._legacy_plans.items() if v[0]() is not None}
            if len(self._legacy_plans) >= 256:
                self._legacy_plans.clear()
        self._legacy_plans[key] = (weakref.ref(idx_in), None if rm_in is None else weakref.ref(rm_in), plan)
        return plan


class TorchDistEngine(CollectiveEngine):
    """``torch.distributed`` engine; ``backend`` in {"nccl", "gloo"}."""

    def __init__(self, dist_backend: Optional[str] = None, ranks_per_graph: int = -1, group=None, **kwargs):
        super().__init__(ranks_per_graph=ranks_per_graph)
        self._group = group
        self._dist_backend = dist_backend
        self.init_process_group(**kwargs)

    def init_process_group(self, **kwargs):
        if not dist.is_initialized():
            backend = self._dist_backend or ("nccl" if torch.cuda.is_available() else "gloo")
            dist.init_process_group(backend=backend, **{k: v for k, v in kwargs.items()
                                                        if k in ("init_method", "rank", "world_size", "timeout", "store")})
        self._dist_backend = str(dist.get_backend(self._group))
        self._rank = dist.get_rank(self._group)
        self._world = dist.get_world_size(self._group)

    def get_rank(self) -> int:
        return self._rank

    def get_world_size(self) -> int:
        return self._world

    def _needs_cuda(self) -> bool:
        return "nccl" in self._dist_backend

    def all_to_all_rows(self, send, in_splits, out_splits, out=None):
        F = send.shape[1] if send.dim() > 1 else 1
        n_out = int(sum(out_splits))
        if out is None or out.shape[0] != n_out or out.dtype != send.dtype or not out.is_contiguous():
            out = torch.empty((n_out, F), dtype=send.dtype, device=send.device)
        if self._world == 1:
            out.copy_(send.view(out.shape))
            return out
        s, o = send.contiguous(), out
        moved = False
        if self._needs_cuda() and not s.is_cuda:
            s, o, moved = s.cuda(), torch.empty_like(out, device="cuda"), True
        dist.all_to_all_single(o.view(-1), s.view(-1), output_split_sizes=[c * F for c in out_splits],
                               input_split_sizes=[c * F for c in in_splits], group=self._group)
        if moved:
            out.copy_(o)
        return out

    def all_reduce(self, tensor, op="sum"):
        if self._world > 1:
            dist.all_reduce(tensor, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX,
                                        "min": dist.ReduceOp.MIN}[op], group=self._group)
        return tensor

    def barrier(self) -> None:
        if self._world > 1:
            dist.barrier(group=self._group)

    def finalize(self) -> None:
        if dist.is_initialized() and self._group is None:
            dist.destroy_process_group()


class _ThreadWorld:
    def __init__(self, world_size: int):
        self.W = world_size
        self.barrier = threading.Barrier(world_size)
        self.mail = [[None] * world_size for _ in range(world_size)]  # mail[src][dst]
        self.red = [None] * world_size


class ThreadEngine(CollectiveEngine):
    """One rank of an in-process emulated world (see :func:`dgraph_b200.parallel.emulator.run_world`)."""

    def __init__(self, world: _ThreadWorld, rank: int, ranks_per_graph: int = -1):
        super().__init__(ranks_per_graph=ranks_per_graph)
        self._w, self._rank = world, rank

    def init_process_group(self, *a, **k):
        return None

    def get_rank(self):
        return self._rank

    def get_world_size(self):
        return self._w.W

    def all_to_all_rows(self, send, in_splits, out_splits, out=None):
        W, r = self._w.W, self._rank
        chunks = torch.split(send.detach(), [int(c) for c in in_splits], dim=0)
        for d in range(W):
            self._w.mail[r][d] = chunks[d].clone()
        self._w.barrier.wait()
        got = [self._w.mail[s][r] for s in range(W)]
        for s in range(W):
            assert got[s].shape[0] == int(out_splits[s]), f"rank {r}: expected {out_splits[s]} rows from {s}, got {got[s].shape[0]}"
        res = torch.cat(got, dim=0) if W > 1 else got[0]
        self._w.barrier.wait()
        if out is not None and out.shape == res.shape and out.dtype == res.dtype:
            out.copy_(res)
            return out
        return res

    def all_reduce(self, tensor, op="sum"):
        W, r = self._w.W, self._rank
        self._w.red[r] = tensor.detach().clone()
        self._w.barrier.wait()
        stack = torch.stack(self._w.red)
        res = {"sum": stack.sum(0), "max": stack.max(0).values, "min": stack.min(0).values}[op]
        self._w.barrier.wait()
        tensor.copy_(res)
        return tensor

    def barrier(self):
        self._w.barrier.wait()

    def finalize(self):
        return None
