"""Legacy NCCL gather / scatter caches and all-to-all-v helpers.

Parity: reference ``DGraph/distributed/nccl/_nccl_cache.py:16-312`` (``NCCLGatherCache`` / ``NCCLScatterCache`` +
generators), ``alltoallv_impl.py:7-180`` (``_nccl_alltoall_v``, ``_nccl_alltoallv_with_dict``,
``torch_alltoallv_with_comm_map``) and ``_indices_utils.py``.  In the reference these feed the legacy
``indices + rank_mappings`` API, which is broken in the snapshot; here the caches are thin views over the same
:class:`NCCLGraphCommPlan` machinery that backs ``comm.gather`` / ``comm.scatter``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from dgraph_b200.parallel.comm_plan import NCCLGraphCommPlan, plan_from_requests


@dataclass
class NCCLGatherCache:
    """What rank ``rank`` needs to run a gather without any metadata exchange."""
    gather_recv_comm_vector: List[int]  # rows received from every rank (unique vertices)
    gather_send_comm_vector: List[int]  # rows sent to every rank
    gather_send_local_placement: Dict[int, torch.Tensor]  # per peer: my local rows to ship
    gather_recv_local_placement: Dict[int, torch.Tensor]  # per peer: edge slots filled from that peer's rows
    gather_local_indices: torch.Tensor  # local rows addressed by local edges
    gather_local_remapped_ranks: Optional[torch.Tensor]
    rank: int
    world_size: int
    num_features: int = 0
    plan: Optional[NCCLGraphCommPlan] = None


@dataclass
class NCCLScatterCache:
    scatter_recv_comm_vector: List[int]
    scatter_send_comm_vector: List[int]
    scatter_send_local_placement: Dict[int, torch.Tensor]
    scatter_recv_local_placement: Dict[int, torch.Tensor]
    scatter_local_indices: torch.Tensor
    num_output_rows: int
    rank: int
    world_size: int
    plan: Optional[NCCLGraphCommPlan] = None


def _split_dict(t: torch.Tensor, splits: List[int]) -> Dict[int, torch.Tensor]:
    out, o = {}, 0
    for r, k in enumerate(splits):
        if k:
            out[r] = t[o:o + k]
        o += k
    return out


def _plan(indices, edge_placement, edge_dest_ranks, num_rows, rank, world_size, exchange):
    idx = indices.reshape(-1).long()
    mine = edge_placement.reshape(-1) == rank
    owner = edge_dest_ranks.reshape(-1).long()[mine]
    return plan_from_requests(rank, world_size, owner, idx[mine] % num_rows, num_rows, exchange)


def NCCLGatherCacheGenerator(indices, edge_placement, edge_dest_ranks, num_input_rows, rank, world_size, exchange=None):
    p = _plan(indices, edge_placement, edge_dest_ranks, num_input_rows, rank, world_size, exchange)
    slots = torch.argsort(p.boundary_edge_buffer_map, stable=True)
    per_slot_counts = torch.bincount(p.boundary_edge_buffer_map, minlength=sum(p.boundary_edge_splits))
    recv_edges = []
    o = 0
    for k in p.boundary_edge_splits:
        n_edges = int(per_slot_counts[o:o + k].sum()) if k else 0
        recv_edges.append(n_edges)
        o += k
    return NCCLGatherCache(
        gather_recv_comm_vector=list(p.boundary_edge_splits), gather_send_comm_vector=list(p.boundary_vertex_splits),
        gather_send_local_placement=_split_dict(p.boundary_vertex_idx, p.boundary_vertex_splits),
        gather_recv_local_placement=_split_dict(p.boundary_edge_idx[slots], recv_edges),
        gather_local_indices=p.local_vertex_idx, gather_local_remapped_ranks=None, rank=rank, world_size=world_size, plan=p)


def NCCLScatterCacheGenerator(indices, edge_placement, edge_dest_ranks, num_output_rows, rank, world_size, exchange=None):
    p = _plan(indices, edge_placement, edge_dest_ranks, num_output_rows, rank, world_size, exchange)
    return NCCLScatterCache(
        scatter_recv_comm_vector=list(p.boundary_vertex_splits), scatter_send_comm_vector=list(p.boundary_edge_splits),
        scatter_send_local_placement=_split_dict(torch.arange(sum(p.boundary_edge_splits)), p.boundary_edge_splits),
        scatter_recv_local_placement=_split_dict(p.boundary_vertex_idx, p.boundary_vertex_splits),
        scatter_local_indices=p.local_vertex_idx, num_output_rows=num_output_rows, rank=rank, world_size=world_size, plan=p)


def _peer_exchange(send: Dict[int, torch.Tensor], recv: Dict[int, torch.Tensor], rank: int) -> None:
    """One ``isend`` / ``irecv`` pair per peer with something to move, batched (``dist.batch_isend_irecv``)."""
    ops = []
    for r in sorted(set(send) | set(recv)):
        if r == rank:
            continue
        if r in send and send[r].numel():
            ops.append(dist.P2POp(dist.isend, send[r].contiguous(), r))
        if r in recv and recv[r].numel():
            ops.append(dist.P2POp(dist.irecv, recv[r], r))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


def _nccl_alltoall_v(local_send_tensor: torch.Tensor, local_recv_tensor: torch.Tensor, indices: torch.Tensor,
                     local_rank_mapping: torch.Tensor, edge_rank_loc: torch.Tensor, src_rank_loc: torch.Tensor, rank: int,
                     world_size: int, cache=None):
    """Remote half of the legacy gather (``alltoallv_impl.py:7-131``): for every edge held by rank ``i != rank`` whose vertex
    lives here, send row ``indices % rows`` of ``local_send_tensor [1, rows, F]`` to ``i``; rows arriving from peer ``i`` land
    in the slots of ``local_recv_tensor [1, E_local, F]`` where ``local_rank_mapping == i`` (in order).  ``edge_rank_loc`` /
    ``src_rank_loc`` describe ALL edges (holder, vertex owner); with a :class:`NCCLGatherCache` the per-peer row lists come from
    the cache and the index tensors are not read."""
    F = local_send_tensor.shape[-1]
    rows = local_send_tensor.shape[-2]
    send2d = local_send_tensor.reshape(-1, F)
    if cache is None:
        holder, owner, idx = edge_rank_loc.reshape(-1), src_rank_loc.reshape(-1), indices.reshape(-1)
        if not bool((holder != owner).any()):
            return local_recv_tensor
        send_rows = {i: idx[(owner == rank) & (holder == i)].long() % rows for i in range(world_size) if i != rank}
        lrm = local_rank_mapping.reshape(-1)
        recv_slots = {i: torch.nonzero(lrm == i, as_tuple=True)[0] for i in range(world_size) if i != rank}
    else:
        send_rows = {i: t.reshape(-1).long() for i, t in cache.gather_send_local_placement.items() if i != rank}
        recv_slots = {i: t.reshape(-1).long() for i, t in cache.gather_recv_local_placement.items() if i != rank}
    send = {i: send2d.index_select(0, r) for i, r in send_rows.items() if r.numel()}
    recv = {i: torch.zeros((s.numel(), F), dtype=send2d.dtype, device=send2d.device) for i, s in recv_slots.items() if s.numel()}
    _peer_exchange(send, recv, rank)
    out = local_recv_tensor.reshape(-1, F)
    for i, buf in recv.items():
        out[recv_slots[i]] = buf
    return local_recv_tensor


def _nccl_alltoallv_with_dict(send_buffer_dict: Dict[int, torch.Tensor], recv_buffer_dict: Dict[int, torch.Tensor], rank: int,
                              world_size: int):
    """Per-peer tensors in, the filled receive dictionary out (``alltoallv_impl.py:134-161``)."""
    _peer_exchange(send_buffer_dict, recv_buffer_dict, rank)
    return recv_buffer_dict


def torch_alltoallv_with_comm_map(contiguous_send_tensor: torch.Tensor, contiguous_recv_tensor: torch.Tensor,
                                  send_comm_map, recv_comm_map, rank: int, world_size: int):
    """``dist.all_to_all`` over the slices of two ``[1, rows, F]`` buffers cut along dim 1 by ``send_comm_map`` / ``recv_comm_map``
    (rows per rank); returns the list of receive slices, which are views of ``contiguous_recv_tensor``
    (``alltoallv_impl.py:164-180``).  gloo has no list all-to-all: one send / receive pair per peer there."""
    send_sizes = [int(x) for x in (send_comm_map.tolist() if torch.is_tensor(send_comm_map) else send_comm_map)]
    recv_sizes = [int(x) for x in (recv_comm_map.tolist() if torch.is_tensor(recv_comm_map) else recv_comm_map)]
    assert len(send_sizes) == world_size and len(recv_sizes) == world_size, "comm maps must have world_size entries"
    send_list = [t.contiguous() for t in torch.split(contiguous_send_tensor, send_sizes, dim=1)]
    recv_views = list(torch.split(contiguous_recv_tensor, recv_sizes, dim=1))
    if dist.get_backend() == "gloo":
        recv_views[rank].copy_(send_list[rank])
        bufs = {r: torch.empty_like(recv_views[r]) for r in range(world_size) if r != rank}
        _peer_exchange({r: send_list[r] for r in range(world_size) if r != rank}, bufs, rank)
        for r, b in bufs.items():
            recv_views[r].copy_(b)
    else:
        bufs = [torch.empty_like(v) for v in recv_views]
        dist.all_to_all(bufs, send_list)
        for v, b in zip(recv_views, bufs):
            v.copy_(b)
    return recv_views


def all_to_all_cache_helper(indices, edge_placement, edge_vertex_ranks, num_rows, rank, world_size):
    """Gather-side bookkeeping: ``(send_comm_vector [W], recv_comm_vector [W], send_local_placement {peer: local rows to ship},
    recv_local_placement {peer: positions among this rank's edges that peer's rows fill})``.  An edge lives on
    ``edge_placement[e]`` and reads a vertex owned by ``edge_vertex_ranks[e]``."""
    from dgraph_b200.parallel import indices_utils as U

    place, owner = edge_placement.reshape(-1), edge_vertex_ranks.reshape(-1)
    cross = place != owner
    send_v = U._get_send_comm_vector(owner[cross], place[cross], rank, world_size)
    recv_v = U._get_recv_comm_vector(owner[cross], place[cross], rank, world_size)
    recv_place = U._get_local_recv_placement(recv_v, owner[place == rank], rank)
    send_place = U._get_local_send_placement(send_v, indices, owner, place, rank, num_rows)
    return send_v, recv_v, send_place, recv_place


def all_to_all_cache_with_local_reduce_helper(indices, edge_placement, edge_vertex_ranks, num_output_rows, rank, world_size):
    """Scatter-side bookkeeping with local pre-reduction: ``(renumbered_indices, num_remote_rows, remapped_ranks,
    receiving_ranks, recv_placement)`` -- this rank's remote-bound edge rows are summed per distinct target (``renumbered``
    addresses a compact ``[num_remote_rows]`` buffer whose rows go to ``remapped_ranks``); ``recv_placement[peer]`` are the
    local rows the peer's pre-reduced rows add to."""
    from dgraph_b200.parallel import indices_utils as U
    from dgraph_b200.parallel.rank_local import RankLocalRenumberingWithMapping

    idx, place, owner = indices.reshape(-1), edge_placement.reshape(-1), edge_vertex_ranks.reshape(-1)
    mine = place == rank
    remote = mine & (owner != rank)
    renumbered, unique_idx, remapped = RankLocalRenumberingWithMapping(idx[remote], owner[remote])
    receiving = torch.unique(owner[remote])
    incoming = (place != owner) & (owner == rank)
    recv_placement = U._get_local_unique_recv_placement(idx, place, incoming, num_output_rows, rank, world_size)
    return renumbered, int(unique_idx.numel()), remapped, receiving, recv_placement
