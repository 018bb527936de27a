"""Halo-exchange communication plans.

What this provides (parity: reference ``DGraph/distributed/commInfo.py:7-207``):

* :class:`CommunicationPattern` -- same fields as the reference dataclass.
* the same builder helper names (``compute_local_vertices`` ... ``build_communication_pattern``).

What is different on purpose:

* Halo vertices are numbered in **receive order** (owner-rank-major, then ascending global id), so
  ``cat([local, halo])[local_edge_list[:, 1]]`` is right for *any* partition (round-robin, METIS,
  contiguous) and any world size.  The reference numbers them by ascending global id, which only
  coincides with the receive order for rank-contiguous ids or W == 2.
* The send list is derived from what the **receiver** needs, so directed (non-symmetric) edge sets and
  bipartite relations (``neighbor_partitioning``) work.
* No collective is required when the global edge list is available (the reference API): every rank
  can compute the whole ``comm_map`` locally.  A shard-based builder that only sees this rank's edges
  (papers100M-scale) lives in :func:`build_communication_pattern_from_shard` and uses two metadata
  collectives instead.
* Builders are device agnostic (the reference forces ``.cuda()`` in ``compute_comm_map``).
"""
from __future__ import annotations

from dataclasses import dataclass, field, fields
from typing import Optional, Tuple, Dict, Any

import torch
import torch.distributed as dist


@dataclass
class CommunicationPattern:
    # --- identity ---
    rank: int
    world_size: int
    # --- vertex counts ---
    num_local_vertices: int
    num_halo_vertices: int
    # --- local subgraph: col 0 < num_local (owned central vertex); col 1 in [0, num_local + num_halo) ---
    local_edge_list: torch.Tensor  # [E_loc, 2]
    # --- send indexing (CSR over destination ranks) ---
    send_local_idx: torch.Tensor  # [total_sends] local row ids
    send_offset: torch.Tensor  # [W + 1]
    # --- receive indexing ---
    recv_offset: torch.Tensor  # [W + 1]
    # --- comm_map[a, b] = rows rank a sends to rank b ---
    comm_map: torch.Tensor  # [W, W]
    # --- one-sided write offsets ---
    put_forward_remote_offset: torch.Tensor  # [W]
    put_backward_remote_offset: torch.Tensor  # [W]
    # --- extras (not in the reference): global ids, useful for debugging / checkpoint re-sharding ---
    local_vertices_global: Optional[torch.Tensor] = None  # [num_local]
    halo_vertices_global: Optional[torch.Tensor] = None  # [num_halo] in receive order
    num_neighbor_local_vertices: Optional[int] = None  # bipartite: rows of the neighbour-side matrix
    halo_owner: Optional[torch.Tensor] = None  # [num_halo] owner rank of each halo row
    halo_remote_row: Optional[torch.Tensor] = None  # [num_halo] row of that vertex in its owner's local matrix
    _cache: Dict[str, Any] = field(default_factory=dict, repr=False, compare=False)

    def to(self, device) -> "CommunicationPattern":
        """Move index tensors used on the hot path to ``device`` (offsets stay on the host)."""
        self.local_edge_list = self.local_edge_list.to(device)
        self.send_local_idx = self.send_local_idx.to(device)
        if self.halo_owner is not None:
            self.halo_owner = self.halo_owner.to(device)
            self.halo_remote_row = self.halo_remote_row.to(device)
        self._cache.clear()
        return self

    @property
    def num_local_edges(self) -> int:
        return int(self.local_edge_list.shape[0])

    def memory_usage(self, unit: str = "MB") -> Dict[str, Any]:
        div = {"B": 1, "KB": 1024, "MB": 1024 ** 2, "GB": 1024 ** 3}.get(unit.upper(), 1024 ** 2)
        cpu = gpu = 0
        for f in fields(self):
            v = getattr(self, f.name)
            if isinstance(v, torch.Tensor):
                n = v.element_size() * v.nelement()
                if v.is_cuda:
                    gpu += n
                else:
                    cpu += n
        return {"cpu": cpu / div, "gpu": gpu / div, "total": (cpu + gpu) / div, "unit": unit.upper()}

    def kernel_plan(self, device=None):
        """CSR / CSC int32 structures consumed by the sm_100a kernels (built once, cached)."""
        from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

        key = f"kp:{device}"
        if key not in self._cache:
            self._cache[key] = GraphKernelPlan.from_pattern(self, device)
        return self._cache[key]

    def __getstate__(self):
        d = dict(self.__dict__)
        d["_cache"] = {}
        return d


# --------------------------------------------------------------------------------------------
# helpers with the reference's names
# --------------------------------------------------------------------------------------------
def compute_local_vertices(partitioning: torch.Tensor, rank: int) -> torch.Tensor:
    """Global ids owned by ``rank`` in ascending order (``commInfo.py:35-38``)."""
    return torch.nonzero(partitioning == rank, as_tuple=True)[0]


def _halo_sorted(edge_list, src_part, rank, dst_part):
    """Unique remote column-1 vertices of the edges owned by ``rank``, in (owner, gid) order."""
    owned = src_part[edge_list[:, 0]] == rank
    nbr = edge_list[owned, 1]
    nbr_rank = dst_part[nbr]
    remote = nbr_rank != rank
    v_total = int(dst_part.numel())
    enc = torch.unique(nbr_rank[remote] * v_total + nbr[remote])  # sorted => rank-major then gid
    return enc % v_total, enc // v_total


def compute_halo_vertices(
    edge_list: torch.Tensor,
    src_partitioning: torch.Tensor,
    rank: int,
    dst_partitioning: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Halo vertex global ids in **receive order** (``commInfo.py:41-61``; ordering fixed, Q8).

    ``src_partitioning`` owns column 0 (the edge owner side), ``dst_partitioning`` column 1 (defaults to the
    same partition).  A column-1 vertex is halo when its owner differs from ``rank``.
    """
    if dst_partitioning is None:
        dst_partitioning = src_partitioning
    gids, _ = _halo_sorted(edge_list, src_partitioning, rank, dst_partitioning)
    return gids


def compute_local_edge_list(
    global_edge_list: torch.Tensor,
    partitioning: torch.Tensor,
    local_vertices_global: torch.Tensor,
    halo_vertices_global: torch.Tensor,
    rank: int,
    neighbor_partitioning: Optional[torch.Tensor] = None,
    neighbor_local_vertices_global: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Owned edges renumbered to ``[local | halo]`` indices (``commInfo.py:64-91``).

    Homogeneous: column 1 indexes ``cat([x_local, x_halo])``.  Bipartite: column 0 indexes the central
    side's local rows; column 1 indexes ``cat([nbr_local, nbr_halo])`` of the neighbour side.
    """
    dev = global_edge_list.device
    owned = partitioning[global_edge_list[:, 0]] == rank
    e = global_edge_list[owned]
    n_loc = int(local_vertices_global.numel())
    g2l_c = torch.full((int(partitioning.numel()),), -1, dtype=torch.long, device=dev)
    g2l_c[local_vertices_global] = torch.arange(n_loc, device=dev)
    if neighbor_partitioning is None:
        g2l_n = g2l_c
        n_nbr_loc = n_loc
    else:
        if neighbor_local_vertices_global is None:
            neighbor_local_vertices_global = compute_local_vertices(neighbor_partitioning, rank)
        n_nbr_loc = int(neighbor_local_vertices_global.numel())
        g2l_n = torch.full((int(neighbor_partitioning.numel()),), -1, dtype=torch.long, device=dev)
        g2l_n[neighbor_local_vertices_global] = torch.arange(n_nbr_loc, device=dev)
    g2l_n = g2l_n.clone() if g2l_n is g2l_c else g2l_n
    g2l_n[halo_vertices_global] = torch.arange(
        n_nbr_loc, n_nbr_loc + int(halo_vertices_global.numel()), device=dev
    )
    return torch.stack([g2l_c[e[:, 0]], g2l_n[e[:, 1]]], dim=1)


def _all_pairs(edge_list, src_part, dst_part):
    """Unique (consumer rank, owner rank, gid) triples over all cross-rank edges, sorted.

    consumer = owner of column 0 (the edge owner), owner = owner of the column-1 vertex.
    Returned encoded as consumer * (W*V) ... decomposed into three tensors.
    """
    cons = src_part[edge_list[:, 0]]
    nbr = edge_list[:, 1]
    own = dst_part[nbr]
    cross = cons != own
    return cons[cross], own[cross], nbr[cross]


def compute_boundary_vertices(
    edge_list: torch.Tensor,
    src_partitioning: torch.Tensor,
    src_local_vertices_global: torch.Tensor,
    rank: int,
    num_ranks: int,
    dst_partitioning: Optional[torch.Tensor] = None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Sender-centric send list, same contract as the reference (``commInfo.py:94-145``).

    For every cross-rank edge ``(u, v)`` with ``u`` owned by ``rank``, ``u`` is sent once to the owner of
    ``v``.  Returns ``(send_local_idx, send_offset)`` sorted by destination rank then global id.
    """
    if dst_partitioning is None:
        dst_partitioning = src_partitioning
    return _send_list(
        edge_list[:, 0], src_partitioning, dst_partitioning[edge_list[:, 1]],
        src_local_vertices_global, rank, num_ranks,
    )


def _send_list(sent_vertex, send_part, dest_rank, send_local_vertices_global, rank, W):
    """Unique ``(dest_rank, sent_vertex)`` pairs with ``sent_vertex`` owned by ``rank`` -> CSR send list."""
    dev = sent_vertex.device
    own = send_part[sent_vertex]
    m = (own == rank) & (dest_rank != rank)
    v_total = int(send_part.numel())
    enc = torch.unique(dest_rank[m] * v_total + sent_vertex[m])  # sorted by destination rank then gid
    dest = enc // v_total
    gid = enc % v_total
    g2l = torch.full((v_total,), -1, dtype=torch.long, device=dev)
    g2l[send_local_vertices_global] = torch.arange(int(send_local_vertices_global.numel()), device=dev)
    send_local_idx = g2l[gid]
    counts = torch.bincount(dest, minlength=W)
    send_offset = torch.zeros(W + 1, dtype=torch.long, device=dev)
    send_offset[1:] = counts.cumsum(0)
    return send_local_idx, send_offset


def _send_list_from_consumers(edge_list, cons_part, send_part, send_local_vertices_global, rank, W):
    """What each consumer (owner of column 0) needs from ``rank`` (owner of column 1): correct for directed
    and bipartite edge sets because it is derived from the receiver's edges."""
    return _send_list(
        edge_list[:, 1], send_part, cons_part[edge_list[:, 0]], send_local_vertices_global, rank, W
    )


def compute_comm_map(send_offset: torch.Tensor, world_size: int, group=None) -> torch.Tensor:
    """All-gather of per-rank send counts -> ``comm_map[W, W]`` (``commInfo.py:148-154``).

    Works on whatever device the active backend needs (CUDA for NCCL, host for gloo); returns a host tensor.
    """
    counts = (send_offset[1:] - send_offset[:-1]).to(torch.long)
    if not (dist.is_available() and dist.is_initialized()) or world_size == 1:
        assert world_size == 1, "compute_comm_map needs an initialised process group for world_size > 1"
        return counts.view(1, 1).cpu()
    dev = _collective_device(group)
    counts = counts.to(dev)
    out = [torch.zeros(world_size, dtype=torch.long, device=dev) for _ in range(world_size)]
    dist.all_gather(out, counts, group=group)
    return torch.stack(out).cpu()


def compute_comm_map_local(edge_list, cons_part, send_part, W) -> torch.Tensor:
    """``comm_map`` for all ranks from the global edge list, without communication."""
    cons, own, nbr = _all_pairs(edge_list, cons_part, send_part)
    v_total = int(send_part.numel())
    enc = torch.unique((own * W + cons) * v_total + nbr)
    pair = enc // v_total
    return torch.bincount(pair, minlength=W * W).view(W, W).cpu()


def compute_recv_offsets(comm_map: torch.Tensor, rank: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(recv_offset[W+1], put_forward_remote_offset[W])`` from ``comm_map`` (``commInfo.py:157-164``)."""
    recv_counts = comm_map[:, rank].cpu()
    recv_offset = torch.zeros(comm_map.size(0) + 1, dtype=torch.long)
    recv_offset[1:] = recv_counts.cumsum(0)
    return recv_offset, comm_map[:rank, :].sum(0)


def _collective_device(group=None):
    backend = dist.get_backend(group)
    if "nccl" in str(backend):
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def build_communication_pattern(
    global_edge_list: torch.Tensor,
    partitioning: torch.Tensor,
    rank: int,
    world_size: int,
    neighbor_partitioning: Optional[torch.Tensor] = None,
) -> CommunicationPattern:
    """Build the halo plan of ``rank`` from the global edge list (``commInfo.py:167-207``).

    Args:
        global_edge_list: ``[E, 2]`` int64.  Column 0 is the central vertex (its owner owns the edge and
            aggregates messages); column 1 the neighbour whose features are needed.
        partitioning: ``[V]`` owner rank of every central-side vertex.
        neighbor_partitioning: optional ``[V_nbr]`` owner ranks of the neighbour side (bipartite
            relations such as GraphCast grid->mesh).  ``None`` means homogeneous.

    No collective is issued: all ranks hold the same global edge list, so the whole ``comm_map`` is computed
    locally (the reference all-gathers it, ``commInfo.py:148-154``).
    """
    W = world_size
    edge_list = global_edge_list.to(torch.long)
    part = partitioning.to(torch.long)
    nbr_part = part if neighbor_partitioning is None else neighbor_partitioning.to(torch.long)
    bip = neighbor_partitioning is not None

    local_verts = compute_local_vertices(part, rank)
    nbr_local_verts = compute_local_vertices(nbr_part, rank) if bip else local_verts
    halo_verts, halo_owner = _halo_sorted(edge_list, part, rank, nbr_part)
    # row of every neighbour-side vertex inside its owner's local matrix (owners number ascending by gid)
    order = torch.argsort(nbr_part, stable=True)
    starts = torch.zeros(W + 1, dtype=torch.long, device=order.device)
    starts[1:] = torch.bincount(nbr_part, minlength=W).cumsum(0)
    row_in_owner = torch.empty_like(order)
    row_in_owner[order] = torch.arange(order.numel(), device=order.device) - starts[nbr_part[order]]
    local_edges = compute_local_edge_list(
        edge_list, part, local_verts, halo_verts, rank,
        neighbor_partitioning=nbr_part if bip else None,
        neighbor_local_vertices_global=nbr_local_verts if bip else None,
    )
    send_idx, send_off = _send_list_from_consumers(edge_list, part, nbr_part, nbr_local_verts, rank, W)
    comm = compute_comm_map_local(edge_list, part, nbr_part, W)
    recv_off, put_fwd = compute_recv_offsets(comm, rank)
    assert int(recv_off[-1]) == int(halo_verts.numel()), "halo plan inconsistent"
    assert torch.equal(comm[rank].cpu(), (send_off[1:] - send_off[:-1]).cpu()), "send plan inconsistent"
    return CommunicationPattern(
        rank=rank,
        world_size=W,
        num_local_vertices=int(local_verts.numel()),
        num_halo_vertices=int(halo_verts.numel()),
        local_edge_list=local_edges,
        send_local_idx=send_idx,
        send_offset=send_off.cpu(),
        recv_offset=recv_off,
        comm_map=comm,
        put_forward_remote_offset=put_fwd,
        put_backward_remote_offset=comm[:, :rank].sum(1),
        local_vertices_global=local_verts,
        halo_vertices_global=halo_verts,
        num_neighbor_local_vertices=int(nbr_local_verts.numel()),
        halo_owner=halo_owner,
        halo_remote_row=row_in_owner[halo_verts],
    )


def build_communication_pattern_from_shard(
    local_edges_global: torch.Tensor,
    vertex_offsets: torch.Tensor,
    rank: int,
    world_size: int,
    group=None,
    neighbor_offsets: Optional[torch.Tensor] = None,
) -> CommunicationPattern:
    """Build the halo plan from **this rank's edge shard only** (papers100M scale).

    Vertices are partitioned contiguously: rank ``r`` owns ``[vertex_offsets[r], vertex_offsets[r+1])``, so no
    O(V_global) table is needed.  ``local_edges_global`` holds the edges whose column-0 vertex this rank
    owns, with global ids.  Two metadata collectives (counts, then ids) tell each owner what to send --
    the role ``all_to_all_single`` plays at ``_NCCLCommPlan.py:238-262`` in the reference.
    """
    W = world_size
    e = local_edges_global.to(torch.long)
    dev = e.device
    off = vertex_offsets.to(torch.long).to(dev)
    noff = off if neighbor_offsets is None else neighbor_offsets.to(torch.long).to(dev)
    my0, my1 = int(off[rank]), int(off[rank + 1])
    n_loc = my1 - my0
    nbr0, nbr1 = int(noff[rank]), int(noff[rank + 1])
    n_nbr_loc = nbr1 - nbr0
    assert e.numel() == 0 or (int(e[:, 0].min()) >= my0 and int(e[:, 0].max()) < my1), "edge shard holds foreign edges"

    nbr = e[:, 1]
    nbr_rank = torch.bucketize(nbr, noff, right=True) - 1
    remote = nbr_rank != rank
    halo_gid = torch.unique(nbr[remote])  # contiguous partition => gid order == (owner, gid) order
    halo_rank = torch.bucketize(halo_gid, noff, right=True) - 1
    need_counts = torch.bincount(halo_rank, minlength=W)

    col1 = torch.where(
        remote,
        torch.searchsorted(halo_gid, nbr) + n_nbr_loc,
        nbr - nbr0,
    )
    local_edges = torch.stack([e[:, 0] - my0, col1], dim=1)

    if W == 1 or not (dist.is_available() and dist.is_initialized()):
        assert W == 1, "shard-based plan build needs an initialised process group"
        send_counts = torch.zeros(1, dtype=torch.long)
        send_gid = torch.empty(0, dtype=torch.long, device=dev)
        comm = torch.zeros(1, 1, dtype=torch.long)
    else:
        cdev = _collective_device(group)
        need_c = need_counts.to(cdev)
        send_c = torch.zeros(W, dtype=torch.long, device=cdev)
        dist.all_to_all_single(send_c, need_c, group=group)
        send_counts = send_c.cpu()
        send_gid = torch.empty(int(send_counts.sum()), dtype=torch.long, device=cdev)
        dist.all_to_all_single(
            send_gid, halo_gid.to(cdev),
            output_split_sizes=send_counts.tolist(), input_split_sizes=need_counts.cpu().tolist(), group=group,
        )
        rows = [torch.zeros(W, dtype=torch.long, device=cdev) for _ in range(W)]
        dist.all_gather(rows, send_c, group=group)
        comm = torch.stack(rows).cpu()
        send_gid = send_gid.to(dev)
    send_off = torch.zeros(W + 1, dtype=torch.long)
    send_off[1:] = send_counts.cumsum(0)
    recv_off, put_fwd = compute_recv_offsets(comm, rank)
    return CommunicationPattern(
        rank=rank,
        world_size=W,
        num_local_vertices=n_loc,
        num_halo_vertices=int(halo_gid.numel()),
        local_edge_list=local_edges,
        send_local_idx=send_gid - nbr0,
        send_offset=send_off,
        recv_offset=recv_off,
        comm_map=comm,
        put_forward_remote_offset=put_fwd,
        put_backward_remote_offset=comm[:, :rank].sum(1),
        local_vertices_global=torch.arange(my0, my1, device=dev),
        halo_vertices_global=halo_gid,
        num_neighbor_local_vertices=n_nbr_loc,
        halo_owner=halo_rank,
        halo_remote_row=halo_gid - noff[halo_rank],
    )


def build_all_patterns(global_edge_list, partitioning, world_size, neighbor_partitioning=None):
    """Plans of every rank in one process (world emulator / tests)."""
    return [
        build_communication_pattern(global_edge_list, partitioning, r, world_size, neighbor_partitioning)
        for r in range(world_size)
    ]
