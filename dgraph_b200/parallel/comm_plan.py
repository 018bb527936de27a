"""Vector-addressed gather / scatter plans ("CommPlan").

Parity: reference ``DGraph/distributed/nccl/_NCCLCommPlan.py:10-333`` -- same dataclasses, field names and
builder signatures (``COO_to_NCCLCommPlan``, ``COO_to_NCCLEdgeConditionedCommPlan``).

One plan answers, for the edges this rank owns and ONE endpoint array ``global_edges_vertex_ids``:
  * which edges index a vertex owned here (``local_edge_idx`` -> ``local_vertex_idx``),
  * which edges index a remote vertex (``boundary_edge_idx``) and which slot of the de-duplicated,
    rank-major receive buffer holds that vertex (``boundary_edge_buffer_map``; ``boundary_edge_splits[q]`` =
    number of distinct vertices needed from rank q),
  * which of this rank's rows other ranks asked for (``boundary_vertex_idx`` split by
    ``boundary_vertex_splits``).

Differences from the reference: device agnostic (no forced ``.cuda()``), works without an initialised process
group when ``world_size == 1``, accepts an explicit ``exchange`` callable (used by the in-process world
emulator), de-duplicated slots are ordered (rank, gid) deterministically, and the plan also carries int32 /
CSR forms for the sm_100a kernels.
"""
from __future__ import annotations

import sys
from dataclasses import dataclass, field, fields
from typing import Any, Callable, Dict, List, Optional, Union

import torch
import torch.distributed as dist


@dataclass
class NCCLGraphCommPlan:
    rank: int
    world_size: int
    num_local_vertices: int
    num_local_edges: int
    local_edge_idx: torch.Tensor
    local_vertex_idx: torch.Tensor
    boundary_edge_idx: torch.Tensor
    boundary_edge_buffer_map: torch.Tensor
    boundary_edge_splits: List[int]
    boundary_vertex_idx: torch.Tensor
    boundary_vertex_splits: List[int]
    # extras for the peer-memory engine (not in the reference): owner rank and row inside the owner's matrix, per edge
    edge_owner: Optional[torch.Tensor] = None
    edge_owner_row: Optional[torch.Tensor] = None
    _cache: Dict[str, Any] = field(default_factory=dict, repr=False, compare=False)

    _TENSORS = ("local_edge_idx", "local_vertex_idx", "boundary_edge_idx", "boundary_edge_buffer_map",
                "boundary_vertex_idx", "edge_owner", "edge_owner_row")

    def to(self, device) -> "NCCLGraphCommPlan":
        for n in self._TENSORS:
            if getattr(self, n) is not None:
                setattr(self, n, getattr(self, n).to(device))
        self._cache.clear()
        return self

    def memory_usage(self, unit: str = "MB") -> Dict[str, Union[float, str]]:
        div = {"B": 1, "KB": 1024, "MB": 1024 ** 2, "GB": 1024 ** 3}.get(unit.upper(), 1024 ** 2)
        cpu = gpu = 0
        for f in fields(self):
            if f.name.startswith("_"):
                continue
            v = getattr(self, f.name)
            if isinstance(v, torch.Tensor):
                n = v.element_size() * v.nelement()
                if v.is_cuda:
                    gpu += n
                else:
                    cpu += n
            elif isinstance(v, list):
                cpu += sys.getsizeof(v) + sum(sys.getsizeof(i) for i in v)
            else:
                cpu += sys.getsizeof(v)
        return {"cpu": cpu / div, "gpu": gpu / div, "total": (cpu + gpu) / div, "unit": unit.upper()}

    # ---- derived forms for the fused kernels -------------------------------------------------
    def edge_to_row(self) -> torch.Tensor:
        """``[E_loc]`` int64: row of ``cat([x_local, recv_buffer])`` that each edge reads / writes."""
        if "e2r" not in self._cache:
            dev = self.local_edge_idx.device
            m = torch.empty(self.num_local_edges, dtype=torch.long, device=dev)
            m[self.local_edge_idx] = self.local_vertex_idx
            m[self.boundary_edge_idx] = self.boundary_edge_buffer_map + self.num_local_vertices
            self._cache["e2r"] = m
        return self._cache["e2r"]

    def __getstate__(self):
        d = dict(self.__dict__)
        d["_cache"] = {}
        return d


@dataclass
class NCCLEdgeConditionedGraphCommPlan:
    rank: int
    world_size: int
    source_graph_plan: NCCLGraphCommPlan
    dest_graph_plan: Optional[NCCLGraphCommPlan] = None

    def to(self, device):
        self.source_graph_plan = self.source_graph_plan.to(device)
        if self.dest_graph_plan is not None:
            self.dest_graph_plan = self.dest_graph_plan.to(device)
        return self

    def reverse(self) -> "NCCLEdgeConditionedGraphCommPlan":
        if self.dest_graph_plan is None:
            raise ValueError("Destination graph plan is None, cannot reverse.")
        return NCCLEdgeConditionedGraphCommPlan(
            rank=self.rank, world_size=self.world_size,
            source_graph_plan=self.dest_graph_plan, dest_graph_plan=self.source_graph_plan,
        )


def compute_edge_slices(dest_ranks, rank, my_dst_global, offset):
    """Split this rank's edges into internal / boundary (``_NCCLCommPlan.py:140-161``)."""
    internal = dest_ranks == rank
    internal_node_idx = my_dst_global[internal] - offset[rank]
    internal_edge_idx = torch.nonzero(internal, as_tuple=True)[0]
    boundary_edge_idx = torch.nonzero(~internal, as_tuple=True)[0]
    return internal_node_idx, internal_edge_idx, my_dst_global[~internal], dest_ranks[~internal], boundary_edge_idx


def fast_2D_unique(indices_1, indices_2):
    """Unique ``(a, b)`` pairs via 64-bit packing, **sorted** by (a, b) (``_NCCLCommPlan.py:164-171``)."""
    packed = (indices_1.to(torch.int64) << 32) | indices_2.to(torch.int64)
    uniq, inv = torch.unique(packed, return_inverse=True, sorted=True)
    return uniq >> 32, uniq & 0xFFFFFFFF, inv


def _dist_exchange_ids(ids: torch.Tensor, send_splits: List[int], group=None):
    """Variable all-to-all of int64 ids through torch.distributed; returns ``(recv_ids, recv_splits)``."""
    from dgraph_b200.parallel.comm_pattern import _collective_device

    W = len(send_splits)
    cdev = _collective_device(group)
    sc = torch.tensor(send_splits, dtype=torch.long, device=cdev)
    rc = torch.zeros(W, dtype=torch.long, device=cdev)
    dist.all_to_all_single(rc, sc, group=group)
    recv_splits = rc.tolist()
    out = torch.empty(sum(recv_splits), dtype=torch.long, device=cdev)
    dist.all_to_all_single(out, ids.to(cdev).contiguous(), output_split_sizes=recv_splits,
                           input_split_sizes=send_splits, group=group)
    return out.to(ids.device), recv_splits


def COO_to_NCCLCommPlan(
    rank: int,
    world_size: int,
    global_edges_vertex_ids: torch.Tensor,
    local_edge_list: torch.Tensor,
    offset: torch.Tensor,
    exchange: Optional[Callable] = None,
) -> NCCLGraphCommPlan:
    """Plan for one endpoint array (``_NCCLCommPlan.py:174-278``).

    Args:
        global_edges_vertex_ids: global vertex id indexed by every edge (either endpoint), ``[E]``; only the
            entries selected by ``local_edge_list`` are read, so a rank may pass its own shard with
            ``local_edge_list = arange(E_loc)``.
        local_edge_list: ids (into ``global_edges_vertex_ids``) of the edges this rank owns.
        offset: ``[W+1]`` contiguous vertex partition.
        exchange: ``f(ids, send_splits) -> (recv_ids, recv_splits)``; defaults to torch.distributed.
    """
    dev = local_edge_list.device
    offset = offset.to(dev).to(torch.long)
    if int(offset[-1]) > 2 ** 32:
        raise ValueError(f"{int(offset[-1])} vertices exceed the 2**32 limit of the packed (rank, id) key")
    mine = global_edges_vertex_ids.to(dev)[local_edge_list].to(torch.long)
    owners = torch.bucketize(mine, offset, right=True) - 1
    return plan_from_requests(rank, world_size, owners, mine - offset[owners],
                              int(offset[rank + 1] - offset[rank]), exchange)


def plan_from_requests(
    rank: int,
    world_size: int,
    owner: torch.Tensor,
    owner_row: torch.Tensor,
    num_local_vertices: int,
    exchange: Optional[Callable] = None,
) -> NCCLGraphCommPlan:
    """Generic plan: edge ``e`` addresses row ``owner_row[e]`` of rank ``owner[e]``'s matrix.

    Used by :func:`COO_to_NCCLCommPlan` (contiguous partitions) and by the legacy
    ``gather(x, indices, rank_mappings)`` API, where ownership comes from an explicit rank map.
    """
    dev = owner.device
    zero_off = torch.zeros(world_size + 1, dtype=torch.long, device=dev)
    loc_v, loc_e, b_row, b_rank, b_e = compute_edge_slices(owner, rank, owner_row, zero_off)
    u_rank, u_row, inv = fast_2D_unique(b_rank, b_row)
    edge_splits = torch.bincount(u_rank, minlength=world_size).tolist()
    if world_size == 1:
        recv_rows, vertex_splits = torch.empty(0, dtype=torch.long, device=dev), [0]
    else:
        ex = exchange if exchange is not None else _dist_exchange_ids
        recv_rows, vertex_splits = ex(u_row, edge_splits)
    return NCCLGraphCommPlan(
        rank=rank,
        world_size=world_size,
        num_local_vertices=int(num_local_vertices),
        num_local_edges=int(owner.numel()),
        local_edge_idx=loc_e,
        local_vertex_idx=loc_v,
        boundary_edge_idx=b_e,
        boundary_edge_buffer_map=inv,
        boundary_edge_splits=[int(x) for x in edge_splits],
        boundary_vertex_idx=recv_rows.to(dev),
        boundary_vertex_splits=[int(x) for x in vertex_splits],
        edge_owner=owner.to(torch.int32),
        edge_owner_row=owner_row.to(torch.int32),
    )


def offline_exchange(global_edge_owner_key: torch.Tensor, global_vertex_ids: torch.Tensor, key_offset: torch.Tensor,
                     vertex_offset: torch.Tensor, rank: int, world_size: int) -> Callable:
    """An ``exchange`` callable that needs no process group: with the *global* edge arrays at hand, what every other rank will
    ask ``rank`` for can be computed locally.  ``global_edge_owner_key`` decides which rank holds an edge (its source vertex with
    ``key_offset``), ``global_vertex_ids`` is the endpoint the plan indexes (partitioned by ``vertex_offset``).  Lets one process
    pre-build and store the plans of all ranks (``experiments/OGB/GenerateCache.py``)."""
    dev = global_vertex_ids.device
    koff, voff = key_offset.to(dev).long(), vertex_offset.to(dev).long()
    holder = torch.bucketize(global_edge_owner_key.to(dev).long().contiguous(), koff, right=True) - 1
    ids = global_vertex_ids.to(dev).long()
    wanted = (ids >= voff[rank]) & (ids < voff[rank + 1]) & (holder != rank)  # edges of other ranks that read my vertices

    def exchange(_ids, _send_splits):
        rows, splits = [], []
        for q in range(world_size):
            if q == rank:
                splits.append(0)
                continue
            u = torch.unique(ids[wanted & (holder == q)]) - voff[rank]  # sorted: the order fast_2D_unique gives the requester
            rows.append(u)
            splits.append(int(u.numel()))
        return (torch.cat(rows) if rows else torch.empty(0, dtype=torch.long, device=dev)), splits

    return exchange


def COO_to_NCCLEdgeConditionedCommPlan(
    rank: int,
    world_size: int,
    global_edges_src: torch.Tensor,
    global_edges_dst: torch.Tensor,
    local_edge_list: torch.Tensor,
    src_offset: torch.Tensor,
    dest_offset: Optional[torch.Tensor],
    exchange: Optional[Callable] = None,
) -> NCCLEdgeConditionedGraphCommPlan:
    """Source- and destination-endpoint plans of the same edge set (``_NCCLCommPlan.py:281-333``)."""
    src_plan = COO_to_NCCLCommPlan(rank, world_size, global_edges_src, local_edge_list, src_offset, exchange)
    if dest_offset is None:
        dest_offset = src_offset
    dst_plan = COO_to_NCCLCommPlan(rank, world_size, global_edges_dst, local_edge_list, dest_offset, exchange)
    return NCCLEdgeConditionedGraphCommPlan(rank=rank, world_size=world_size, source_graph_plan=src_plan,
                                            dest_graph_plan=dst_plan)
