"""Peer-memory engine: the sm_100a communication path (NVLink loads issued from inside the kernels).

Replaces all three reference back-ends on a B200 box: NCCL ``put`` = ``all_to_all_single``
(``nccl/NCCLBackendEngine.py:281-306``), MPI RMA windows and the NVSHMEM extension
(``nvshmem/NVSHMEMBackendEngine.py:22-160``, ``csrc/torch_nvshmem_p2p.cu``).  ``torch.distributed`` (NCCL) is kept
for bootstrap, plan-build metadata and as the generic fallback for arbitrary ``put`` calls; every hot exchange is a
pull kernel over a **symmetric arena**:

* one ``cudaMalloc`` per rank at start-up, exported through CUDA IPC and mapped by all peers (``csrc/peer.cu``);
* named persistent *slots* inside the arena at identical offsets on all ranks, so a peer address is
  ``peer_base[r] + offset`` and nothing is allocated or registered on the hot path;
* a 64-entry signal pad for epoch-flag barriers that run as tiny kernels **on the compute stream** (no host sync).

Producers write straight into slots (the GEMM's ``out=`` is the slot that peers read; features live in ``symmetric_rows``
buffers), so an exchange is a barrier plus the consumer: either the aggregation kernel reads remote rows from the owners while it
runs, or ONE pull kernel lands every halo row behind the owned rows of the same slot (``pull_mode``).  Backward: the CSC walk writes
halo-row gradients into a slot, barrier, ONE deterministic pull-reduce kernel on the owners, in place.  Gradient all-reduce + Adam
= ONE kernel reading all peers' flat gradient buffers.  Every decision that changes which slots exist or how large they are is
taken from all-gathered totals: slot offsets must be identical on all ranks.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _pending
from dgraph_b200.parallel.collective_engine import TorchDistEngine

_ALIGN = 256


class _DevMem:
    """Wraps a raw device range for ``torch.as_tensor`` (the tensor keeps ``owner`` alive)."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr: int, shape, dtype, device, owner=None) -> torch.Tensor:
    n = 1
    for s in shape:
        n *= int(s)
    nbytes = n * torch.empty((), dtype=dtype).element_size()
    if nbytes == 0:
        return torch.empty(shape, dtype=dtype, device=device)
    t = torch.as_tensor(_DevMem(ptr, nbytes, owner), device=device)
    return t.view(dtype).view(shape)


class SymmetricArena:
    """One device allocation per rank, mapped into every peer; bump-allocated named slots at symmetric offsets.

    ``members`` (optional): the ranks of ``group`` (the ranks that map one another) that form this engine's *graph replica*, in
    replica-rank order.  ``ptr_table`` then lists only those -- halo traffic stays inside the replica -- while ``ptr_table_all`` lists
    every mapped rank (the data-parallel gradient sum of the hybrid DP x graph-partition layout reads all of them)."""

    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device, group=None, local_ptrs: Optional[List[int]] = None,
                 members: Optional[List[int]] = None):
        self.rank, self.world, self.device, self.nbytes = rank, world, device, int(nbytes)
        self.members = list(members) if members is not None else list(range(world))
        self._slots: Dict[str, Tuple[int, int]] = {}
        self._top = 2 * _ALIGN  # first 256 bytes: signal pad of the replica's barriers; next 256: pad of the whole-world barriers
        lib = N.lib()
        if local_ptrs is not None:  # in-process emulation: all "ranks" are arenas of this process
            self.base = local_ptrs[rank]
            self.peer_bases = list(local_ptrs)
            self._owned = False
        else:
            p = ctypes.c_void_p()
            N.check(lib.dgb_symm_alloc(ctypes.c_size_t(self.nbytes), ctypes.byref(p)), "dgb_symm_alloc")
            self.base = p.value
            self._owned = True
            handle = (ctypes.c_ubyte * 64)()
            N.check(lib.dgb_symm_export(ctypes.c_void_p(self.base), handle), "dgb_symm_export")
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
            allh = [torch.zeros(64, dtype=torch.uint8, device=device) for _ in range(world)]
            dist.all_gather(allh, mine, group=group)
            self.peer_bases = []
            for r in range(world):
                if r == rank:
                    self.peer_bases.append(self.base)
                    continue
                hb = (ctypes.c_ubyte * 64)(*allh[r].cpu().tolist())
                q = ctypes.c_void_p()
                N.check(lib.dgb_symm_import(hb, ctypes.byref(q)), f"dgb_symm_import(rank {r})")
                self.peer_bases.append(q.value)
        self._ptr_tables: Dict[Tuple, torch.Tensor] = {}
        self.pad_table = self.ptr_table(0)
        self.pad_table_all = self.ptr_table_all(_ALIGN)

    def ptr_table(self, offset: int) -> torch.Tensor:
        """Device array of ``peer_base[r] + offset`` (int64 == pointer) over the replica's ranks, in replica-rank order."""
        t = self._ptr_tables.get((offset, 0))
        if t is None:
            t = torch.tensor([self.peer_bases[r] + offset for r in self.members], dtype=torch.int64, device=self.device)
            self._ptr_tables[(offset, 0)] = t
        return t

    def ptr_table_all(self, offset) -> torch.Tensor:
        """Same over every mapped rank; ``offset`` may be a per-rank list (slots need not be symmetric across replicas)."""
        key = (tuple(offset) if isinstance(offset, (list, tuple)) else offset, 1)
        t = self._ptr_tables.get(key)
        if t is None:
            offs = list(offset) if isinstance(offset, (list, tuple)) else [offset] * len(self.peer_bases)
            t = torch.tensor([b + o for b, o in zip(self.peer_bases, offs)], dtype=torch.int64, device=self.device)
            self._ptr_tables[key] = t
        return t

    def slot(self, name: str, nbytes: int) -> int:
        """Offset of the named slot (allocated on first use; ``nbytes`` must be the same on all ranks)."""
        s = self._slots.get(name)
        if s is not None and s[1] >= nbytes:
            return s[0]
        if s is not None:
            raise RuntimeError(f"symmetric slot {name} would have to grow from {s[1]} to {nbytes} bytes")
        off = (self._top + _ALIGN - 1) // _ALIGN * _ALIGN
        if off + nbytes > self.nbytes:
            raise RuntimeError(
                f"symmetric arena exhausted ({self.nbytes >> 20} MiB): set DGRAPH_B200_ARENA_MB (need {nbytes >> 20} MiB more)")
        self._slots[name] = (off, nbytes)
        self._top = off + nbytes
        return off

    def tensor(self, offset: int, shape, dtype, guard=None) -> torch.Tensor:
        """Tensor over arena memory.  ``guard`` (any object) lives exactly as long as the tensor's storage, i.e. as long as any view
        of it -- autograd's saved copies included."""
        return tensor_from_ptr(self.base + offset, shape, dtype, self.device, owner=(self, guard))

    def close(self):
        if self._owned and self.base:
            lib = N.lib()
            for r, b in enumerate(self.peer_bases):
                if r != self.rank and b:
                    lib.dgb_symm_unmap(ctypes.c_void_p(b))
            lib.dgb_symm_free(ctypes.c_void_p(self.base))
            self.base = 0


class _PatternPeerInfo:
    """Per-pattern device tables for the pull kernels (built once)."""

    def __init__(self, engine: "PeerMemoryEngine", pat):
        dev = engine.device
        W, r = engine.get_world_size(), engine.get_rank()
        self.n_local = pat.num_local_vertices
        self.n_nbr_local = pat.num_neighbor_local_vertices if pat.num_neighbor_local_vertices is not None else pat.num_local_vertices
        self.n_halo = pat.num_halo_vertices
        self.halo_owner = pat.halo_owner.to(dev, torch.int32).contiguous()
        self.halo_row = pat.halo_remote_row.to(dev, torch.int32).contiguous()
        # start pulling at the segment of owner (rank + 1) % W: staggers the ranks over the owners (no incast on owner 0)
        self.rot = int(pat.recv_offset[(r + 1) % W]) if W > 1 else 0
        if self.rot >= max(self.n_halo, 1):
            self.rot = 0
        remote_edges = int((pat.local_edge_list[:, 1] >= self.n_nbr_local).sum()) if self.n_halo else 0
        counts = engine.all_gather_ints([self.n_nbr_local, self.n_halo, int(pat.send_offset[-1]), remote_edges])
        # how often the remote edges of the whole graph re-read a halo row on average: the input of the (collective) pull-route choice
        self.halo_reuse = sum(c[3] for c in counts) / max(sum(c[1] for c in counts), 1)
        self.nbr_local_all = [c[0] for c in counts]
        self.max_nbr_local = max(self.nbr_local_all)
        self.max_aug = max(c[0] + c[1] for c in counts)
        self.max_halo = max(max(c[1] for c in counts), 1)
        # remote rows are addressed as (owner << peer_shift) | row inside the pull kernels (GraphKernelPlan.col_peer)
        self.peer_shift = max(1, (max(self.max_nbr_local, 2) - 1).bit_length())
        assert (W << self.peer_shift) <= (1 << 31), "per-rank partition too large for the 32-bit (owner, row) neighbour code"
        # backward: boundary rows of mine and where their halo copies live on the consumers
        so = pat.send_offset.tolist()
        send_idx = pat.send_local_idx.to("cpu", torch.long)
        fwd_off = pat.put_forward_remote_offset.tolist()  # my segment's start inside rank d's halo range
        ranks, srows = [], []
        for d in range(W):
            k = so[d + 1] - so[d]
            if k:
                ranks.append(torch.full((k,), d, dtype=torch.long))
                srows.append(torch.arange(k, dtype=torch.long) + fwd_off[d])  # row inside d's halo-gradient slot
        if ranks:
            ranks, srows = torch.cat(ranks), torch.cat(srows)
            order = torch.argsort(send_idx, stable=True)
            rows_sorted = send_idx[order]
            uniq, cnt = torch.unique_consecutive(rows_sorted, return_counts=True)
            ptr = torch.zeros(uniq.numel() + 1, dtype=torch.long)
            ptr[1:] = cnt.cumsum(0)
            self.bnd_row = uniq.to(dev, torch.int32)
            self.bnd_ptr = ptr.to(dev, torch.int32)
            self.src_rank = ranks[order].to(dev, torch.int32)
            self.src_row = srows[order].to(dev, torch.int32)
        else:
            z = torch.zeros(0, dtype=torch.int32, device=dev)
            self.bnd_row, self.src_rank, self.src_row = z, z, z
            self.bnd_ptr = torch.zeros(1, dtype=torch.int32, device=dev)
        self.n_bnd = int(self.bnd_row.numel())


class _PlanPeerInfo:
    """Per-CommPlan tables for the pull kernels: where every edge's row lives, and -- for the transposed direction -- which
    slots of which peers hold partial sums for my rows."""

    def __init__(self, engine: "PeerMemoryEngine", plan):
        from dgraph_b200.parallel.kernel_plan import GraphKernelPlan

        dev = engine.device
        W, r = engine.get_world_size(), engine.get_rank()
        self.n_local, self.n_edges = plan.num_local_vertices, plan.num_local_edges
        self.n_remote = int(sum(plan.boundary_edge_splits))  # distinct remote rows my edges address (= my partial-sum slots)
        self.edge_owner = plan.edge_owner.to(dev, torch.int32).contiguous()
        self.edge_owner_row = plan.edge_owner_row.to(dev, torch.int32).contiguous()
        allv = engine.all_gather_ints([self.n_local, self.n_edges, self.n_remote] + [int(x) for x in plan.boundary_edge_splits])
        self.max_local = max(a[0] for a in allv)
        self.max_remote = max(max(a[2] for a in allv), 1)
        # incidence (row of [local | remote-slot] <- edge id): atomic-free pre-reduction of edge rows
        e2r = plan.edge_to_row().to(dev)
        inc = torch.stack([e2r, torch.arange(self.n_edges, device=dev)], dim=1)
        self.inc_plan = GraphKernelPlan.from_edges(inc, self.n_local + self.n_remote, self.n_edges, device=dev,
                                                   num_nbr_local=self.n_edges)
        # my rows requested by rank q (q's order) live in q's partial-sum slot at offset sum_{p < me} q.edge_splits[p]
        ranks, srows, rows = [], [], []
        o = 0
        bvi = plan.boundary_vertex_idx.to("cpu", torch.long)
        for q, k in enumerate(plan.boundary_vertex_splits):
            if k:
                q_off = sum(allv[q][3 + p] for p in range(r))
                ranks.append(torch.full((k,), q, dtype=torch.long))
                srows.append(torch.arange(k, dtype=torch.long) + q_off)
                rows.append(bvi[o:o + k])
            o += k
        if ranks:
            ranks, srows, rows = torch.cat(ranks), torch.cat(srows), torch.cat(rows)
            order = torch.argsort(rows, stable=True)
            uniq, cnt = torch.unique_consecutive(rows[order], return_counts=True)
            ptr = torch.zeros(uniq.numel() + 1, dtype=torch.long)
            ptr[1:] = cnt.cumsum(0)
            self.bnd_row, self.bnd_ptr = uniq.to(dev, torch.int32), ptr.to(dev, torch.int32)
            self.src_rank, self.src_row = ranks[order].to(dev, torch.int32), srows[order].to(dev, torch.int32)
        else:
            z = torch.zeros(0, dtype=torch.int32, device=dev)
            self.bnd_row, self.src_rank, self.src_row = z, z, z
            self.bnd_ptr = torch.zeros(1, dtype=torch.int32, device=dev)
        self.n_bnd = int(self.bnd_row.numel())


def _plan_gather_peer(engine, x2d, plan):
    """``out[e] = X@owner[e][row[e]]``: publish + barrier + ONE pull kernel (replaces index + all-to-all + index)."""
    info = engine.plan_info(plan)
    F = x2d.shape[1]
    pub, off = engine.slot_tensor(("plan_pub", info.uid), (info.max_local, F), x2d.dtype, ring=False)
    pub[: info.n_local].copy_(x2d)
    engine.barrier_on_stream()
    out = torch.empty((info.n_edges, F), dtype=x2d.dtype, device=x2d.device)
    rb = F * x2d.element_size()
    N.check(N.lib().dgb_gather_rows_peer(N.ptr(out), N.i64(rb), N.i64(info.n_edges), N.i32(rb),
                                         N.ptr(engine.arena.ptr_table(off)), N.i64(rb), N.ptr(info.edge_owner),
                                         N.ptr(info.edge_owner_row), N.stream()), "dgb_gather_rows_peer")
    N.count_launch(2)
    engine.mark_read(off)
    return out


def _plan_scatter_peer(engine, e2d, plan):
    """``Y[row[e]] += E[e]`` on the owners: local CSR pre-reduction, publish the partial sums of remote rows, barrier, ONE
    deterministic pull-reduce kernel on the owners (replaces 3 scatter kernels + all-to-all)."""
    from dgraph_b200.ops.segment import _native_seg_sum

    info = engine.plan_info(plan)
    F = e2d.shape[1]
    ip = info.inc_plan
    acc = _native_seg_sum(ip.rowptr, ip.col, ip.num_local, e2d.contiguous(), None, False)  # [n_local + n_remote, F] bf16
    part, off = engine.slot_tensor(("plan_part", info.uid), (info.max_remote, F), e2d.dtype, ring=False)
    part[: info.n_remote].copy_(acc[info.n_local:])
    engine.barrier_on_stream()
    out = acc[: info.n_local].clone()
    N.check(N.lib().dgb_halo_grad_reduce(N.ptr(out), N.i64(F), N.ptr(acc), N.i64(F), N.i32(F), N.i64(info.n_bnd),
                                         N.ptr(info.bnd_row), N.ptr(info.bnd_ptr), N.ptr(info.src_rank), N.ptr(info.src_row),
                                         N.ptr(engine.arena.ptr_table(off)), N.i64(F), N.stream()), "dgb_halo_grad_reduce")
    N.count_launch(2)
    engine.mark_read(off)
    return out


class _PlanGatherPeer(Function):
    @staticmethod
    def forward(ctx, x, plan, engine):
        ctx.plan, ctx.engine, ctx.dim = plan, engine, x.dim()
        x2d = x[0] if x.dim() == 3 else x
        out = _plan_gather_peer(engine, x2d.contiguous(), plan)
        return out.unsqueeze(0) if x.dim() == 3 else out

    @staticmethod
    def backward(ctx, g):
        g2d = g[0] if g.dim() == 3 else g
        out = _plan_scatter_peer(ctx.engine, g2d.contiguous(), ctx.plan)
        return (out.unsqueeze(0) if ctx.dim == 3 else out), None, None


class _PlanScatterPeer(Function):
    @staticmethod
    def forward(ctx, e, plan, engine):
        ctx.plan, ctx.engine, ctx.dim = plan, engine, e.dim()
        e2d = e[0] if e.dim() == 3 else e
        out = _plan_scatter_peer(engine, e2d.contiguous(), plan)
        return out.unsqueeze(0) if e.dim() == 3 else out

    @staticmethod
    def backward(ctx, g):
        g2d = g[0] if g.dim() == 3 else g
        out = _plan_gather_peer(ctx.engine, g2d.contiguous(), ctx.plan)
        return (out.unsqueeze(0) if ctx.dim == 3 else out), None, None


class _SlotGuard:
    """Frees a ring entry when the last tensor that aliases it dies.  The guard rides in the storage owner of the tensor that
    ``slot_tensor(..., track=True)`` hands out, so it lives exactly as long as any view of the buffer -- including the copies that
    autograd saved for a backward pass -- no matter which layer saved it or whether its producer's backward ever runs."""

    def __init__(self, engine, key, off):
        self.engine, self.key, self.off = engine, key, off

    def release(self):
        if self.engine is not None:
            self.engine.release_slot(self.key, self.off)
            self.engine = None

    def __del__(self):
        try:
            self.release()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass


def _halo_grad_reduce(engine, pat, g_aug):
    """Backward of any halo read: gradients of halo rows are published in a symmetric slot and pull-reduced by the owners (in
    place, on the rows of ``g_aug`` that this rank owns)."""
    info = engine.pattern_info(pat)
    F = g_aug.shape[1]
    g_aug = g_aug.contiguous()
    assert g_aug.dtype == torch.bfloat16, "halo gradient reduce kernel is bf16 (fp32 accumulate)"
    gh, off = engine.slot_tensor(("halo_bwd", info.uid), (info.max_halo, F), g_aug.dtype, ring=False)
    gh[: info.n_halo].copy_(g_aug[info.n_nbr_local:])
    engine.barrier_on_stream()
    gx = g_aug[: info.n_nbr_local]
    engine.halo_grad_reduce_inplace(info, gx, F, F, off, F)
    engine.mark_read(off)  # the next writer of this slot inserts a barrier only if none happened in between
    return gx


class _FusedPullAggregate(Function):
    """``out[i] = (1/deg) sum_{j in N(i)} x[j]`` where remote neighbour rows are read **directly from the owner GPU's matrix**
    inside the aggregation kernel (no halo buffer, no exchange kernel): publish + barrier + ONE segment kernel."""

    @staticmethod
    def forward(ctx, x, engine, pat, kp, mean):
        from dgraph_b200.ops.segment import _native_seg_sum

        info = engine.pattern_info(pat)
        F = x.shape[1]
        pub, off = engine.published(x, info)  # no copy when x already lives in a symmetric slot (engine.symmetric_rows)
        engine.barrier_on_stream()
        out = _native_seg_sum(kp.rowptr, kp.col_peer(info.peer_shift), kp.num_local, pub, None, mean,
                              peer=(engine.arena.ptr_table(off), engine.get_world_size(), info.peer_shift, pub.stride(0)))
        engine.mark_read(off)
        ctx.engine, ctx.pat, ctx.kp, ctx.mean = engine, pat, kp, mean
        return out

    @staticmethod
    def backward(ctx, g):
        from dgraph_b200.ops.segment import _mean_bwd_weights, _native_seg_sum

        kp, engine = ctx.kp, ctx.engine
        info = engine.pattern_info(ctx.pat)
        F = g.shape[1]
        w = _mean_bwd_weights(kp) if ctx.mean else None
        # the CSC walk writes the halo rows' gradients straight into the symmetric slot; the owners reduce them in place
        gh, off = engine.slot_tensor(("halo_bwd", info.uid), (info.max_halo, F), torch.bfloat16, ring=False)
        gx = _native_seg_sum(kp.colptr, kp.row, kp.num_aug, g.contiguous(), w, False, out2=gh, out_split=info.n_nbr_local)
        engine.barrier_on_stream()
        engine.halo_grad_reduce_inplace(info, gx, F, F, off, F)
        engine.mark_read(off)
        return gx, None, None, None, None


class _HaloAug(Function):
    """``aug = cat([x, halo(x)])`` through the symmetric arena; backward pull-reduces halo gradients on the owners."""

    @staticmethod
    def forward(ctx, x, engine, pat, defer=False, track=False):
        info = engine.pattern_info(pat)
        F = x.shape[1]
        assert x.dim() == 2 and (F * x.element_size()) % 16 == 0, "feature rows must be multiples of 16 bytes"
        own = engine.symmetric_slot_of(x, info.max_aug)
        if own is not None:  # the rows already sit at the head of an augmented symmetric buffer (engine.symmetric_rows): no copy
            off = own
            aug = engine.arena.tensor(off, (info.max_aug, F), x.dtype)[: info.n_nbr_local + info.n_halo]
        else:
            # `track` (decided by the caller: grad mode is always off in here) keeps this ring entry out of circulation while any
            # view of the buffer is alive, e.g. saved for a backward pass (see _SlotGuard)
            aug, off = engine.slot_tensor(("halo_fwd", info.uid), (info.max_aug, F), x.dtype, ring=True, track=track)
            aug = aug[: info.n_nbr_local + info.n_halo]
            aug[: info.n_nbr_local].copy_(x)
        engine.barrier_on_stream()
        rb = F * x.element_size()

        def pull():
            N.check(N.lib().dgb_halo_pull(N.ptr(aug[info.n_nbr_local:]), N.i64(rb), N.i64(info.n_halo), N.i32(rb),
                                          N.ptr(engine.arena.ptr_table(off)), N.i64(rb), N.ptr(info.halo_owner),
                                          N.ptr(info.halo_row), N.i64(info.rot), N.stream()), "dgb_halo_pull")

        if defer and info.n_halo > 0:
            # pull on the side stream, ordered after the barrier; the consumer overlaps local work with it (ops/_pending.py)
            main = torch.cuda.current_stream()
            side = engine.side_stream()
            side.wait_stream(main)
            with torch.cuda.stream(side):
                pull()
                ev = torch.cuda.Event()
                ev.record(side)
            _pending.put(aug, ev)
        else:
            pull()
        N.count_launch(2)
        engine.mark_read(off)
        ctx.engine, ctx.pat = engine, pat
        return aug

    @staticmethod
    def backward(ctx, g_aug):
        return _halo_grad_reduce(ctx.engine, ctx.pat, g_aug), None, None, None, None


class PeerMemoryEngine(TorchDistEngine):
    """``group`` (optional): the process group of one graph replica in the hybrid data-parallel x graph-partition layout
    (``Communicator.init_process_group(..., ranks_per_graph=k)``, reference arithmetic ``nccl/NCCLBackendEngine.py:56-64``).  Halo /
    plan traffic and its barriers then stay inside the replica, while every rank of the job still maps every other rank's arena so
    that the fused gradient all-reduce + Adam kernel sums over the whole world."""

    def __init__(self, ranks_per_graph: int = -1, arena_mb: Optional[int] = None, group=None, **kwargs):
        super().__init__(dist_backend="nccl", ranks_per_graph=ranks_per_graph, group=group, **kwargs)
        self.device = torch.device("cuda", torch.cuda.current_device())
        N.lib()
        self._epoch = 0
        self._epoch_all = 0
        self._job_rank = dist.get_rank() if dist.is_initialized() else 0
        self._job_world = dist.get_world_size() if dist.is_initialized() else 1
        self._uid = 0
        self._rings: Dict[Tuple, List] = {}
        self._by_off: Dict[int, dict] = {}
        self._sym_rows: Dict[int, Tuple] = {}
        self._flat_offsets: Dict[int, List[int]] = {}
        self._ar_parity = 0
        self._timeout_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._timeout_host = self._timeout_ev = None
        self._spins = ctypes.c_longlong(int(os.environ.get("DGRAPH_B200_BARRIER_SPINS", str(1 << 26))))  # read once
        # DGRAPH_B200_DEVICE_EPOCH=1 (what `bench.py` sets for its multi-GPU CUDA-graph runs): barrier epochs counted in device memory, so a training step
        # with exchanges can be captured in a CUDA graph.  Fixed at construction: the two barrier forms must never be mixed.
        self.device_epoch = os.environ.get("DGRAPH_B200_DEVICE_EPOCH", "0") == "1"
        self._epoch_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.arena: Optional[SymmetricArena] = None
        self._epoch_all_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        if self._job_world > 1:
            mb = arena_mb or int(os.environ.get("DGRAPH_B200_ARENA_MB", "8192"))
            members = None
            if group is not None and self._world != self._job_world:  # replica = a contiguous block of job ranks
                members = dist.get_process_group_ranks(group)
            self.arena = SymmetricArena(mb << 20, self._job_rank, self._job_world, self.device, group=None, members=members)
            dist.barrier()

    # ---- small host-side helpers ---------------------------------------------------------------
    def all_gather_ints(self, vals: List[int]) -> List[List[int]]:
        if self._world == 1:
            return [list(vals)]
        t = torch.tensor(vals, dtype=torch.long, device=self.device)
        out = [torch.zeros_like(t) for _ in range(self._world)]
        dist.all_gather(out, t, group=self._group)
        return [o.tolist() for o in out]

    def pattern_info(self, pat) -> _PatternPeerInfo:
        """Per-pattern tables live on the pattern object itself (``id()`` values are recycled after garbage collection)."""
        i = pat._cache.get("peer_info")
        if i is None or i.engine_ref is not self:
            i = _PatternPeerInfo(self, pat)
            i.engine_ref = self
            self._uid += 1
            i.uid = self._uid
            pat._cache["peer_info"] = i
        return i

    def plan_info(self, plan) -> _PlanPeerInfo:
        i = plan._cache.get("peer_info")
        if i is None or i.engine_ref is not self:
            i = _PlanPeerInfo(self, plan)
            i.engine_ref = self
            self._uid += 1
            i.uid = self._uid
            plan._cache["peer_info"] = i
        return i

    # ---- CommPlan gather / scatter over peer memory -----------------------------------------------------
    def _peer_plan_ok(self, t, plan) -> bool:
        F = t.shape[-1]
        return (self._world > 1 and plan is not None and getattr(plan, "edge_owner", None) is not None and t.is_cuda
                and t.dtype == torch.bfloat16 and (F * 2) % 16 == 0 and t.dim() in (2, 3) and (t.dim() == 2 or t.shape[0] == 1))

    def gather(self, src_tensor=None, indices=None, rank_mappings=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, _ = self._reference_keywords(src_tensor, rank_mappings, None, kwargs)
        plan = comm_plan if comm_plan is not None else (indices if hasattr(indices, "boundary_edge_idx") else None)
        if self._peer_plan_ok(src_tensor, plan):
            return _PlanGatherPeer.apply(src_tensor, plan, self)
        return super().gather(src_tensor, indices, rank_mappings, comm_plan=comm_plan, **kwargs)

    def scatter(self, src_tensor=None, indices=None, rank_mappings=None, output_size=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, output_size = self._reference_keywords(src_tensor, rank_mappings, output_size, kwargs)
        plan = comm_plan if comm_plan is not None else (indices if hasattr(indices, "boundary_edge_idx") else None)
        if self._peer_plan_ok(src_tensor, plan):
            return _PlanScatterPeer.apply(src_tensor, plan, self)
        return super().scatter(src_tensor, indices, rank_mappings, output_size, comm_plan=comm_plan, **kwargs)

    # ---- symmetric slots -----------------------------------------------------------------------
    def slot_tensor(self, key, shape, dtype, ring: bool = False, track: bool = False):
        """Persistent symmetric tensor for ``key``.  ``ring=True``: rotate over entries so that a buffer saved for
        backward is never overwritten by the next forward call with the same key (entries are added on demand; all
        ranks execute the same program, so they grow their rings in lock-step and offsets stay symmetric)."""
        nbytes = int(torch.empty((), dtype=dtype).element_size())
        for s in shape:
            nbytes *= int(s)
        rk = (key, tuple(shape), dtype)
        entries = self._rings.setdefault(rk, [])
        pick = None
        if ring:
            for e in entries:
                if not e["busy"]:
                    pick = e
                    break
        elif entries:
            pick = entries[0]
        if pick is None:
            off = self.arena.slot(f"{rk}#{len(entries)}", max(nbytes, 16))
            pick = {"off": off, "busy": False, "read_epoch": -1}
            entries.append(pick)
            self._check_symmetric(rk, off, nbytes)
        guard = None
        if ring and track:
            pick["busy"] = True
            guard = _SlotGuard(self, key, pick["off"])
        self._by_off[pick["off"]] = pick
        self.note_write(pick["off"])
        return self.arena.tensor(pick["off"], shape, dtype, guard), pick["off"]

    def _check_symmetric(self, rk, off: int, nbytes: int) -> None:
        """``DGRAPH_B200_VALIDATE=1``: every new slot is compared across the ranks of the replica (one small all-gather per
        allocation, i.e. during the first steps only).  Peers address a slot as ``peer_base + offset``, so a rank that allocated a
        different slot at this point -- a data-dependent branch, a ring that grew on one rank only -- would make every later
        exchange read the wrong memory without any error."""
        from dgraph_b200.utils.debug import validate_enabled

        if not validate_enabled() or self._world == 1 or not getattr(self.arena, "_owned", False):
            return
        mine = torch.tensor([off, nbytes], dtype=torch.long, device=self.device)
        allv = [torch.zeros_like(mine) for _ in range(self._world)]
        dist.all_gather(allv, mine, group=self._group)
        seen = [tuple(v.tolist()) for v in allv]
        if any(v != seen[0] for v in seen):
            raise RuntimeError(f"symmetric slot {rk} differs between ranks (offset, bytes per rank: {seen}): the ranks did not "
                               "allocate the same slots in the same order")

    def note_write(self, off):
        """This rank is about to overwrite slot ``off``.  Peers read it during epoch ``read_epoch``; they are provably done once any
        later barrier has completed, so a barrier is inserted only when none happened since."""
        e = self._by_off.get(off)
        if e is not None and e["read_epoch"] == self._epoch:
            self.barrier_on_stream()

    def mark_read(self, off):
        """Record that peers read slot ``off`` after the latest barrier (see ``note_write``)."""
        self._by_off[off]["read_epoch"] = self._epoch

    # ---- zero-copy publication --------------------------------------------------------------------
    def symmetric_rows(self, pat, F: int, dtype=torch.bfloat16, tag=0) -> torch.Tensor:
        """``[num_local, F]`` tensor at the head of an augmented symmetric buffer of its own.  Features written here (host -> device
        copy, a producing kernel's ``out=``) are exchanged without the copy into a slot: ``halo_exchange_aug`` pulls the halo rows
        in right behind them and returns the same memory as ``cat([x, halo])``; ``aggregate_pull`` lets peers read it in place.
        Peers read these rows between the exchange's barrier and the next barrier on the stream, so refill the buffer only in stream
        order behind a later barrier -- e.g. between training steps (the optimizer step ends with one)."""
        if self._world == 1:
            return torch.empty((pat.num_local_vertices, F), dtype=dtype, device=self.device)
        info = self.pattern_info(pat)
        buf, off = self.slot_tensor(("rows", info.uid, F, tag), (info.max_aug, F), dtype, ring=False)
        self._sym_rows[buf.data_ptr()] = (off, info.max_aug, F, dtype)
        return buf[: info.n_nbr_local]

    def symmetric_slot_of(self, x: torch.Tensor, min_rows: int):
        """Arena offset of ``x`` when it is a ``symmetric_rows`` buffer with at least ``min_rows`` rows of the same width."""
        if self.arena is None:
            return None
        hit = self._sym_rows.get(x.data_ptr())
        if hit is None or not x.is_contiguous():
            return None
        off, rows, F, dtype = hit
        return off if (rows >= min_rows and x.dim() == 2 and x.shape[1] == F and x.dtype == dtype) else None

    def published(self, x: torch.Tensor, info):
        """``(tensor peers may read, arena offset)`` holding the rows of ``x``: ``x`` itself when it already lives in a symmetric
        buffer, else a copy into the pattern's publication slot."""
        own = self.symmetric_slot_of(x, info.n_nbr_local)
        if own is not None:
            return x, own
        pub, off = self.slot_tensor(("pub", info.uid, x.shape[1]), (info.max_nbr_local, x.shape[1]), x.dtype, ring=False)
        pub[: info.n_nbr_local].copy_(x)
        return pub, off

    # ---- what travels and how ------------------------------------------------------------------------
    def pull_mode(self, info, kp) -> str:
        """How a consumer fetches the remote rows it needs.  ``"kernel"``: the consuming kernel reads them from the owners while
        it runs (each remote edge costs one NVLink row read, nothing is staged).  ``"buffer"``: one pull kernel copies every halo
        row once into local memory first.  Peer loads bypass the local L2, so re-reads cross NVLink again: ``auto`` takes the
        in-kernel route only while the remote edges re-read a halo row fewer than ``DGRAPH_B200_PULL_REUSE`` (1.6) times on
        average OVER ALL RANKS.  The choice must be the same on every rank -- the two routes use differently sized symmetric slots,
        and a rank that chose differently would shift every later slot offset -- so it is made from the all-gathered totals."""
        mode = os.environ.get("DGRAPH_B200_PULL", "auto")
        if mode in ("kernel", "buffer"):
            return mode
        c = getattr(info, "_pull_mode", None)
        if c is None:
            c = info._pull_mode = "kernel" if info.halo_reuse < float(os.environ.get("DGRAPH_B200_PULL_REUSE", "1.6")) else "buffer"
        return c

    def halo_grad_reduce_inplace(self, info, gx, ld_gx, H, slot_off, ld_slot, plan=None):
        """``gx[b] += sum over the ranks holding b as a halo row of their slot rows`` for my boundary rows ``b`` (ONE deterministic
        pull-reduce kernel; ``gx`` may be a strided view such as the ``dQ`` half of a ``[n, 2H]`` buffer).  ``plan``: a restricted
        ``(bnd_row, bnd_ptr, src_rank, src_row)`` from :meth:`restricted_reduce_plan`."""
        bnd_row, bnd_ptr, src_rank, src_row = plan if plan is not None else (info.bnd_row, info.bnd_ptr, info.src_rank, info.src_row)
        N.check(N.lib().dgb_halo_grad_reduce(N.ptr(gx), N.i64(ld_gx), N.ptr(gx), N.i64(ld_gx), N.i32(H), N.i64(bnd_row.numel()),
                                             N.ptr(bnd_row), N.ptr(bnd_ptr), N.ptr(src_rank), N.ptr(src_row),
                                             N.ptr(self.arena.ptr_table(slot_off)), N.i64(ld_slot), N.stream()),
                "dgb_halo_grad_reduce")
        N.count_launch()

    def restricted_reduce_plan(self, info, kp, rows, colptr_r):
        """Pull-reduce plan limited to the halo copies that can carry a non-zero gradient when the incoming gradient is zero outside
        the central rows ``rows``: a halo row of rank q matters only if the restricted CSC (``colptr_r`` =
        ``kp.restricted_csc(rows)[0]``) gives it an edge.  Every rank publishes one flag per halo row, the owners gather the flags of
        their copies over NVLink (one barrier + one ``gather_rows_peer_kernel``, once per label set; cached like the restricted
        CSC itself) and compact their boundary tables.  COLLECTIVE: all ranks must call it in the same order."""
        cache = info.__dict__.setdefault("_restricted_reduce", {})
        key = (rows.data_ptr(), rows.numel(), rows._version)
        hit = cache.get(key)
        if hit is not None and hit[0] is rows:
            return hit[1]
        dev = self.device
        n, n_halo = info.n_nbr_local, info.n_halo
        flags, off = self.slot_tensor(("nzflags", info.uid), (info.max_halo, 4), torch.int32)
        if n_halo:
            cp = colptr_r.long()
            flags[:n_halo].zero_()
            flags[:n_halo, 0] = (cp[n + 1:n + n_halo + 1] > cp[n:n + n_halo]).to(torch.int32)
        self.barrier_on_stream()
        n_src = int(info.src_row.numel())
        got = torch.zeros((max(n_src, 1), 4), dtype=torch.int32, device=dev)
        if n_src:
            N.check(N.lib().dgb_gather_rows_peer(N.ptr(got), N.i64(16), N.i64(n_src), N.i32(16), N.ptr(self.arena.ptr_table(off)),
                                                 N.i64(16), N.ptr(info.src_rank), N.ptr(info.src_row), N.stream()),
                    "dgb_gather_rows_peer")
            N.count_launch()
        self.mark_read(off)
        keep = got[:n_src, 0] != 0
        pref = torch.zeros(n_src + 1, dtype=torch.long, device=dev)
        torch.cumsum(keep, 0, out=pref[1:])
        cnt = pref[info.bnd_ptr.long()[1:]] - pref[info.bnd_ptr.long()[:-1]]
        live = cnt > 0
        ptr = torch.zeros(int(live.sum()) + 1, dtype=torch.int32, device=dev)
        ptr[1:] = cnt[live].cumsum(0).to(torch.int32)
        plan = (info.bnd_row[live].contiguous(), ptr, info.src_rank[keep].contiguous(), info.src_row[keep].contiguous())
        if len(cache) >= 4:
            cache.clear()
        cache[key] = (rows, plan)
        return plan

    def release_slot(self, key, off):
        for rk, entries in self._rings.items():
            if rk[0] == key:
                for e in entries:
                    if e["off"] == off:
                        e["busy"] = False

    def reset_arena(self):
        """Forget every slot and start allocating from the bottom of the arena again (the arena is a bump allocator: slots are
        never freed one by one).  COLLECTIVE and only legal when no tensor handed out by ``slot_tensor`` / ``symmetric_rows`` is in
        use any more -- e.g. between the points of a bandwidth sweep.  Ends with a barrier on the stream."""
        if self.arena is None:
            return
        self.barrier_on_stream("world")
        self._rings.clear()
        self._by_off.clear()
        self._sym_rows.clear()
        self._flat_offsets.clear()
        self.arena._slots.clear()
        self.arena._top = 2 * _ALIGN

    def new_step(self):
        """Mark every ring entry reusable (call once per training step when running without autograd)."""
        for entries in self._rings.values():
            for e in entries:
                e["busy"] = False

    def allocate_buffer(self, size, dtype, device):
        return torch.empty(size, dtype=dtype, device=device)

    # ---- ordering --------------------------------------------------------------------------------
    def barrier_on_stream(self, scope: str = "graph"):
        """All ranks' streams have reached this point (epoch-flag barrier kernel; no host synchronisation).  ``scope="graph"``: the
        ranks of this graph replica; ``scope="world"``: every rank of the job (hybrid layout; identical when there is one replica)."""
        whole = scope == "world" and self._job_world != self._world
        if (self._job_world if whole else self._world) == 1:
            return
        _pending.wait_all()  # no rank may run ahead of a pull that still reads its peers' slots
        self._epoch += 1  # a world barrier orders the replica too: slots read before it may be rewritten after it
        if whole:
            self._epoch_all += 1
            pads, rank, world, epoch, epoch_dev = (self.arena.pad_table_all, self._job_rank, self._job_world, self._epoch_all,
                                                   self._epoch_all_dev)
        else:
            pads, rank, world, epoch, epoch_dev = self.arena.pad_table, self._rank, self._world, self._epoch, self._epoch_dev
        if self.device_epoch:  # the kernel takes the epoch from (and advances) the device counter: capturable
            N.check(N.lib().dgb_peer_barrier_dev(N.ptr(pads), N.i32(rank), N.i32(world), N.ptr(epoch_dev),
                                                 N.ptr(self._timeout_flag), self._spins, N.stream()), "dgb_peer_barrier_dev")
            return
        N.check(N.lib().dgb_peer_barrier(N.ptr(pads), N.i32(rank), N.i32(world), ctypes.c_uint(epoch & 0xFFFFFFFF),
                                         N.ptr(self._timeout_flag), self._spins, N.stream()), "dgb_peer_barrier")

    def check_timeouts(self):
        """Raise if any barrier since the last call gave up waiting (a rank did not arrive within ``DGRAPH_B200_BARRIER_SPINS`` polls of
        ~64 ns each -- default 2^26, about 5 s; the kernels behind a timed-out barrier read incomplete peer data, so the step that
        contains it is garbage).  Synchronises with the device."""
        if int(self._timeout_flag.item()) != 0:
            raise RuntimeError("peer barrier timed out (a rank did not arrive): results since the last check are invalid")

    def poll_timeouts(self):
        """Cheap per-step form of :meth:`check_timeouts` (called by the fused optimizer step): copies the flag into pinned host memory
        on the stream and inspects the copy made one step earlier, so it never waits for the device -- a timeout surfaces one step
        late instead of never."""
        if self.arena is None or torch.cuda.is_current_stream_capturing():
            return
        if self._timeout_host is None:
            self._timeout_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._timeout_ev = torch.cuda.Event()
        elif self._timeout_ev.query() and int(self._timeout_host[0]) != 0:
            raise RuntimeError("peer barrier timed out (a rank did not arrive): the last optimizer steps used incomplete peer data")
        self._timeout_host.copy_(self._timeout_flag, non_blocking=True)
        self._timeout_ev.record()

    # ---- fused halo exchange ---------------------------------------------------------------------
    def side_stream(self) -> "torch.cuda.Stream":
        s = getattr(self, "_side_stream", None)
        if s is None:
            s = self._side_stream = torch.cuda.Stream(device=self.device)
        return s

    def halo_exchange_aug(self, x: torch.Tensor, pat, defer: bool = False) -> torch.Tensor:
        """``cat([x, halo])`` as one symmetric buffer (rows ``[0, n_local)`` = ``x``).  ``defer=True``: the halo rows arrive
        on a side stream; hand the buffer only to a consumer that waits for ``ops._pending.take(buffer)``."""
        if self._world == 1:
            return x
        # grad mode is off inside autograd.Function.forward, so whether the buffer may be saved for a backward pass is decided here
        return _HaloAug.apply(x, self, pat, bool(defer) and os.environ.get("DGRAPH_B200_OVERLAP", "0") != "0",
                              torch.is_grad_enabled())

    def halo_exchange(self, x: torch.Tensor, pat) -> torch.Tensor:
        if self._world == 1:
            return x.new_zeros((0, x.shape[1]))
        info = self.pattern_info(pat)
        return self.halo_exchange_aug(x, pat)[info.n_nbr_local:]

    def aggregate_pull(self, x: torch.Tensor, pat, mean: bool = True) -> torch.Tensor:
        """Neighbourhood sum / mean with remote rows pulled inside the aggregation kernel (F1a + F7 fused)."""
        kp = pat.kernel_plan(x.device)
        if self._world == 1:
            from dgraph_b200.ops.segment import segment_sum

            return segment_sum(x, kp, mean=mean)
        return _FusedPullAggregate.apply(x, self, pat, kp, mean)

    def publish(self, x: torch.Tensor, pat):
        """Make ``x`` readable by the peers (in place when it is a ``symmetric_rows`` buffer, else one copy into a slot) and order it;
        returns ``(local tensor, (peer pointer table, halo_owner, halo_row, n_local))`` -- the ``peers`` entry that
        ``ops.gemm_native.gather_gemm_bf16`` needs to pull remote rows itself."""
        info = self.pattern_info(pat)
        pub, off = self.published(x, info)
        self.barrier_on_stream()
        self.mark_read(off)
        return pub[: info.n_nbr_local], (self.arena.ptr_table(off), info.halo_owner, info.halo_row, info.n_nbr_local)

    # ---- small all-reduces over peer memory --------------------------------------------------------
    def all_reduce(self, tensor, op="sum"):
        """Small fp32 sums (BatchNorm statistics, loss scalars; up to 64 K elements) are published in a symmetric slot and summed by
        ONE kernel that reads every peer's copy (``dgb_allreduce_sum_f32``; same summation order on every rank, so the results are
        bit-identical across ranks): flag barrier + kernel instead of an NCCL launch, and capturable with the device-epoch barriers.
        Two slots alternate so that back-to-back reductions need no extra barrier before the slot is rewritten.
        ``DGRAPH_B200_PEER_ALLREDUCE=0`` sends everything through NCCL, as does anything larger or not a sum."""
        if (self._world > 1 and op == "sum" and os.environ.get("DGRAPH_B200_PEER_ALLREDUCE", "1") == "1" and tensor.is_cuda
                and tensor.dtype == torch.float32 and tensor.is_contiguous() and 0 < tensor.numel() <= (1 << 16)):
            n = tensor.numel()
            n4 = (n + 3) // 4 * 4
            self._ar_parity ^= 1
            buf, off = self.slot_tensor(("allreduce", n4, self._ar_parity), (n4,), torch.float32, ring=False)
            if n4 != n:
                buf[n:].zero_()
            buf[:n].copy_(tensor.reshape(-1))
            self.barrier_on_stream()
            flat = tensor.reshape(-1)
            out = flat if n4 == n else torch.empty(n4, dtype=torch.float32, device=tensor.device)
            N.check(N.lib().dgb_allreduce_sum_f32(N.ptr(out), N.ptr(self.arena.ptr_table(off)), N.i32(self._world), N.i64(n4),
                                                  N.stream()), "dgb_allreduce_sum_f32")
            N.count_launch(2)
            self.mark_read(off)
            if n4 != n:
                flat.copy_(out[:n])
            return tensor
        return super().all_reduce(tensor, op)

    # ---- fused gradient all-reduce + Adam ----------------------------------------------------------
    def symmetric_flat(self, key: str, numel: int, dtype=torch.float32) -> Tuple[torch.Tensor, int]:
        n = (numel + 3) // 4 * 4
        mx = max(c[0] for c in self.all_gather_ints([n]))
        t, off = self.slot_tensor(("flat", key), (mx,), dtype)
        if self._job_world != self._world:  # hybrid layout: replicas need not have allocated their slots in the same order
            offs = torch.tensor([off], dtype=torch.long, device=self.device)
            allo = [torch.zeros_like(offs) for _ in range(self._job_world)]
            dist.all_gather(allo, offs)
            self._flat_offsets[off] = [int(o) for o in allo]
        return t, off

    def _grad_table(self, flat_g_off):
        """Pointers to the flat gradient buffer of every rank that holds a copy of the weights (the whole job)."""
        if self._job_world != self._world:
            return self.arena.ptr_table_all(self._flat_offsets[flat_g_off]), self._job_world
        return self.arena.ptr_table(flat_g_off), self._world

    def adam_allreduce(self, flat_p, flat_g_off, flat_m, flat_v, n, lr, b1, b2, eps, wd, grad_scale, step):
        table, w = self._grad_table(flat_g_off)
        self.barrier_on_stream("world")  # every rank finished writing its gradients
        N.check(N.lib().dgb_adam_allreduce(N.ptr(flat_p), N.ptr(table), N.i32(w), N.ptr(flat_m), N.ptr(flat_v), N.ptr(None),
                                           N.i64(n), N.f32(lr), N.f32(b1), N.f32(b2), N.f32(eps), N.f32(wd), N.f32(grad_scale),
                                           N.i32(step), N.stream()), "dgb_adam_allreduce")
        N.count_launch(2)
        self.barrier_on_stream("world")  # nobody zeroes its gradient buffer while a peer is still reading it
        self.poll_timeouts()

    def adam_allreduce_capturable(self, flat_p, flat_g_off, flat_m, flat_v, n, b1, b2, eps, wd, grad_scale, step_dev, dev_state):
        """``adam_allreduce`` with the step counter and the learning rate in device memory (CUDA-graph capture; needs the
        device-epoch barriers)."""
        assert self.device_epoch, "capturable multi-GPU steps need DGRAPH_B200_DEVICE_EPOCH=1 (set before the engine is created)"
        table, w = self._grad_table(flat_g_off)
        self.barrier_on_stream("world")
        N.check(N.lib().dgb_adam_allreduce_capturable(N.ptr(flat_p), N.ptr(table), N.i32(w), N.ptr(flat_m), N.ptr(flat_v),
                                                      N.ptr(None), N.i64(n), N.f32(b1), N.f32(b2), N.f32(eps), N.f32(wd),
                                                      N.f32(grad_scale), N.ptr(step_dev), N.ptr(dev_state), N.stream()),
                "dgb_adam_allreduce_capturable")
        N.count_launch(3)
        self.barrier_on_stream("world")
        self.poll_timeouts()

    def finalize(self) -> None:
        if self.arena is not None:
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier()
            self.arena.close()
            self.arena = None


def emulated_engines(patterns, arena_bytes: int = 64 << 20) -> List["PeerMemoryEngine"]:
    """``W = len(patterns)`` peer-memory engines inside ONE process on the current GPU: every "rank's" symmetric arena is an ordinary
    device buffer (a peer pointer is just a pointer), barriers are no-ops, and the caller runs every phase rank after rank on one
    stream -- so no kernel ever waits for another one.  What a rank reads from its peers is whatever they wrote before, hence run
    each phase twice when the result matters (the first pass fills the slots).  For tests, single-GPU profiling of the peer kernels
    (`benchmarks/kernel_zoo.py`) and the smoke test; real runs use one process per GPU."""
    W = len(patterns)
    counts = [[p.num_local_vertices if p.num_neighbor_local_vertices is None else p.num_neighbor_local_vertices, p.num_halo_vertices,
               int(p.send_offset[-1]), 0] for p in patterns]
    bufs = [torch.zeros(arena_bytes, dtype=torch.uint8, device="cuda") for _ in range(W)]
    ptrs = [b.data_ptr() for b in bufs]

    class _Emulated(PeerMemoryEngine):
        def __init__(self, rank):  # noqa: D401 - no process group: everything the kernels need is set by hand
            self._rank, self._world, self._group = rank, W, None
            self._job_rank, self._job_world = rank, W
            self._ranks_per_graph = -1
            self.device = torch.device("cuda", torch.cuda.current_device())
            self._epoch, self._epoch_all, self._uid, self._ar_parity = 0, 0, 0, 0
            self._rings, self._by_off, self._sym_rows, self._flat_offsets = {}, {}, {}, {}
            self._timeout_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._timeout_host = self._timeout_ev = None
            self.device_epoch = False
            self.arena = SymmetricArena(arena_bytes, rank, W, self.device, local_ptrs=ptrs)
            self._keep = bufs

        def get_rank(self):
            return self._rank

        def get_world_size(self):
            return self._world

        def all_gather_ints(self, vals):
            assert len(vals) == 4, "emulation knows the pattern counts only"
            return counts

        def barrier_on_stream(self, scope: str = "graph"):
            self._epoch += 1

    return [_Emulated(r) for r in range(W)]
