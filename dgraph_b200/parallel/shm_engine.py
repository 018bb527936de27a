"""One-sided engine for CPU ranks of one node: windows in POSIX shared memory (``csrc/host/shm_window.cpp``).

Role: the reference's MPI engine (``DGraph/distributed/mpi/MPIBackendEngine.py:33-530``) -- vector-addressed ``gather`` / ``scatter``
as remote-memory-access epochs (``Win.Fence`` -> one ``Get`` / ``Accumulate(SUM)`` per row -> ``Win.Fence``) -- without MPI, which
this image does not have.  What is kept: the call signatures and tensor conventions (``[1, N, F]`` data, ``[1, E]`` indices, ``[1, E]``
owner map, ``rank_mapping=None`` meaning an even contiguous split, the row inside the owner being ``index % rows``), gather and
scatter being each other's backward, and the one-sided ``put`` that honours ``remote_offsets`` (``Engine.py:67-86``).  What is
different: the window is created once and grown on demand instead of ``Win.Create`` / ``Win.Free`` per call, the per-row loop runs in
C++ over a few threads instead of one Python iteration per row, duplicate targets need no local pre-reduction (the accumulate is a
row-locked add), and fences are bounded so that a dead rank raises instead of hanging the job.

``comm_plan=`` gather / scatter run as window epochs too (every edge addresses ``(owner, row)``; the plan's send lists are not
needed).  Collectives (plan building, all-reduce, barrier) go through gloo, inherited from :class:`TorchDistEngine`.  On a GPU box the one-sided engine is :class:`PeerMemoryEngine` (NVLink loads and stores issued from inside
the kernels); this one is what ``backend="mpi"`` selects on a CPU-only box.
"""
from __future__ import annotations

import ctypes
import os
import time
import weakref
from typing import Optional

import torch
import torch.distributed as dist
from torch.autograd import Function

from dgraph_b200.parallel.collective_engine import TorchDistEngine, _splits

_ERRORS = {1: "timed out waiting for a peer", 2: "row outside the owner's window", 3: "shared-memory system call failed",
           4: "bad argument", 5: "another rank reported a failure"}
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        from dgraph_b200.build import build_host

        lib = ctypes.CDLL(str(build_host()))
        lib.dgbh_win_open.restype = ctypes.c_void_p
        lib.dgbh_win_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_int]
        lib.dgbh_win_connect.argtypes = [ctypes.c_void_p, ctypes.c_double]
        lib.dgbh_win_base.restype = ctypes.c_void_p
        lib.dgbh_win_base.argtypes = [ctypes.c_void_p]
        lib.dgbh_win_bytes.restype = ctypes.c_uint64
        lib.dgbh_win_bytes.argtypes = [ctypes.c_void_p]
        lib.dgbh_win_epoch.restype = ctypes.c_uint64
        lib.dgbh_win_epoch.argtypes = [ctypes.c_void_p]
        lib.dgbh_win_grow.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        lib.dgbh_win_fence.argtypes = [ctypes.c_void_p, ctypes.c_double]
        lib.dgbh_win_get_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                          ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
        lib.dgbh_win_put.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64]
        lib.dgbh_win_accumulate_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                                 ctypes.c_int, ctypes.c_int64, ctypes.c_void_p]
        lib.dgbh_win_set_lock_offset.argtypes = [ctypes.c_void_p, ctypes.c_int64]
        lib.dgbh_win_aborted.argtypes = [ctypes.c_void_p]
        lib.dgbh_win_close.argtypes = [ctypes.c_void_p]
        lib.dgbh_win_close.restype = None
        _LIB = lib
    return _LIB


def available() -> bool:
    """True when the host library (a C++ compiler at first use) is there."""
    try:
        _lib()
        return True
    except Exception:  # noqa: BLE001
        return False


class ShmWindow:
    """This rank's window plus the mappings of its peers' windows.  ``expose`` / ``fence`` / ``get_rows`` / ``accumulate_rows`` /
    ``put`` are the access-epoch vocabulary of an MPI window; sizes need not agree between ranks."""

    def __init__(self, job: str, rank: int, world: int, nbytes: int = 1 << 20, threads: int = 1, timeout_ms: float = 120_000.0):
        self.rank, self.world, self.timeout_ms = rank, world, float(timeout_ms)
        self._h = _lib().dgbh_win_open(job.encode(), rank, world, int(nbytes), int(threads))
        if not self._h:
            raise RuntimeError(f"cannot create the shared-memory window {job!r} (see stderr)")
        self._view = None
        self._locks_published = False
        self._finalizer = weakref.finalize(self, _lib().dgbh_win_close, self._h)

    def _check(self, rc: int, what: str) -> None:
        if rc:
            raise RuntimeError(f"shared-memory window, rank {self.rank}: {what}: {_ERRORS.get(rc, rc)}")

    def connect(self) -> None:
        self._check(_lib().dgbh_win_connect(self._h, self.timeout_ms), "connect")

    @property
    def nbytes(self) -> int:
        return int(_lib().dgbh_win_bytes(self._h))

    @property
    def epoch(self) -> int:
        return int(_lib().dgbh_win_epoch(self._h))

    def bytes_view(self) -> torch.Tensor:
        """uint8 tensor over this rank's window (re-made after the window grew)."""
        if self._view is None or self._view.numel() != self.nbytes:
            buf = (ctypes.c_char * self.nbytes).from_address(_lib().dgbh_win_base(self._h))
            self._view = torch.frombuffer(buf, dtype=torch.uint8)
        return self._view

    def reserve(self, nbytes: int) -> None:
        """At least ``nbytes`` of window (outside an access epoch only; contents are not kept when it grows)."""
        if nbytes > self.nbytes:
            self._check(_lib().dgbh_win_grow(self._h, int(nbytes * 1.25)), "grow")
            self._view = None

    def region(self, rows: int, F: int, dtype: torch.dtype) -> torch.Tensor:
        """``[rows, F]`` view of the start of this rank's window, valid until the window grows (withdraws a lock array published by :meth:`accumulator`: the bytes
        it occupied are data again)."""
        n = rows * F * torch.empty((), dtype=dtype).element_size()
        self.reserve(n)
        if self._locks_published:
            self._check(_lib().dgbh_win_set_lock_offset(self._h, -1), "set_lock_offset")
            self._locks_published = False
        return self.bytes_view()[:n].view(dtype).view(rows, F)

    def fence(self) -> None:
        """Closes the current access epoch and opens the next.  A fence that timed out raises on this rank and marks the window set
        aborted: every later fence on any rank raises at once ("another rank reported a failure") -- build a new engine to go on."""
        self._check(_lib().dgbh_win_fence(self._h, self.timeout_ms), "fence")

    def get_rows(self, rows: torch.Tensor, owners: torch.Tensor, out: torch.Tensor, out_rows: Optional[torch.Tensor] = None) -> None:
        rows, owners = rows.to(torch.int64).contiguous(), owners.to(torch.int32).contiguous()
        assert out.is_contiguous() and out.dim() == 2
        dr = None if out_rows is None else out_rows.to(torch.int64).contiguous()
        self._check(_lib().dgbh_win_get_rows(self._h, rows.data_ptr(), owners.data_ptr(), 0 if dr is None else dr.data_ptr(), rows.numel(),
                                             out.shape[1] * out.element_size(), 0, out.data_ptr()), "get")

    def accumulate_rows(self, rows: torch.Tensor, owners: torch.Tensor, src: torch.Tensor) -> None:
        """``window(owners[i])[rows[i]] += src[i]``; rows of owners that opened the epoch with :meth:`accumulator` are locked one at a
        time and added with plain vector adds, others are added element by element with compare-and-swap."""
        rows, owners = rows.to(torch.int64).contiguous(), owners.to(torch.int32).contiguous()
        assert src.is_contiguous() and src.dim() == 2 and src.dtype in (torch.float32, torch.float64) and src.shape[0] == rows.numel()
        self._check(_lib().dgbh_win_accumulate_rows(self._h, rows.data_ptr(), owners.data_ptr(), rows.numel(), src.shape[1],
                                                    src.element_size(), 0, src.data_ptr()), "accumulate")

    def accumulator(self, rows: int, F: int, dtype: torch.dtype) -> torch.Tensor:
        """Zeroed ``[rows, F]`` region at the head of this rank's window with one zeroed lock byte per row behind it, published to the
        peers (call before the epoch's opening fence)."""
        n = rows * F * torch.empty((), dtype=dtype).element_size()
        lock_off = (n + 63) // 64 * 64
        self.reserve(lock_off + rows)
        view = self.bytes_view()
        view[lock_off:lock_off + rows].zero_()
        region = view[:n].view(dtype).view(rows, F)
        region.zero_()
        self._check(_lib().dgbh_win_set_lock_offset(self._h, lock_off), "set_lock_offset")
        self._locks_published = True
        return region

    def put(self, peer: int, dst_offset_bytes: int, src: torch.Tensor) -> None:
        assert src.is_contiguous()
        self._check(_lib().dgbh_win_put(self._h, int(peer), int(dst_offset_bytes), src.data_ptr(), src.numel() * src.element_size()), "put")

    def aborted(self) -> bool:
        return bool(_lib().dgbh_win_aborted(self._h))

    def close(self) -> None:
        self._view = None
        self._finalizer()


def _owners_and_rows(engine, indices, rank_mappings, n_rows_local: int):
    """``(rows inside the owner, owner rank inside the replica, mask of the entries this rank handles or None)``; conventions of the
    reference engines: ``[1, E]`` / ``[E]`` = owner per entry (MPI, NVSHMEM), ``[2, E]`` = (rank handling the edge, owner) (NCCL),
    ``None`` = even contiguous split (``MPIBackendEngine.py:458-459,507-508``)."""
    idx = indices.reshape(-1).to(torch.long).cpu()
    W, r = engine.get_world_size(), engine.get_rank()
    mine = None
    if rank_mappings is None:
        owner = torch.div(idx, n_rows_local, rounding_mode="floor").clamp_(max=W - 1)
    else:
        rm = rank_mappings.cpu()
        if rm.dim() == 2 and rm.shape[0] == 2 and rm.shape[1] == idx.numel():
            owner, mine = rm[1].to(torch.long), rm[0] == r
            if bool(mine.all()):
                mine = None
        else:
            owner = rm.reshape(-1).to(torch.long)
    assert owner.numel() == idx.numel(), f"indices {tuple(indices.shape)} and rank mapping {tuple(rank_mappings.shape)} disagree"
    if mine is not None:
        idx, owner = idx[mine], owner[mine]
    return idx % n_rows_local, owner, mine


def _win_gather(engine, x2d: torch.Tensor, rows: torch.Tensor, owners: torch.Tensor) -> torch.Tensor:
    """``out[i] = x2d@owners[i][rows[i]]``: expose, fence, get, fence."""
    win = engine.window
    src = x2d.detach()
    host = src if not src.is_cuda else src.cpu()
    win.region(host.shape[0], host.shape[1], host.dtype).copy_(host)
    win.fence()
    out = torch.empty((rows.numel(), host.shape[1]), dtype=host.dtype)
    win.get_rows(rows, owners, out)
    win.fence()
    return out.to(x2d.device) if src.is_cuda else out


def _win_scatter_add(engine, v2d: torch.Tensor, rows: torch.Tensor, owners: torch.Tensor, n_rows_local: int) -> torch.Tensor:
    """``out@owners[i][rows[i]] += v2d[i]`` and return this rank's ``[n_rows_local, F]``: zero, fence, accumulate, fence, read."""
    win = engine.window
    src = v2d.detach()
    host = src if not src.is_cuda else src.cpu()
    acc_dtype = host.dtype if host.dtype in (torch.float32, torch.float64) else torch.float32
    host = host.to(acc_dtype).contiguous()
    region = win.accumulator(n_rows_local, host.shape[1], acc_dtype)  # zeroed rows + one lock byte per row, offset published
    win.fence()
    win.accumulate_rows(rows, owners, host)
    win.fence()
    out = region.clone().to(src.dtype)
    return out.to(v2d.device) if src.is_cuda else out


class _WindowGather(Function):
    @staticmethod
    def forward(ctx, x, rows, owners, engine):
        x2d = x[0] if x.dim() == 3 else x
        ctx.rows, ctx.owners, ctx.engine, ctx.n_rows, ctx.dim = rows, owners, engine, x2d.shape[0], x.dim()
        out = _win_gather(engine, x2d.contiguous(), rows, owners)
        return out.unsqueeze(0) if x.dim() == 3 else out

    @staticmethod
    def backward(ctx, g):
        g2d = g[0] if g.dim() == 3 else g
        out = _win_scatter_add(ctx.engine, g2d.contiguous(), ctx.rows, ctx.owners, ctx.n_rows)
        return (out.unsqueeze(0) if ctx.dim == 3 else out), None, None, None


class _WindowScatter(Function):
    @staticmethod
    def forward(ctx, v, rows, owners, n_rows_local, engine):
        v2d = v[0] if v.dim() == 3 else v
        ctx.rows, ctx.owners, ctx.engine, ctx.dim = rows, owners, engine, v.dim()
        out = _win_scatter_add(engine, v2d.contiguous(), rows, owners, int(n_rows_local))
        return out.unsqueeze(0) if v.dim() == 3 else out

    @staticmethod
    def backward(ctx, g):
        g2d = g[0] if g.dim() == 3 else g
        out = _win_gather(ctx.engine, g2d.contiguous(), ctx.rows, ctx.owners)
        return (out.unsqueeze(0) if ctx.dim == 3 else out), None, None, None, None


class ShmWindowEngine(TorchDistEngine):
    """``torch.distributed`` (gloo) for collectives + shared-memory windows for the one-sided operations."""

    one_sided = True
    _current = None  # weak reference to the engine created last (the static window-management calls of the MPI engine use it)

    def __init__(self, ranks_per_graph: int = -1, group=None, window_bytes: Optional[int] = None, **kwargs):
        super().__init__(dist_backend="gloo", ranks_per_graph=ranks_per_graph, group=group, **kwargs)
        nbytes = int(window_bytes if window_bytes is not None else float(os.environ.get("DGRAPH_B200_SHM_WINDOW_MB", "16")) * (1 << 20))
        threads = int(os.environ.get("DGRAPH_B200_SHM_THREADS", str(max(1, min(4, (os.cpu_count() or 1) // max(1, self._world))))))
        timeout = float(os.environ.get("DGRAPH_B200_SHM_TIMEOUT_MS", "120000"))
        name = [f"dgb{os.getpid():x}{time.monotonic_ns() & 0xFFFFFF:x}" if self._rank == 0 else None]
        if self._world > 1:
            src = dist.get_global_rank(group, 0) if group is not None else 0
            dist.broadcast_object_list(name, src=src, group=group)
        self.window = ShmWindow(name[0], self._rank, self._world, nbytes, threads, timeout)
        self.barrier()  # every control block exists before anyone maps it (connect would wait for it anyway)
        self.window.connect()
        ShmWindowEngine._current = weakref.ref(self)

    # ---- window management names of the reference's MPI engine (MPIBackendEngine.py:343-365) -------------------------------------
    @staticmethod
    def Malloc(size, device="cpu"):  # noqa: N802
        """``(window, tensor)``: a zeroed fp32 buffer; the window is the persistent one of the engine created last (data is staged
        into it per operation, so any tensor can take part in one-sided calls and nothing has to be attached first)."""
        eng = ShmWindowEngine._current() if ShmWindowEngine._current is not None else None
        return (None if eng is None else eng.window), torch.zeros(int(size), dtype=torch.float32, device=device)

    @staticmethod
    def Attach(tensor: torch.Tensor):  # noqa: N802
        eng = ShmWindowEngine._current() if ShmWindowEngine._current is not None else None
        return (None if eng is None else eng.window), tensor

    @staticmethod
    def Detach(win) -> None:  # noqa: N802
        return None

    def _plan_one_sided(self, t: torch.Tensor, plan) -> bool:
        """Plans that know where every edge's row lives (``edge_owner`` / ``edge_owner_row``, set by the plan builders) are served
        with window epochs; the choice depends on the plan type and the tensor's rank only, so every rank makes the same one."""
        return (self._world > 1 and getattr(plan, "edge_owner", None) is not None and not t.is_cuda
                and (t.dim() == 2 or (t.dim() == 3 and t.shape[0] == 1)))

    # ---- vector-addressed one-sided gather / scatter ----------------------------------------------------------------------------------
    def gather(self, src_tensor=None, indices=None, rank_mappings=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, _ = self._reference_keywords(src_tensor, rank_mappings, None, kwargs)
        plan = comm_plan if comm_plan is not None else (indices if hasattr(indices, "boundary_edge_idx") else None)
        if plan is not None:
            if self._plan_one_sided(src_tensor, plan):  # every edge reads its row from the owner's window: no send lists needed
                return _WindowGather.apply(src_tensor, plan.edge_owner_row, plan.edge_owner, self)
            return super().gather(src_tensor, indices, rank_mappings, comm_plan=comm_plan, **kwargs)
        assert indices.dtype in (torch.long, torch.int), f"Indices must be long or int, found {indices.dtype}"
        assert src_tensor.dim() == 2 or src_tensor.shape[0] == 1, "one graph per call: [1, N, F] or [N, F]"
        rows, owners, _ = _owners_and_rows(self, indices, rank_mappings, src_tensor.shape[-2])
        return _WindowGather.apply(src_tensor, rows, owners, self)

    def scatter(self, src_tensor=None, indices=None, rank_mappings=None, output_size=None, *, comm_plan=None, **kwargs):
        src_tensor, rank_mappings, output_size = self._reference_keywords(src_tensor, rank_mappings, output_size, kwargs)
        plan = comm_plan if comm_plan is not None else (indices if hasattr(indices, "boundary_edge_idx") else None)
        if plan is not None:
            if self._plan_one_sided(src_tensor, plan):
                return _WindowScatter.apply(src_tensor, plan.edge_owner_row, plan.edge_owner, int(plan.num_local_vertices), self)
            return super().scatter(src_tensor, indices, rank_mappings, output_size, comm_plan=comm_plan, **kwargs)
        # the reference's MPI engine takes (tensor, indices, num_output_rows, rank_mapping=None): a row count in third position
        if isinstance(rank_mappings, int) and (output_size is None or isinstance(output_size, torch.Tensor)):
            rank_mappings, output_size = output_size, rank_mappings
        assert output_size is not None, "scatter needs the number of local output rows"
        assert indices.dtype in (torch.long, torch.int), f"Indices must be long or int, found {indices.dtype}"
        assert src_tensor.dim() == 2 or src_tensor.shape[0] == 1, "one graph per call: [1, E, F] or [E, F]"
        rows, owners, mine = _owners_and_rows(self, indices, rank_mappings, int(output_size))
        if mine is not None:
            src_tensor = src_tensor[..., mine.to(src_tensor.device), :]
        return _WindowScatter.apply(src_tensor, rows, owners, int(output_size), self)

    # ---- halo primitive: one-sided writes at the receivers' offsets ---------------------------------------------------------------------
    def put(self, send_buffer, recv_buffer, send_offsets, recv_offsets, remote_offsets=None) -> None:
        if remote_offsets is None or self._world == 1 or send_buffer.is_cuda:
            return super().put(send_buffer, recv_buffer, send_offsets, recv_offsets, remote_offsets)
        F = 1
        for d in send_buffer.shape[1:]:
            F *= int(d)
        s2 = send_buffer.detach().reshape(send_buffer.shape[0], F).contiguous()
        r2 = recv_buffer.view(recv_buffer.shape[0], F)
        row_bytes = F * s2.element_size()
        win = self.window
        region = win.region(r2.shape[0], F, s2.dtype)  # the image of my receive buffer; peers write their chunks into it
        win.fence()
        offs = send_offsets.tolist() if isinstance(send_offsets, torch.Tensor) else list(send_offsets)
        rem = remote_offsets.tolist() if isinstance(remote_offsets, torch.Tensor) else list(remote_offsets)
        for p, n in enumerate(_splits(offs)):
            if n:
                win.put(p, int(rem[p]) * row_bytes, s2[int(offs[p]):int(offs[p]) + n])
        win.fence()
        r2.copy_(region)

    def finalize(self) -> None:
        self.window.close()
        super().finalize()
