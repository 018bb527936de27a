"""Race / bounds / ordering detection (SURVEY §5 "Race detection / sanitizers": the reference has none beyond ``CUDACHECK``).

Three layers, cheapest first:

1. **Plan validators** (:func:`validate_kernel_plan`, :func:`validate_pattern`, :func:`validate_comm_plan`) -- every index the
   kernels will dereference is range-checked on the host once per plan, so the kernels themselves carry no bounds branches.
   ``DGRAPH_B200_VALIDATE=1`` makes the plan constructors call them.
2. **Sync-checked launches** (:func:`enable_sync_checks` / ``DGRAPH_B200_DEBUG_SYNC=1``) -- after every native launch the
   stream is synchronised and the sticky error read, so an asynchronous fault is attributed to the launch that caused it.
3. **compute-sanitizer** (``tools/sanitize.sh``: memcheck / racecheck / synccheck / initcheck over the kernel tests; one tool
   per ``gpurun`` call, 1 GPU).  Peer-memory kernels are exercised through the single-process rank emulation of
   ``tests/test_peer_kernels_gpu.py`` because kernels that wait on one another must not run as separate launches on one GPU.
"""
from __future__ import annotations

import os

import torch

from dgraph_b200.ops import _native as N

_sync_hook = None


def enable_sync_checks(on: bool = True):
    global _sync_hook
    if on and _sync_hook is None:
        def _hook(what: str):
            try:
                torch.cuda.synchronize()
            except RuntimeError as e:  # attribute the sticky error to this launch
                raise RuntimeError(f"asynchronous CUDA fault surfaced after native launch '{what}': {e}") from e

        _sync_hook = _hook
        N._post_launch_hooks.append(_hook)
    elif not on and _sync_hook is not None:
        N._post_launch_hooks.remove(_sync_hook)
        _sync_hook = None


if os.environ.get("DGRAPH_B200_DEBUG_SYNC", "0") == "1":
    enable_sync_checks(True)


def validate_enabled() -> bool:
    return os.environ.get("DGRAPH_B200_VALIDATE", "0") == "1"


def _monotone_csr(ptr: torch.Tensor, nnz: int, what: str):
    p = ptr.to("cpu", torch.int64)
    assert p.numel() >= 1 and int(p[0]) == 0, f"{what}: rowptr must start at 0"
    assert int(p[-1]) == nnz, f"{what}: rowptr ends at {int(p[-1])}, expected {nnz}"
    assert bool((p[1:] >= p[:-1]).all()), f"{what}: rowptr not monotone"


def _in_range(idx: torch.Tensor, hi: int, what: str):
    if idx.numel():
        lo_v, hi_v = int(idx.min()), int(idx.max())
        assert lo_v >= 0 and hi_v < hi, f"{what}: index range [{lo_v}, {hi_v}] outside [0, {hi})"


def validate_kernel_plan(plan) -> None:
    """CSR/CSC structure of a ``GraphKernelPlan``: monotone pointers, column/row ids in range, edge-id maps are permutations."""
    E = int(plan.col.numel())
    n_dst, n_src = int(plan.rowptr.numel()) - 1, int(plan.colptr.numel()) - 1
    _monotone_csr(plan.rowptr, E, "csr")
    _monotone_csr(plan.colptr, E, "csc")
    _in_range(plan.col, n_src, "csr.col")
    _in_range(plan.row, n_dst, "csc.row")
    for name in ("csr_eid", "csc_eid", "csc_to_csr"):
        t = getattr(plan, name, None)
        if t is not None and t.numel():
            assert t.numel() == E and int(torch.unique(t).numel()) == E, f"{name} is not a permutation of the edges"
            _in_range(t, E, name)


def validate_pattern(pat) -> None:
    """Halo plan invariants (the reference checks these only in its 2-GPU tests, ``tests/test_comm_info.py:441-689``)."""
    W = pat.world_size
    assert pat.send_offset.numel() == W + 1 and pat.recv_offset.numel() == W + 1
    so, ro = pat.send_offset.cpu(), pat.recv_offset.cpu()
    assert bool((so[1:] >= so[:-1]).all()) and bool((ro[1:] >= ro[:-1]).all()), "offsets not monotone"
    assert int(so[-1]) == pat.send_local_idx.numel(), "send_offset[-1] != len(send_local_idx)"
    assert int(ro[-1]) == pat.num_halo_vertices, "recv_offset[-1] != num_halo_vertices"
    assert int(so[pat.rank + 1] - so[pat.rank]) == 0 and int(ro[pat.rank + 1] - ro[pat.rank]) == 0, "self-send must be empty"
    _in_range(pat.send_local_idx, pat.num_local_vertices, "send_local_idx")
    e = pat.local_edge_list
    if e.numel():
        _in_range(e[:, 0], pat.num_local_vertices, "edge central column")
        _in_range(e[:, 1], pat.num_neighbor_local_vertices + pat.num_halo_vertices, "edge neighbour column")
    for p in range(W):  # a vertex is sent at most once per destination
        seg = pat.send_local_idx[int(so[p]):int(so[p + 1])]
        assert int(torch.unique(seg).numel()) == seg.numel(), f"duplicate sends to rank {p}"


def validate_comm_plan(plan) -> None:
    E = plan.num_local_edges
    _in_range(plan.local_edge_idx, E, "local_edge_idx")
    _in_range(plan.boundary_edge_idx, E, "boundary_edge_idx")
    _in_range(plan.local_vertex_idx, plan.num_local_vertices, "local_vertex_idx")
    _in_range(plan.boundary_vertex_idx, plan.num_local_vertices, "boundary_vertex_idx")
    assert plan.local_edge_idx.numel() + plan.boundary_edge_idx.numel() == E, "local + boundary edges must cover the edges"
    assert plan.local_edge_idx.numel() == plan.local_vertex_idx.numel()
    n_recv = int(sum(int(v) for v in plan.boundary_edge_splits))
    _in_range(plan.boundary_edge_buffer_map, max(n_recv, 1), "boundary_edge_buffer_map")
    assert int(sum(int(v) for v in plan.boundary_vertex_splits)) == plan.boundary_vertex_idx.numel()


def check_finite(named_tensors, what: str = "tensor") -> None:
    """Raise naming the first non-finite tensor (one fused reduction per tensor, one host read at the end)."""
    bad = []
    flags = []
    for name, t in named_tensors:
        if t is not None and t.is_floating_point():
            flags.append((name, (~torch.isfinite(t)).any()))
    if flags:
        stacked = torch.stack([f for _, f in flags]).cpu()
        bad = [n for (n, _), b in zip(flags, stacked.tolist()) if b]
    if bad:
        raise FloatingPointError(f"non-finite values in {what}: {bad[:8]}")
