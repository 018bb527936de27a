"""Peer-memory kernels on ONE GPU: W ranks emulated as W buffers of one process (pointers are just device pointers), each
rank's kernels launched back to back on one stream -- no kernel ever waits for another (B200_PROFILING.md rule)."""
import ctypes

import pytest
import torch

from tests._graphs import random_graph, partition

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]


class _FakeEngine:
    def __init__(self, rank, W, counts):
        self.device = torch.device("cuda")
        self._rank, self._W, self._counts = rank, W, counts

    def get_rank(self):
        return self._rank

    def get_world_size(self):
        return self._W

    def all_gather_ints(self, vals):
        return self._counts


@pytest.mark.parametrize("W", [2, 4])
@pytest.mark.parametrize("F", [8, 104, 256])
def test_halo_pull_and_grad_reduce_emulated(W, F):
    from dgraph_b200.ops import _native as N
    from dgraph_b200.parallel import build_all_patterns
    from dgraph_b200.parallel.peer_engine import _PatternPeerInfo

    torch.manual_seed(0)
    V = 300
    e = random_graph(V, 6, seed=W, symmetric=False)
    part = partition(V, W, "random")
    pats = build_all_patterns(e, part, W)
    counts = [[p.num_local_vertices, p.num_halo_vertices, int(p.send_offset[-1]), 0] for p in pats]
    infos = [_PatternPeerInfo(_FakeEngine(r, W, counts), pats[r]) for r in range(W)]
    max_aug = max(c[0] + c[1] for c in counts)
    xg = torch.randn(V, F, device="cuda").bfloat16()
    # "symmetric slots": one buffer per rank, same shape
    slots = [torch.zeros(max_aug, F, device="cuda", dtype=torch.bfloat16) for _ in range(W)]
    table = torch.tensor([s.data_ptr() for s in slots], dtype=torch.int64, device="cuda")
    for r in range(W):  # phase 1: every rank publishes its rows
        slots[r][: counts[r][0]] = xg[pats[r].local_vertices_global.cuda()]
    rb = F * 2
    for r in range(W):  # phase 2: every rank pulls its halo
        i = infos[r]
        N.check(N.lib().dgb_halo_pull(N.ptr(slots[r][i.n_nbr_local:]), N.i64(rb), N.i64(i.n_halo), N.i32(rb), N.ptr(table),
                                      N.i64(rb), N.ptr(i.halo_owner), N.ptr(i.halo_row), N.i64(i.rot), N.stream()), "pull")
    for r in range(W):
        want = xg[torch.cat([pats[r].local_vertices_global, pats[r].halo_vertices_global]).cuda()]
        assert torch.equal(slots[r][: want.shape[0]], want)

    # backward: g_aug per rank -> owners sum the halo parts (oracle: index_add over global ids in fp32)
    g_aug = [torch.randn(counts[r][0] + counts[r][1], F, device="cuda").bfloat16() for r in range(W)]
    max_halo = max(c[1] for c in counts)
    gslots = [torch.zeros(max(max_halo, 1), F, device="cuda", dtype=torch.bfloat16) for _ in range(W)]
    gtable = torch.tensor([s.data_ptr() for s in gslots], dtype=torch.int64, device="cuda")
    for r in range(W):
        gslots[r][: counts[r][1]] = g_aug[r][counts[r][0]:]
    want = torch.zeros(V, F, device="cuda")
    for r in range(W):
        ids = torch.cat([pats[r].local_vertices_global, pats[r].halo_vertices_global]).cuda()
        want.index_add_(0, ids, g_aug[r].float())
    for r in range(W):
        i = infos[r]
        gx = g_aug[r][: i.n_nbr_local].clone()
        N.check(N.lib().dgb_halo_grad_reduce(N.ptr(gx), N.i64(F), N.ptr(g_aug[r]), N.i64(F), N.i32(F), N.i64(i.n_bnd),
                                             N.ptr(i.bnd_row), N.ptr(i.bnd_ptr), N.ptr(i.src_rank), N.ptr(i.src_row),
                                             N.ptr(gtable), N.i64(F), N.stream()), "grad_reduce")
        assert torch.allclose(gx.float(), want[pats[r].local_vertices_global.cuda()], rtol=2e-2, atol=3e-2)


def test_adam_allreduce_and_allreduce_sum_emulated():
    from dgraph_b200.ops import _native as N

    torch.manual_seed(1)
    W, n = 4, 1000
    g = [torch.randn(n, device="cuda") for _ in range(W)]
    table = torch.tensor([x.data_ptr() for x in g], dtype=torch.int64, device="cuda")
    out = torch.empty(n, device="cuda")
    N.check(N.lib().dgb_allreduce_sum_f32(N.ptr(out), N.ptr(table), N.i32(W), N.i64(n), N.stream()), "allreduce")
    assert torch.allclose(out, sum(g), atol=1e-5)
    p0 = torch.randn(n, device="cuda")
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=1e-2)
    ps = [p0.clone() for _ in range(W)]
    ms = [torch.zeros(n, device="cuda") for _ in range(W)]
    vs = [torch.zeros(n, device="cuda") for _ in range(W)]
    for step in (1, 2):
        ref_p.grad = sum(g) / W
        opt.step()
        for r in range(W):
            N.check(N.lib().dgb_adam_allreduce(N.ptr(ps[r]), N.ptr(table), N.i32(W), N.ptr(ms[r]), N.ptr(vs[r]), N.ptr(None),
                                               N.i64(n), N.f32(1e-2), N.f32(0.9), N.f32(0.999), N.f32(1e-8), N.f32(0.0),
                                               N.f32(1.0 / W), N.i32(step), N.stream()), "adam_allreduce")
    for r in range(W):
        assert torch.allclose(ps[r], ref_p.detach(), atol=1e-5)
        assert torch.equal(ps[r], ps[0])  # replicas stay bit-identical


@pytest.mark.parametrize("W", [2, 3])
def test_fused_pull_kernels_emulated(W):
    """Segment-mean with in-kernel halo pulls and the gather-GEMM with peer rows == the same ops on a materialised [local | halo]
    matrix (ranks emulated as buffers of one process)."""
    from dgraph_b200.ops import _native as N
    from dgraph_b200.ops.gemm_native import gather_gemm_bf16, pack_segment_weight
    from dgraph_b200.ops.segment import _native_seg_sum
    from dgraph_b200.parallel import build_all_patterns
    from dgraph_b200.parallel.peer_engine import _PatternPeerInfo

    torch.manual_seed(0)
    V, F, H = 400, 64, 128
    e = random_graph(V, 7, seed=W + 5, symmetric=False)
    part = partition(V, W, "random")
    pats = build_all_patterns(e, part, W)
    counts = [[p.num_local_vertices, p.num_halo_vertices, int(p.send_offset[-1]), 0] for p in pats]
    infos = [_PatternPeerInfo(_FakeEngine(r, W, counts), pats[r]) for r in range(W)]
    xg = torch.randn(V, F, device="cuda").bfloat16()
    nmax = max(c[0] for c in counts)
    pubs = [torch.zeros(nmax, F, device="cuda", dtype=torch.bfloat16) for _ in range(W)]
    for r in range(W):
        pubs[r][: counts[r][0]] = xg[pats[r].local_vertices_global.cuda()]
    table = torch.tensor([p.data_ptr() for p in pubs], dtype=torch.int64, device="cuda")
    w = (torch.randn(H, 2 * F, device="cuda") * 0.1).bfloat16()
    wp = pack_segment_weight(w, [F, F])
    for r in range(W):
        i, pat = infos[r], pats[r]
        kp = pat.kernel_plan("cuda")
        aug = xg[torch.cat([pat.local_vertices_global, pat.halo_vertices_global]).cuda()].contiguous()
        want = _native_seg_sum(kp.rowptr, kp.col, kp.num_local, aug, None, True)
        got = _native_seg_sum(kp.rowptr, kp.col_peer(i.peer_shift), kp.num_local, pubs[r], None, True,
                              peer=(table, W, i.peer_shift, F))
        assert torch.equal(got, want)
        # gather-GEMM over cat([x[src], x_aug[dst]]): dst rows beyond n_local are pulled from the owner
        src = pat.local_edge_list[:, 0].cuda().int()
        dst = pat.local_edge_list[:, 1].cuda().int()
        E = int(src.numel())
        ref = gather_gemm_bf16([aug, aug], [src, dst], wp, None, E, [F, F], act="silu")
        out = gather_gemm_bf16([pubs[r], pubs[r]], [src, dst], wp, None, E, [F, F], act="silu",
                               peers=[None, (table, i.halo_owner, i.halo_row, i.n_nbr_local)])
        assert torch.equal(out, ref)
