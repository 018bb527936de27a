"""One-sided shared-memory window engine (`parallel/shm_engine.py` over `csrc/host/shm_window.cpp`): real processes, gloo control
plane, CPU only.  Oracle = the same operation on the replicated global tensor (what the reference's MPI tests do with mpirun,
`tests/test_mpi_backend.py` there)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests._graphs import partition, random_graph

pytestmark = pytest.mark.skipif(not __import__("dgraph_b200.parallel.shm_engine", fromlist=["available"]).available(),
                                reason="host library cannot be built (no C++ compiler)")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _requests(rank, V, E, seed=0):
    g = torch.Generator().manual_seed(seed * 100 + rank)
    return torch.randint(0, V, (E,), generator=g), torch.randn(E, generator=g, dtype=torch.float64)


def _worker(rank, W, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W),
                      DGRAPH_B200_SHM_WINDOW_MB="0.0625", DGRAPH_B200_SHM_THREADS="2")
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from dgraph_b200.parallel.shm_engine import ShmWindowEngine

    comm = Communicator.init_process_group("mpi")
    eng = comm.engine
    checks = {"engine": isinstance(eng, ShmWindowEngine)}
    n, F, E = 40, 6, 150
    V = n * W
    xg = torch.randn(V, F, generator=torch.Generator().manual_seed(7), dtype=torch.float64)
    reqs = [_requests(r, V, E) for r in range(W)]
    idx, wts = reqs[rank]
    lo = rank * n

    # gather: [1, N, F] data, [1, E] global indices, [1, E] owner map; backward is the one-sided accumulate
    x = xg[lo:lo + n].clone().unsqueeze(0).requires_grad_(True)
    out = comm.gather(x, idx.unsqueeze(0), (idx // n).unsqueeze(0))
    checks["gather"] = out.shape == (1, E, F) and torch.equal(out[0], xg[idx])
    (out[0] * wts.unsqueeze(1)).sum().backward()
    want = torch.zeros(V, F, dtype=torch.float64)
    for i, w in reqs:
        want.index_add_(0, i, w.unsqueeze(1).expand(-1, F).contiguous())
    checks["gather_bwd"] = torch.allclose(x.grad[0], want[lo:lo + n], atol=1e-12)
    # no owner map = even contiguous split; [2, E] map = (rank that handles the entry, owner), entries of other ranks dropped
    checks["gather_default_map"] = torch.equal(eng.gather(send_tensor=x.detach(), indices=idx.unsqueeze(0))[0], xg[idx])
    handled = torch.arange(E) % W
    got = comm.gather(x.detach(), idx.unsqueeze(0), torch.stack([handled, idx // n]))
    checks["gather_2row_map"] = torch.equal(got[0], xg[idx[handled == rank]])

    # scatter: every rank adds its rows into the owners' windows (many duplicates, from several ranks at once)
    v = (torch.randn(E, F, generator=torch.Generator().manual_seed(50 + rank), dtype=torch.float64)).unsqueeze(0).requires_grad_(True)
    s = comm.scatter(v, idx.unsqueeze(0), (idx // n).unsqueeze(0), n)
    want = torch.zeros(V, F, dtype=torch.float64)
    for r in range(W):
        want.index_add_(0, reqs[r][0], torch.randn(E, F, generator=torch.Generator().manual_seed(50 + r), dtype=torch.float64))
    checks["scatter"] = s.shape == (1, n, F) and torch.allclose(s[0], want[lo:lo + n], atol=1e-12)
    g_out = torch.randn(V, F, generator=torch.Generator().manual_seed(9), dtype=torch.float64)
    (s[0] * g_out[lo:lo + n]).sum().backward()
    checks["scatter_bwd"] = torch.allclose(v.grad[0], g_out[idx], atol=1e-12)
    # the reference's MPI argument order: (tensor, indices, num_output_rows, rank_mapping)
    s2 = eng.scatter(v.detach(), idx.unsqueeze(0), n, (idx // n).unsqueeze(0))
    checks["scatter_mpi_order"] = torch.allclose(s2, s.detach(), atol=1e-12)

    # other dtypes: bytes for the gather, fp32 accumulation for a bf16 scatter; a tensor larger than the window grows it
    before = eng.window.nbytes
    big = torch.randn(V, 2048, generator=torch.Generator().manual_seed(3)).bfloat16()
    got = comm.gather(big[lo:lo + n].unsqueeze(0), idx.unsqueeze(0), (idx // n).unsqueeze(0))
    checks["gather_bf16_grow"] = torch.equal(got[0], big[idx]) and eng.window.nbytes > before
    ones = torch.ones(1, E, 8, dtype=torch.bfloat16)
    cnt = comm.scatter(ones, idx.unsqueeze(0), (idx // n).unsqueeze(0), n)
    total = torch.zeros(V)
    for i, _ in reqs:
        total += torch.bincount(i, minlength=V)
    checks["scatter_bf16"] = cnt.dtype == torch.bfloat16 and torch.equal(cnt[0, :, 0].float(), total[lo:lo + n])

    # halo exchange through the one-sided put (remote offsets of the plan), forward and backward
    e = random_graph(V, 6, seed=4, symmetric=False)
    part = partition(V, W, "random")
    pat = build_communication_pattern(e, part, rank, W)
    xl = xg[pat.local_vertices_global].clone().requires_grad_(True)
    epoch0 = eng.window.epoch
    halo = HaloExchange(comm)(xl, pat)
    checks["halo_put"] = torch.equal(halo, xg[pat.halo_vertices_global]) and eng.window.epoch > epoch0
    halo.sum().backward()
    cross = part[e[:, 0]] != part[e[:, 1]]
    pairs = torch.unique(torch.stack([part[e[cross, 0]], e[cross, 1]], 1), dim=0)
    need = torch.zeros(V, dtype=torch.float64).index_add_(0, pairs[:, 1], torch.ones(pairs.shape[0], dtype=torch.float64))
    checks["halo_put_bwd"] = torch.allclose(xl.grad[:, 0], need[pat.local_vertices_global])

    # comm_plan= gather / scatter as window epochs (every edge reads / adds at (owner, row)); forward and backward vs the global graph
    from dgraph_b200.parallel.comm_plan import COO_to_NCCLCommPlan

    off = torch.tensor([n * r for r in range(W + 1)])
    mine = torch.nonzero((e[:, 0] >= off[rank]) & (e[:, 0] < off[rank + 1]), as_tuple=True)[0]
    plan = COO_to_NCCLCommPlan(rank, W, e[:, 1], mine, off)
    xp = xg[lo:lo + n].clone().unsqueeze(0).requires_grad_(True)
    epoch0 = eng.window.epoch
    got = comm.gather(xp, comm_plan=plan)
    checks["plan_gather"] = torch.equal(got[0], xg[e[mine, 1]]) and eng.window.epoch == epoch0 + 2
    got.sum().backward()
    indeg = torch.bincount(e[:, 1], minlength=V).to(torch.float64)
    checks["plan_gather_bwd"] = torch.allclose(xp.grad[0, :, 0], indeg[lo:lo + n])
    ev = torch.randn(e.shape[0], F, generator=torch.Generator().manual_seed(21), dtype=torch.float64)
    sc = comm.scatter(ev[mine].unsqueeze(0), comm_plan=plan)
    checks["plan_scatter"] = torch.allclose(sc[0], torch.zeros(V, F, dtype=torch.float64).index_add_(0, e[:, 1], ev)[lo:lo + n], atol=1e-12)

    # collectives still work (gloo), and a fence that a rank never reaches is reported instead of hanging
    t = torch.ones(1, dtype=torch.float64) * (rank + 1)
    comm.all_reduce(t)
    checks["all_reduce"] = float(t) == W * (W + 1) / 2
    comm.barrier()
    if rank == 0:
        eng.window.timeout_ms = 300.0
        try:
            eng.window.fence()
            checks["fence_timeout"] = False
        except RuntimeError as err:
            checks["fence_timeout"] = "timed out" in str(err) and eng.window.aborted()
    ret[rank] = checks
    dist.barrier()
    comm.destroy()
    eng.window.close()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
@pytest.mark.parametrize("W", [2, 3])
def test_window_engine_matches_global_oracle(W):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(W, _free_port(), ret), nprocs=W, join=True)
    for r in range(W):
        bad = [k for k, ok in ret[r].items() if not ok]
        assert not bad, f"rank {r}: {bad}"
    assert "fence_timeout" in ret[0]


def _hybrid_worker(rank, W, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.parallel.shm_engine import ShmWindowEngine

    comm = Communicator.init_process_group("mpi", ranks_per_graph=2)  # 2 replicas x 2 ranks: one window set per replica
    eng = comm.engine
    ok = isinstance(eng, ShmWindowEngine) and comm.get_world_size() == 2 and comm.get_num_replicas() == 2
    r, n, F = comm.get_rank(), 16, 4
    xg = torch.randn(2 * n, F, generator=torch.Generator().manual_seed(comm.get_replica_id()), dtype=torch.float32)  # per-replica data
    idx = torch.randint(0, 2 * n, (64,), generator=torch.Generator().manual_seed(rank))
    got = comm.gather(xg[r * n:(r + 1) * n].unsqueeze(0), idx.unsqueeze(0), (idx // n).unsqueeze(0))
    ok = ok and torch.equal(got[0], xg[idx]) and eng.to_global_rank(r) == rank
    t = torch.ones(1)
    comm.all_reduce(t, scope="world")
    ret[rank] = bool(ok and float(t) == W)
    dist.barrier()
    eng.window.close()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_window_engine_per_replica_windows():
    W = 4
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_hybrid_worker, args=(W, _free_port(), ret), nprocs=W, join=True)
    assert all(ret[r] for r in range(W))


def test_window_single_rank_roundtrip():
    from dgraph_b200.parallel.shm_engine import ShmWindow

    win = ShmWindow(f"dgbtest{os.getpid():x}", 0, 1, nbytes=4096)
    win.connect()
    x = torch.arange(48, dtype=torch.float32).view(12, 4)
    win.region(12, 4, torch.float32).copy_(x)
    win.fence()
    rows = torch.tensor([3, 3, 11, 0])
    out = torch.empty(4, 4)
    win.get_rows(rows, torch.zeros(4, dtype=torch.int32), out)
    assert torch.equal(out, x[rows])
    win.accumulate_rows(rows, torch.zeros(4, dtype=torch.int32), torch.ones(4, 4))
    win.fence()
    want = x.clone().index_add_(0, rows, torch.ones(4, 4))
    assert torch.equal(win.region(12, 4, torch.float32), want)
    # the same accumulate through an accumulator (zeroed rows + published lock bytes: row-locked vector adds), then through a plain
    # region again (lock array withdrawn: the bytes behind the rows are data, element-wise compare-and-swap adds)
    acc = win.accumulator(12, 4, torch.float32)
    win.fence()
    win.accumulate_rows(rows, torch.zeros(4, dtype=torch.int32), torch.ones(4, 4))
    win.fence()
    assert torch.equal(acc, torch.zeros(12, 4).index_add_(0, rows, torch.ones(4, 4)))
    big = win.region(64, 4, torch.float32)  # covers the bytes the lock array used
    big.fill_(1.0)
    win.fence()
    win.accumulate_rows(torch.arange(64), torch.zeros(64, dtype=torch.int32), torch.ones(64, 4))
    win.fence()
    assert torch.equal(big, torch.full((64, 4), 2.0))
    with pytest.raises(RuntimeError, match="outside"):
        win.get_rows(torch.tensor([1 << 40]), torch.zeros(1, dtype=torch.int32), torch.empty(1, 4))
    win.close()


def _train(model, opt, x, y, pat, V, steps, sync=None):
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(x, pat), y, reduction="sum") / V
        loss.backward()
        if sync is not None:
            for p in model.parameters():
                sync(p.grad)
        opt.step()
        losses.append(float(loss.detach()))
    return losses


def _train_worker(rank, W, port, e, V, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    comm = Communicator.init_process_group("mpi")
    part = partition(V, W, "random")
    pat = build_communication_pattern(e, part, rank, W)
    g = torch.Generator().manual_seed(5)
    xg = torch.randn(V, 6, generator=g, dtype=torch.float64)
    yg = torch.randint(0, 4, (V,), generator=g)
    torch.manual_seed(0)
    model = CommAwareGCN(6, 8, 4, HaloExchange(comm), comm).double()
    gids = pat.local_vertices_global
    epoch0 = comm.engine.window.epoch
    losses = _train(model, torch.optim.SGD(model.parameters(), lr=0.05), xg[gids], yg[gids], pat, V, 3, sync=comm.all_reduce)
    total = torch.tensor(losses, dtype=torch.float64)
    dist.all_reduce(total)
    ret[rank] = (total.tolist(), [p.detach().clone() for p in model.parameters()], comm.engine.window.epoch - epoch0)
    dist.barrier()
    comm.engine.window.close()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_gcn_training_over_windows_equals_single_process():
    """Two ranks, every halo exchange (forward and backward) a one-sided `put` epoch: same losses and weights as one process."""
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from dgraph_b200.parallel.local_engine import LocalEngine

    W, V = 2, 80
    e = random_graph(V, 6, seed=11, symmetric=True)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_train_worker, args=(W, _free_port(), e, V, ret), nprocs=W, join=True)
    comm = Communicator._from_engine(LocalEngine(), backend="single")
    pat = build_communication_pattern(e, torch.zeros(V, dtype=torch.long), 0, 1)
    g = torch.Generator().manual_seed(5)
    xg = torch.randn(V, 6, generator=g, dtype=torch.float64)
    yg = torch.randint(0, 4, (V,), generator=g)
    torch.manual_seed(0)
    model = CommAwareGCN(6, 8, 4, HaloExchange(comm), comm).double()
    want = _train(model, torch.optim.SGD(model.parameters(), lr=0.05), xg, yg, pat, V, 3)
    for r in range(W):
        losses, params, fences = ret[r]
        assert all(abs(a - b) < 1e-10 for a, b in zip(losses, want)), (losses, want)
        assert all(torch.allclose(a, b, atol=1e-10) for a, b in zip(params, model.parameters()))
        assert fences == 3 * (2 + 1) * 2  # steps x (two forward exchanges + one backward: the features need no gradient) x two fences
