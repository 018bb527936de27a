"""Real multi-process CPU run over gloo, world_size 2 (BASELINE config 1 plumbing)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests._graphs import random_graph, partition


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, W, port, e, part, xg, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from dgraph_b200.parallel.comm_plan import COO_to_NCCLCommPlan

    comm = Communicator.init_process_group("mpi")  # CPU box: mpi/nccl/nvshmem all map to gloo
    try:
        Communicator.init_process_group("nccl")
        raise AssertionError("second init must raise")
    except RuntimeError:
        pass
    pat = build_communication_pattern(e, part, rank, W)
    x = xg[pat.local_vertices_global].clone().requires_grad_(True)
    halo = HaloExchange(comm)(x, pat)
    ok = torch.allclose(halo, xg[pat.halo_vertices_global])
    halo.sum().backward()
    cnt = torch.zeros(xg.shape[0], dtype=torch.float64)
    cross = part[e[:, 0]] != part[e[:, 1]]
    pairs = torch.unique(torch.stack([part[e[cross, 0]], e[cross, 1]], 1), dim=0)
    cnt.index_add_(0, pairs[:, 1], torch.ones(pairs.shape[0], dtype=torch.float64))
    ok = ok and torch.allclose(x.grad[:, 0], cnt[pat.local_vertices_global])
    # CommPlan through real collectives
    V = xg.shape[0]
    off = torch.tensor([V * r // W for r in range(W + 1)])
    mine = torch.nonzero((e[:, 0] >= off[rank]) & (e[:, 0] < off[rank + 1]), as_tuple=True)[0]
    plan = COO_to_NCCLCommPlan(rank, W, e[:, 1], mine, off)
    got = comm.gather(xg[off[rank]:off[rank + 1]].unsqueeze(0), comm_plan=plan)
    ok = ok and torch.allclose(got[0], xg[e[mine, 1]])
    t = torch.ones(2, dtype=torch.float64) * (rank + 1)
    comm.all_reduce(t)
    ok = ok and float(t[0]) == sum(range(1, W + 1))
    comm.barrier()
    ret[rank] = bool(ok)
    comm.destroy()
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_gloo_world2_halo_and_plan():
    W = 2
    e = random_graph(50, 5, seed=4, symmetric=False)
    part = partition(50, W, "random")
    xg = torch.randn(50, 4, dtype=torch.float64)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(W, _free_port(), e, part, xg, ret), nprocs=W, join=True)
    assert all(ret[r] for r in range(W))
