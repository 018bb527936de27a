"""Shard-based plan build (each rank only sees its own edges) == global-edge-list plan build, over real gloo collectives."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests._graphs import random_graph


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, W, port, e, V, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200.parallel import build_communication_pattern, build_communication_pattern_from_shard

    off = torch.tensor([V * r // W for r in range(W + 1)])
    part = torch.bucketize(torch.arange(V), off, right=True) - 1
    ref = build_communication_pattern(e, part, rank, W)
    mine = e[part[e[:, 0]] == rank]
    got = build_communication_pattern_from_shard(mine, off, rank, W)
    ok = (got.num_local_vertices == ref.num_local_vertices and got.num_halo_vertices == ref.num_halo_vertices
          and torch.equal(got.halo_vertices_global, ref.halo_vertices_global)
          and torch.equal(got.send_local_idx, ref.send_local_idx) and torch.equal(got.send_offset, ref.send_offset)
          and torch.equal(got.recv_offset, ref.recv_offset) and torch.equal(got.comm_map, ref.comm_map)
          and torch.equal(got.halo_owner, ref.halo_owner) and torch.equal(got.halo_remote_row, ref.halo_remote_row)
          and torch.equal(torch.unique(got.local_edge_list, dim=0), torch.unique(ref.local_edge_list, dim=0)))
    # training on the shard-built plan == training on the plan built from the global edge list (same losses, same weights)
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.parallel import HaloExchange

    Communicator._is_initialized = False
    comm = Communicator.init_process_group("gloo")
    g = torch.Generator().manual_seed(5)
    xg = torch.randn(V, 6, generator=g, dtype=torch.float64)
    yg = torch.randint(0, 4, (V,), generator=g)
    runs = []
    for pat in (ref, got):
        torch.manual_seed(0)
        model = CommAwareGCN(6, 8, 4, HaloExchange(comm), comm).double()
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        gids = pat.local_vertices_global
        losses = []
        for _ in range(3):
            opt.zero_grad()
            loss = torch.nn.functional.cross_entropy(model(xg[gids], pat), yg[gids], reduction="sum") / V
            loss.backward()
            for p_ in model.parameters():
                dist.all_reduce(p_.grad)
            opt.step()
            losses.append(float(loss))
        runs.append((losses, [p_.detach().clone() for p_ in model.parameters()]))
    ok = ok and all(abs(a - b) < 1e-12 for a, b in zip(runs[0][0], runs[1][0]))
    ok = ok and all(torch.allclose(a, b, atol=1e-12) for a, b in zip(runs[0][1], runs[1][1]))
    comm.destroy()
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_shard_plan_equals_global_plan():
    W, V = 3, 90
    e = random_graph(V, 6, seed=8, symmetric=False)  # directed: the plan derives the send lists from the receivers' needs
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(W, _free_port(), e, V, ret), nprocs=W, join=True)
    assert all(ret[r] for r in range(W))
