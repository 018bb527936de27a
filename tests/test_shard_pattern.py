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
    ret[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_shard_plan_equals_global_plan():
    W, V = 3, 90
    e = random_graph(V, 6, seed=8, symmetric=False)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(W, _free_port(), e, V, ret), nprocs=W, join=True)
    assert all(ret[r] for r in range(W))
