"""Hybrid DP x graph-partition layout over gloo (4 processes, ranks_per_graph=2); C header / source sync; sampler."""
import glob
import os
import re
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests._graphs import random_graph, partition

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_declares_every_entry_point():
    src = set()
    for f in glob.glob(os.path.join(ROOT, "csrc", "*.cu")):
        src |= set(re.findall(r"^DGB_API\s+[\w\s\*]+?\b(dgb_\w+)\s*\(", open(f).read(), flags=re.M))
    hdr = set(re.findall(r"\b(dgb_\w+)\s*\(", open(os.path.join(ROOT, "include", "dgraph_b200.h")).read()))
    assert src and not (src - hdr), f"missing from include/dgraph_b200.h: {sorted(src - hdr)}"
    import shutil
    import subprocess

    if shutil.which("gcc"):  # the header is plain C (ctypes / dlopen users): it must compile on its own
        r = subprocess.run(["gcc", "-fsyntax-only", "-x", "c", os.path.join(ROOT, "include", "dgraph_b200.h")],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[:2000]


def test_sampler_gives_replicas_disjoint_samples():
    from dgraph_b200.utils.dist_utils import CommAwareDistributedSampler, SingleProcessDummyCommunicator

    ds = list(range(10))
    a = list(CommAwareDistributedSampler(ds, num_replicas=2, replica_id=0, shuffle=False))
    b = list(CommAwareDistributedSampler(ds, num_replicas=2, replica_id=1, shuffle=False))
    assert a == [0, 2, 4, 6, 8] and b == [1, 3, 5, 7, 9]
    c = SingleProcessDummyCommunicator()
    assert c.get_world_size() == 1 and c.get_num_replicas() == 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, W, port, e, xg, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    Communicator._is_initialized = False
    comm = Communicator.init_process_group("nccl", ranks_per_graph=2)
    ok = comm.get_world_size() == 2 and comm.get_global_world_size() == 4 and comm.get_num_replicas() == 2
    ok = ok and comm.get_replica_id() == rank // 2 and comm.get_rank() == rank % 2
    part = partition(xg.shape[0], 2, "random")
    pat = build_communication_pattern(e, part, comm.get_rank(), 2)
    scale = float(comm.get_replica_id() + 1)  # every replica works on its own sample
    x = xg[pat.local_vertices_global] * scale
    halo = HaloExchange(comm)(x, pat)
    ok = ok and torch.allclose(halo, xg[pat.halo_vertices_global] * scale)
    t = torch.ones(1, dtype=torch.float64)
    comm.all_reduce(t)  # graph scope: 2 ranks
    u = torch.ones(1, dtype=torch.float64)
    comm.all_reduce(u, scope="world")
    ok = ok and float(t) == 2.0 and float(u) == 4.0
    ret[rank] = bool(ok)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_hybrid_dp_graph_partition_gloo():
    W = 4
    e = random_graph(40, 5, seed=1, symmetric=False)
    xg = torch.randn(40, 3, dtype=torch.float64)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(W, _free_port(), e, xg, ret), nprocs=W, join=True)
    assert all(ret[r] for r in range(W))


def _train_worker(rank, W, port, e, xs, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    Communicator._is_initialized = False
    comm = Communicator.init_process_group("nccl", ranks_per_graph=2)
    V = xs[0].shape[0]
    part = partition(V, 2, "contiguous")
    pat = build_communication_pattern(e, part, comm.get_rank(), 2)
    torch.manual_seed(0)
    model = CommAwareGCN(xs[0].shape[1], 8, 3, HaloExchange(comm), comm)
    opt = FusedAdam([p for p in model.parameters()], lr=1e-2, comm=comm)
    x = xs[comm.get_replica_id()][pat.local_vertices_global]  # every replica trains on its own sample
    opt.zero_grad()
    (model(x, pat) ** 2).sum().backward()
    opt.sync_grads()  # SUM over all four ranks: two graph partitions x two replicas
    ret[rank] = opt.flat_g.clone()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_hybrid_gradients_equal_sum_over_samples_and_partitions():
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern
    from dgraph_b200.utils.dist_utils import SingleProcessDummyCommunicator

    V = 30
    e = random_graph(V, 4, seed=2, symmetric=True)
    xs = [torch.randn(V, 4, generator=torch.Generator().manual_seed(s)) for s in (10, 11)]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_train_worker, args=(4, _free_port(), e, xs, ret), nprocs=4, join=True)
    comm = SingleProcessDummyCommunicator()
    pat = build_communication_pattern(e, torch.zeros(V, dtype=torch.long), 0, 1)
    torch.manual_seed(0)
    model = CommAwareGCN(4, 8, 3, HaloExchange(comm), comm)
    opt = FusedAdam([p for p in model.parameters()], lr=1e-2)
    opt.zero_grad()
    for x in xs:  # gradients accumulate over the two samples
        (model(x, pat) ** 2).sum().backward()
    for r in range(4):
        assert torch.allclose(ret[r], opt.flat_g, rtol=1e-4, atol=1e-4 * float(opt.flat_g.abs().max())), r


def test_docs_pages_follow_the_sources():
    """``docs/kernels.md`` is generated from the C header and every ``:::`` target of ``docs/reference.md`` is importable."""
    import importlib
    import re
    import subprocess
    import sys

    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_kernel_docs.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for ln in open(os.path.join(ROOT, "docs", "reference.md")):
        m = re.match(r"::: (\S+)", ln)
        if not m:
            continue
        parts = m.group(1).split(".")
        for i in range(len(parts), 0, -1):
            try:
                obj = importlib.import_module(".".join(parts[:i]))
                break
            except ImportError:
                continue
        else:
            raise AssertionError(f"cannot import any prefix of {m.group(1)}")
        for a in parts[i:]:
            obj = getattr(obj, a)
