"""Hybrid data-parallel x graph-partition layout on the peer-memory engine (needs >= 4 GPUs; skipped otherwise): every replica
splits its own sample's graph over ``ranks_per_graph`` GPUs (halo traffic over NVLink inside the replica), gradients are summed
over the WHOLE job by the fused all-reduce + Adam kernel.  Oracle: the same run on the ``torch.distributed`` (NCCL all-to-all +
NCCL all-reduce) engine.  Reference arithmetic: ``DGraph/distributed/nccl/NCCLBackendEngine.py:56-64``; loss scaling
``experiments/GraphCast/train_graphcast.py:29-34``."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from tests._graphs import random_graph, partition

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu, pytest.mark.timeout(300)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, W, rpg, port, engine, e, xg, yg, state, ret):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W),
                      DGRAPH_B200_ARENA_MB="256", DGRAPH_B200_ENGINE=engine)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=W, device_id=dev)
    from dgraph_b200 import Communicator
    from dgraph_b200.models.gcn import CommAwareGCN
    from dgraph_b200.optim import FusedAdam
    from dgraph_b200.parallel import HaloExchange, build_communication_pattern

    comm = Communicator.init_process_group("nccl", ranks_per_graph=rpg)
    assert comm.get_world_size() == rpg and comm.get_global_world_size() == W and comm.get_num_replicas() == W // rpg
    rep, r = comm.get_replica_id(), comm.get_rank()
    part = partition(xg.shape[0], rpg, "contiguous")
    pat = build_communication_pattern(e.to(dev), part.to(dev), r, rpg)
    gids = pat.local_vertices_global
    # every replica trains on its own sample: same graph, features shifted by the replica id
    xl = (xg.to(dev)[gids] + 0.25 * rep).bfloat16()
    yl = ((yg.to(dev)[gids] + rep) % 5)
    torch.manual_seed(0)
    model = CommAwareGCN(xg.shape[1], 64, 5, HaloExchange(comm), comm, compute_dtype=torch.bfloat16).to(dev)
    model.load_state_dict(state)
    opt = FusedAdam(model.parameters(), lr=1e-3, comm=comm)
    losses = []
    for _ in range(3):
        opt.zero_grad()
        out = model(xl, pat)
        # sum over the replica's ranks = mean over the sample; FusedAdam averages over all W ranks -> scale by ranks_per_graph
        loss = torch.nn.functional.cross_entropy(out.float(), yl, reduction="sum") / xg.shape[0] * rpg
        loss.backward()
        opt.step()
        l = loss.detach().clone() / rpg
        dist.all_reduce(l)
        losses.append(float(l) / (W // rpg))
    if hasattr(comm.engine, "check_timeouts"):
        comm.engine.check_timeouts()
    ret[(engine, rank)] = (losses, {k: v.detach().cpu() for k, v in model.state_dict().items()}, type(comm.engine).__name__)
    comm.destroy()
    if hasattr(comm.engine, "arena"):
        comm.engine.finalize()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("W,rpg", [(4, 2), (8, 4), (8, 2), (4, 1)])
def test_hybrid_replicas_on_peer_engine_match_nccl_engine(W, rpg):
    if torch.cuda.device_count() < W:
        pytest.skip(f"needs {W} GPUs")
    from dgraph_b200.models.gcn import CommAwareGCN

    torch.manual_seed(0)
    V, F = 1500, 32
    e = random_graph(V, 8, seed=5)
    xg = torch.randn(V, F)
    yg = torch.randint(0, 5, (V,))
    state = {k: v.clone() for k, v in CommAwareGCN(F, 64, 5, None, None).state_dict().items()}
    mgr = mp.Manager()
    ret = mgr.dict()
    for engine in ("torch", "peer"):
        mp.spawn(_worker, args=(W, rpg, _free_port(), engine, e, xg, yg, state, ret), nprocs=W, join=True)
    assert ret[("peer", 0)][2] == "PeerMemoryEngine" and ret[("torch", 0)][2] == "TorchDistEngine"
    for r in range(W):
        lp, sp, _ = ret[("peer", r)]
        lt, st, _ = ret[("torch", r)]
        assert torch.allclose(torch.tensor(lp), torch.tensor(lt), rtol=2e-2, atol=2e-2), (lp, lt)
        for k in sp:
            d = (sp[k] - st[k]).abs()
            assert d.mean() < 4e-4 and d.max() <= 6.5e-3, (k, float(d.mean()), float(d.max()))
            assert torch.equal(sp[k], ret[("peer", 0)][1][k])  # replicated weights stay bit-identical across the whole job
