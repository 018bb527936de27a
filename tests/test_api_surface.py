"""Less travelled parts of the public API: ``Communicator.put`` contract, rank slices, global scalars, legacy caches and
all-to-all-v helpers, rank-local aggregation, hetero dataset accessors, TimingReport, the cached OGB wrapper."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dgraph_b200.parallel.emulator import run_world
from tests._graphs import random_graph


def test_put_contract_and_slices_and_global_val():
    """``put``: chunk i of the send buffer (rows ``send_offsets[i]:send_offsets[i+1]``) lands in rank i's receive buffer at
    ``recv_offsets[me]`` (``Engine.py:67-86``)."""
    from dgraph_b200.nn.dist_bn import GetGlobalVal

    W = 3
    counts = torch.tensor([[0, 2, 1], [3, 0, 2], [1, 1, 0]])  # counts[a, b] rows from a to b

    def job(comm):
        r = comm.get_rank()
        send_off = torch.cat([torch.zeros(1, dtype=torch.long), counts[r].cumsum(0)])
        recv_off = torch.cat([torch.zeros(1, dtype=torch.long), counts[:, r].cumsum(0)])
        send = torch.stack([torch.full((2,), 100.0 * r + 10 * d + k) for d in range(W) for k in range(int(counts[r, d]))]) \
            if int(send_off[-1]) else torch.zeros(0, 2)
        recv = comm.alloc_buffer((int(recv_off[-1]), 2), torch.float32, "cpu")
        comm.put(send, recv, send_off, recv_off)
        t = torch.arange(12.0).view(2, 6)
        return recv, comm.get_local_rank_slice(t, dim=1), GetGlobalVal(r + 1, comm, device="cpu")

    for r, (recv, sl, total) in enumerate(run_world(W, job)):
        want = [100.0 * s + 10 * r + k for s in range(W) for k in range(int(counts[s, r]))]
        assert recv[:, 0].tolist() == want
        assert torch.equal(sl, torch.arange(12.0).view(2, 6)[:, 2 * r:2 * r + 2]) and total == 6.0


def test_rank_local_aggregate_and_hetero_accessors():
    from dgraph_b200.data.hetero import RELATIONS, SyntheticHeterogeneousDataset
    from dgraph_b200.parallel.rank_local import LocalAggregateWithRemapping, RankLocalRenumberingWithMapping

    data = torch.arange(10.0).view(1, 5, 2)
    idx = torch.tensor([7, 3, 7, 9, 3])
    summed, uniq = LocalAggregateWithRemapping(data, idx, 2)
    assert uniq.tolist() == [3, 7, 9] and summed.shape == (1, 3, 2)
    summed_r, owners = LocalAggregateWithRemapping(data, idx, torch.tensor([1, 0, 1, 2, 0]), 2, torch.device("cpu"))  # reference form
    assert torch.equal(summed_r, summed) and owners.tolist() == [0, 1, 2]
    assert torch.equal(summed[0], torch.stack([data[0, 1] + data[0, 4], data[0, 0] + data[0, 2], data[0, 3]]))
    renum, uniq2, ranks = RankLocalRenumberingWithMapping(idx, torch.tensor([1, 0, 1, 2, 0]))
    assert renum.tolist() == [1, 0, 1, 2, 0] and uniq2.tolist() == [3, 7, 9] and ranks.tolist() == [0, 1, 2]
    ds = SyntheticHeterogeneousDataset(1, 2, num_papers=64, num_authors=32, num_institutions=4, num_features=6)
    assert ds.num_relations == len(RELATIONS) == 5 and ds.num_features == 6
    lo, hi = ds.local_range(0)
    assert ds.features[0].shape == (hi - lo, 6) and ds.labels.shape == (hi - lo,)
    m = ds.get_mask("train")
    assert bool(((m >= 0) & (m < hi - lo)).all()) and ds.get_target("train").shape == m.shape
    ds.add_batch_dimension()
    assert ds.features[0].shape == (1, hi - lo, 6)
    ds.prepare_plans(fused=True, reference_layout=False)
    assert len(ds.patterns) == 5 and ds.patterns[0].num_local_vertices == hi - lo


def test_timing_report_scopes_and_dump(tmp_path):
    from dgraph_b200.utils import TimingReport
    from dgraph_b200.utils.dist_utils import SingleProcessDummyCommunicator

    TimingReport.init(SingleProcessDummyCommunicator())
    with TimingReport("a"):
        torch.ones(10).sum()
    TimingReport.start("b")
    TimingReport.stop("b")
    TimingReport.add_time("c", 1.5)
    t = TimingReport.get_timers()
    assert set(t) == {"a", "b", "c"} and t["c"] == [1.5] and all(v[0] >= 0 for v in t.values())
    TimingReport.dump(str(tmp_path / "t.json"))
    assert os.path.getsize(tmp_path / "t.json") > 10
    TimingReport.enabled = False
    with TimingReport("off"):
        pass
    assert "off" not in TimingReport.get_timers()
    TimingReport.reset()


def test_distributed_ogb_wrapper_caches_and_slices(tmp_path):
    from dgraph_b200.data.ogbn_datasets import DistributedOGBWrapper

    def job(comm):
        ds = DistributedOGBWrapper("tiny", comm, dir_name=str(tmp_path))
        feats, edges, rank_maps, labels = ds[0]
        return feats.shape, edges.shape, rank_maps.shape, labels.shape, ds.graph_obj.get_nodes_per_rank()

    res = run_world(2, job)
    assert os.path.exists(tmp_path / "tiny_graph_data_2.pt")  # written once by rank 0, loaded by everyone
    total = 0
    for r, (fs, es, rs, ls, per_rank) in enumerate(res):
        assert fs[0] == ls[0] == int(per_rank[r]) and es[0] == 2 and rs == (2, es[1])
        total += fs[0]
    assert total == 10_000


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _a2av_worker(rank, W, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200.parallel.legacy_cache import (NCCLGatherCacheGenerator, NCCLScatterCacheGenerator, _nccl_alltoall_v,
                                                   _nccl_alltoallv_with_dict, torch_alltoallv_with_comm_map)

    other = 1 - rank
    # every rank owns 3 rows; all edges: holder, owner of the addressed vertex, global vertex id (row = id % 3)
    holder = torch.tensor([[0, 0, 0, 1, 1]])
    owner = torch.tensor([[1, 0, 1, 0, 1]])
    ids = torch.tensor([[4, 1, 5, 2, 3]])
    x = torch.arange(6, dtype=torch.float32).view(1, 3, 2) + 100 * rank
    mine = holder[0] == rank
    recv = torch.zeros(1, int(mine.sum()), 2)
    _nccl_alltoall_v(x, recv, ids, owner[:, mine], holder, owner, rank, W)
    got_v = recv.clone()
    d = _nccl_alltoallv_with_dict({other: torch.full((3,), float(rank))}, {other: torch.zeros(3)}, rank, W)
    # rank r sends (r + 1) rows to the other rank and keeps 1
    send_counts = [1, rank + 1] if rank == 0 else [rank + 1, 1]
    recv_counts = [1, 2] if rank == 0 else [1, 1]
    send = torch.arange(sum(send_counts) * 2, dtype=torch.float32).view(1, -1, 2) + 100 * rank
    rbuf = torch.zeros(1, sum(recv_counts), 2)
    views = torch_alltoallv_with_comm_map(send, rbuf, torch.tensor(send_counts), torch.tensor(recv_counts), rank, W)
    assert len(views) == W and views[0].data_ptr() == rbuf.data_ptr()
    cat = rbuf
    # legacy caches: 8 vertices, 4 per rank; every rank owns the edges whose source it owns
    e = torch.tensor([[0, 5], [1, 2], [2, 7], [3, 4], [4, 1], [5, 6], [6, 0], [7, 3]])
    place, dest = e[:, 0] // 4, e[:, 1] // 4
    gc = NCCLGatherCacheGenerator(e[:, 1], place, dest, 4, rank, W)
    sc = NCCLScatterCacheGenerator(e[:, 1], place, dest, 4, rank, W)
    ret[rank] = (got_v, d[other], cat, gc.gather_send_comm_vector, gc.gather_recv_comm_vector, sc.scatter_send_comm_vector,
                 sc.scatter_recv_comm_vector)
    dist.barrier()
    dist.destroy_process_group()


def test_legacy_alltoallv_helpers_and_caches_gloo():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_a2av_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    r0, r1 = ret[0], ret[1]
    # gather halves: rank 0 holds edges to vertices 4, 1, 5 -> its remote slots 0 and 2 get rank 1's rows 1 and 2; slot 1 is local
    # (left for the local gather); rank 1 holds edges to 2, 3 -> slot 0 gets rank 0's row 2
    assert torch.equal(r0[0][0], torch.tensor([[102., 103.], [0., 0.], [104., 105.]]))
    assert torch.equal(r1[0][0], torch.tensor([[4., 5.], [0., 0.]]))
    assert r0[1].tolist() == [1.0] * 3 and r1[1].tolist() == [0.0] * 3
    # list all-to-all: rank 0 keeps its row 0 and receives rank 1's first two rows; rank 1 receives rank 0's row 1 and keeps its last
    assert torch.equal(r0[2][0], torch.tensor([[0., 1.], [100., 101.], [102., 103.]]))
    assert torch.equal(r1[2][0], torch.tensor([[2., 3.], [104., 105.]]))
    # what one rank sends for a gather is what the other receives, and scatter is the mirror image
    assert r0[3][1] == r1[4][0] and r1[3][0] == r0[4][1]
    assert r0[5] == r0[4] and r0[6] == r0[3]


def test_benchmark_graph_utils_all_pairs_through_legacy_api():
    """The reference's benchmark data generators (``experiments/Benchmarks/graph_utils.py``) drive the legacy gather / scatter
    API: with one vertex per rank every rank gathers the vertices of all other ranks, and scattering the edge rows back sums, per
    vertex, the rows of all edges that point at it."""
    import importlib.util

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("_bench_graph_utils", os.path.join(root, "experiments", "Benchmarks", "graph_utils.py"))
    gu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gu)
    W, F = 4, 3
    gd = gu.get_nccl_gather_benchmark_data(F, W, "cpu")
    sd = gu.get_nccl_scatter_benchmark_data(F, W, "cpu")
    assert gd.edge_indices.shape == (1, W * (W - 1)) and sd.vertex_data.shape == (1, W * (W - 1), F)

    def job(comm):
        r = comm.get_rank()
        x = gd.vertex_data[r:r + 1].unsqueeze(0)  # [1, 1, F]: this rank's vertex
        rm = torch.cat([gd.edge_rank_placement, gd.edge_src_rank], 0)
        got = comm.gather(x, gd.edge_indices, rm)  # rows for the edges held by this rank
        mine = sd.edge_rank_placement[0] == r
        back = comm.scatter(sd.vertex_data[:, mine], sd.edge_indices, torch.cat([sd.edge_rank_placement, sd.edge_dest_rank], 0),
                            sd.num_local_vertices)
        return got[0], back[0]

    for r, (got, back) in enumerate(run_world(W, job)):
        others = [j for j in range(W) if j != r]
        assert torch.allclose(got, gd.vertex_data[others])
        to_me = sd.edge_indices[0] == r
        assert torch.allclose(back[0], sd.vertex_data[0, to_me].sum(0))


def test_indices_utils_counts_and_placements():
    from dgraph_b200.parallel import indices_utils as U
    from dgraph_b200.parallel.legacy_cache import all_to_all_cache_helper, all_to_all_cache_with_local_reduce_helper
    from dgraph_b200.parallel.rank_local import RankLocalGather, RankLocalReNumbering

    W = 3
    # edge e lives on place[e] and addresses vertex idx[e] owned by owner[e]; 4 vertices per rank
    place = torch.tensor([0, 0, 0, 1, 1, 2, 2, 2])
    idx = torch.tensor([[1, 5, 9, 4, 0, 8, 1, 5]])
    owner = idx[0] // 4
    send, recv = U._get_send_recv_comm_vectors(owner, place, 0, W)  # data flows owner -> place
    assert send.tolist() == [0, 1, 1] and recv.tolist() == [0, 1, 1]
    sp = U._get_local_send_placement(send, idx, owner, place, 0, 4)
    assert sp[1].tolist() == [0] and sp[2].tolist() == [1]
    lrm = owner[place == 0]
    rp = U._get_local_recv_placement(recv, lrm, 0)
    assert rp[1].reshape(-1).tolist() == [1] and rp[2].reshape(-1).tolist() == [2]
    bufs, rp2 = U._get_local_recv_buffer_w_placement(recv, lrm, 6, 0, "cpu")
    assert bufs[1].shape == (1, 1, 6) and set(rp2) == {1, 2}
    assert U._generate_local_rank_mapping(torch.zeros(7), 3).tolist() == [0, 0, 0, 1, 1, 1, 2]
    sv, rv, s_place, r_place = all_to_all_cache_helper(idx, place, owner, 4, 0, W)
    assert sv.tolist() == send.tolist() and rv.tolist() == recv.tolist() and s_place[1].tolist() == [0]
    # scatter direction: rank 2 holds edges 5..7 whose targets live on ranks 2, 0, 1
    ren, n_remote, remapped, receiving, recv_pl = all_to_all_cache_with_local_reduce_helper(idx, place, owner, 4, 2, W)
    assert n_remote == 2 and sorted(receiving.tolist()) == [0, 1] and ren.tolist() == [0, 1] and remapped.tolist() == [0, 1]
    assert recv_pl[0].tolist() == [1]  # rank 0's edge 2 (vertex 9) adds to local row 1 of rank 2
    r, u = RankLocalReNumbering(torch.tensor([7, 3, 7, 9]))
    assert r.tolist() == [1, 0, 1, 2] and u.tolist() == [3, 7, 9]
    src = torch.arange(12.0).view(6, 2)
    assert torch.equal(RankLocalGather(src, torch.tensor([5, 0, 2]), torch.tensor([1, 0, 1]), 1), src[[5, 2]])


def test_engine_calls_accept_the_reference_engines_argument_names_and_orders():
    """The three reference engines name and order the gather / scatter arguments differently (NCCL ``local_send_tensor``; MPI
    ``send_tensor, indices, num_output_rows, rank_mapping``; NVSHMEM ``input_tensor, indices, rank_mappings, num_output_rows``):
    every form reaches the same plan."""
    W, F = 3, 2
    idx_all = torch.tensor([[0, 1, 2, 3, 4, 5]])  # rank r addresses rows 2r, 2r+1 of a 2-rows-per-rank matrix ... of everyone

    def job(comm):
        r = comm.get_rank()
        x = (torch.arange(2 * F, dtype=torch.float64).view(1, 2, F) + 100 * r)
        own = torch.div(idx_all, 2, rounding_mode="floor")
        a = comm.gather(x, idx_all, own)
        b = comm.gather(local_send_tensor=x, indices=idx_all, rank_mappings=own)
        c = comm.gather(send_tensor=x, indices=idx_all, rank_mapping=own)
        e = torch.ones(1, idx_all.shape[1], F, dtype=torch.float64) * (r + 1)
        s0 = comm.scatter(e, idx_all, own, 2)
        s1 = comm.scatter(e, idx_all, 2, own)  # MPI order
        s2 = comm.scatter(input_tensor=e, indices=idx_all, rank_mappings=own, num_output_rows=2)
        s3 = comm.scatter(e, idx_all, 2)  # MPI order, owners from the even split
        return a, b, c, s0, s1, s2, s3

    for r, (a, b, c, s0, s1, s2, s3) in enumerate(run_world(W, job)):
        assert torch.equal(a, b) and torch.equal(a, c) and a.shape == (1, 6, F)
        assert torch.equal(a[0, 2 * 1], torch.arange(F, dtype=torch.float64) + 100)  # row 0 of rank 1
        for s in (s1, s2, s3):
            assert torch.equal(s0, s)
        assert torch.equal(s0, torch.full((1, 2, F), float(sum(range(1, W + 1))), dtype=torch.float64))
