"""CPU ranks of one node: the one-sided window engine (`parallel/shm_engine.py`) against the gloo all-to-all engine on the same
operations -- the halo `put` and the vector-addressed gather / scatter of the reference's MPI engine.  Wall-clock per call (median
of `--iters`, max over ranks), no GPU.

    python benchmarks/bench_shm_window.py [--world 2] [--rows 200000] [--feat 128] [--iters 20]
"""
import argparse
import json
import os
import socket
import sys
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _median_max(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        dist.barrier()
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3


def _worker(rank, W, port, a, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W))
    torch.set_num_threads(max(1, (os.cpu_count() or 2) // W))
    dist.init_process_group("gloo", rank=rank, world_size=W)
    from dgraph_b200.parallel.collective_engine import TorchDistEngine
    from dgraph_b200.parallel.shm_engine import ShmWindowEngine

    engines = {"gloo all-to-all": TorchDistEngine(dist_backend="gloo"), "shm windows": ShmWindowEngine()}
    n, F = a.rows, a.feat
    g = torch.Generator().manual_seed(rank)
    x = torch.randn(n, F, generator=g)
    # halo put: every rank sends n / W rows to every rank
    k = n // W
    send_off = torch.arange(W + 1) * k
    recv = torch.empty(k * W, F)
    remote = torch.full((W,), rank * k, dtype=torch.long)  # my chunk lands at row rank * k of every receiver
    idx = torch.randint(0, n * W, (n,), generator=g)
    owners = (idx // n).unsqueeze(0)
    out = {}
    for name, eng in engines.items():
        res = {"put_ms": _median_max(lambda: eng.put(x[: k * W], recv, send_off, send_off, remote_offsets=remote), a.iters)}
        res["put_GBps_per_rank"] = k * W * F * 4 / (res["put_ms"] * 1e-3) / 1e9
        res["gather_ms"] = _median_max(lambda: eng.gather(x.unsqueeze(0), idx.unsqueeze(0), owners), a.iters)
        res["scatter_ms"] = _median_max(lambda: eng.scatter(x.unsqueeze(0), idx.unsqueeze(0), owners, n), a.iters)
        out[name] = {k_: round(v, 3) for k_, v in res.items()}
    if rank == 0:
        ret["res"] = out
    dist.barrier()
    engines["shm windows"].window.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--rows", type=int, default=200_000)
    ap.add_argument("--feat", type=int, default=128)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(a.world, port, a, ret), nprocs=a.world, join=True)
    print(json.dumps({"world": a.world, "rows_per_rank": a.rows, "feat": a.feat, "dtype": "fp32", "iters": a.iters,
                      "cpus": os.cpu_count(), "timing": "wall clock, median, max over ranks", **ret["res"]}))


if __name__ == "__main__":
    main()
