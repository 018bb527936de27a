"""Peer (NVLink / NVSwitch) bandwidth between two GPUs of one process: the denominator of the halo-exchange roofline.

For message sizes 1 MiB .. 1 GiB it times (CUDA events on the reading GPU's stream, 3 warm-up + 10 timed, L2 irrelevant: the source
is remote) (a) `tensor.copy_` from GPU 1 to GPU 0 (cudaMemcpyPeerAsync) and (b) `halo_pull_kernel` reading GPU 1's rows with 16-byte
loads (row widths 128 B .. 1536 B, identity row order) -- the instruction stream every pull kernel of this library uses.
One JSON line per point; `python benchmarks/p2p_sweep.py > profiles/p2p_sweep.jsonl` (needs 2 visible GPUs)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dgraph_b200.ops import _native as N  # noqa: E402


def timed(fn, warm=3, it=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def main():
    assert torch.cuda.device_count() >= 2, "needs two GPUs"
    torch.cuda.set_device(0)
    N.check(N.lib().dgb_enable_peer_access(1), "enable_peer_access")
    for mib in (1, 4, 16, 64, 256, 1024):
        nbytes = mib << 20
        src = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1").random_(0, 255)
        dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
        torch.cuda.synchronize()
        ms = timed(lambda: dst.copy_(src, non_blocking=True))
        print(json.dumps({"what": "memcpy_peer", "MiB": mib, "ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1)}), flush=True)
        for row_bytes in (128, 512, 1536):
            rows = nbytes // row_bytes
            owner = torch.zeros(rows, dtype=torch.int32, device="cuda:0")
            ridx = torch.arange(rows, dtype=torch.int32, device="cuda:0")
            table = torch.tensor([src.data_ptr()], dtype=torch.int64, device="cuda:0")

            def pull():
                N.check(N.lib().dgb_halo_pull(N.ptr(dst), N.i64(row_bytes), N.i64(rows), N.i32(row_bytes), N.ptr(table),
                                              N.i64(row_bytes), N.ptr(owner), N.ptr(ridx), N.i64(0), N.stream()), "dgb_halo_pull")

            ms = timed(pull)
            print(json.dumps({"what": "halo_pull_kernel", "MiB": mib, "row_bytes": row_bytes, "ms": round(ms, 4),
                              "GBps": round(rows * row_bytes / ms / 1e6, 1)}), flush=True)
        assert torch.equal(dst[: 1 << 20].cpu(), src[: 1 << 20].cpu())


if __name__ == "__main__":
    main()
