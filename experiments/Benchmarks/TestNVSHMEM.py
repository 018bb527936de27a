"""The one-sided arm of the communication benchmark (parity: reference ``experiments/Benchmarks/TestNVSHMEM.py:19-197``).  On B200
``backend="nvshmem"`` selects the NVLink peer-memory engine, so the sweeps and the all-pairs protocol are those of
``TestNCCL.py`` with another default backend; this module adds the reference's benchmark object for the one-sided calling
convention (every rank passes only the edges it holds, with the owner rank of each addressed row).

``torchrun --nproc-per-node N experiments/Benchmarks/TestNVSHMEM.py --message_size 128 --num_iters 1000 --log_dir logs``
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import TestNCCL as _nccl  # noqa: E402
from TestNCCL import _time, run, run_legacy  # noqa: F401,E402


class NVSHMEMBenchmark:
    """Per-iteration device times of ``comm.gather`` / ``comm.scatter`` in the one-sided convention: ``edge_indices [1, E_local]``
    are the global rows this rank's edges address and ``edge_*_rank [1, E_local]`` the ranks that own them."""

    def __init__(self, comm_object, *args, **kwargs) -> None:
        self.comm_object = comm_object
        self.rank, self.world_size = comm_object.get_rank(), comm_object.get_world_size()

    print = _nccl.NCCLBenchmark.print
    save_np = _nccl.NCCLBenchmark.save_np

    def benchmark_gather(self, data, edge_src_rank, edge_indices, num_iters: int = 1000):
        self.comm_object.barrier()
        return _time(lambda: self.comm_object.gather(data, edge_indices, edge_src_rank), num_iters, data.device)

    def benchmark_scatter(self, data, edge_dest_rank, edge_indices, num_local_vertices, num_iters: int = 1000):
        self.comm_object.barrier()
        return _time(lambda: self.comm_object.scatter(data, edge_indices, edge_dest_rank, num_local_vertices), num_iters,
                     data.device)


def _mine(t, placement, rank):
    """This rank's columns of an all-ranks ``[1, E]`` (or ``[1, E, F]``) tensor; a tensor that is already local passes through."""
    if t.shape[1] != placement.shape[1]:
        return t.contiguous()
    return t[:, placement[0] == rank].contiguous()


def _local_rows(graph_data, mapping, rank):
    v = graph_data.vertex_data
    F = v.shape[-1]
    if v.dim() == 3 and (mapping.dim() == 1 or v.shape[1] != mapping.shape[-1]):
        return v.contiguous()  # already this rank's [1, n, F]
    return v[mapping == rank].view(1, -1, F).contiguous()


def run_gather_benchmark(benchmark: NVSHMEMBenchmark, num_iters: int, gather_graph_data):
    """``gather_graph_data``: a ``GatherGraphData`` -- this rank's share (``get_nvshmem_gather_benchmark_data``) or the record of
    all ranks' edges (``get_nccl_gather_benchmark_data``), from which this rank's share is cut."""
    r, g = benchmark.rank, gather_graph_data
    pl = g.edge_rank_placement
    return benchmark.benchmark_gather(_local_rows(g, g.vertex_rank_mapping, r), _mine(g.edge_src_rank, pl, r),
                                      _mine(g.edge_indices, pl, r), num_iters=num_iters)


def run_scatter_benchmark(benchmark: NVSHMEMBenchmark, num_iters: int, scatter_graph_data):
    r, g = benchmark.rank, scatter_graph_data
    pl = g.edge_rank_placement
    return benchmark.benchmark_scatter(_mine(g.vertex_data, pl, r), _mine(g.edge_dest_rank, pl, r), _mine(g.edge_indices, pl, r),
                                       g.num_local_vertices, num_iters=num_iters)


def main():
    if not any(a == "--backend" or a.startswith("--backend=") for a in sys.argv[1:]):
        sys.argv += ["--backend", "nvshmem"]
    _nccl.main()


if __name__ == "__main__":
    main()
