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
    return t[:, placement[0] == rank].contiguous()


def run_gather_benchmark(benchmark: NVSHMEMBenchmark, num_iters: int, graph_data):
    """``graph_data``: a ``GatherGraphData`` holding the edges of ALL ranks (``graph_utils.py``); this rank's share is cut out."""
    r = benchmark.rank
    F = graph_data.vertex_data.shape[-1]
    local = graph_data.vertex_data[graph_data.vertex_rank_mapping == r, :].view(1, -1, F).contiguous()
    pl = graph_data.edge_rank_placement
    return benchmark.benchmark_gather(local, _mine(graph_data.edge_src_rank, pl, r), _mine(graph_data.edge_indices, pl, r),
                                      num_iters=num_iters)


def run_scatter_benchmark(benchmark: NVSHMEMBenchmark, num_iters: int, graph_data):
    r = benchmark.rank
    F = graph_data.vertex_data.shape[-1]
    pl = graph_data.edge_rank_placement
    local = graph_data.vertex_data[graph_data.data_rank_mapping == r].view(1, -1, F).contiguous()
    return benchmark.benchmark_scatter(local, _mine(graph_data.edge_dest_rank, pl, r), _mine(graph_data.edge_indices, pl, r),
                                       graph_data.num_local_vertices, num_iters=num_iters)


def main():
    if not any(a == "--backend" or a.startswith("--backend=") for a in sys.argv[1:]):
        sys.argv += ["--backend", "nvshmem"]
    _nccl.main()


if __name__ == "__main__":
    main()
