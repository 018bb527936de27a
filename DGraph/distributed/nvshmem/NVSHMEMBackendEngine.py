"""Reference module path ``DGraph/distributed/nvshmem/NVSHMEMBackendEngine.py``: the peer-memory engine (CUDA-IPC symmetric arena,
kernels that read / reduce rows of other GPUs over NVLink) in the role of the NVSHMEM engine."""
from DGraph.distributed.nvshmem import NVSHMEMBackendEngine  # noqa: F401
from dgraph_b200.parallel.functional import CommPlan_GatherFunction as NVSHMEMGatherFunction  # noqa: F401
from dgraph_b200.parallel.functional import CommPlan_ScatterFunction as NVSHMEMScatterFunction  # noqa: F401


def _nvshmmem_gather(send_tensor, indices, rank_mappings):
    """One-sided gather of the reference (``NVSHMEMBackendEngine.py:22-63``) through the current communicator's engine."""
    from DGraph.distributed.nccl._torch_func_impl import _default_engine

    return _default_engine().gather(send_tensor, indices, rank_mappings)


def _nvshmem_scatter(input_tensor, indices, rank_mappings, num_output_rows):
    from DGraph.distributed.nccl._torch_func_impl import _default_engine

    return _default_engine().scatter(input_tensor, indices, rank_mappings, int(num_output_rows))
