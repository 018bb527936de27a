"""``import DGraph`` keeps working for users of the reference (module-level re-exports)."""


def test_reference_import_paths_resolve():
    import DGraph
    from DGraph import Communicator, CommunicatorBase
    from DGraph.Communicator import SUPPORTED_BACKENDS
    from DGraph.distributed import HaloExchange, DGraphMessagePassing, CommunicationPattern, build_communication_pattern
    from DGraph.distributed.commInfo import compute_halo_vertices, compute_comm_map
    from DGraph.distributed.nccl import (NCCLBackendEngine, TIMINGS, NCCLGraphCommPlan, NCCLEdgeConditionedGraphCommPlan,
                                         COO_to_NCCLCommPlan, COO_to_NCCLEdgeConditionedCommPlan)
    from DGraph.distributed.RankLocalOps import OptimizedLocalScatterGather, RankLocalRenumberingWithMapping
    from DGraph.distributed.Engine import BackendEngine
    from DGraph.data.datasets import DistributedOGBWrapper
    from DGraph.data.graph import DistributedGraph, get_round_robin_node_rank_map
    from DGraph.utils import largest_split
    from DGraph.utils.TimingReport import TimingReport
    from DGraph.checks import check_dist_initialized

    assert SUPPORTED_BACKENDS == ["nccl", "mpi", "nvshmem"]
    assert issubclass(Communicator, CommunicatorBase) and largest_split(10, 4) == 3
    assert DGraph.__version__


def test_communicator_contract():
    import pytest
    from DGraph import Communicator

    Communicator._is_initialized = False
    c = Communicator.init_process_group("single")
    with pytest.raises(RuntimeError, match="already initialized"):
        Communicator.init_process_group("nccl")
    assert c.get_rank() == 0 and c.get_world_size() == 1
    assert hasattr(c, "_Communicator__backend_engine")  # reference models reach into the name-mangled attribute
    import torch

    t = torch.arange(12.0).view(3, 4)
    assert torch.equal(c.get_local_tensor(t, torch.tensor([0, 1, 0, 0]), dim=1), t[:, [0, 2, 3]])
    assert c.alloc_buffer((2, 3), torch.float32, "cpu").shape == (2, 3)
    c.destroy()
    assert Communicator._is_initialized is False
