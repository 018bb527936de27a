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


def test_entry_points_keep_the_reference_flags():
    """every CLI flag of the reference entry points is still accepted (extra flags are additions)."""
    import importlib.util
    import inspect
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    want = {
        "experiments/OGB/main.py": ["backend", "dataset", "epochs", "lr", "runs", "hidden_dims", "log_dir",
                                    "node_rank_placement_file", "root_dir"],
        "experiments/OGB-LSC/main.py": ["comm_type", "dataset", "num_papers", "num_authors", "num_institutions",
                                        "optimized_graph_structure", "paper_rank_mapping_file", "author_rank_mapping_file",
                                        "institution_rank_mapping_file", "data_dir"],
        "experiments/GraphCast/train_graphcast.py": ["batch_size", "is_distributed", "test_run", "use_synthetic_data",
                                                     "benchmark", "backend", "procs_per_graph"],
    }
    saved_path, saved_mods = list(sys.path), set(sys.modules)
    try:
        for i, (rel, flags) in enumerate(want.items()):
            path = os.path.join(root, rel)
            for m in ("utils", "config", "Trainer", "_common"):  # the experiment folders reuse these short module names
                sys.modules.pop(m, None)
            sys.path[:0] = [os.path.dirname(path), os.path.join(root, "experiments")]
            spec = importlib.util.spec_from_file_location(f"_entry_{i}", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            params = inspect.signature(mod.main).parameters
            assert not [f for f in flags if f not in params], (rel, [f for f in flags if f not in params])
    finally:
        sys.path[:] = saved_path
        for m in set(sys.modules) - saved_mods:
            if m in ("utils", "config", "Trainer", "_common", "GCN", "RGAT"):
                sys.modules.pop(m, None)


def test_torch_local_module_matches_reference_semantics():
    """Same cases as the reference's ``tests/test_local_kernels.py:19-154`` (CPU formulation here; the CUDA kernel behind the
    same functions is covered by ``test_peer_kernels_gpu.py`` / ``test_segment_ops_gpu.py``)."""
    import torch
    from DGraph.torch_local import (local_masked_gather, local_masked_scatter, local_masked_scatter_add_gather,
                                    local_masked_scatter_gather)
    import DGraph.torch_nvshmem_p2p as nv

    assert all(hasattr(nv.NVSHMEMP2P, n) for n in ("init", "finalize", "dist_put", "dist_get", "allocate_symmetric_memory",
                                                   "register_memory", "deregister_memory", "clone_tensor",
                                                   "padded_clone_tensor", "barrier", "barrier_stream"))
    torch.manual_seed(0)
    src = torch.randn(1, 4, 4)
    idx = torch.tensor([0, 0, 2, 1, 1, 2, 3, 3])
    rm = torch.tensor([0, 0, 1, 0, 0, 1, 1, 1])
    out = torch.zeros(1, 8, 4)
    local_masked_gather(src, idx, rm, out, 1, 4, 4, 8, 0)
    want = torch.zeros(1, 8, 4)
    for j in range(8):
        if rm[j] == 0:
            want[0, j] = src[0, idx[j]]
    assert torch.equal(out, want)
    vals = torch.randn(1, 8, 4)
    acc = torch.zeros(1, 4, 4)
    local_masked_scatter(vals, idx, rm, acc, 1, 8, 4, 4, 1)
    want = torch.zeros(1, 4, 4)
    for j in range(8):
        if rm[j] == 1:
            want[0, idx[j]] += vals[0, j]
    assert torch.allclose(acc, want)
    src8 = torch.randn(1, 8, 4)
    out = torch.zeros(1, 8, 4)
    local_masked_scatter_gather(src8, torch.tensor([0, 3, 2, 1]), torch.tensor([1, 3, 5, 7]), out, 1, 4, 4, 8)
    want = torch.zeros(1, 8, 4)
    for s, d in zip([0, 3, 2, 1], [1, 3, 5, 7]):
        want[0, d] = src8[0, s]
    assert torch.equal(out, want)
    out = torch.zeros(1, 8, 4)
    local_masked_scatter_add_gather(src8, torch.tensor([0, 3, 2, 1, 3]), torch.tensor([1, 3, 5, 7, 3]), out, 1, 5, 4, 8)
    want = torch.zeros(1, 8, 4)
    for s, d in zip([0, 3, 2, 1, 3], [1, 3, 5, 7, 3]):
        want[0, d] += src8[0, s]
    assert torch.allclose(out, want)
