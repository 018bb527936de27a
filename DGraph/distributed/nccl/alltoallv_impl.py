from dgraph_b200.parallel.legacy_cache import _nccl_alltoall_v, _nccl_alltoallv_with_dict, torch_alltoallv_with_comm_map  # noqa: F401
