from dgraph_b200.checks import check_dist_initialized, check_nccl_availability  # noqa: F401
