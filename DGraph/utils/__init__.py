from dgraph_b200.utils.data_splitting import largest_split, split_per_rank, try_barrier  # noqa: F401
