from dgraph_b200.utils.data_splitting import largest_split  # noqa: F401
