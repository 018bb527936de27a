from dgraph_b200.parallel.engine import BackendEngine  # noqa: F401
