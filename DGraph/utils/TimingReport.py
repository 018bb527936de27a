from dgraph_b200.utils.timing import TimingReport  # noqa: F401
