from dgraph_b200.data.ogbn_datasets import DistributedOGBWrapper  # noqa: F401
