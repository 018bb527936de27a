from dgraph_b200.data.ogbn_datasets import *  # noqa: F401,F403
