from dgraph_b200.data.preprocess import *  # noqa: F401,F403
