from dgraph_b200.parallel.legacy_cache import *  # noqa: F401,F403
