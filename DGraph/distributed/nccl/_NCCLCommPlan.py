from dgraph_b200.parallel.comm_plan import *  # noqa: F401,F403
from dgraph_b200.parallel.comm_plan import compute_edge_slices, fast_2D_unique  # noqa: F401
