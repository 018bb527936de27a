from dgraph_b200.parallel.comm_pattern import *  # noqa: F401,F403
from dgraph_b200.parallel.comm_pattern import (  # noqa: F401
    CommunicationPattern, compute_local_vertices, compute_halo_vertices, compute_local_edge_list, compute_boundary_vertices,
    compute_comm_map, compute_recv_offsets, build_communication_pattern,
)
