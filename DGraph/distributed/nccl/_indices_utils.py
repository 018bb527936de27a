"""Reference module path ``DGraph/distributed/nccl/_indices_utils.py`` -> ``dgraph_b200.parallel.indices_utils``."""
from dgraph_b200.parallel.indices_utils import (  # noqa: F401
    _allocate_local_recv_buffers, _generate_local_rank_mapping, _get_local_recv_buffer_w_placement, _get_local_recv_placement,
    _get_local_send_placement, _get_local_unique_recv_placement, _get_recv_comm_vector, _get_send_comm_vector,
    _get_send_recv_comm_vectors)
