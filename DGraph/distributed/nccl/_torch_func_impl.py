from dgraph_b200.parallel.functional import CommPlan_GatherFunction, CommPlan_ScatterFunction  # noqa: F401


import os as _os

import torch as _torch


def clear_cached_memory():
    """``DGRAPH_CLEAR_BUFFER_CACHE=1``: release the caching allocator's unused blocks (and synchronise) between exchanges --
    slower, but keeps peak memory down (reference ``_torch_func_impl.py:16-24``)."""
    if _os.environ.get("DGRAPH_CLEAR_BUFFER_CACHE", "0") == "1" and _torch.cuda.is_available():
        _torch.cuda.empty_cache()
        _torch.cuda.synchronize()


def _default_engine():
    from dgraph_b200.parallel.collective_engine import TorchDistEngine

    eng = getattr(_default_engine, "_eng", None)
    if eng is None:
        eng = _default_engine._eng = TorchDistEngine()
    return eng


class GatherFunction:
    """Legacy vector-addressed gather (``_torch_func_impl.py:355-579``; it raises ``NameError`` in the reference snapshot).
    ``GatherFunction.apply(x [1, N_loc, F], indices [1, E], edge_rank_loc [E], edge_dest_ranks [E], rank, world_size)`` returns
    the rows for the edges placed on ``rank``; differentiable."""

    @staticmethod
    def apply(local_send_tensor, indices, edge_rank_loc, edge_dest_ranks, rank, world_size):
        rm = _torch.stack([edge_rank_loc.reshape(-1), edge_dest_ranks.reshape(-1)])
        out = _default_engine().gather(local_send_tensor, indices, rm)
        clear_cached_memory()
        return out


class ScatterFunction:
    """Legacy vector-addressed scatter-sum: ``ScatterFunction.apply(x [1, E_loc, F], indices, edge_src_ranks, edge_dest_ranks,
    num_local_output_rows, rank, world_size)`` (``_torch_func_impl.py:582-782``)."""

    @staticmethod
    def apply(send_tensor, indices, edge_src_ranks, edge_dest_ranks, num_local_output_rows, rank, world_size):
        rm = _torch.stack([edge_src_ranks.reshape(-1), edge_dest_ranks.reshape(-1)])
        out = _default_engine().scatter(send_tensor, indices, rm, int(num_local_output_rows))
        clear_cached_memory()
        return out
