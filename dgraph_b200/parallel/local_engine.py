"""Single-process engine (world size 1): the role of the reference's ``SingleProcessDummyCommunicator``
(``experiments/GraphCast/dist_utils.py:8-46``, which cannot be instantiated in the snapshot)."""
from __future__ import annotations

import torch

from dgraph_b200.parallel.collective_engine import CollectiveEngine


class LocalEngine(CollectiveEngine):
    def init_process_group(self, *a, **k):
        return None

    def get_rank(self) -> int:
        return 0

    def get_world_size(self) -> int:
        return 1

    def all_to_all_rows(self, send, in_splits, out_splits, out=None):
        if out is not None and out.shape == send.shape and out.dtype == send.dtype:
            out.copy_(send)
            return out
        return send.clone()

    def all_reduce(self, tensor, op="sum"):
        return tensor

    def barrier(self) -> None:
        return None

    def finalize(self) -> None:
        return None
