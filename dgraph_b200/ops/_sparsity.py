"""Row-support hints between adjacent backward ops.

A loss over a subset ``S`` of the output rows (the labelled vertices: 8-10 % of an OGB graph) makes ``dLogits`` -- and the
gradient of every row-wise op between the loss and the last message-passing layer -- structurally zero outside ``S``.  Autograd
hands dense tensors from one backward to the next, so the producer leaves a one-slot note here ("the tensor at this address is
zero except rows ``idx``; here are those rows, compact") and the consumer picks it up *iff it receives exactly that tensor*.
The consumer then skips the zero rows: gathered weight-gradient GEMM, compact dX GEMM, CSC walk restricted to edges that start
in ``S``.  Results are bit-identical to the dense computation (only additions of exact zeros are dropped).  Anything in between
(another op, gradient accumulation) produces a different tensor -> no match -> dense path.

``DGRAPH_B200_ROW_SPARSITY=0`` disables the mechanism.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

ENABLED = os.environ.get("DGRAPH_B200_ROW_SPARSITY", "1") == "1"
STATS = {"put": 0, "hit": 0, "miss": 0}
# (tensor, version at put time, idx, compact).  The tensor itself is kept alive: while the note exists its memory cannot be handed
# to another tensor, so "same address + same shape / dtype + unchanged version" really means "the very data the note describes".
_slot = None


def put(t: torch.Tensor, idx: torch.Tensor, compact: torch.Tensor) -> None:
    global _slot
    if not ENABLED:
        return
    _slot = (t, t._version, idx, compact)
    STATS["put"] += 1


def take(t: torch.Tensor) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """Hint for exactly this tensor (same storage address, shape, dtype, not modified since) or None.  One-shot."""
    global _slot
    s, _slot = _slot, None
    if s is None:
        return None
    ref = s[0]
    if (ENABLED and ref.data_ptr() == t.data_ptr() and ref.shape == t.shape and ref.dtype == t.dtype
            and ref.stride() == t.stride() and t._version == s[1] and ref._version == s[1]):
        STATS["hit"] += 1
        return s[2], s[3]
    STATS["miss"] += 1
    return None


def clear() -> None:
    global _slot
    _slot = None
