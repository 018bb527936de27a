"""Fused masked cross-entropy: ``mean_{i in idx} CE(logits[i], labels[i])`` in one kernel each way (``csrc/loss_ops.cu``).

Parity: the reference trainer's ``criterion(out[train_mask], y[train_mask])`` (``experiments/OGB/main.py:115,140-148``) -- an
ATen chain of ~8 launches.  ``idx`` is an int64 index tensor (no boolean masks: they force a device->host sync).  The backward
leaves a row-support hint (``ops/_sparsity.py``) so the classifier's backward works on the ``len(idx)`` non-zero rows only.
The gradient rows are produced in bf16 (they feed bf16 tensor-core GEMMs) also when the logits are fp32.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F
from torch.autograd import Function

from dgraph_b200.ops import _native as N
from dgraph_b200.ops import _sparsity


def masked_cross_entropy_ref(logits: torch.Tensor, labels: torch.Tensor, idx: Optional[torch.Tensor]) -> torch.Tensor:
    up = (lambda t: t) if logits.dtype == torch.float64 else (lambda t: t.float())
    if idx is None:
        return F.cross_entropy(up(logits), labels.reshape(-1))
    return F.cross_entropy(up(logits.index_select(0, idx)), labels.reshape(-1).index_select(0, idx))


def _row_support(idx: torch.Tensor):
    """``None`` when ``idx`` lists every row at most once, else ``(unique rows, position of every entry among them)``.  One
    device-to-host read per index tensor and version; the result is cached on the tensor, so the unique-row tensor is the same object
    every step (consumers of the row-support hint cache per-label-set plans on its identity)."""
    tag = (idx.data_ptr(), idx.numel(), idx._version)
    hit = getattr(idx, "_dgb_support", None)
    if hit is not None and hit[0] == tag:
        return hit[1]
    sup = None
    if idx.numel() and bool(torch.bincount(idx, minlength=1).max() > 1):
        uniq, inv = torch.unique(idx, return_inverse=True)
        sup = (uniq.contiguous(), inv.contiguous())
    try:
        idx._dgb_support = (tag, sup)
    except Exception:  # noqa: BLE001
        pass
    return sup


def _merge_duplicate_rows(dsel: torch.Tensor, sup) -> torch.Tensor:
    """Compact gradient rows of the unique rows: a row listed k times counts k times in the loss, so its k rows add up (fp32 sum,
    rounded once)."""
    uniq, inv = sup
    acc = torch.zeros((uniq.numel(), dsel.shape[1]), dtype=torch.float32, device=dsel.device)
    acc.index_add_(0, inv, dsel.float())
    return acc.to(dsel.dtype)


class _MaskedCE(Function):
    @staticmethod
    def forward(ctx, logits, labels, idx):
        V, C = logits.shape
        n_sel = V if idx is None else idx.numel()
        assert logits.stride(1) == 1 and logits.dtype in (torch.bfloat16, torch.float32)
        loss = torch.zeros((), dtype=torch.float32, device=logits.device)
        lse = torch.empty(n_sel, dtype=torch.float32, device=logits.device)
        N.check(N.lib().dgb_masked_ce_fwd(N.ptr(logits), N.i64(logits.stride(0)), N.i32(logits.dtype == torch.float32),
                                          N.ptr(labels), N.ptr(idx), N.i64(n_sel), N.i32(C), N.ptr(loss), N.ptr(lse),
                                          N.stream()), "dgb_masked_ce_fwd")
        N.count_launch()
        ctx.save_for_backward(logits, labels, idx, lse)
        return loss

    @staticmethod
    def backward(ctx, gout):
        logits, labels, idx, lse = ctx.saved_tensors
        V, C = logits.shape
        n_sel = V if idx is None else idx.numel()
        ld_o = (C + 7) // 8 * 8  # 16-byte rows for the GEMMs that consume the compact gradient
        dsel = torch.empty((n_sel, ld_o), dtype=torch.bfloat16, device=logits.device)
        gout = gout.to(torch.float32).contiguous()
        N.check(N.lib().dgb_masked_ce_bwd(N.ptr(logits), N.i64(logits.stride(0)), N.i32(logits.dtype == torch.float32),
                                          N.ptr(labels), N.ptr(idx), N.i64(n_sel), N.i32(C), N.ptr(lse), N.ptr(gout),
                                          N.ptr(dsel), N.i64(ld_o), N.stream()), "dgb_masked_ce_bwd")
        N.count_launch()
        if idx is None:
            return dsel[:, :C].to(logits.dtype), None, None
        dlogits = torch.zeros((V, C), dtype=logits.dtype, device=logits.device)
        # The hint is left on every call, whatever the index tensor holds: in a distributed step the consumers choose collective
        # code paths by its presence, so it must not depend on one rank's data (duplicates on one rank only, an empty label set).
        sup = _row_support(idx)
        if sup is not None:
            idx, dsel = sup[0], _merge_duplicate_rows(dsel, sup)
        dlogits.index_copy_(0, idx, dsel[:, :C].to(logits.dtype))
        if logits.dtype == torch.bfloat16:
            _sparsity.put(dlogits, idx, dsel)  # padded compact rows [n_sel, ld_o]
        return dlogits, None, None


def masked_cross_entropy(logits: torch.Tensor, labels: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Mean cross-entropy over rows ``idx`` of ``logits [V, C]`` with ``labels [V]`` (int64).  ``idx=None`` = all rows."""
    labels = labels.reshape(-1)
    if logits.is_cuda and N.use_native(logits) and labels.dtype == torch.int64 and (idx is None or idx.dtype == torch.int64):
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        return _MaskedCE.apply(logits, labels.contiguous(), None if idx is None else idx.contiguous())
    return masked_cross_entropy_ref(logits, labels, idx)


class MaskedCrossEntropyLoss(torch.nn.Module):
    def forward(self, logits, labels, idx=None):
        return masked_cross_entropy(logits, labels, idx)
