"""Packed bf16 copies of fp32 master weights, refreshed once per optimizer step instead of once per call.

The fused layers consume re-laid-out bf16 operands (``[W_i ; W_j]`` stacked and zero-padded, its transpose, the padded bias).
They depend only on the parameter values, which change once per step, so the packed tensors are cached on the parameter and
rebuilt when the parameter changed: either its autograd version counter moved (stock optimizers update in place) or
:func:`bump` was called (``FusedAdam`` writes the flat parameter arena from a native kernel, which PyTorch cannot see).
"""
from __future__ import annotations

_EPOCH = 0


def bump() -> None:
    """Some parameters changed behind PyTorch's back (native optimizer kernel, checkpoint load into the flat arena)."""
    global _EPOCH
    _EPOCH += 1


def get(param, key, build):
    """``build()`` result cached on ``param`` under ``key`` until the parameter changes."""
    tag = (_EPOCH, param._version, param.data_ptr(), key)
    hit = getattr(param, "_dgb_packed", None)
    if hit is not None and hit[0] == tag:
        return hit[1]
    val = build()
    try:
        param._dgb_packed = (tag, val)
    except Exception:  # noqa: BLE001 - tensors that refuse attributes simply repack every call
        pass
    return val
