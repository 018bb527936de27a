"""Tiny ``fire``-compatible CLI shim: ``run_cli(main)`` maps ``--flag value`` to keyword arguments of ``main``.

The reference's entry points use ``python-fire`` (``experiments/OGB/main.py:362``); ``fire`` is used when importable,
otherwise an ``argparse`` parser is derived from the function signature (types from defaults / annotations).
"""
from __future__ import annotations

import argparse
import inspect


def _str2bool(v: str) -> bool:
    return str(v).lower() in ("1", "true", "yes", "y", "t")


def run_cli(fn):
    try:
        import fire  # type: ignore

        return fire.Fire(fn)
    except ImportError:
        pass
    sig = inspect.signature(fn)
    ap = argparse.ArgumentParser(description=(fn.__doc__ or "").strip().split("\n")[0])
    for name, p in sig.parameters.items():
        default = None if p.default is inspect._empty else p.default
        ann = p.annotation if p.annotation is not inspect._empty else (type(default) if default is not None else str)
        if ann is bool or isinstance(default, bool):
            ap.add_argument(f"--{name}", type=_str2bool, nargs="?", const=True, default=default)
        else:
            typ = ann if ann in (int, float, str) else (type(default) if default is not None else str)
            ap.add_argument(f"--{name}", type=typ, default=default, required=p.default is inspect._empty)
    return fn(**vars(ap.parse_args()))
