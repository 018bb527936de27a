"""Checkpoint save **and resume** (the reference only saves ``model.state_dict()``, ``experiments/GraphCast/
train_graphcast.py:150-151``; SURVEY §5 "no load/resume path anywhere").

Format: the model file stays a plain ``state_dict`` (reference key names, loadable by the reference's modules); a side
file carries optimizer / scheduler / step / RNG state.  Replicated weights are identical on all ranks, so rank 0 writes.
"""
from __future__ import annotations

import os
from typing import Optional

import torch


def save_checkpoint(path: str, model, optimizer=None, scheduler=None, step: int = 0, rank: int = 0, extra=None):
    if rank != 0:
        return
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    sd = model.state_dict()
    torch.save({k: v.detach().cpu() for k, v in sd.items()}, path)
    aux = {"step": int(step), "rng_cpu": torch.get_rng_state(), "extra": extra}
    if torch.cuda.is_available():
        aux["rng_cuda"] = torch.cuda.get_rng_state()
    if optimizer is not None:
        aux["optimizer"] = optimizer.state_dict()
    if scheduler is not None:
        aux["scheduler"] = scheduler.state_dict()
    torch.save(aux, path + ".train_state")


def load_checkpoint(path: str, model, optimizer=None, scheduler=None, map_location="cpu", restore_rng: bool = False) -> int:
    sd = torch.load(path, map_location=map_location, weights_only=True)
    sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}  # DDP-prefixed reference files
    model.load_state_dict(sd)
    step = 0
    aux_path = path + ".train_state"
    if os.path.exists(aux_path):
        aux = torch.load(aux_path, map_location=map_location, weights_only=False)
        step = int(aux.get("step", 0))
        if optimizer is not None and "optimizer" in aux:
            optimizer.load_state_dict(aux["optimizer"])
        if scheduler is not None and "scheduler" in aux:
            scheduler.load_state_dict(aux["scheduler"])
        if restore_rng:
            torch.set_rng_state(aux["rng_cpu"])
            if torch.cuda.is_available() and "rng_cuda" in aux:
                torch.cuda.set_rng_state(aux["rng_cuda"])
    return step
