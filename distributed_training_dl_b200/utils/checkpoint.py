"""Checkpoint / resume in torch ``state_dict`` format.

Covers the three save spellings of the reference -- bare keys (pytorch/single_gpu.py:81-84),
``module.``-prefixed wrapper keys (pytorch/data_parallel.py:87-90), per-rank DDP files
(pytorch/distributed_data_parallel.py:103-115) -- and the full-trainer resume of the Chainer
examples (chainer/train_mnist.py:92-93,120-122): model + optimizer + scheduler + epoch.
Fixes the reference's race (non-recursive mkdir from every rank): the
directory is created with ``exist_ok`` and a rank-0-only save is one call.
"""
from __future__ import annotations

import os

import torch

from .logging import get_rank


def strip_module_prefix(sd: dict) -> dict:
    return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}


def add_module_prefix(sd: dict) -> dict:
    return {(k if k.startswith("module.") else "module." + k): v for k, v in sd.items()}


def _cpu(sd):
    return {k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in sd.items()}


def save_model(model, path, rank0_only=True, per_rank=False):
    """``torch.save(model.state_dict(), path)``.  ``per_rank=True`` appends ``_{rank}`` like the
    reference's DDP script; ``rank0_only`` skips other ranks."""
    rank = get_rank()
    if per_rank:
        root, ext = os.path.splitext(path)
        path = f"{root}_{rank}{ext}"
    elif rank0_only and rank != 0:
        return None
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(_cpu(model.state_dict()), path)
    return path


def load_model(model, path, strict=True):
    """Loads either key spelling into either a wrapped or a bare model."""
    sd = torch.load(path, map_location="cpu")
    own = model.state_dict().keys()
    wants_prefix = any(k.startswith("module.") for k in own)
    sd = add_module_prefix(sd) if wants_prefix else strip_module_prefix(sd)
    return model.load_state_dict(sd, strict=strict)


def save_snapshot(path, model, optimizer=None, scheduler=None, epoch=0, iteration=0, extra=None,
                  rank0_only=True):
    """Everything needed to resume (Chainer ``extensions.snapshot``)."""
    if rank0_only and get_rank() != 0:
        if optimizer is not None and hasattr(optimizer, "_gather_state"):
            optimizer.state_dict()  # collective on sharded optimizer state: all ranks take part
        return None
    bare = model.module if hasattr(model, "module") else model
    snap = {"model": _cpu(bare.state_dict()), "epoch": epoch, "iteration": iteration, "extra": extra or {}}
    if optimizer is not None:
        snap["optimizer"] = optimizer.state_dict()
    if scheduler is not None:
        snap["scheduler"] = scheduler.state_dict()
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(snap, tmp)
    os.replace(tmp, path)
    return path


def load_snapshot(path, model, optimizer=None, scheduler=None):
    snap = torch.load(path, map_location="cpu", weights_only=False)
    bare = model.module if hasattr(model, "module") else model
    bare.load_state_dict(snap["model"])
    if optimizer is not None and "optimizer" in snap:
        optimizer.load_state_dict(snap["optimizer"])
    if scheduler is not None and "scheduler" in snap:
        scheduler.load_state_dict(snap["scheduler"])
    return snap
