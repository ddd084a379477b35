"""Rank-aware logging and running meters.

Keeps the reference's human-readable line -- ``Epoch[e]: [i/n]| loss | acc | batch time`` with a
``From Rank: r,`` prefix under DDP (pytorch/single_gpu.py:114-116;
pytorch/distributed_data_parallel.py:144-148) -- and adds images/sec.  Loss / accuracy counters
stay on the device and are read back only at the log interval, instead of the reference's
three ``.item()`` host syncs per step.
"""
from __future__ import annotations

import datetime
import json
import sys
import time

import torch


def get_rank() -> int:
    from ..parallel import rendezvous
    c = rendezvous.current()
    return c.rank if c else 0


def rank0_only(fn):
    def wrapper(*a, **k):
        if get_rank() == 0:
            return fn(*a, **k)
    return wrapper


class RankLogger:
    def __init__(self, rank=None, all_ranks=True, stream=None):
        self.rank = get_rank() if rank is None else rank
        self.all_ranks, self.stream = all_ranks, stream or sys.stdout

    def info(self, msg, force=False):
        if self.all_ranks or self.rank == 0 or force:
            print(f"From Rank: {self.rank}, {msg}", file=self.stream, flush=True)


class TrainMeter:
    """Running mean loss / accuracy with device-side accumulation."""

    def __init__(self, device="cpu"):
        self.device = torch.device(device)
        self.reset()

    def reset(self):
        self.loss_sum = torch.zeros((), device=self.device)
        self.correct = torch.zeros((), device=self.device)
        self.total = 0
        self.steps = 0
        self.t0 = time.time()

    def update(self, loss, outputs=None, targets=None, correct=None):
        self.loss_sum += loss.detach()
        self.steps += 1
        if correct is not None:
            self.correct += correct
        elif outputs is not None:
            self.correct += (outputs.detach().argmax(1) == targets).sum()
        if targets is not None:
            self.total += targets.shape[0]

    def summary(self):
        """One host sync for everything."""
        vals = torch.stack([self.loss_sum.float(), self.correct.float()]).cpu()
        loss = vals[0].item() / max(self.steps, 1)
        acc = 100.0 * vals[1].item() / max(self.total, 1)
        return loss, acc

    def line(self, epoch, batch_idx, n_batches, batch_time):
        loss, acc = self.summary()
        ips = self.total / max(time.time() - self.t0, 1e-9)
        return (f"Epoch:[{epoch}][{batch_idx}/{n_batches}]| loss: {loss:.3f} | acc: {acc:.3f} | "
                f"batch time: {batch_time:.3f}s | {ips:.0f} img/s")


def format_elapsed(seconds: float) -> str:
    return str(datetime.timedelta(seconds=seconds))


def format_bucket_report(rows) -> str:
    """Table of ``Reducer.bucket_report()`` rows: per-bucket collective time and achieved bus
    bandwidth, averaged over the steps that were timed (SURVEY.md section 5.5 asks for it next to
    the reference's loss / accuracy line, pytorch/single_gpu.py:114-116)."""
    agg = {}
    for r in rows:
        a = agg.setdefault(r["bucket"], {"bytes": r["bytes"], "kind": r["kind"], "world": r["world"], "ms": 0.0, "n": 0})
        a["ms"] += r["ms"]
        a["n"] += 1
    lines = ["bucket |     bytes | variant        |    ms | bus GB/s"]
    for b in sorted(agg):
        a = agg[b]
        ms = a["ms"] / max(a["n"], 1)
        w = a["world"]
        bus = 2.0 * (w - 1) / w * a["bytes"] / (ms * 1e-3) / 1e9 if ms > 0 else float("inf")
        lines.append(f"{b:6d} | {a['bytes']:9d} | {a['kind']:14s} | {ms:5.3f} | {bus:8.1f}")
    return "\n".join(lines)


class JsonLog:
    """Append-only JSON log (Chainer ``LogReport``, chainer/train_mnist.py:96)."""

    def __init__(self, path=None):
        self.path, self.entries = path, []

    def append(self, entry: dict):
        self.entries.append(entry)
        if self.path and get_rank() == 0:
            with open(self.path, "w") as f:
                json.dump(self.entries, f, indent=1)
