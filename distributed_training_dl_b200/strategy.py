"""Keras-shaped front end: strategies, ``compile`` / ``fit`` / ``evaluate``, callbacks.

Covers what the reference's TensorFlow 2 scripts do (tensorflow2/mnist_single.py:65-92,
mnist_mirror_strategy.py:12,68-79, mnist_multi_worker_strategy.py:14-29,85-102):

    strategy = MirroredStrategy()                      # all GPUs of this process
    with strategy.scope():
        model = Model(MnistCNN())
        model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    model.fit(x, y, batch_size=64, epochs=10, callbacks=[ModelCheckpoint("dir/ckpt_{epoch}")],
              validation_data=(xv, yv))

``batch_size`` is the GLOBAL batch, split across replicas; every replica applies the identical
update after the gradient all-reduce (here: fused into it).  ``MultiWorkerMirroredStrategy``
takes the reference's ``worker_hosts`` / ``task_index`` spelling (its ``TF_CONFIG``) or any
rendezvous ``init()`` understands, one process per worker.
"""
from __future__ import annotations

import contextlib
import glob
import os
import re
import time

import torch

from . import engine
from .parallel import rendezvous
from .parallel.data_parallel import DataParallel, DataParallelStep
from .parallel.optim import FusedAdam, FusedSGD
from .parallel.wrap import wrap
from .utils.logging import get_rank

_scope_stack = []


class _Strategy:
    num_replicas_in_sync = 1

    @contextlib.contextmanager
    def scope(self):
        _scope_stack.append(self)
        try:
            yield self
        finally:
            _scope_stack.pop()


class OneDeviceStrategy(_Strategy):
    """Single device (tensorflow2/mnist_single.py)."""

    def __init__(self, device=None):
        self.device = torch.device(device or ("cuda:0" if torch.cuda.is_available() else "cpu"))

    def build(self, net, make_opt):
        comm = rendezvous.current() or rendezvous.init(
            backend="b200" if self.device.type == "cuda" else "gloo", device=self.device)
        net = wrap(net.to(self.device), comm=comm)
        opt = make_opt(net.parameters())
        return net, opt, engine.TrainStep(net, opt, engine.CrossEntropyLoss()), self.device


class MirroredStrategy(_Strategy):
    """One process, all local GPUs (tensorflow2/mnist_mirror_strategy.py:12)."""

    def __init__(self, devices=None):
        if devices is None:
            n = torch.cuda.device_count()
            devices = [f"cuda:{i}" for i in range(n)] if n else ["cpu"]
        self.devices = [torch.device(d) for d in devices]
        self.num_replicas_in_sync = len(self.devices)

    def build(self, net, make_opt):
        if len(self.devices) == 1:
            return OneDeviceStrategy(self.devices[0]).build(net, make_opt)
        dp = DataParallel(net, device_ids=[d.index for d in self.devices])
        opt = make_opt(dp.parameters())
        step = DataParallelStep(dp, opt)
        return dp, opt, step, self.devices[0]


class MultiWorkerMirroredStrategy(_Strategy):
    """One process per worker (tensorflow2/mnist_multi_worker_strategy.py:14-29)."""

    def __init__(self, worker_hosts=None, task_index=None, job_name="worker", backend=None, **init_kw):
        if job_name != "worker":
            raise ValueError("only job type 'worker' takes part in synchronous training "
                             "(tensorflow2/mnist_multi_worker_strategy.py:15-16)")
        self.comm = rendezvous.init(backend=backend, worker_hosts=worker_hosts, task_index=task_index, **init_kw)
        self.num_replicas_in_sync = self.comm.size

    def build(self, net, make_opt):
        net = wrap(net.to(self.comm.device), comm=self.comm)
        opt = make_opt(net.parameters())
        return net, opt, engine.TrainStep(net, opt, engine.CrossEntropyLoss()), self.comm.device


# --------------------------------------------------------------------------------------
# callbacks
# --------------------------------------------------------------------------------------
class Callback:
    def set_model(self, model):
        self.model = model

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass


class ModelCheckpoint(Callback):
    """``filepath`` may contain ``{epoch}``; weights only, every epoch, rank 0 writes
    (the reference lets every worker write the same files: SURVEY.md Q14)."""

    def __init__(self, filepath, save_weights_only=True):
        self.filepath = filepath

    def on_epoch_end(self, epoch, logs=None):
        self.model.save_weights(self.filepath.format(epoch=epoch + 1))


class PrintLR(Callback):
    """The callback the reference defines but never registers (tensorflow2/mnist_single.py:50-56)."""

    def on_epoch_end(self, epoch, logs=None):
        if get_rank() == 0:
            print(f"\nLearning rate for epoch {epoch + 1} is {self.model.optimizer.param_groups[0]['lr']}")


class TensorBoard(Callback):
    """``tf.keras.callbacks.TensorBoard(log_dir, histogram_freq=1)`` as all three TF2 scripts register
    it (tensorflow2/mnist_single.py:72-76, mnist_multi_worker_strategy.py:79-83): per-epoch scalars
    (loss / accuracy / val_*) and, every ``histogram_freq`` epochs, a histogram of every weight
    tensor.  Event files come from ``torch.utils.tensorboard`` when the ``tensorboard`` package is
    importable; otherwise the same records go to ``<log_dir>/events.jsonl`` (one JSON object per line,
    histograms as 30-bin counts).  Only rank 0 writes (the reference lets every worker write into the
    same directory: SURVEY.md Q14)."""

    def __init__(self, log_dir="./logs", histogram_freq=0, bins=30):
        self.log_dir, self.histogram_freq, self.bins = log_dir, int(histogram_freq), bins
        self._writer = self._jsonl = None

    def _open(self):
        if self._writer is not None or self._jsonl is not None:
            return
        os.makedirs(self.log_dir, exist_ok=True)
        try:
            from torch.utils.tensorboard import SummaryWriter
            self._writer = SummaryWriter(log_dir=self.log_dir)
        except Exception:  # noqa: BLE001 - tensorboard not installed: plain JSON lines instead
            self._jsonl = open(os.path.join(self.log_dir, "events.jsonl"), "a")

    def on_epoch_end(self, epoch, logs=None):
        if get_rank() != 0:
            return
        self._open()
        step = epoch + 1
        scalars = {k: float(v) for k, v in (logs or {}).items() if k != "time"}
        hist = {}
        if self.histogram_freq and step % self.histogram_freq == 0:
            for name, p in self.model._bare().named_parameters():
                hist[name] = p.detach().float().cpu()
        if self._writer is not None:
            for k, v in scalars.items():
                self._writer.add_scalar("epoch_" + k, v, step)
            for name, t in hist.items():
                self._writer.add_histogram(name, t, step)
            self._writer.flush()
        else:
            import json
            rec = {"epoch": step, "scalars": scalars, "histograms": {}}
            for name, t in hist.items():
                lo, hi = float(t.min()), float(t.max())
                rec["histograms"][name] = {"min": lo, "max": hi,
                                           "counts": torch.histc(t, bins=self.bins, min=lo, max=hi).int().tolist()}
            self._jsonl.write(json.dumps(rec) + "\n")
            self._jsonl.flush()

    def close(self):
        if self._writer is not None:
            self._writer.close()
        if self._jsonl is not None:
            self._jsonl.close()
        self._writer = self._jsonl = None


class History(Callback):
    def __init__(self):
        self.history = {}

    def on_epoch_end(self, epoch, logs=None):
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


def latest_checkpoint(directory):
    """``tf.train.latest_checkpoint``: newest ``ckpt_<n>`` file in ``directory``."""
    best, best_n = None, -1
    for path in glob.glob(os.path.join(directory, "ckpt_*")):
        m = re.search(r"ckpt_(\d+)", os.path.basename(path))
        if m and int(m.group(1)) > best_n:
            best, best_n = path, int(m.group(1))
    return best


# --------------------------------------------------------------------------------------
# model
# --------------------------------------------------------------------------------------
class Model:
    def __init__(self, net):
        self.net = net
        self.strategy = _scope_stack[-1] if _scope_stack else None
        self.compiled = False

    def compile(self, optimizer="adam", loss="sparse_categorical_crossentropy", metrics=("accuracy",),
                learning_rate=None):
        if loss not in ("sparse_categorical_crossentropy", "softmax_cross_entropy"):
            raise ValueError(f"unsupported loss {loss!r}")
        strategy = self.strategy or (_scope_stack[-1] if _scope_stack else OneDeviceStrategy())
        self.strategy = strategy

        def make_opt(params):
            if callable(optimizer):
                return optimizer(params)
            if optimizer == "adam":
                return FusedAdam(params, lr=learning_rate or 1e-3, keras=True)
            if optimizer == "sgd":
                return FusedSGD(params, lr=learning_rate or 0.01)
            raise ValueError(f"unsupported optimizer {optimizer!r}")

        self.module, self.optimizer, self.step, self.device = strategy.build(self.net, make_opt)
        self.metrics = list(metrics)
        self.compiled = True
        return self

    # -- data ---------------------------------------------------------------------------
    def _local_view(self, x, y):
        """Multi-worker: every worker trains on its own strided shard of the arrays."""
        comm = getattr(self.strategy, "comm", None)
        if comm is None or comm.size == 1:
            return x, y, 1
        return x[comm.rank::comm.size], y[comm.rank::comm.size], comm.size

    @staticmethod
    def _as_tensors(x, y):
        x = torch.as_tensor(x, dtype=torch.float32)
        y = torch.as_tensor(y).long()
        return x, y

    def fit(self, x, y, batch_size=32, epochs=1, callbacks=(), validation_data=None, shuffle=True, verbose=1):
        assert self.compiled, "call compile() first"
        x, y = self._as_tensors(x, y)
        x, y, workers = self._local_view(x, y)
        local_batch = max(1, batch_size // workers)
        history = History()
        cbs = list(callbacks) + [history]
        for cb in cbs:
            cb.set_model(self)
        pin = self.device.type == "cuda"
        self.module.train()
        for epoch in range(epochs):
            for cb in cbs:
                cb.on_epoch_begin(epoch)
            t0 = time.time()
            perm = torch.randperm(x.shape[0]) if shuffle else torch.arange(x.shape[0])
            stats_sum, seen = torch.zeros(2, device=self.device), 0
            for i in range(0, x.shape[0], local_batch):
                idx = perm[i: i + local_batch]
                xb, yb = x[idx], y[idx]
                if pin:
                    xb, yb = xb.pin_memory(), yb.pin_memory()
                out = self.step(xb, yb)
                stats = out if out.numel() == 2 else self.step.last_stats
                if stats is not None:
                    stats_sum += stats.detach() * torch.tensor([float(len(idx)), 1.0], device=stats.device)
                else:
                    stats_sum[0] += out.detach() * len(idx)
                seen += len(idx)
            tot = stats_sum.cpu()
            logs = {"loss": tot[0].item() / max(seen, 1), "accuracy": tot[1].item() / max(seen, 1)}
            if validation_data is not None:
                vl, va = self.evaluate(*validation_data, batch_size=max(local_batch, 256), verbose=0)
                logs.update(val_loss=vl, val_accuracy=va)
                self.module.train()
            logs["time"] = time.time() - t0
            if verbose and get_rank() == 0:
                print(f"Epoch {epoch + 1}/{epochs} - " + " - ".join(
                    f"{k}: {v:.4f}" for k, v in logs.items()), flush=True)
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
        return history

    @torch.no_grad()
    def evaluate(self, x, y, batch_size=256, verbose=0):
        x, y = self._as_tensors(x, y)
        net = self.module
        net.eval()
        loss_sum, correct = 0.0, 0
        for i in range(0, x.shape[0], batch_size):
            xb, yb = x[i: i + batch_size].to(self.device), y[i: i + batch_size].to(self.device)
            out = net(xb)
            loss_sum += torch.nn.functional.cross_entropy(out, yb, reduction="sum").item()
            correct += (out.argmax(1) == yb).sum().item()
        n = max(x.shape[0], 1)
        if verbose and get_rank() == 0:
            print(f"{x.shape[0]}/{x.shape[0]} - loss: {loss_sum / n:.4f} - accuracy: {correct / n:.4f}")
        return loss_sum / n, correct / n

    @torch.no_grad()
    def predict(self, x, batch_size=256):
        self.module.eval()
        x = torch.as_tensor(x, dtype=torch.float32)
        return torch.cat([self.module(x[i: i + batch_size].to(self.device)).softmax(1).cpu()
                          for i in range(0, x.shape[0], batch_size)])

    # -- weights --------------------------------------------------------------------------
    def _bare(self):
        return self.module.module if hasattr(self.module, "module") else self.module

    def save_weights(self, path):
        if get_rank() != 0:
            return
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        torch.save({k: v.detach().cpu() for k, v in self._bare().state_dict().items()}, path)

    def load_weights(self, path):
        sd = torch.load(path, map_location="cpu")
        self._bare().load_state_dict(sd)
        dp = self.module if isinstance(self.module, DataParallel) else None
        if dp is not None:  # copy replica 0's weights to the others
            for a in dp.arenas[1:]:
                a.param_flat.copy_(dp.arenas[0].param_flat)
            for b in dp.buffer_arenas[1:]:
                b.bytes.copy_(dp.buffer_arenas[0].bytes)
        mods = [m for lst in dp.packed for m in lst] if dp is not None else list(self.module.modules())
        for m in mods:
            if hasattr(m, "_packed_version"):
                m._packed_version = None

    def summary(self):
        n = sum(p.numel() for p in self._bare().parameters())
        print(self._bare())
        print(f"Total params: {n:,}")
