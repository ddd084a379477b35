"""Chainer-shaped front end: ``Classifier``, updaters, ``Trainer`` and its extensions.

Covers what the reference's Chainer scripts do (chainer/train_mnist.py:62-125,
chainer/train_mnist_gpu.py:49-93, chainer/train_mnist_multi.py:51-119):

    comm = create_communicator("pure_nccl")                     # rank / size / intra_rank
    model = Classifier(MLP(1000, 10)).to(comm.device)
    optimizer = create_multi_node_optimizer(Adam(), comm).setup(model)
    train = scatter_dataset(train, comm, shuffle=True)
    updater = StandardUpdater(SerialIterator(train, 100), optimizer, device=comm.device)
    trainer = Trainer(updater, (20, "epoch"), out="result")
    trainer.extend(create_multi_node_evaluator(Evaluator(test_iter, model), comm))
    if comm.rank == 0:
        trainer.extend(LogReport()); trainer.extend(PrintReport([...]))
    trainer.extend(snapshot(), trigger=(5, "epoch"))
    load_npz(args.resume, trainer)                               # full resume
    trainer.run()
"""
from __future__ import annotations

import json
import os
import time

import torch
import torch.nn as nn

from . import engine
from .parallel import rendezvous
from .parallel.data_parallel import DataParallel, DataParallelStep
from .parallel.optim import FusedAdam, FusedSGD
from .parallel.wrap import wrap
from .utils.data import scatter_dataset  # noqa: F401  (re-export, ChainerMN spelling)

create_communicator = rendezvous.create_communicator


class Classifier(nn.Module):
    """``L.Classifier``: wraps a predictor; reports softmax-CE loss and accuracy (train_mnist.py:62)."""

    def __init__(self, predictor):
        super().__init__()
        self.predictor = predictor

    def forward(self, x):
        return self.predictor(x)


class SerialIterator:
    """Batches of a dataset; ``repeat=False`` stops after one pass (train_mnist.py:74-76)."""

    def __init__(self, dataset, batch_size, repeat=True, shuffle=True, seed=0):
        self.dataset, self.batch_size, self.repeat, self.shuffle = dataset, batch_size, repeat, shuffle
        self.gen = torch.Generator().manual_seed(seed)
        self.epoch, self.is_new_epoch, self.pos = 0, False, 0
        self._order()

    def _order(self):
        n = len(self.dataset)
        self.order = torch.randperm(n, generator=self.gen).tolist() if self.shuffle else list(range(n))

    def reset(self):
        self.epoch, self.is_new_epoch, self.pos = 0, False, 0
        self._order()

    def __iter__(self):
        return self

    def __next__(self):
        n = len(self.dataset)
        if not self.repeat and self.pos >= n:
            raise StopIteration
        idx = self.order[self.pos: self.pos + self.batch_size]
        self.pos += self.batch_size
        self.is_new_epoch = self.pos >= n
        if self.is_new_epoch and self.repeat:
            self.epoch += 1
            self.pos = 0
            self._order()
        items = [self.dataset[i] for i in idx]
        xs = torch.stack([torch.as_tensor(a[0], dtype=torch.float32) for a in items])
        ts = torch.as_tensor([int(a[1]) for a in items])
        return xs, ts

    @property
    def epoch_detail(self):
        return self.epoch + self.pos / max(len(self.dataset), 1)


# --------------------------------------------------------------------------------------
# optimizers (Chainer's ``optimizer.setup(model)`` two-step construction)
# --------------------------------------------------------------------------------------
class _Deferred:
    def __init__(self, kind, comm=None, **kw):
        self.kind, self.kw, self.comm = kind, kw, comm
        self.target = self.inner = None

    def setup(self, model):
        self.target = model
        return self

    def materialise(self, wrapped):
        cls = FusedAdam if self.kind == "adam" else FusedSGD
        self.inner = cls(wrapped.parameters(), **self.kw)
        return self.inner


def Adam(alpha=1e-3, beta1=0.9, beta2=0.999, eps=1e-8):
    """``chainer.optimizers.Adam()`` defaults (train_mnist.py:69)."""
    return _Deferred("adam", lr=alpha, betas=(beta1, beta2), eps=eps)


def MomentumSGD(lr=0.01, momentum=0.9):
    return _Deferred("sgd", lr=lr, momentum=momentum)


def create_multi_node_optimizer(optimizer, comm):
    """Gradients are all-reduced over ``comm`` before every update (train_mnist_multi.py:81-82);
    here the update is fused into that all-reduce."""
    optimizer.comm = comm
    return optimizer


# --------------------------------------------------------------------------------------
# updaters
# --------------------------------------------------------------------------------------
class StandardUpdater:
    """One device per process (train_mnist.py:80-81, train_mnist_multi.py:97)."""

    def __init__(self, iterator, optimizer, device=None):
        self.iterator, self.deferred = iterator, optimizer
        comm = optimizer.comm or rendezvous.current()
        if device is None or (isinstance(device, int) and device < 0):
            device = torch.device("cpu")
        elif isinstance(device, int):
            device = torch.device("cuda", device)
        self.device = torch.device(device)
        if comm is None:
            comm = rendezvous.init(backend="b200" if self.device.type == "cuda" else "gloo", device=self.device)
        self.comm = comm
        self.model = optimizer.target
        self.wrapped = wrap(self.model.to(self.device), comm=comm)
        self.optimizer = optimizer.materialise(self.wrapped)
        self.step = engine.TrainStep(self.wrapped, self.optimizer, engine.CrossEntropyLoss())
        self.iteration = 0

    def update(self):
        x, t = next(self.iterator)
        if self.device.type == "cuda":
            x, t = x.pin_memory(), t.pin_memory()
        loss = self.step(x, t)
        stats = self.step.last_stats
        self.iteration += 1
        return loss, stats, x.shape[0]

    @property
    def epoch(self):
        return self.iterator.epoch

    @property
    def is_new_epoch(self):
        return self.iterator.is_new_epoch


class ParallelUpdater(StandardUpdater):
    """One process, several GPUs: the minibatch is split across ``devices`` and gradients are
    combined before the update (train_mnist_gpu.py:52-62,87-93).  Unlike Chainer's
    main/slave scheme no device is special: every replica applies the fused update."""

    def __init__(self, iterator, optimizer, devices):
        self.iterator, self.deferred = iterator, optimizer
        ids = [int(v) for v in (devices.values() if isinstance(devices, dict) else devices)]
        self.device = torch.device("cuda", ids[0])
        self.comm = rendezvous.current()
        self.model = optimizer.target
        self.wrapped = DataParallel(self.model, device_ids=ids)
        self.optimizer = optimizer.materialise(self.wrapped)
        self._dp_step = DataParallelStep(self.wrapped, self.optimizer)
        self.iteration = 0

    def update(self):
        x, t = next(self.iterator)
        stats = self._dp_step(x.pin_memory(), t.pin_memory())
        self.iteration += 1
        return stats[0], stats, x.shape[0]


# --------------------------------------------------------------------------------------
# trainer + extensions
# --------------------------------------------------------------------------------------
class Trainer:
    def __init__(self, updater, stop_trigger=(20, "epoch"), out="result"):
        self.updater, self.stop_trigger, self.out = updater, stop_trigger, out
        self.extensions = []
        self.observation = {}
        self.start_epoch = 0
        self.elapsed_time = 0.0

    def extend(self, ext, trigger=(1, "epoch"), name=None):
        self.extensions.append((ext, trigger))
        return self

    def _stop(self):
        n, unit = self.stop_trigger
        return (self.updater.epoch >= n) if unit == "epoch" else (self.updater.iteration >= n)

    def run(self):
        os.makedirs(self.out, exist_ok=True)
        up = self.updater
        up.iterator.epoch = max(up.iterator.epoch, self.start_epoch)
        t0 = time.time() - self.elapsed_time
        acc = torch.zeros(2, device=up.device)
        seen = 0
        up.wrapped.train()
        while not self._stop():
            loss, stats, n = up.update()
            if stats is not None:
                acc += stats.detach().to(acc.device) * torch.tensor([float(n), 1.0], device=acc.device)
            else:
                acc[0] += loss.detach() * n
            seen += n
            if up.is_new_epoch:
                tot = acc.cpu()
                self.observation = {"epoch": up.epoch, "iteration": up.iteration,
                                    "main/loss": tot[0].item() / max(seen, 1),
                                    "main/accuracy": tot[1].item() / max(seen, 1)}
                acc.zero_()
                seen = 0
                self.elapsed_time = time.time() - t0
                self.observation["elapsed_time"] = self.elapsed_time
                for ext, (k, unit) in self.extensions:
                    if unit == "epoch" and up.epoch % k == 0:
                        ext(self)
                up.wrapped.train()

    # whole-trainer (de)serialisation = Chainer's snapshot / load_npz(resume, trainer)
    def state_dict(self):
        bare = self.updater.model
        return {"model": {k: v.detach().cpu() for k, v in bare.state_dict().items()},
                "optimizer": self.updater.optimizer.state_dict(), "epoch": self.updater.epoch,
                "iteration": self.updater.iteration, "elapsed_time": self.elapsed_time}

    def load_state_dict(self, sd):
        self.updater.model.load_state_dict(sd["model"])
        self.updater.optimizer.load_state_dict(sd["optimizer"])
        self.start_epoch = sd["epoch"]
        self.updater.iterator.epoch = sd["epoch"]
        self.updater.iteration = sd["iteration"]
        self.elapsed_time = sd.get("elapsed_time", 0.0)
        w = self.updater.wrapped
        if isinstance(w, DataParallel):
            for a in w.arenas[1:]:
                a.param_flat.copy_(w.arenas[0].param_flat)
            for b in w.buffer_arenas[1:]:
                b.bytes.copy_(w.buffer_arenas[0].bytes)
        mods = [m for lst in w.packed for m in lst] if isinstance(w, DataParallel) else list(w.modules())
        for m in mods:
            if hasattr(m, "_packed_version"):
                m._packed_version = None


class Evaluator:
    """Per-epoch evaluation on a non-repeating iterator (train_mnist.py:85)."""

    def __init__(self, iterator, model, device=None):
        self.iterator, self.model, self.comm = iterator, model, None

    @torch.no_grad()
    def evaluate(self, device):
        self.iterator.reset()
        self.model.eval()
        loss_sum, correct, n = 0.0, 0, 0
        for x, t in self.iterator:
            x, t = x.to(device), t.to(device)
            out = self.model(x)
            loss_sum += torch.nn.functional.cross_entropy(out, t, reduction="sum").item()
            correct += (out.argmax(1) == t).sum().item()
            n += x.shape[0]
        res = {"validation/main/loss": loss_sum / max(n, 1), "validation/main/accuracy": correct / max(n, 1)}
        if self.comm is not None and self.comm.size > 1:
            res = {k: self.comm.allreduce_mean(v) for k, v in res.items()}
        return res

    def __call__(self, trainer):
        trainer.observation.update(self.evaluate(trainer.updater.device))


def create_multi_node_evaluator(evaluator, comm):
    """Averages the evaluation metrics over ranks (train_mnist_multi.py:102-103)."""
    evaluator.comm = comm
    return evaluator


class LogReport:
    def __init__(self, log_name="log"):
        self.log_name, self.log = log_name, []

    def __call__(self, trainer):
        self.log.append(dict(trainer.observation))
        with open(os.path.join(trainer.out, self.log_name), "w") as f:
            json.dump(self.log, f, indent=1)


class PrintReport:
    def __init__(self, entries):
        self.entries, self._header = entries, False

    def __call__(self, trainer):
        if not self._header:
            print("  ".join(f"{e:>24s}" for e in self.entries), flush=True)
            self._header = True
        row = []
        for e in self.entries:
            v = trainer.observation.get(e, "")
            row.append(f"{v:>24.6g}" if isinstance(v, (int, float)) else f"{str(v):>24s}")
        print("  ".join(row), flush=True)


class snapshot:  # noqa: N801  (Chainer's lower-case extension name)
    def __init__(self, filename="snapshot_iter_{iteration}"):
        self.filename = filename

    def __call__(self, trainer):
        sd = trainer.state_dict()      # collective when optimizer state is sharded: all ranks call
        comm = trainer.updater.comm
        if comm is None or comm.rank == 0:
            path = os.path.join(trainer.out, self.filename.format(iteration=trainer.updater.iteration))
            torch.save(sd, path + ".tmp")
            os.replace(path + ".tmp", path)


class dump_graph:  # noqa: N801
    """Chainer writes the computational graph once (train_mnist.py:89); here: the module tree."""

    def __init__(self, root_name="main/loss", out_name="cg.txt"):
        self.out_name, self.done = out_name, False

    def __call__(self, trainer):
        if not self.done:
            with open(os.path.join(trainer.out, self.out_name), "w") as f:
                f.write(repr(trainer.updater.model))
            self.done = True


def load_npz(path, trainer):
    """``chainer.serializers.load_npz(args.resume, trainer)`` (train_mnist.py:120-122)."""
    trainer.load_state_dict(torch.load(path, map_location="cpu", weights_only=False))
