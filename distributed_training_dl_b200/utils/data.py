"""Data sharding and synthetic datasets.

``DistributedSampler`` reproduces the semantics the reference relies on
(pytorch/distributed_data_parallel.py:87; torch/utils/data/distributed.py): a seeded
permutation of the dataset, padded by wrap-around to a multiple of the world size and
sub-sampled with stride ``world_size`` starting at ``rank``.  The reference never calls
``set_epoch`` so every epoch reuses one order; ``set_epoch`` is provided.
``scatter_dataset`` is ChainerMN's spelling (chainer/train_mnist_multi.py:91-92).
"""
from __future__ import annotations

import math

import torch
from torch.utils.data import Dataset, Sampler, Subset


class DistributedSampler(Sampler):
    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False):
        if num_replicas is None or rank is None:
            from ..parallel import rendezvous
            comm = rendezvous.current()
            num_replicas = num_replicas if num_replicas is not None else (comm.size if comm else 1)
            rank = rank if rank is not None else (comm.rank if comm else 0)
        if not 0 <= rank < num_replicas:
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.dataset, self.num_replicas, self.rank = dataset, num_replicas, rank
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        n = len(dataset)
        if drop_last and n % num_replicas:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def indices(self):
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(n, generator=g).tolist()
        else:
            idx = list(range(n))
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad > 0:
                idx += (idx * math.ceil(pad / max(len(idx), 1)))[:pad]
        else:
            idx = idx[: self.total_size]
        return idx[self.rank: self.total_size: self.num_replicas]

    def __iter__(self):
        return iter(self.indices())

    def __len__(self):
        return self.num_samples


def scatter_dataset(dataset, comm, shuffle=False, seed=0):
    """Every rank gets an equal-sized shard; ``dataset`` may be ``None`` on ranks other than 0,
    in which case rank 0's dataset object is broadcast first (ChainerMN semantics)."""
    if comm.size > 1:
        dataset = comm.broadcast_object(dataset if comm.rank == 0 else None, src=0)
    sampler = DistributedSampler(dataset, num_replicas=comm.size, rank=comm.rank, shuffle=shuffle, seed=seed)
    return Subset(dataset, sampler.indices())


class SyntheticDataset(Dataset):
    """Random images / labels of a named shape, generated once from a seed.

    Shapes: ``"cifar"`` 3x32x32, ``"mnist"`` 1x28x28, ``"mnist_flat"`` 784, ``"imagenet"`` 3x224x224.
    """
    SHAPES = {"cifar": (3, 32, 32), "mnist": (1, 28, 28), "mnist_flat": (784,), "imagenet": (3, 224, 224)}

    def __init__(self, kind="cifar", length=1024, num_classes=10, seed=0):
        g = torch.Generator().manual_seed(seed)
        shape = self.SHAPES[kind] if isinstance(kind, str) else tuple(kind)
        self.x = torch.randn((length,) + shape, generator=g)
        self.y = torch.randint(0, num_classes, (length,), generator=g)

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i], self.y[i]


class PinnedPrefetcher:
    """Feeds ``(input, target)`` batches to a training step through a ring of reusable page-locked
    buffers filled by a background thread (the real-data input path: pytorch/single_gpu.py:96-97
    copies every batch from pageable memory, synchronously).

    ``tensor.pin_memory()`` per batch allocates page-locked memory every step; here ``depth``
    buffers per field are allocated once (re-allocated only if a batch is larger than any before)
    and the copy into them overlaps the previous step.  A slot is refilled only after the work
    the consumer enqueued while it held that slot has finished on the GPU: when the consumer asks
    for the next batch, an event is recorded on its current stream and the worker waits for it
    (on its own thread) before overwriting the slot.  Without CUDA the buffers are ordinary
    memory and there is nothing to wait for.

        for x, t in PinnedPrefetcher(loader):
            loss = step(x, t)          # TrainStep copies x, t to the device asynchronously
    """

    def __init__(self, loader, depth: int = 3, device=None):
        import queue
        import threading
        if depth < 2:
            raise ValueError("depth must be at least 2 (one slot in use, one being filled)")
        self.loader, self.depth = loader, depth
        self.cuda = torch.cuda.is_available()
        self.device = device
        self._queue_mod, self._threading = queue, threading
        self._slots = [{"bufs": None, "event": None} for _ in range(depth)]

    def __len__(self):
        return len(self.loader)

    def _stage(self, slot, batch):
        """Copies the fields of ``batch`` into the slot's buffers; returns views of the right size."""
        fields = list(batch) if isinstance(batch, (tuple, list)) else [batch]
        if slot["bufs"] is None:
            slot["bufs"] = [None] * len(fields)
        views = []
        for i, f in enumerate(fields):
            f = torch.as_tensor(f)
            buf = slot["bufs"][i]
            if buf is None or buf.dtype != f.dtype or buf.numel() < f.numel():
                buf = torch.empty(f.numel(), dtype=f.dtype, pin_memory=self.cuda)
                slot["bufs"][i] = buf
            view = buf[: f.numel()].view(f.shape)
            view.copy_(f)
            views.append(view)
        return tuple(views)

    def _worker(self, free_q, ready_q, stop):
        try:
            for batch in self.loader:
                idx = free_q.get()
                if stop.is_set():
                    return
                slot = self._slots[idx]
                if slot["event"] is not None:
                    slot["event"].synchronize()     # the consumer's copies out of this slot are done
                    slot["event"] = None
                ready_q.put((idx, self._stage(slot, batch)))
            ready_q.put((None, None))
        except BaseException as e:  # noqa: BLE001 - surfaces in the consumer thread
            ready_q.put((None, e))

    def __iter__(self):
        free_q, ready_q = self._queue_mod.Queue(), self._queue_mod.Queue()
        for i in range(self.depth):
            free_q.put(i)
        stop = self._threading.Event()
        t = self._threading.Thread(target=self._worker, args=(free_q, ready_q, stop), daemon=True)
        t.start()
        held = None
        try:
            while True:
                idx, item = ready_q.get()
                if held is not None:
                    # everything the consumer enqueued for the previous batch is in its stream now
                    if self.cuda:
                        ev = torch.cuda.Event()
                        ev.record(torch.cuda.current_stream(self.device))
                        self._slots[held]["event"] = ev
                    free_q.put(held)
                    held = None
                if idx is None:
                    if isinstance(item, BaseException):
                        raise item
                    return
                held = idx
                yield item
        finally:
            stop.set()
            free_q.put(0)       # wake the worker if it is waiting for a slot
