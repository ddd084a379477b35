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
