"""Process launchers.

``spawn(fn, nprocs, args)`` has the signature the reference calls
(``mp.spawn(main_worker, nprocs=ngpus_per_node, args=(ngpus_per_node, args))``,
pytorch/distributed_data_parallel.py:56) and additionally exports torchrun-style environment
variables to each worker so ``init()`` with no arguments works.  Fixes the reference's
"all spawned workers share one --rank" bug: worker ``i`` gets global rank
``node_rank * nprocs + i``.
"""
from __future__ import annotations

import os
import socket

import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(i, fn, args, nprocs, node_rank, nnodes, master_addr, master_port, set_env):
    if set_env:
        os.environ["LOCAL_RANK"] = str(i)
        os.environ["RANK"] = str(node_rank * nprocs + i)
        os.environ["WORLD_SIZE"] = str(nnodes * nprocs)
        os.environ["MASTER_ADDR"] = master_addr
        os.environ["MASTER_PORT"] = str(master_port)
    fn(i, *args)


def spawn(fn, nprocs=1, args=(), node_rank=0, nnodes=1, master_addr="127.0.0.1", master_port=None,
          join=True, set_env=True, start_method="spawn"):
    master_port = master_port or free_port()
    return mp.start_processes(_entry, args=(fn, args, nprocs, node_rank, nnodes, master_addr,
                                            master_port, set_env),
                              nprocs=nprocs, join=join, start_method=start_method)
