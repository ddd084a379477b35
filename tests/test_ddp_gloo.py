"""Multi-process plumbing on CPU (gloo, world_size 2): SURVEY.md section 4, test plan item 1."""
import os

import pytest

import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.utils import free_port

from _workers import ddp_equivalence, ddp_pyramidnet, fd_channel, multi_worker_fit, slurm_and_hosts_init


@pytest.mark.parametrize("init_kind", ["tcp", "env"])
def test_ddp_equivalence_gloo(tmp_path, init_kind):
    world = 2
    dtdl.spawn(ddp_equivalence, nprocs=world, args=(world, init_kind, free_port(), str(tmp_path)),
               set_env=(init_kind == "env"))
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def test_ddp_pyramidnet_gloo(tmp_path):
    world = 2
    dtdl.spawn(ddp_pyramidnet, nprocs=world, args=(world, free_port(), str(tmp_path)), set_env=False)
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


@pytest.mark.parametrize("world", [2, 4])
def test_fd_channel(tmp_path, world):
    """The descriptor all-gather that carries cuMem / multicast handles between ranks."""
    dtdl.spawn(fd_channel, nprocs=world, args=(world, free_port(), str(tmp_path)), set_env=False)
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


@pytest.mark.parametrize("kind", ["slurm", "hosts"])
def test_init_from_slurm_env_and_host_list(tmp_path, kind):
    world = 2
    dtdl.spawn(slurm_and_hosts_init, nprocs=world, args=(world, free_port(), str(tmp_path), kind), set_env=False)
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def test_multi_worker_mirrored_strategy_fit(tmp_path):
    """DP-d: Model.fit under MultiWorkerMirroredStrategy with two workers."""
    world = 2
    dtdl.spawn(multi_worker_fit, nprocs=world, args=(world, free_port(), str(tmp_path)), set_env=False)
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))
