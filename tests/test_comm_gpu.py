"""Peer-memory collectives and DDP-style training on >= 2 GPUs (skipped on a 1-GPU box), plus the
single-replica variants of the single-process engine and front-end checks that a 1-GPU box can run."""
import os

import pytest
import torch

import distributed_training_dl_b200 as dtdl

from _gpu_workers import bf16_gradient_transport, collectives, collectives_sweep, ddp_training, ddp_variants

pytestmark = pytest.mark.gpu


def _world():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    return min(n, int(os.environ.get("DTDL_TEST_WORLD", "2")))


def test_peer_collectives(tmp_path):
    world = _world()
    dtdl.spawn(collectives, nprocs=world, args=(world, str(tmp_path)))
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def _done_or_skipped(tmp_path, world):
    if all(os.path.exists(tmp_path / f"skip_{i}") for i in range(world)):
        pytest.skip("this box has no NVLS multicast support")
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def test_nvls_collectives(tmp_path):
    """multimem.ld_reduce / multimem.st all-reduce, fused reduce+update and broadcast over
    cuMem symmetric buffers bound to an NVSwitch multicast object."""
    world = _world()
    dtdl.spawn(collectives, nprocs=world, args=(world, str(tmp_path), True))
    _done_or_skipped(tmp_path, world)


@pytest.mark.parametrize("nvls", [False, True])
def test_collectives_sweep(tmp_path, nvls):
    """Every variant x 1 KiB .. 64 MiB (1 GiB with DTDL_TEST_MAX_BYTES) x 10 repetitions, bitwise exact."""
    world = _world()
    dtdl.spawn(collectives_sweep, nprocs=world, args=(world, str(tmp_path), nvls))
    if nvls:
        _done_or_skipped(tmp_path, world)
    else:
        assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


@pytest.mark.parametrize("nvls", [False, True])
def test_bf16_gradient_transport(tmp_path, nvls):
    world = _world()
    dtdl.spawn(bf16_gradient_transport, nprocs=world, args=(world, str(tmp_path), nvls))
    if nvls:
        _done_or_skipped(tmp_path, world)
    else:
        assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def test_nvls_ddp_training(tmp_path):
    world = _world()
    dtdl.spawn(ddp_training, nprocs=world, args=(world, str(tmp_path), True, True))
    _done_or_skipped(tmp_path, world)


@pytest.mark.parametrize("use_graph", [False, True])
def test_ddp_training(tmp_path, use_graph):
    world = _world()
    dtdl.spawn(ddp_training, nprocs=world, args=(world, str(tmp_path), use_graph))
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


@pytest.mark.parametrize("nvls", ["0", "1"])
def test_data_parallel_engine(tmp_path, nvls):
    """nvls=1: the single-process flavour of the cuMem + multicast allocator (handles shared
    directly, no descriptor exchange)."""
    _world()
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = f"import sys; sys.path.insert(0, {here!r}); import _gpu_workers as w; w.data_parallel_engine({str(tmp_path)!r})"
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                         env=dict(os.environ, DTDL_NVLS=nvls))
    if nvls == "1" and "cannot join an NVLS multicast team" in res.stderr:
        pytest.skip("this box has no NVLS multicast support")
    assert res.returncode == 0 and os.path.exists(tmp_path / "ok_dp"), res.stdout[-2000:] + res.stderr[-4000:]


def test_data_parallel_engine_one_gpu(tmp_path):
    """The same engine checks with ONE visible GPU (a single replica): runs on a 1-GPU box, where the
    multi-GPU variants above are skipped -- eager call contract, live-graph refusal, graph fast path,
    equivalence with torch SGD, optimizer state format."""
    if torch.cuda.device_count() < 1:
        pytest.skip("needs a GPU")
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = f"import sys; sys.path.insert(0, {here!r}); import _gpu_workers as w; w.data_parallel_engine({str(tmp_path)!r})"
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                         env=dict(os.environ, CUDA_VISIBLE_DEVICES="0", DTDL_NVLS="0"))
    assert res.returncode == 0 and os.path.exists(tmp_path / "ok_dp"), res.stdout[-2000:] + res.stderr[-4000:]


def test_frontends_one_gpu(tmp_path):
    """The Keras- and Chainer-shaped front ends on ONE visible GPU (what a 1-GPU box can run of
    ``test_frontends_on_all_gpus``): MirroredStrategy.fit with the fused Keras-Adam, ModelCheckpoint,
    load_weights into a model whose step graph is already captured, ParallelUpdater in a Trainer."""
    if torch.cuda.device_count() < 1:
        pytest.skip("needs a GPU")
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = f"import sys; sys.path.insert(0, {here!r}); import _gpu_workers as w; w.frontends_multi_gpu({str(tmp_path)!r})"
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                         env=dict(os.environ, CUDA_VISIBLE_DEVICES="0", DTDL_NVLS="0"))
    assert res.returncode == 0 and os.path.exists(tmp_path / "ok_fe"), res.stdout[-2000:] + res.stderr[-4000:]


@pytest.mark.parametrize("backend", ["b200", "nccl"])
def test_ddp_backends_optimizers_no_sync(tmp_path, backend):
    world = _world()
    dtdl.spawn(ddp_variants, nprocs=world, args=(world, str(tmp_path), backend))
    assert all(os.path.exists(tmp_path / f"ok_{i}") for i in range(world))


def test_frontends_on_all_gpus(tmp_path):
    """MirroredStrategy.fit, load_weights replica re-sync and ParallelUpdater on >= 2 GPUs."""
    _world()
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    code = f"import sys; sys.path.insert(0, {here!r}); import _gpu_workers as w; w.frontends_multi_gpu({str(tmp_path)!r})"
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(min(torch.cuda.device_count(), 2)))
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900, env=env)
    assert res.returncode == 0 and os.path.exists(tmp_path / "ok_fe"), res.stdout[-2000:] + res.stderr[-4000:]
