"""CPU unit tests: rendezvous spec parsing, sampler, checkpoints, fused optimizers, models."""
import copy
import os

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200 import parallel as par
from distributed_training_dl_b200.models import MLP, MnistCNN, pyramidnet, pyramid_widths


def test_slurm_nodelist():
    assert par.expand_slurm_nodelist("gpu[01-03,07],login1") == ["gpu01", "gpu02", "gpu03", "gpu07", "login1"]
    assert par.expand_slurm_nodelist("c1") == ["c1"]
    assert par.expand_slurm_nodelist("n[9-11]") == ["n9", "n10", "n11"]


def test_rendezvous_specs():
    s = par.resolve_spec("tcp://c1:20201", rank=3, world_size=4, env={})
    assert (s.master_addr, s.master_port, s.rank, s.world_size) == ("c1", 20201, 3, 4)
    with pytest.raises(ValueError):
        par.resolve_spec("tcp://c1:20201", env={})
    env = {"RANK": "5", "WORLD_SIZE": "8", "MASTER_ADDR": "10.0.0.1", "MASTER_PORT": "1234", "LOCAL_RANK": "1"}
    s = par.resolve_spec("env://", env=env)
    assert (s.rank, s.world_size, s.local_rank, s.init_method) == (5, 8, 1, "tcp://10.0.0.1:1234")
    assert par.resolve_spec(None, env=env).source == "env"
    senv = {"SLURM_PROCID": "9", "SLURM_NTASKS": "16", "SLURM_LOCALID": "1", "SLURM_JOB_NODELIST": "gpu[3-4]",
            "SLURM_JOB_ID": "77"}
    s = par.resolve_spec(None, env=senv)
    assert (s.source, s.rank, s.world_size, s.local_rank, s.master_addr) == ("slurm", 9, 16, 1, "gpu3")
    s = par.resolve_spec(worker_hosts="a:2222,a:2223,b:2222", task_index=2, env={})
    assert (s.rank, s.world_size, s.master_addr, s.master_port, s.local_rank) == (2, 3, "a", 2222, 0)
    assert par.resolve_spec(None, env={}).world_size == 1


def test_sampler_matches_torch():
    from torch.utils.data.distributed import DistributedSampler as TorchSampler
    ds = list(range(103))
    for world in (1, 2, 4):
        for rank in range(world):
            for shuffle in (False, True):
                a = dtdl.DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=3)
                b = TorchSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=3)
                a.set_epoch(2)
                b.set_epoch(2)
                assert list(a) == list(b) and len(a) == len(b)


def test_pyramidnet_structure():
    w = pyramid_widths()
    assert w[0] == 21 and w[-1] == 271 and len(w) == 51
    m = pyramidnet()
    assert sum(p.numel() for p in m.parameters()) == 24253410
    assert len(list(m.parameters())) == 415 and len(list(m.buffers())) == 465
    assert sum(p.numel() for p in MnistCNN().parameters()) == 93322
    assert sum(p.numel() for p in MLP().parameters()) == 1796010


def test_pyramidnet_matches_reference_definition():
    """Same keys / shapes as the reference's model.py and identical output for its weights."""
    import importlib.util
    path = None
    for cand in ("baseline/_ref/pytorch/model.py", "/root/reference/pytorch/model.py"):
        if os.path.exists(cand):
            path = cand
            break
    if path is None:
        pytest.skip("reference model.py not available")
    spec = importlib.util.spec_from_file_location("ref_model", path)
    ref_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_mod)
    torch.manual_seed(0)
    ref = ref_mod.pyramidnet()
    mine = pyramidnet()
    assert list(ref.state_dict().keys()) == list(mine.state_dict().keys())
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(3, 3, 32, 32)
    ref.train()
    mine.train()
    assert torch.allclose(ref(x), mine(x), atol=1e-5)
    ref.eval()
    mine.eval()
    assert torch.allclose(ref(x), mine(x), atol=1e-5)


def _tiny():
    return nn.Sequential(nn.Linear(12, 16), nn.ReLU(), nn.Linear(16, 5))


@pytest.mark.parametrize("kind", ["sgd", "sgd_damp", "adam", "adam_wd"])
def test_fused_optimizers_match_torch(kind):
    torch.manual_seed(0)
    a, b = _tiny(), None
    b = copy.deepcopy(a)
    if kind.startswith("sgd"):
        kw = dict(lr=0.1, momentum=0.9, weight_decay=1e-4, dampening=0.1 if kind == "sgd_damp" else 0.0)
        oa, ob = dtdl.FusedSGD(a.parameters(), **kw), torch.optim.SGD(b.parameters(), **kw)
    else:
        kw = dict(lr=1e-2, weight_decay=1e-2 if kind == "adam_wd" else 0.0)
        oa, ob = dtdl.FusedAdam(a.parameters(), **kw), torch.optim.Adam(b.parameters(), **kw)
    sched_a = torch.optim.lr_scheduler.StepLR(oa, step_size=2, gamma=0.1)
    sched_b = torch.optim.lr_scheduler.StepLR(ob, step_size=2, gamma=0.1)
    for step in range(5):
        x, t = torch.randn(8, 12), torch.randint(0, 5, (8,))
        for m, o, s in ((a, oa, sched_a), (b, ob, sched_b)):
            loss = F.cross_entropy(m(x), t)
            o.zero_grad()
            loss.backward()
            o.step()
            s.step()
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa, pb, atol=1e-6), kind
    sa, sb = oa.state_dict(), ob.state_dict()
    key = "momentum_buffer" if kind.startswith("sgd") else "exp_avg"
    assert set(sa["state"].keys()) == set(sb["state"].keys())
    for i in sb["state"]:
        assert torch.allclose(sa["state"][i][key], sb["state"][i][key], atol=1e-6)
    # resume: a fresh optimizer continues identically from the state dict
    c = copy.deepcopy(a)
    oc = type(oa)(c.parameters(), **kw)
    oc.load_state_dict(sa)
    x, t = torch.randn(8, 12), torch.randint(0, 5, (8,))
    for m, o in ((a, oa), (c, oc)):
        loss = F.cross_entropy(m(x), t)
        o.zero_grad()
        loss.backward()
        o.step()
    for pa, pc in zip(a.parameters(), c.parameters()):
        assert torch.allclose(pa, pc, atol=1e-7)


def test_keras_adam_first_step():
    p = nn.Parameter(torch.tensor([1.0, -2.0]))
    o = dtdl.FusedAdam([p], lr=1e-3, keras=True)
    p.grad = torch.tensor([0.5, -0.25])
    o.step()
    # first Adam step moves every coordinate by ~lr against the gradient sign
    assert torch.allclose(p.detach(), torch.tensor([1.0 - 1e-3, -2.0 + 1e-3]), atol=1e-6)


def test_wrap_single_process_cpu(tmp_path):
    """world_size 1: wrapper keys, fused step == torch SGD, checkpoint round trips."""
    torch.manual_seed(0)
    base = MLP(n_units=32)
    ref = copy.deepcopy(base)
    net = dtdl.wrap(base)
    assert list(net.state_dict().keys()) == ["module." + k for k in ref.state_dict().keys()]
    assert list(net.module.state_dict().keys()) == list(ref.state_dict().keys())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    for _ in range(3):
        x, t = torch.randn(16, 784), torch.randint(0, 10, (16,))
        out = net(x)
        loss = F.cross_entropy(out, t)
        opt.zero_grad()          # between forward and backward, as the reference does (R11)
        loss.backward()
        opt.step()
        rl = F.cross_entropy(ref(x), t)
        ropt.zero_grad()
        rl.backward()
        ropt.step()
    for a, b in zip(net.module.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    # checkpoints: wrapper spelling and bare spelling load into either kind of model
    p1 = dtdl.save_model(net, str(tmp_path / "a" / "wrapped.pth"))
    p2 = dtdl.save_model(net.module, str(tmp_path / "bare.pth"))
    fresh = MLP(n_units=32)
    dtdl.load_model(fresh, p1)
    dtdl.load_model(net, p2)
    for a, b in zip(fresh.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    # full snapshot resume
    sched = torch.optim.lr_scheduler.StepLR(opt, 2, 0.1)
    snap = dtdl.save_snapshot(str(tmp_path / "snap.pt"), net, opt, sched, epoch=3)
    net2 = dtdl.wrap(MLP(n_units=32))
    opt2 = dtdl.FusedSGD(net2.parameters(), lr=0.5)
    sched2 = torch.optim.lr_scheduler.StepLR(opt2, 2, 0.1)
    info = dtdl.load_snapshot(snap, net2, opt2, sched2)
    assert info["epoch"] == 3 and opt2.param_groups[0]["momentum"] == 0.9
    for a, b in zip(net2.module.parameters(), net.module.parameters()):
        assert torch.equal(a, b)


def test_wrap_with_torch_optimizer_cpu():
    torch.manual_seed(0)
    base = MnistCNN()
    ref = copy.deepcopy(base)
    net = dtdl.wrap(base)
    opt = torch.optim.Adam(net.parameters())
    ropt = torch.optim.Adam(ref.parameters())
    for _ in range(2):
        x, t = torch.randn(4, 1, 28, 28), torch.randint(0, 10, (4,))
        loss = F.cross_entropy(net(x), t)
        opt.zero_grad()
        loss.backward()
        opt.step()
        rl = F.cross_entropy(ref(x), t)
        ropt.zero_grad()
        rl.backward()
        ropt.step()
    for a, b in zip(net.module.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6)


def test_native_library_resolves_every_symbol():
    """dlopen with RTLD_NOW: an unresolved symbol in libdtdl_b200.so must fail here, on the build
    machine, not on the GPU box; and every prototype of the header must exist in the library."""
    import ctypes
    from distributed_training_dl_b200 import _native as nv
    if not nv.available():
        pytest.skip("library not built")
    handle = ctypes.CDLL(nv.LIB_PATH, mode=ctypes.RTLD_GLOBAL | 2)   # 2 = RTLD_NOW
    for name in nv._parse_header(nv.HEADER_PATH):
        assert hasattr(handle, name), name


def _bare_peer(world, use_nvls=False, nvls_min=0):
    """A PeerComm without a GPU behind it: enough state for the pure policy methods."""
    from distributed_training_dl_b200.parallel.peer import PeerComm
    pc = object.__new__(PeerComm)
    pc.rank, pc.world, pc.use_nvls, pc.nvls_min_bytes = 0, world, use_nvls, nvls_min
    pc.two_shot_min_bytes = 256 * 1024
    pc._local_group = None
    return pc


def test_collective_variant_policy(monkeypatch):
    """one-shot below 256 KiB, two-shot above; multicast (2) for every size once NVLS is on;
    the auto policy turns NVLS on from 4 ranks and the forced mode refuses a box without it."""
    from distributed_training_dl_b200 import _native as nv
    pc = _bare_peer(2)
    assert pc.choose_two_shot(1024) == 0 and pc.choose_two_shot(1 << 20) == 1
    assert _bare_peer(1).choose_two_shot(1 << 30) == 0          # a single rank never communicates
    pn = _bare_peer(8, use_nvls=True)
    assert pn.choose_two_shot(1024) == 2 and pn.choose_two_shot(1 << 28) == 2
    pm = _bare_peer(8, use_nvls=True, nvls_min=1 << 20)
    assert pm.choose_two_shot(1024) == 0 and pm.choose_two_shot(1 << 19) == 1 and pm.choose_two_shot(1 << 21) == 2

    monkeypatch.delenv("DTDL_NVLS", raising=False)
    for world, ok, want in [(2, True, False), (4, True, True), (8, True, True), (8, False, False)]:
        p = _bare_peer(world)
        p._decide_nvls(ok)
        assert p.use_nvls is want, (world, ok)
    monkeypatch.setenv("DTDL_NVLS", "1")
    p = _bare_peer(2)
    p._decide_nvls(True)
    assert p.use_nvls is True                                     # forced on below the auto threshold
    with pytest.raises(nv.NativeError):
        _bare_peer(2)._decide_nvls(False)
    monkeypatch.setenv("DTDL_NVLS", "0")
    p = _bare_peer(8)
    p._decide_nvls(True)
    assert p.use_nvls is False


def test_bucket_report_formatting():
    from distributed_training_dl_b200.utils import format_bucket_report
    rows = [{"bucket": 0, "bytes": 1 << 20, "ms": 0.02, "bus_GBps": 0.0, "kind": "fused nvls", "world": 8},
            {"bucket": 1, "bytes": 25 << 20, "ms": 0.10, "bus_GBps": 0.0, "kind": "fused nvls", "world": 8},
            {"bucket": 0, "bytes": 1 << 20, "ms": 0.04, "bus_GBps": 0.0, "kind": "fused nvls", "world": 8}]
    text = format_bucket_report(rows).splitlines()
    assert len(text) == 3 and "bus GB/s" in text[0]
    # bucket 0: mean of 0.02 and 0.04 ms -> 2*(7/8)*1 MiB / 0.03 ms = 61.2 GB/s
    assert abs(float(text[1].split("|")[-1]) - 2 * 7 / 8 * (1 << 20) / 0.03e-3 / 1e9) < 0.1


def test_pinned_prefetcher_order_reuse_and_errors():
    """Batches arrive in order and intact, the staging buffers are a fixed ring, a smaller last
    batch fits the same buffers, loader errors surface in the consumer, early exit does not hang."""
    from distributed_training_dl_b200.utils import PinnedPrefetcher
    g = torch.Generator().manual_seed(0)
    batches = [(torch.randn(8 if i < 9 else 3, 4, generator=g), torch.arange(8 if i < 9 else 3) + i) for i in range(10)]
    ptrs, got = set(), []
    for x, t in PinnedPrefetcher(batches, depth=3):
        ptrs.add(x.data_ptr())
        got.append((x.clone(), t.clone()))
    assert len(got) == 10 and len(ptrs) <= 3
    for (x, t), (rx, rt) in zip(got, batches):
        assert torch.equal(x, rx) and torch.equal(t, rt)
    assert len(PinnedPrefetcher(batches)) == 10

    def broken():
        yield batches[0]
        raise RuntimeError("disk on fire")

    it = iter(PinnedPrefetcher(broken(), depth=2))
    next(it)
    with pytest.raises(RuntimeError, match="disk on fire"):
        next(it)
    for i, _ in enumerate(PinnedPrefetcher(batches, depth=2)):     # leaving early must not block
        if i == 2:
            break
    with pytest.raises(ValueError):
        PinnedPrefetcher(batches, depth=1)


def _write_idx(path, arr):
    import gzip
    import struct
    with gzip.open(path, "wb") as f:
        f.write(struct.pack(">HBB", 0, 0x08, arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape))
        f.write(arr.astype("uint8").tobytes())


def test_mnist_idx_files_npz_cache_and_preprocessing(tmp_path):
    """The reference's chainer/mnist_dataset.py + mnist_helper.py: four idx-gz files -> mnist.npz
    cache -> (float32 image scaled to [0, scale], int label) pairs, flat or 1x28x28."""
    import numpy as np
    from distributed_training_dl_b200.utils import mnist
    rng = np.random.default_rng(0)
    xtr, ytr = rng.integers(0, 256, (20, 28, 28)), rng.integers(0, 10, 20)
    xte, yte = rng.integers(0, 256, (7, 28, 28)), rng.integers(0, 10, 7)
    for key, arr in (("train_x", xtr), ("train_y", ytr), ("test_x", xte), ("test_y", yte)):
        _write_idx(str(tmp_path / mnist.FILES[key]), arr)
    assert np.array_equal(mnist.read_idx(str(tmp_path / mnist.FILES["train_x"])), xtr.astype("uint8"))
    (a, b), (c, d) = mnist.load_mnist(str(tmp_path), synthetic=False)
    assert (tmp_path / "mnist.npz").exists()                       # cache written on first load
    assert a.shape == (20, 28, 28) and np.array_equal(b, ytr) and np.array_equal(c, xte) and len(d) == 7
    for key in mnist.FILES:                                        # second load needs only the cache
        (tmp_path / mnist.FILES[key]).unlink()
    train, test = mnist.get_mnist(str(tmp_path), ndim=1, scale=1.0, synthetic=False)
    x0, y0 = train[0]
    assert x0.shape == (784,) and x0.dtype == np.float32 and 0.0 <= x0.min() and x0.max() <= 1.0
    assert np.allclose(x0, xtr[0].reshape(-1) / 255.0) and int(y0) == int(ytr[0]) and len(test) == 7
    train3, _ = mnist.get_mnist(str(tmp_path), ndim=3, scale=255.0, synthetic=False)
    assert train3[0][0].shape == (1, 28, 28) and np.allclose(train3[0][0][0], xtr[0])
    with pytest.raises(FileNotFoundError):
        mnist.load_mnist(str(tmp_path / "nothing_here"), synthetic=False)
    (st, _), _ = mnist.load_mnist(str(tmp_path / "nothing_here"), synthetic=True)   # no network: random data
    assert st.shape[1:] == (28, 28)
    bad = tmp_path / "bad.gz"
    import gzip
    with gzip.open(bad, "wb") as f:
        f.write(b"\\x00\\x00\\x0d\\x01" + b"\\x00" * 8)
    with pytest.raises(ValueError):
        mnist.read_idx(str(bad))


def test_train_meter_log_line_and_rank_prefix(capsys):
    """The reference's only observable output (pytorch/single_gpu.py:114-116 and the
    'From Rank: r,' prefix of pytorch/distributed_data_parallel.py:144-148)."""
    from distributed_training_dl_b200.utils import RankLogger, TrainMeter
    m = TrainMeter("cpu")
    logits = torch.tensor([[2.0, 0.0], [0.0, 3.0], [1.0, 0.0], [0.0, 1.0]])
    targets = torch.tensor([0, 1, 1, 1])
    m.update(torch.tensor(0.5), outputs=logits, targets=targets)
    m.update(torch.tensor(1.5), targets=targets, correct=torch.tensor(4.0))
    loss, acc = m.summary()
    assert abs(loss - 1.0) < 1e-6 and abs(acc - 100.0 * 7 / 8) < 1e-4
    line = m.line(3, 20, 782, 0.0123)
    assert line.startswith("Epoch:[3][20/782]| loss: 1.000 | acc: 87.500 | batch time: 0.012s")
    RankLogger().info("hello")
    assert "hello" in capsys.readouterr().out


def test_bucket_plan_keeps_the_last_bucket_small():
    """The final bucket (earliest layers, ready only when backward ends) is capped at 4 MiB; buckets
    tile the arena without gaps and every offset is 32-byte aligned."""
    from distributed_training_dl_b200.models import pyramidnet
    from distributed_training_dl_b200.parallel.arena import ParamArena
    arena = ParamArena(pyramidnet().parameters(), "cpu")
    sizes = [b.numel * 4 for b in arena.buckets]
    assert sizes[-1] <= 4 << 20 and len(sizes) >= 4
    off = 0
    for b in arena.buckets:
        assert b.offset == off and b.offset % 8 == 0 and b.numel % 8 == 0
        off += b.numel
    assert off == arena.numel
    assert sorted(id(p) for b in arena.buckets for p in b.params) == sorted(id(p) for p in arena.params)


def test_set_precision_marks_every_layer():
    import torch
    from distributed_training_dl_b200.models import pyramidnet, resnet50
    from distributed_training_dl_b200.ops import InputLayout, set_precision
    from distributed_training_dl_b200.ops.modules import _PackedWeightMixin
    for make in (pyramidnet, resnet50):
        m = set_precision(make(), "tf32")
        layers = [x for x in m.modules() if isinstance(x, (_PackedWeightMixin, InputLayout))]
        assert layers and all(x.act_dtype == torch.float32 for x in layers)
        ops = [x._op for x in m.modules() if isinstance(getattr(x, "_op", None), _PackedWeightMixin)]
        assert all(o.act_dtype == torch.float32 for o in ops)
        set_precision(m, "bf16")
        assert all(x.act_dtype == torch.bfloat16 for x in layers)
    import pytest
    with pytest.raises(ValueError):
        set_precision(pyramidnet(), "fp8")


def test_data_parallel_shards_are_balanced():
    from distributed_training_dl_b200.parallel.data_parallel import DataParallel
    dp = DataParallel.__new__(DataParallel)       # only the arithmetic: no GPUs needed
    dp.world = 8
    assert dp.shard_bounds(41) == [(0, 6), (6, 11), (11, 16), (16, 21), (21, 26), (26, 31), (31, 36), (36, 41)]
    assert dp.shard_bounds(64)[-1] == (56, 64)
    import pytest
    with pytest.raises(ValueError):
        dp.shard_bounds(7)


def test_collective_variant_table(tmp_path, monkeypatch):
    """Variant choice per message size comes from a table (measured sweeps), overridable from a file."""
    import json
    from distributed_training_dl_b200.parallel.peer import PeerComm
    pc = PeerComm.__new__(PeerComm)
    pc.two_shot_min_bytes, pc.nvls_min_bytes = 256 * 1024, 0
    PeerComm._variant_table = None
    pc.world, pc.use_nvls = 2, False
    assert [pc.choose_two_shot(n) for n in (1024, 256 << 10, (256 << 10) + 16, 64 << 20)] == [0, 0, 1, 1]
    pc.world, pc.use_nvls = 8, True
    assert pc.choose_two_shot(1024) == 2 and pc.choose_two_shot(1 << 30) == 2
    pc.use_nvls = False                      # multicast rows fall back to the peer-pointer variants
    assert pc.choose_two_shot(1024) == 0 and pc.choose_two_shot(1 << 24) == 1
    pc.world = 1
    assert pc.choose_two_shot(1 << 24) == 0
    path = tmp_path / "table.json"
    path.write_text(json.dumps({"8": [[65536, 0], [None, 2]]}))
    monkeypatch.setenv("DTDL_VARIANT_TABLE", str(path))
    PeerComm._variant_table = None
    pc.world, pc.use_nvls = 8, True
    assert pc.choose_two_shot(4096) == 0 and pc.choose_two_shot(1 << 20) == 2
    PeerComm._variant_table = None


def test_train_step_returns_a_detached_loss():
    """The step has already run backward; a caller that keeps the loss must not keep the graph."""
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    torch.manual_seed(0)
    net = MLP(32)
    opt = dtdl.FusedSGD(net.parameters(), lr=0.1)
    step = TrainStep(net, opt, CrossEntropyLoss())
    loss = step(torch.randn(8, 784), torch.randint(0, 10, (8,)))
    assert loss.grad_fn is None and not loss.requires_grad and loss.item() > 0
