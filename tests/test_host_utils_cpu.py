"""Host-side pieces that the larger tests only touch in passing: iterators, the Chainer-shaped
optimizers, logging helpers, checkpoint key spellings, the communicator's object collectives at
world size 1, a stock torch optimizer behind ``wait_gradients()``.  All on CPU."""
import copy
import json

import torch
import torch.nn.functional as F

import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200 import trainer as ch
from distributed_training_dl_b200.models import MLP
from distributed_training_dl_b200.utils import checkpoint as ck
from distributed_training_dl_b200.utils import logging as lg


def _toy(n=40, seed=0):
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    x = torch.randn(n, 784, generator=g) * 0.1 + (y.float().view(-1, 1) - 4.5) / 3
    return list(zip(x, y))


def test_serial_iterator_epochs_detail_and_reset():
    data = _toy(10)
    it = ch.SerialIterator(data, 4, repeat=True, shuffle=False)
    sizes, new_epoch, detail = [], [], []
    for _ in range(4):
        xs, ts = next(it)
        sizes.append(len(ts))
        new_epoch.append(it.is_new_epoch)
        detail.append(round(it.epoch_detail, 3))
    assert sizes == [4, 4, 2, 4]                      # the short tail batch closes the epoch
    assert new_epoch == [False, False, True, False]
    assert detail == [0.4, 0.8, 1.0, 1.4] and it.epoch == 1
    it.reset()
    assert (it.epoch, it.pos, it.is_new_epoch) == (0, 0, False)
    once = ch.SerialIterator(data, 4, repeat=False, shuffle=False)
    assert [len(t) for _, t in once] == [4, 4, 2]    # stops after one pass
    a = ch.SerialIterator(data, 10, shuffle=True, seed=3)
    b = ch.SerialIterator(data, 10, shuffle=True, seed=3)
    assert torch.equal(next(a)[1], next(b)[1])       # the order is a function of the seed


def test_momentum_sgd_trainer_matches_torch_sgd(tmp_path):
    """Chainer's two-step construction (``MomentumSGD().setup(model)``) + StandardUpdater on CPU:
    the same numbers as torch.optim.SGD over the same batches."""
    data = _toy(32)
    torch.manual_seed(0)
    clf = ch.Classifier(MLP(16, 10))
    ref = copy.deepcopy(clf.predictor)
    opt = ch.MomentumSGD(lr=0.05, momentum=0.9).setup(clf)
    updater = ch.StandardUpdater(ch.SerialIterator(data, 8, shuffle=False), opt, device=-1)
    tr = ch.Trainer(updater, (1, "epoch"), out=str(tmp_path / "out"))
    tr.extend(ch.LogReport())
    tr.run()
    assert updater.iteration == 4 and updater.is_new_epoch
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    it = ch.SerialIterator(data, 8, shuffle=False)
    for _ in range(4):
        x, t = next(it)
        ropt.zero_grad()
        F.cross_entropy(ref(x), t).backward()
        ropt.step()
    for p, q in zip(clf.predictor.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5)
    log = json.load(open(tmp_path / "out" / "log"))
    assert log[-1]["epoch"] == 1 and log[-1]["iteration"] == 4 and "main/loss" in log[-1]


def test_logging_helpers(tmp_path):
    assert lg.get_rank() == 0
    calls = []
    lg.rank0_only(lambda v: calls.append(v))(7)
    assert calls == [7]
    assert lg.format_elapsed(3725.5) == "1:02:05.500000"
    path = tmp_path / "log.json"
    jl = lg.JsonLog(str(path))
    jl.append({"epoch": 1, "loss": 2.0})
    jl.append({"epoch": 2, "loss": 1.0})
    assert [e["epoch"] for e in json.load(open(path))] == [1, 2]
    assert lg.JsonLog().append({"a": 1}) is None      # no path: memory only
    meter = lg.TrainMeter()
    meter.update(torch.tensor(2.0), correct=torch.tensor(3.0), targets=torch.zeros(8))
    assert meter.summary() == (2.0, 37.5)
    meter.reset()
    assert (meter.steps, meter.total, meter.summary()) == (0, 0, (0.0, 0.0))


def test_module_prefix_spellings_round_trip():
    sd = MLP(8, 10).state_dict()
    wrapped = ck.add_module_prefix(sd)
    assert list(wrapped) == ["module." + k for k in sd]
    assert list(ck.add_module_prefix(wrapped)) == list(wrapped)         # idempotent
    bare = ck.strip_module_prefix(wrapped)
    assert list(bare) == list(sd) and all(torch.equal(bare[k], sd[k]) for k in sd)
    assert list(ck.strip_module_prefix(sd)) == list(sd)


def test_communicator_object_collectives_and_torch_optimizer_single_process():
    comm = dtdl.init(backend="gloo")
    try:
        assert comm.size == 1 and comm.rank == 0
        assert comm.broadcast_object({"k": [1, 2]}, src=0) == {"k": [1, 2]}
        assert comm.allgather_object("x") == ["x"]
        assert abs(comm.allreduce_mean(3.5) - 3.5) < 1e-12
        torch.manual_seed(0)
        net = dtdl.wrap(MLP(8, 10))
        ref = copy.deepcopy(net.module)
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        ropt = torch.optim.SGD(ref.parameters(), lr=0.1)
        x, t = torch.randn(6, 784), torch.randint(0, 10, (6,))
        for _ in range(2):
            opt.zero_grad()
            F.cross_entropy(net(x), t).backward()
            net.wait_gradients()                      # stock optimizer: gradients are final after this
            opt.step()
            ropt.zero_grad()
            F.cross_entropy(ref(x), t).backward()
            ropt.step()
        for p, q in zip(net.module.parameters(), ref.parameters()):
            assert torch.allclose(p, q, atol=1e-6)
    finally:
        dtdl.shutdown()
