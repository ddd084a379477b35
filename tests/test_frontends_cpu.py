"""Keras-shaped (strategy / fit) and Chainer-shaped (trainer / extensions) front ends on CPU."""
import os

import torch

from distributed_training_dl_b200 import strategy as ks
from distributed_training_dl_b200 import trainer as ch
from distributed_training_dl_b200.models import MLP, MnistCNN


def _separable(n, shape, seed=0):
    """Tiny learnable problem: the label is encoded in the mean of the image."""
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    x = torch.randn((n,) + shape, generator=g) * 0.1 + (y.float().view(-1, *[1] * len(shape)) - 4.5) / 3
    return x, y


def test_keras_style_fit_checkpoint_evaluate(tmp_path):
    x, y = _separable(256, (1, 28, 28))
    strategy = ks.MirroredStrategy(devices=["cpu"])
    assert strategy.num_replicas_in_sync == 1
    with strategy.scope():
        model = ks.Model(MnistCNN())
        model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    ckpt = os.path.join(tmp_path, "ckpt_{epoch}")
    tb = ks.TensorBoard(log_dir=os.path.join(tmp_path, "tb"), histogram_freq=1)
    hist = model.fit(x, y, batch_size=64, epochs=3, callbacks=[tb, ks.ModelCheckpoint(ckpt), ks.PrintLR()],
                     validation_data=(x[:64], y[:64]), verbose=0)
    tb.close()
    assert os.listdir(os.path.join(tmp_path, "tb")), "the TensorBoard callback wrote nothing"
    assert hist.history["loss"][-1] < hist.history["loss"][0]
    assert "val_accuracy" in hist.history
    latest = ks.latest_checkpoint(str(tmp_path))
    assert latest.endswith("ckpt_3")
    before = model.evaluate(x, y)
    with strategy.scope():
        fresh = ks.Model(MnistCNN()).compile()
    fresh.load_weights(latest)
    after = fresh.evaluate(x, y)
    assert abs(before[0] - after[0]) < 1e-5 and before[1] == after[1]
    assert fresh.predict(x[:5]).shape == (5, 10)


def test_chainer_style_trainer_snapshot_resume(tmp_path):
    xs, ys = _separable(200, (784,))
    data = list(zip(xs, ys))
    comm = ch.create_communicator("naive")
    assert (comm.rank, comm.size, comm.intra_rank) == (0, 1, 0)
    train = ch.scatter_dataset(data, comm, shuffle=True)

    def build(out):
        model = ch.Classifier(MLP(32, 10))
        opt = ch.create_multi_node_optimizer(ch.Adam(), comm).setup(model)
        updater = ch.StandardUpdater(ch.SerialIterator(train, 50), opt, device=-1)
        tr = ch.Trainer(updater, (2, "epoch"), out=out)
        ev = ch.create_multi_node_evaluator(
            ch.Evaluator(ch.SerialIterator(train, 100, repeat=False, shuffle=False), model), comm)
        tr.extend(ev)
        tr.extend(ch.dump_graph("main/loss"))
        tr.extend(ch.LogReport())
        tr.extend(ch.PrintReport(["epoch", "main/loss", "validation/main/loss", "main/accuracy",
                                  "validation/main/accuracy", "elapsed_time"]))
        tr.extend(ch.snapshot(filename="snap_{iteration}"), trigger=(1, "epoch"))
        return tr

    torch.manual_seed(0)
    t1 = build(str(tmp_path / "a"))
    t1.run()
    assert os.path.exists(tmp_path / "a" / "log") and os.path.exists(tmp_path / "a" / "cg.txt")
    assert t1.observation["epoch"] == 2 and "validation/main/accuracy" in t1.observation
    snaps = sorted(os.listdir(tmp_path / "a"))
    assert "snap_4" in snaps and "snap_8" in snaps
    # resume from the epoch-1 snapshot: the trainer continues at epoch 1 and stops at 2
    t2 = build(str(tmp_path / "b"))
    ch.load_npz(str(tmp_path / "a" / "snap_4"), t2)
    assert t2.updater.iteration == 4 and t2.updater.optimizer._step == 4
    t2.run()
    assert t2.updater.epoch == 2 and t2.updater.iteration == 8
