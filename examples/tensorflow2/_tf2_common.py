"""Shared pieces of the Keras-style MNIST examples (flags of tensorflow2/mnist_single.py:95-115)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_training_dl_b200 import strategy as ks  # noqa: E402,F401
from distributed_training_dl_b200.models import MnistCNN  # noqa: E402,F401
from distributed_training_dl_b200.utils.mnist import load_mnist  # noqa: E402


def flags(description, multi_worker=False):
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--train_dir", "-td", default="./train_dir")
    p.add_argument("--dataset_dir", "-dd", default="./data/mnist")
    p.add_argument("--batch_size", "-b", type=int, default=64)
    p.add_argument("--test_batchsize", "-tb", type=int, default=1000)
    p.add_argument("--epochs", "-e", type=int, default=10)
    p.add_argument("--gpu_nums", "-g", type=int, default=0)
    p.add_argument("--cpu_nums", "-c", type=int, default=0)
    p.add_argument("--learning_rate", "-lr", type=float, default=None, help="Adam's default when unset")
    p.add_argument("--log_interval", type=int, default=10)
    p.add_argument("--save_model", "-sm", action="store_true")
    if multi_worker:
        p.add_argument("--worker_hosts", default=None, help="comma separated host:port list")
        p.add_argument("--job_name", default="worker")
        p.add_argument("--task_index", type=int, default=0)
    return p


def dataset(args):
    (xt, yt), (xv, yv) = load_mnist(args.dataset_dir)
    prep = lambda x: (x.astype(np.float32) / 255.0).reshape(len(x), 1, 28, 28)  # noqa: E731
    return (prep(xt), yt.astype(np.int64)), (prep(xv), yv.astype(np.int64))


def train_and_eval(args, strategy):
    (xt, yt), (xv, yv) = dataset(args)
    ckpt = os.path.join(args.train_dir, "ckpt_{epoch}")
    with strategy.scope():
        model = ks.Model(MnistCNN())
        model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"],
                      learning_rate=args.learning_rate)
    model.summary()
    model.fit(xt, yt, batch_size=args.batch_size, epochs=args.epochs,
              callbacks=[ks.TensorBoard(log_dir=args.train_dir, histogram_freq=1),
                         ks.ModelCheckpoint(filepath=ckpt, save_weights_only=True), ks.PrintLR()],
              validation_data=(xv, yv))
    latest = ks.latest_checkpoint(args.train_dir)
    if latest:
        model.load_weights(latest)
    loss, acc = model.evaluate(xv, yv, batch_size=args.test_batchsize, verbose=2)
    print("Eval loss: {}, Eval Accuracy: {}".format(loss, acc))
