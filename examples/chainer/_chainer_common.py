"""Shared pieces of the Chainer-style MNIST examples (flags of chainer/train_mnist.py:30-47)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_training_dl_b200 import trainer as ch  # noqa: E402,F401
from distributed_training_dl_b200.models import MLP  # noqa: E402,F401
from distributed_training_dl_b200.utils.mnist import get_mnist  # noqa: E402,F401


def flags(description):
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--batchsize", "-b", type=int, default=100)
    p.add_argument("--epoch", "-e", type=int, default=20)
    p.add_argument("--frequency", "-f", type=int, default=-1, help="snapshot every N epochs")
    p.add_argument("--out", "-o", default="result")
    p.add_argument("--resume", "-r", default="")
    p.add_argument("--unit", "-u", type=int, default=1000)
    p.add_argument("--dataset-dir", default="./data/mnist")
    return p


def add_reporting(trainer, args, rank0=True):
    if rank0:
        trainer.extend(ch.dump_graph("main/loss"))
        trainer.extend(ch.LogReport())
        trainer.extend(ch.PrintReport(["epoch", "main/loss", "validation/main/loss", "main/accuracy",
                                       "validation/main/accuracy", "elapsed_time"]))
    frequency = args.epoch if args.frequency == -1 else max(1, args.frequency)
    trainer.extend(ch.snapshot(), trigger=(frequency, "epoch"))
    if args.resume:
        ch.load_npz(args.resume, trainer)
