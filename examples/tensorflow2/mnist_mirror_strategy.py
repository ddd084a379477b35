#!/usr/bin/env python
"""Keras-style MNIST CNN mirrored over all GPUs of one process
(the job of tensorflow2/mnist_mirror_strategy.py)."""
from _tf2_common import flags, ks, train_and_eval

if __name__ == "__main__":
    args = flags("MNIST CNN, MirroredStrategy").parse_args()
    devices = [f"cuda:{i}" for i in range(args.gpu_nums)] if args.gpu_nums else None
    train_and_eval(args, ks.MirroredStrategy(devices))
