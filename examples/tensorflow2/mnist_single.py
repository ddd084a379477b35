#!/usr/bin/env python
"""Keras-style MNIST CNN on one device (the job of tensorflow2/mnist_single.py)."""
from _tf2_common import flags, ks, train_and_eval

if __name__ == "__main__":
    train_and_eval(flags("MNIST CNN, single device").parse_args(), ks.OneDeviceStrategy())
