"""The two MNIST workloads of the reference on the framework's op layer.

* ``MnistCNN``: the Keras CNN used by all three TensorFlow 2 scripts
  (tensorflow2/mnist_single.py:16-26): Conv3x3(1->32, valid)+ReLU, MaxPool2, Conv3x3(32->64)
  +ReLU, MaxPool2, Conv3x3(64->64)+ReLU, Flatten, Dense 64+ReLU, Dense 10.  93,322 parameters.
  The reference applies softmax in the model and a probability-space cross entropy
  (:26,79); here the model returns logits and the loss is softmax cross entropy, which is
  the same function without the round trip through probabilities.
* ``MLP``: the Chainer multilayer perceptron (chainer/train_mnist.py:13-26;
  chainer/train_mnist_multi.py:17-28): 784 -> n_units -> n_units -> 10 with ReLU.
"""
from __future__ import annotations

import torch.nn as nn

from ..ops import modules as L


class MnistCNN(nn.Module):
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.input = L.InputLayout()
        self.conv1 = L.Conv2d(1, 32, 3, activation="relu", first_layer=True)
        self.pool1 = L.MaxPool2d(2)
        self.conv2 = L.Conv2d(32, 64, 3, activation="relu")
        self.pool2 = L.MaxPool2d(2)
        self.conv3 = L.Conv2d(64, 64, 3, activation="relu")
        self.flatten = L.Flatten(64)
        self.fc1 = L.Linear(3 * 3 * 64, 64, activation="relu")
        self.fc2 = L.Linear(64, num_classes)
        self.output = L.OutputLayout(num_classes)

    def forward(self, x):
        if x.dim() == 4 and x.shape[-1] == 1 and x.shape[1] != 1:
            x = x.permute(0, 3, 1, 2)  # accept Keras' NHWC [N,28,28,1] batches as well
        x = self.pool1(self.conv1(self.input(x)))
        x = self.pool2(self.conv2(x))
        x = self.flatten(self.conv3(x))
        return self.output(self.fc2(self.fc1(x)))


class MLP(nn.Module):
    def __init__(self, n_units: int = 1000, n_out: int = 10, n_in: int = 784):
        super().__init__()
        self.input = L.InputLayout()
        self.l1 = L.Linear(n_in, n_units, activation="relu", first_layer=True)
        self.l2 = L.Linear(n_units, n_units, activation="relu")
        self.l3 = L.Linear(n_units, n_out)
        self.output = L.OutputLayout(n_out)

    def forward(self, x):
        x = x.reshape(x.shape[0], -1)
        return self.output(self.l3(self.l2(self.l1(self.input(x)))))
