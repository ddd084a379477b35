"""PyramidNet-110 (alpha=270) for 32x32 inputs on the framework's op layer.

Same function, parameter names, shapes and initialisation as the reference model
(pytorch/model.py:24-118), so checkpoints are interchangeable key-for-key; the execution is
different.  Per block the reference runs BN -> conv -> BN -> ReLU -> conv -> BN, F.pad of the
shortcut and an in-place add as nine separate ops; here ReLU, the residual add and the zero
padding of the shortcut are folded into the BatchNorm kernels, every convolution is a tcgen05
implicit GEMM on NHWC bf16, and the head (BN+ReLU, 8x8 average pool, FC) stays in that layout
until the fp32 logits.
"""
from __future__ import annotations

import torch.nn as nn

from ..ops import modules as L


def pyramid_widths(num_layers: int = 18, alpha: float = 270.0, base: int = 16):
    """Channel widths after each of the 3*(num_layers-1) blocks: 21, 26, ..., 271.

    Widths grow by alpha / (3*num_layers) per block and are rounded per block
    (pytorch/model.py:60,87-97); note the reference builds num_layers-1 blocks per stage.
    """
    rate = alpha / (3.0 * num_layers)
    widths, w = [], float(base)
    for _ in range(3 * (num_layers - 1)):
        w += rate
        widths.append(int(round(w)))
    return widths


class PyramidBlock(nn.Module):
    """Pre-activation residual block with a zero-padded identity shortcut."""

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.bn1 = L.BatchNorm2d(cin)
        self.conv1 = L.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)
        self.bn2 = L.BatchNorm2d(cout)
        self.conv2 = L.Conv2d(cout, cout, 3, stride=1, padding=1, bias=False)
        self.bn3 = L.BatchNorm2d(cout)
        # average-pool the shortcut on down-sampling blocks (pytorch/model.py:10-11,19-20)
        self.pool = L.AvgPool2d(2, stride=2, ceil_mode=True) if stride == 2 else None
        self.next_bn = []   # [BatchNorm2d that consumes this block's output]; a list so that it is
        #                     not registered as a sub-module (it belongs to the next block)

    def forward(self, x):
        if self.pool is None:
            h, shortcut = self.bn1(x, fork_input=True)   # shortcut == x; its gradient joins bn1's
        else:
            h, shortcut = self.bn1(x), self.pool(x)
        out = self.conv1(h, stats_for=self.bn2)
        out = self.conv2(self.bn2(out, relu=True), stats_for=self.bn3)
        # bn3 + shortcut on the first cin channels == out += F.pad(shortcut); the BatchNorm that
        # reads the block's output next (set by PyramidNet) gets its statistics from this pass
        return self.bn3(out, residual=shortcut, stats_for=self.next_bn[0] if self.next_bn else None)


class PyramidNet(nn.Module):
    def __init__(self, num_layers: int = 18, alpha: float = 270.0, num_classes: int = 10):
        super().__init__()
        widths = pyramid_widths(num_layers, alpha)
        per_stage = num_layers - 1
        self.input = L.InputLayout()
        self.conv1 = L.Conv2d(3, 16, 3, stride=1, padding=1, bias=False, first_layer=True)
        self.bn1 = L.BatchNorm2d(16)
        cin = 16
        stages = []
        for s in range(3):
            blocks = []
            for b in range(per_stage):
                cout = widths[s * per_stage + b]
                blocks.append(PyramidBlock(cin, cout, stride=2 if (s > 0 and b == 0) else 1))
                cin = cout
            stages.append(nn.Sequential(*blocks))
        self.layer1, self.layer2, self.layer3 = stages
        self.out_channels = cin
        self.bn_out = L.BatchNorm2d(cin)
        chain = [b for stage in stages for b in stage]
        for blk, nxt in zip(chain, chain[1:]):
            blk.next_bn = [nxt.bn1]
        chain[-1].next_bn = [self.bn_out]
        self.avgpool = L.AvgPool2d(8, stride=1)
        self.fc_out = L.Linear(cin, num_classes)
        self.output = L.OutputLayout(num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def forward(self, x):
        x = self.bn1(self.conv1(self.input(x), stats_for=self.bn1), stats_for=self.layer1[0].bn1)
        x = self.layer3(self.layer2(self.layer1(x)))
        x = self.avgpool(self.bn_out(x, relu=True))
        if not L.is_native(x):
            x = x.flatten(1)
        return self.output(self.fc_out(x))


def pyramidnet(num_classes: int = 10) -> PyramidNet:
    """The reference's configuration: num_layers=18, alpha=270 (pytorch/model.py:115-118)."""
    return PyramidNet(num_layers=18, alpha=270.0, num_classes=num_classes)
