"""ResNet-50 for 224x224 inputs on the framework's op layer (BASELINE.json config 4).

ResNet-50 is not part of the reference repository; it is the compute-heavy
scaling workload named by the north star.  Parameter / buffer names and shapes are
torchvision's ``resnet50`` (``conv1``, ``bn1``, ``layer1.0.conv1`` ... ``layer4.2.bn3``,
``downsample.0/1``, ``fc``), so its checkpoints load key-for-key.

Execution: every convolution is the tcgen05 implicit-GEMM kernel on NHWC bf16.  The 7x7 /
stride-2 stem runs as a 4x4 / stride-1 convolution over a 2x2 space-to-depth view of the image
(12 instead of 3 input channels: K = 16 taps x 16 channels, no 4x over-computation); the
stride-2 1x1 shortcut convolutions sub-sample first; the three stride-2 3x3 convolutions use
the stride-1 kernel + sub-sample.  BatchNorm + ReLU (+ residual add) are single fused kernels.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _native as nv
from ..ops import functional as fn
from ..ops import modules as L
from ..ops.modules import _PackedWeightMixin


class _StemOperand(_PackedWeightMixin):
    """Packed-weight holder for the stem's 4x4x12 kernel (not an nn.Module)."""

    def __init__(self, cout):
        self._init_pack_meta(cout, 12, 4, 4, pad=2, stride=1, need_dgrad=False)
        self.bias = None
        self.weight = None            # fp32 [O,12,4,4] copy of the 7x7 parameter, persistent (stable pointer)
        self.out_hw = None
        self._table = {}


def s2d_weight(w7):
    """[O,3,7,7] -> [O,12,4,4]: pad the kernel to 8x8 at the low side, split the 2x2 phases."""
    O = w7.shape[0]
    w8 = F.pad(w7, (1, 0, 1, 0))                             # [O,3,8,8], index 0 is the zero tap
    w = w8.view(O, 3, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4)   # [O, bh, bw, c, a, b]
    return w.reshape(O, 12, 4, 4)


def s2d_weight_adjoint(g4):
    """Gradient of ``s2d_weight``: [O,12,4,4] -> [O,3,7,7]."""
    O = g4.shape[0]
    g8 = g4.view(O, 2, 2, 3, 4, 4).permute(0, 3, 4, 1, 5, 2).reshape(O, 3, 8, 8)
    return g8[:, :, 1:, 1:]


class _StemFn(torch.autograd.Function):
    """7x7/2 convolution as 4x4/1 over the space-to-depth image; the weight gradient is mapped
    back to 7x7 and written into the parameter's gradient buffer like every other layer's."""

    @staticmethod
    def forward(ctx, x, weight, op):
        N, C, H, W = x.shape
        if op.weight is None or op.weight.device != x.device or op._packs_stale():
            op.weight = torch.empty((op.Cout, 12, 4, 4), dtype=torch.float32, device=x.device)
            op._alloc_packs(x.device)
            op._table = {}
        op.weight.copy_(s2d_weight(weight.detach()))
        L.pack_weights([op], op._table)
        xs = fn.to_s2d(x, dtype=op.act_dtype)
        y = fn.conv_fprop_raw(xs, op.wf, op.Cout_p, 4, 4, 2, out_hw=(H // 2, W // 2))
        ctx.op = op
        ctx.save_for_backward(xs, weight)
        return y

    @staticmethod
    def backward(ctx, dy):
        op = ctx.op
        xs, weight = ctx.saved_tensors
        dy = dy.contiguous()
        N, Hs, Ws, Cin_p = xs.shape
        Ho, Wo = dy.shape[1], dy.shape[2]
        g4 = torch.empty((op.Cout, 12, 4, 4), dtype=torch.float32, device=dy.device)
        act = fn.act_of(xs)
        need = nv.lib().dtdl_conv_wgrad_workspace_bytes_ex(N, Ho, Wo, Cin_p, op.Cout_p, 4, 4, 1, act)
        ws = fn.scratch(xs.device, need)
        nv.call("conv_wgrad", nv.ptr(xs), nv.ptr(dy), nv.ptr(g4), nv.ptr(ws), ws.numel(), N, Hs, Ws, 12, Cin_p,
                Ho, Wo, op.Cout, op.Cout_p, 4, 4, 2, 2, 0, 1, act, nv.stream_ptr())
        gw, acc = fn.grad_buffer(weight)
        g7 = s2d_weight_adjoint(g4)
        if acc:
            gw.add_(g7)
        else:
            gw.copy_(g7)
        fn._grad_ready(weight)
        return None, None, None


class StemConv7x7s2(nn.Conv2d):
    """``Conv2d(3, 64, 7, stride=2, padding=3, bias=False)`` computed by space-to-depth."""

    def __init__(self, in_channels=3, out_channels=64):
        super().__init__(in_channels, out_channels, 7, stride=2, padding=3, bias=False)
        assert in_channels == 3
        self._op = _StemOperand(out_channels)

    s2d_weight = staticmethod(s2d_weight)

    def forward(self, x):
        if not L.is_native(x):
            return super().forward(x)
        return _StemFn.apply(x, self.weight, self._op)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, width, stride=1, downsample=False):
        super().__init__()
        cout = width * self.expansion
        self.conv1 = L.Conv2d(cin, width, 1, bias=False)
        self.bn1 = L.BatchNorm2d(width)
        self.conv2 = L.Conv2d(width, width, 3, stride=stride, padding=1, bias=False)
        self.bn2 = L.BatchNorm2d(width)
        self.conv3 = L.Conv2d(width, cout, 1, bias=False)
        self.bn3 = L.BatchNorm2d(cout)
        self.downsample = None
        if downsample:
            self.downsample = nn.Sequential(L.Conv2d(cin, cout, 1, stride=stride, bias=False), L.BatchNorm2d(cout))

    def forward(self, x):
        # every stride-1 GEMM also sums its output for the BatchNorm that follows (stats_for)
        if self.downsample is None:
            identity = x
        else:
            ds_conv, ds_bn = self.downsample
            identity = ds_bn(ds_conv(x, stats_for=ds_bn))
        out = self.bn1(self.conv1(x, stats_for=self.bn1), relu=True)
        out = self.bn2(self.conv2(out, stats_for=self.bn2), relu=True)
        return self.bn3(self.conv3(out, stats_for=self.bn3), residual=identity, relu=True)   # relu(bn3 + identity)


class ResNet50(nn.Module):
    def __init__(self, num_classes=1000, layers=(3, 4, 6, 3)):
        super().__init__()
        self.conv1 = StemConv7x7s2(3, 64)
        self.bn1 = L.BatchNorm2d(64)
        self.maxpool = L.MaxPool2d(3, stride=2, padding=1)
        cin = 64
        stages = []
        for i, (n, width) in enumerate(zip(layers, (64, 128, 256, 512))):
            blocks = []
            for b in range(n):
                stride = 2 if (i > 0 and b == 0) else 1
                blocks.append(Bottleneck(cin, width, stride, downsample=(b == 0)))
                cin = width * 4
            stages.append(nn.Sequential(*blocks))
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.avgpool = L.GlobalAvgPool2d()
        self.fc = L.Linear(cin, num_classes)
        self.output = L.OutputLayout(num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = self.maxpool(self.bn1(self.conv1(x), relu=True))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x)
        if not L.is_native(x):
            x = x.flatten(1)
        return self.output(self.fc(x))


def resnet50(num_classes=1000) -> ResNet50:
    return ResNet50(num_classes)
