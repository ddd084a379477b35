"""Layer modules backed by the sm_100a kernels.

They subclass the torch modules of the same name, so parameter / buffer names, shapes,
initialisation and ``state_dict`` layout are exactly torch's (the reference checkpoints
``model.state_dict()``: pytorch/single_gpu.py:81-84).  On CUDA tensors ``forward`` runs the
hand-written kernels on NHWC activations (bf16, or fp32 with TF32 tensor-core math after
``set_precision(model, "tf32")``); on CPU tensors it runs the stock fp32 torch op
(NCHW), which is also the numerical oracle of the test-suite.  There is no PyTorch fallback
on a GPU: a missing native library raises.
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _native as nv
from . import functional as fn
from .functional import pad8, pad16


def is_native(x: torch.Tensor) -> bool:
    return x.is_cuda


# --------------------------------------------------------------------------------------
# weight packing (fp32 master -> bf16 tcgen05 operand layouts)
# --------------------------------------------------------------------------------------
class _PackedWeightMixin:
    """Owns the fprop / dgrad operand packs of a conv / linear weight (bf16, or fp32 in tf32 mode)."""

    act_dtype = torch.bfloat16      # storage type of the packs = of the activations this layer sees

    def _init_pack_meta(self, Cout, Cin, KH, KW, pad, stride, need_dgrad=True):
        self.Cout, self.Cin, self.KH, self.KW = Cout, Cin, KH, KW
        self.conv_pad, self.conv_stride = pad, stride
        self.Cin_p, self.Cout_p = pad8(Cin), pad8(Cout)
        self.need_dgrad = need_dgrad
        self.fuse_relu = False
        self.wf = None
        self.wd = None
        self._packed_version = None
        self._packed_ptr = None

    def _alloc_packs(self, device):
        taps = self.KH * self.KW
        self.wf = torch.zeros(pad16(self.Cout_p) * taps * self.Cin_p, dtype=self.act_dtype, device=device)
        self.wd = (torch.zeros(pad16(self.Cin_p) * taps * self.Cout_p, dtype=self.act_dtype, device=device)
                   if self.need_dgrad else None)

    def _packs_stale(self):
        w = self.weight
        return self.wf is None or self.wf.device != w.device or self.wf.dtype != self.act_dtype

    def _ensure_packed(self, dtype=None):
        """``dtype``: storage type of the activation about to be multiplied (a layer called with a
        type other than the model's configured one simply re-packs for it)."""
        w = self.weight
        if dtype is not None and dtype != self.act_dtype:
            self.act_dtype = dtype
        if self._packs_stale():
            self._alloc_packs(w.device)
            self._packed_version = None
        if self._packed_version != w._version or self._packed_ptr != w.data_ptr():
            pack_weights([self])


def pack_weights(modules, table_cache=None):
    """Packs the weights of ``modules`` (conv / linear layers on one device) in one launch.

    ``table_cache``: a dict that keeps the device-side layer table between calls.  It is keyed by the
    addresses it describes, so re-allocated packs (a precision switch, a move to another device) or
    re-pointed weights rebuild it instead of sending the kernel to freed memory."""
    modules = [m for m in modules if isinstance(m, _PackedWeightMixin)]
    if not modules:
        return None
    for m in modules:
        if m._packs_stale():
            m._alloc_packs(m.weight.device)

    def src_of(m):
        src = getattr(m, "pack_src", None)
        return src if src is not None else m.weight

    key = tuple((src_of(m).data_ptr(), m.wf.data_ptr(), m.wd.data_ptr() if m.wd is not None else 0) for m in modules)
    if table_cache is not None and table_cache.get("table") is not None and table_cache.get("key") == key:
        table, n, blocks = table_cache["table"], table_cache["n"], table_cache["blocks"]
    else:
        esz = nv.lib().dtdl_pack_entry_size()
        host = (ctypes.c_char * (esz * len(modules)))()
        blocks = 0
        for i, m in enumerate(modules):
            entry = ctypes.addressof(host) + i * esz
            blocks += nv.lib().dtdl_pack_entry_fill(
                entry, nv.ptr(src_of(m)), nv.ptr(m.wf), nv.ptr(m.wd), m.Cout, m.Cin, m.KH * m.KW,
                m.Cin_p, m.Cout_p, blocks, fn.act_of(m.wf))
        dev = modules[0].weight.device
        table = torch.frombuffer(bytearray(bytes(host)), dtype=torch.uint8).to(dev)
        n = len(modules)
        if table_cache is not None:
            table_cache.update(table=table, n=n, blocks=blocks, key=key)
    nv.call("pack_weights", nv.ptr(table), n, blocks, nv.stream_ptr())
    for m in modules:
        m._packed_version = m.weight._version
        m._packed_ptr = m.weight.data_ptr()
    return table


# --------------------------------------------------------------------------------------
# layers
# --------------------------------------------------------------------------------------
class Conv2d(nn.Conv2d, _PackedWeightMixin):
    """3x3 / 1x1 / KxK convolution, stride 1 or 2, as tcgen05 implicit GEMM (K-O1..K-O4, K-O20)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True,
                 activation=None, first_layer=False):
        super().__init__(in_channels, out_channels, kernel_size, stride=stride, padding=padding, bias=bias)
        k = self.kernel_size
        assert k[0] == k[1] and self.padding[0] == self.padding[1], "square kernels only"
        assert self.stride[0] == self.stride[1] and self.stride[0] in (1, 2), "stride 1 or 2"
        assert self.dilation == (1, 1) and self.groups == 1
        # a stride-2 1x1 convolution only ever looks at every other pixel: sub-sample first
        self.pre_subsample = k[0] == 1 and self.stride[0] == 2
        self._init_pack_meta(out_channels, in_channels, k[0], k[1], self.padding[0],
                             1 if self.pre_subsample else self.stride[0], need_dgrad=not first_layer)
        self.fuse_relu = activation == "relu"
        assert activation in (None, "relu")

    def forward(self, x, stats_for=None):
        """``stats_for``: the BatchNorm2d applied next to the result (see fn.conv)."""
        if not is_native(x):
            y = super().forward(x)
            return F.relu(y) if self.fuse_relu else y
        self._ensure_packed(x.dtype)
        if self.pre_subsample:
            x = fn.subsample2(x)
        return fn.conv(x, self, stats_for=stats_for)


class Linear(nn.Linear, _PackedWeightMixin):
    """Fully connected layer = 1x1 "convolution" over [N,1,1,F] activations (K-O12)."""

    def __init__(self, in_features, out_features, bias=True, activation=None, first_layer=False):
        super().__init__(in_features, out_features, bias=bias)
        self._init_pack_meta(out_features, in_features, 1, 1, 0, 1, need_dgrad=not first_layer)
        self.fuse_relu = activation == "relu"

    def forward(self, x):
        if not is_native(x):
            y = super().forward(x)
            return F.relu(y) if self.fuse_relu else y
        self._ensure_packed(x.dtype)
        return fn.conv(x, self)


class BatchNorm2d(nn.BatchNorm2d):
    """BatchNorm with the following residual add and ReLU fused in (K-O5..K-O9).

    ``forward(x, residual=None, relu=False)`` computes ``relu(bn(x) + pad(residual))`` where a
    residual with fewer channels is added to the leading channels only -- PyramidNet's
    zero-padded identity shortcut (pytorch/model.py:17-20,49) without materialising the pad.
    """

    def __init__(self, num_features, eps=1e-5, momentum=0.1):
        super().__init__(num_features, eps=eps, momentum=momentum)
        self._stat = None
        self._fwd_dirty = False
        self._stats_from_producer = False   # set by the kernel that sums our input for us
        self._fused_bwd_fits = {}           # (P, Cp, act) -> the one-kernel backward applies
        self._bar = None                    # its grid-barrier words (device memory, zero-initialised)

    def _grid_barrier(self, device):
        if self._bar is None or self._bar.device != device:
            self._bar = torch.zeros(2, dtype=torch.int32, device=device)
        return self._bar

    def _stat_buffer(self, device, Cp):
        if self._stat is None or self._stat.device != device or self._stat.numel() != 4 * Cp:
            self._stat = torch.zeros(4 * Cp, dtype=torch.float32, device=device)
            self._fwd_dirty = False
        return self._stat

    def forward(self, x, residual=None, relu=False, stats_for=None, fork_input=False):
        """``fork_input=True`` returns ``(y, x')``: use ``x'`` as the block's identity shortcut and the
        shortcut's gradient is added to this layer's input gradient inside the backward kernel
        instead of by a separate autograd add."""
        if not is_native(x):
            if fork_input:
                return self.forward(x, residual=residual, relu=relu), x
            y = super().forward(x)
            if residual is not None:
                c = residual.shape[1]
                if c == y.shape[1]:
                    y = y + residual
                else:
                    y = torch.cat([y[:, :c] + residual, y[:, c:]], dim=1)
            return F.relu(y) if relu else y
        return fn.batch_norm(x, self, residual=residual, relu=relu, stats_for=stats_for, fork_input=fork_input)


class AvgPool2d(nn.AvgPool2d):
    def forward(self, x):
        if not is_native(x):
            return super().forward(x)
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        s = self.stride if isinstance(self.stride, int) else self.stride[0]
        p = self.padding if isinstance(self.padding, int) else self.padding[0]
        if x.shape[1] == k and x.shape[2] == k and p == 0:
            return fn.global_avg_pool(x)
        return fn.avg_pool(x, k, s, p, self.ceil_mode)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x):
        if not is_native(x):
            return super().forward(x)
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        s = self.stride if isinstance(self.stride, int) else self.stride[0]
        p = self.padding if isinstance(self.padding, int) else self.padding[0]
        return fn.max_pool(x, k, s, p, self.ceil_mode)


class GlobalAvgPool2d(nn.Module):
    def forward(self, x):
        if not is_native(x):
            return x.mean(dim=(2, 3), keepdim=True)
        return fn.global_avg_pool(x)


class InputLayout(nn.Module):
    """Entry of a model: NCHW / [N,F] fp32 -> the framework's NHWC layout (GPU only), stored as
    ``act_dtype`` (bf16 by default; ``set_precision(model, "tf32")`` switches to fp32)."""

    act_dtype = torch.bfloat16

    def forward(self, x):
        if not is_native(x):
            return x
        return fn.to_nhwc(x, dtype=self.act_dtype)


class Flatten(nn.Module):
    """Flattens to [N, features].  ``channels`` is the real channel count of the input.

    The feature order is NHWC on both paths (Keras semantics, tensorflow2/mnist_single.py:24),
    so CPU and GPU agree for the same weights.
    """

    def __init__(self, channels):
        super().__init__()
        self.channels = channels

    def forward(self, x):
        if not is_native(x):
            if x.dim() == 4:
                return x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
            return x.reshape(x.shape[0], -1)
        return fn.flatten_nhwc(x, self.channels)


class OutputLayout(nn.Module):
    """Exit of a model: [N,1,1,Cp] bf16 -> [N,C] fp32 logits."""

    def __init__(self, features):
        super().__init__()
        self.features = features

    def forward(self, x):
        if not is_native(x):
            return x.reshape(x.shape[0], -1)
        if getattr(self, "_bypass", False):
            return x  # the single-process engine converts while gathering to GPU 0
        return fn.to_f32(x, self.features)


def set_precision(model: nn.Module, precision="bf16"):
    """Selects the activation / operand storage type of every layer of ``model``:

    * ``"bf16"`` (default): bf16 activations and weight operands, fp32 accumulation, statistics and
      master weights -- the fastest mode;
    * ``"tf32"``: fp32 activations and weight operands, TF32 tensor-core convolutions, fp32
      everything else -- numerically what the reference's fp32 model does under torch's cuDNN default
      (pytorch/single_gpu.py:65-72).

    Returns the model.  Packs are rebuilt lazily at the next forward."""
    dt = fn.act_dtype(precision)
    for m in model.modules():
        if isinstance(m, (_PackedWeightMixin, InputLayout)) or hasattr(m, "act_dtype"):
            m.act_dtype = dt
        op = getattr(m, "_op", None)        # packed-weight holders that are not modules (ResNet stem)
        if isinstance(op, _PackedWeightMixin):
            op.act_dtype = dt
    model._dtdl_precision = "tf32" if dt == torch.float32 else "bf16"
    return model


def packed_layers(model: nn.Module):
    return [m for m in model.modules() if isinstance(m, _PackedWeightMixin)]
