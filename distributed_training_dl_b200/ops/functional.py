"""Autograd functions over the sm_100a kernels.

Activation layout on the GPU: NHWC, channel count padded to a multiple of 8 (``Cp``), pad
channels exactly zero.  A 2-D activation (Linear) is ``[N,1,1,Cp]``.  Two storage types,
chosen per model (``set_precision`` / ``wrap(..., dtype=)``): ``torch.bfloat16`` (bf16
tensor-core math) and ``torch.float32`` ("tf32" mode: fp32 activations and weights, TF32
tensor-core convolutions, everything else fp32 -- the precision the reference itself trains at,
pytorch/single_gpu.py:65-72 under torch's cuDNN default).  Every op takes its type from the
tensor it is given.
Parameter gradients are written by the kernels straight into the parameter's
gradient buffer (a view of the reducer's flat arena when the model is wrapped),
so autograd never materialises or accumulates a weight gradient itself.

What the reference gets from torch.nn / cuDNN and where it is here (SURVEY.md 2.5 b):
``nn.Conv2d`` + its autograd (pytorch/model.py:28-32,62-63) -> ``_Conv``; ``nn.BatchNorm2d`` +
``nn.ReLU`` + ``out += shortcut`` with the zero-padded identity (pytorch/model.py:27-50) ->
``_BatchNorm``; ``nn.AvgPool2d`` (pytorch/model.py:11,76), Keras ``MaxPooling2D``
(tensorflow2/mnist_single.py:19,21) -> ``_Pool`` / ``_GlobalAvgPool``; ``nn.CrossEntropyLoss`` and the
argmax / eq / sum accuracy arithmetic (pytorch/single_gpu.py:99,105-108) -> ``_SoftmaxCE``.
"""
from __future__ import annotations

import os

import torch

from .. import _native as nv


# DTDL_FUSE_BN_STATS=0 keeps the separate statistics pass (A/B measurements)
FUSE_BN_STATS = os.environ.get("DTDL_FUSE_BN_STATS", "1") != "0"

ACT_DTYPES = {"bf16": torch.bfloat16, "tf32": torch.float32, "fp32": torch.float32,
              torch.bfloat16: torch.bfloat16, torch.float32: torch.float32}


def act_dtype(spec):
    """'bf16' | 'tf32' | torch dtype -> the activation storage dtype."""
    try:
        return ACT_DTYPES[spec]
    except KeyError:
        raise ValueError(f"unknown precision {spec!r}: use 'bf16' or 'tf32'") from None


def act_of(t) -> int:
    """The ``act`` argument of the C ABI for a tensor / dtype: 0 = bf16, 1 = fp32 (TF32 math)."""
    dt = t if isinstance(t, torch.dtype) else t.dtype
    if dt == torch.bfloat16:
        return 0
    if dt == torch.float32:
        return 1
    raise TypeError(f"activations must be bfloat16 or float32, got {dt}")


def pad8(c: int) -> int:
    return (c + 7) // 8 * 8


def pad16(c: int) -> int:
    return (c + 15) // 16 * 16


# --------------------------------------------------------------------------------------
# gradient buffers
# --------------------------------------------------------------------------------------
def grad_buffer(p: torch.Tensor):
    """Returns (fp32 gradient tensor of ``p``, accumulate flag) for a kernel to write into.

    * wrapped model: ``p._dtdl_grad`` is a view of the flat gradient arena;
    * bare module:   the buffer is created on first use.

    ``accumulate`` is 0 for the first write after ``zero_grad`` and 1 afterwards, which
    reproduces torch's "grads add up until zeroed" contract without a memset.
    """
    g = getattr(p, "_dtdl_grad", None)
    if g is None:
        g = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
        p._dtdl_grad = g
    fresh = p.grad is None or getattr(p, "_dtdl_fresh", False)
    if p.grad is not g:
        p.grad = g
    p._dtdl_fresh = False
    p._dtdl_native = True
    return g, (0 if fresh else 1)


def mark_fresh(params):
    """``zero_grad`` without touching memory: the next backward overwrites."""
    for p in params:
        p._dtdl_fresh = True


_ready_hooks = []


def add_ready_hook(fn):
    """``fn(param)`` is called when a kernel has enqueued the final gradient of ``param``."""
    _ready_hooks.append(fn)
    return fn


def remove_ready_hook(fn):
    if fn in _ready_hooks:
        _ready_hooks.remove(fn)


def _grad_ready(p):
    for fn in _ready_hooks:
        fn(p)


# --------------------------------------------------------------------------------------
# weight gradients on a second branch of the captured graph
# --------------------------------------------------------------------------------------
# A layer's weight gradient and data gradient both need only dy.  While a step is being captured
# into a CUDA graph the weight-gradient kernels go to a side stream: in the graph they become a
# second chain that runs next to the data-gradient / BatchNorm-backward chain, and two
# latency-bound kernels share the SMs (measured 7-15 % less time per layer, tools/bench_kernels.py
# OVERLAP=1).  Eager execution is CPU-launch-bound and stays on one stream.  The chains are
# joined by an autograd end-of-backward callback, and the reducer waits for both before a bucket's
# collective.  DTDL_WGRAD_FORK=0 keeps everything on one stream.
WGRAD_FORK = os.environ.get("DTDL_WGRAD_FORK", "1") != "0"
_side = {}   # device index -> {"stream", "forked", "main"}


def wgrad_stream(device):
    """Side stream for this layer's weight-gradient kernels (ordered after everything enqueued on
    the current stream so far), or None when not capturing."""
    if not WGRAD_FORK or device.type != "cuda" or not torch.cuda.is_current_stream_capturing():
        return None
    st = _side.get(device.index)
    if st is None:
        st = _side[device.index] = {"stream": torch.cuda.Stream(device=device), "forked": False, "main": None}
    main = torch.cuda.current_stream(device)
    st["stream"].wait_stream(main)
    if not st["forked"]:
        st["forked"], st["main"] = True, main
        torch.autograd.Variable._execution_engine.queue_callback(join_side_streams)
    return st["stream"]


def forked_side_stream(device):
    """The side stream if this backward pass has put work on it (the reducer must wait for it)."""
    st = _side.get(device.index) if device.type == "cuda" else None
    return st["stream"] if st is not None and st["forked"] else None


def join_side_streams():
    """The main chain continues only after every weight gradient has been written."""
    for st in _side.values():
        if st["forked"]:
            st["main"].wait_stream(st["stream"])
            st["forked"] = False


# --------------------------------------------------------------------------------------
# scratch
# --------------------------------------------------------------------------------------
_scratch = {}


def scratch(device, nbytes: int, tag: str = "") -> torch.Tensor:
    """Per-device scratch buffer, grown on demand (stream-ordered reuse; one per ``tag`` so that
    kernels on different streams never share one)."""
    key = (device.type, device.index, tag)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        nbytes = max(nbytes, 32 << 20)
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


# --------------------------------------------------------------------------------------
# layout
# --------------------------------------------------------------------------------------
def to_nhwc(x: torch.Tensor, Cp: int | None = None, dtype=torch.bfloat16) -> torch.Tensor:
    """NCHW (or [N,F]) fp32 -> NHWC padded, stored as ``dtype``.  No gradient flows to the input.

    The result lives on the CURRENT device even when ``x`` is a (peer-accessible) tensor of
    another GPU: the kernel then reads its batch slice over NVLink, which is how the
    single-process engine folds ``scatter`` into the first layer (K-C4)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    act = act_of(dtype)
    if x.dim() == 2:
        N, C = x.shape
        Cp = Cp or pad8(C)
        y = torch.empty((N, 1, 1, Cp), dtype=dtype, device=dev)
        nv.call("pad_cast", nv.ptr(x.contiguous().float()), nv.ptr(y), N, C, Cp, act, nv.stream_ptr())
        return y
    N, C, H, W = x.shape
    Cp = Cp or pad8(C)
    x = x.detach()
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.contiguous().float()
    y = torch.empty((N, H, W, Cp), dtype=dtype, device=dev)
    nv.call("nchw_to_nhwc", nv.ptr(x), nv.ptr(y), N, C, H, W, Cp, act, nv.stream_ptr())
    return y


class _ToF32(torch.autograd.Function):
    """[N,1,1,Cp] activations -> [N,C] fp32 (e.g. logits handed to a user criterion)."""

    @staticmethod
    def forward(ctx, x, C):
        N, Cp = x.shape[0], x.shape[-1]
        ctx.meta = (N, C, Cp, x.dtype)
        y = torch.empty((N, C), dtype=torch.float32, device=x.device)
        nv.call("unpad_cast", nv.ptr(x), nv.ptr(y), N, C, Cp, act_of(x), nv.stream_ptr())
        return y

    @staticmethod
    def backward(ctx, dy):
        N, C, Cp, dtype = ctx.meta
        dy = dy.contiguous().float()
        dx = torch.empty((N, 1, 1, Cp), dtype=dtype, device=dy.device)
        nv.call("pad_cast", nv.ptr(dy), nv.ptr(dx), N, C, Cp, act_of(dtype), nv.stream_ptr())
        return dx, None


def to_f32(x, C):
    return _ToF32.apply(x, C)


# --------------------------------------------------------------------------------------
# convolution / linear
# --------------------------------------------------------------------------------------
def _conv_out(n, k, pad, stride=1):
    return (n + 2 * pad - k) // stride + 1


def conv_fprop_raw(x, wpack, Cout_p, KH, KW, pad, bias=None, relu=False, out_f32=False, out_hw=None,
                   stats=0, stride=1):
    """``stats``: device address of a zeroed fp32 [2*Cout_p] buffer that receives the per-channel
    sum / sum of squares of the output (the following BatchNorm's batch statistics).  The output
    has the storage type of ``x`` (bf16 inputs may ask for an fp32 output with ``out_f32``)."""
    N, H, W, Cin_p = x.shape
    Ho, Wo = out_hw or (_conv_out(H, KH, pad, stride), _conv_out(W, KW, pad, stride))
    act = act_of(x)
    y = torch.empty((N, Ho, Wo, Cout_p), dtype=torch.float32 if (out_f32 or act) else torch.bfloat16,
                    device=x.device)
    flags = (1 if relu else 0) | (2 if out_f32 else 0)
    nv.call("conv_fprop_stats", nv.ptr(x), nv.ptr(wpack), nv.ptr(y), nv.ptr(bias), stats, N, H, W,
            Cin_p, Ho, Wo, Cout_p, KH, KW, pad, pad, flags, stride, act, nv.stream_ptr())
    return y


def _claim_bn_stats(bn, device, Cp):
    """Hand the BatchNorm's statistics buffer to the producer of its input: returns the address
    the producer accumulates into, and marks the module so its forward skips the stats pass."""
    stat = bn._stat_buffer(device, Cp)
    if bn._fwd_dirty:
        stat[: 2 * Cp].zero_()
    bn._fwd_dirty = True
    bn._stats_from_producer = True
    return stat.data_ptr()


def subsample2_raw(x, adjoint=False, full_hw=None):
    act = act_of(x)
    if not adjoint:
        N, H, W, Cp = x.shape
        y = torch.empty((N, (H + 1) // 2, (W + 1) // 2, Cp), dtype=x.dtype, device=x.device)
        nv.call("subsample2", nv.ptr(x), nv.ptr(y), N, H, W, Cp, 0, act, nv.stream_ptr())
        return y
    N, _, _, Cp = x.shape
    H, W = full_hw
    y = torch.empty((N, H, W, Cp), dtype=x.dtype, device=x.device)
    nv.call("subsample2", nv.ptr(x), nv.ptr(y), N, H, W, Cp, 1, act, nv.stream_ptr())
    return y


# fp32 mode: Linear layers up to this many weights run in full fp32 on CUDA cores, like the
# reference's nn.Linear (torch keeps TF32 off for matmuls); larger ones use the TF32 tensor-core path
LINEAR_F32_MAX_WEIGHTS = 1 << 16


def _use_linear_f32(x, mod):
    return (x.dtype == torch.float32 and mod.KH == 1 and mod.KW == 1 and x.shape[1] == 1 and x.shape[2] == 1
            and mod.conv_stride == 1 and mod.Cout * mod.Cin <= LINEAR_F32_MAX_WEIGHTS)


class _LinearF32(torch.autograd.Function):
    """Small fully connected layer in full fp32 (fp32 activation mode; see LINEAR_F32_MAX_WEIGHTS)."""

    @staticmethod
    def forward(ctx, x, weight, bias, mod):
        x = x.contiguous()
        N, Kp = x.shape[0], x.shape[-1]
        y = torch.empty((N, 1, 1, mod.Cout_p), dtype=torch.float32, device=x.device)
        y2, y2_ld = getattr(mod, "_gather_dst", None) or (0, 0)
        nv.call("linear_f32_fwd", nv.ptr(x), nv.ptr(weight), nv.ptr(bias), nv.ptr(y), y2, y2_ld, N, mod.Cin, Kp,
                mod.Cout, mod.Cout_p, 1 if mod.fuse_relu else 0, nv.stream_ptr())
        ctx.mod = mod
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, weight, bias, y if mod.fuse_relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        mod = ctx.mod
        x, weight, bias, y = ctx.saved_tensors
        dy = dy.contiguous()
        st = nv.stream_ptr()
        N, Kp = x.shape[0], x.shape[-1]
        if mod.fuse_relu:
            g = torch.empty_like(dy)
            nv.call("relu_bwd", nv.ptr(dy), nv.ptr(y), nv.ptr(g), dy.numel(), 1, st)
            dy = g
        if ctx.has_bias:
            gb, acc = grad_buffer(bias)
            if not acc:
                gb.zero_()
            nv.call("colsum", nv.ptr(dy), nv.ptr(gb), N, mod.Cout, mod.Cout_p, 1, st)
            _grad_ready(bias)
        gw, acc = grad_buffer(weight)
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        nv.call("linear_f32_bwd", nv.ptr(dy), nv.ptr(x), nv.ptr(weight), nv.ptr(dx), nv.ptr(gw), N, mod.Cin,
                Kp, mod.Cout, mod.Cout_p, acc, st)
        _grad_ready(weight)
        return dx, None, None, None


class _Conv(torch.autograd.Function):
    """Convolution (stride 1 or 2) / Linear with optional fused bias + ReLU epilogue.

    Stride 2 is native: the forward and weight-gradient kernels sample the input through TMA
    element strides, the data gradient runs in gather form over the four output phases
    (``conv_dgrad_s2``), so a stride-2 layer does 1x the arithmetic of its definition.
    """

    @staticmethod
    def forward(ctx, x, weight, bias, mod, stats_for=None):
        x = x.contiguous()
        stats = 0
        if (stats_for is not None and FUSE_BN_STATS
                and (stats_for.training or stats_for.running_mean is None)):
            stats = _claim_bn_stats(stats_for, x.device, mod.Cout_p)
        gather = getattr(mod, "_gather_dst", None)
        if gather is not None and mod.conv_stride == 1 and not stats:
            # last layer of a replica in the single-process engine: the epilogue also stores fp32 logits
            # into GPU 0's output buffer (no gather pass)
            N, H, W, Cin_p = x.shape
            Ho, Wo = _conv_out(H, mod.KH, mod.conv_pad), _conv_out(W, mod.KW, mod.conv_pad)
            y = torch.empty((N, Ho, Wo, mod.Cout_p), dtype=x.dtype, device=x.device)
            nv.call("conv_fprop_gather", nv.ptr(x), nv.ptr(mod.wf), nv.ptr(y), nv.ptr(bias), gather[0], gather[1],
                    mod.Cout, N, H, W, Cin_p, Ho, Wo, mod.Cout_p, mod.KH, mod.KW, mod.conv_pad, mod.conv_pad,
                    1 if mod.fuse_relu else 0, act_of(x), nv.stream_ptr())
        else:
            y = conv_fprop_raw(x, mod.wf, mod.Cout_p, mod.KH, mod.KW, mod.conv_pad, bias=bias,
                               relu=mod.fuse_relu, out_f32=False, out_hw=getattr(mod, "out_hw", None),
                               stats=stats, stride=mod.conv_stride)
        ctx.mod = mod
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, weight, bias, y if mod.fuse_relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        mod = ctx.mod
        x, weight, bias, y = ctx.saved_tensors
        dy = dy.contiguous()
        st = nv.stream_ptr()
        act = act_of(x)
        stride = mod.conv_stride
        if mod.fuse_relu:
            g = torch.empty_like(dy)
            nv.call("relu_bwd", nv.ptr(dy), nv.ptr(y), nv.ptr(g), dy.numel(), act, st)
            dy = g
        if ctx.has_bias:
            gb, acc = grad_buffer(bias)
            if not acc:
                gb.zero_()
            P = dy.shape[0] * dy.shape[1] * dy.shape[2]
            nv.call("colsum", nv.ptr(dy), nv.ptr(gb), P, mod.Cout, mod.Cout_p, act, st)
            _grad_ready(bias)
        N, H, W, Cin_p = x.shape
        Ho, Wo = dy.shape[1], dy.shape[2]
        # weight gradient: straight into the parameter's gradient buffer, or -- when the weight
        # is a computed tensor (ResNet stem's space-to-depth view) -- returned to autograd
        returned = getattr(mod, "return_wgrad", False)
        need = nv.lib().dtdl_conv_wgrad_workspace_bytes_ex(N, Ho, Wo, Cin_p, mod.Cout_p, mod.KH, mod.KW,
                                                           stride, act)
        side = None if returned else wgrad_stream(x.device)
        if side is not None:
            # x and dy are read on the side stream: keep the allocator from recycling them early
            x.record_stream(side)
            dy.record_stream(side)
            with torch.cuda.stream(side):
                gw, acc = grad_buffer(weight)
                ws = scratch(x.device, need, "side")
                nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, H, W,
                        mod.Cin, Cin_p, Ho, Wo, mod.Cout, mod.Cout_p, mod.KH, mod.KW, mod.conv_pad,
                        mod.conv_pad, acc, stride, act, side.cuda_stream)
                _grad_ready(weight)
        else:
            if returned:
                gw, acc = torch.empty(weight.shape, dtype=torch.float32, device=weight.device), 0
            else:
                gw, acc = grad_buffer(weight)
            ws = scratch(x.device, need)
            nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, H, W,
                    mod.Cin, Cin_p, Ho, Wo, mod.Cout, mod.Cout_p, mod.KH, mod.KW, mod.conv_pad, mod.conv_pad,
                    acc, stride, act, st)
            if not returned:
                _grad_ready(weight)
        # data gradient
        dx = None
        if ctx.needs_input_grad[0]:
            if stride == 1:
                dx = conv_fprop_raw(dy, mod.wd, Cin_p, mod.KH, mod.KW, mod.KH - 1 - mod.conv_pad)
            else:
                dx = torch.empty_like(x)
                nv.call("conv_dgrad_s2", nv.ptr(dy), nv.ptr(mod.wd), nv.ptr(dx), N, Ho, Wo, mod.Cout_p, H, W,
                        Cin_p, mod.KH, mod.KW, mod.conv_pad, mod.conv_pad, act, st)
        return dx, (gw if returned else None), None, None, None


def conv(x, mod, stats_for=None):
    """``stats_for``: the training-mode BatchNorm2d that consumes this convolution's output; its
    batch statistics then come out of the GEMM epilogue."""
    if _use_linear_f32(x, mod):
        return _LinearF32.apply(x, mod.weight, mod.bias, mod)
    return _Conv.apply(x, mod.weight, mod.bias, mod, stats_for)


class _Subsample(torch.autograd.Function):
    """y[n,ho,wo] = x[n,2ho,2wo]: a stride-2 1x1 convolution is this followed by a stride-1 one."""

    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        ctx.full_hw = (x.shape[1], x.shape[2])
        return subsample2_raw(x)

    @staticmethod
    def backward(ctx, dy):
        return subsample2_raw(dy.contiguous(), adjoint=True, full_hw=ctx.full_hw)


def subsample2(x):
    return _Subsample.apply(x)


def to_s2d(x: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """NCHW fp32 -> 2x2 space-to-depth NHWC [N,H/2,W/2,pad8(4C)] (ResNet stem input)."""
    N, C, H, W = x.shape
    Cp = pad8(4 * C)
    x = x.detach()
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.contiguous().float()
    y = torch.empty((N, H // 2, W // 2, Cp), dtype=dtype,
                    device=torch.device("cuda", torch.cuda.current_device()))
    nv.call("nchw_to_s2d", nv.ptr(x), nv.ptr(y), N, C, H, W, Cp, act_of(dtype), nv.stream_ptr())
    return y


# --------------------------------------------------------------------------------------
# batch norm (+ residual) (+ ReLU)
# --------------------------------------------------------------------------------------
class _BatchNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, mod, relu, stats_for=None, fork_input=False):
        x = x.contiguous()
        N, H, W, Cp = x.shape
        P = N * H * W
        C = mod.num_features
        st = nv.stream_ptr()
        stat = mod._stat_buffer(x.device, Cp)
        y = torch.empty_like(x)
        training = mod.training or mod.running_mean is None
        if residual is not None:
            residual = residual.contiguous()
        Cres_p = residual.shape[-1] if residual is not None else 0
        if training and getattr(mod, "_stats_from_producer", False):
            mod._stats_from_producer = False  # the kernel that wrote x already summed it
        elif training:
            if mod._fwd_dirty:
                stat[: 2 * Cp].zero_()
            nv.call("bn_stats", nv.ptr(x), nv.ptr(stat), P, Cp, act_of(x), st)
            mod._fwd_dirty = True
        save = torch.empty((2, Cp), dtype=torch.float32, device=x.device)
        mom = mod.momentum if mod.momentum is not None else 0.1
        track = mod.track_running_stats and mod.running_mean is not None
        out_sums = 0
        if (stats_for is not None and FUSE_BN_STATS and stats_for is not mod
                and (stats_for.training or stats_for.running_mean is None)):
            out_sums = _claim_bn_stats(stats_for, x.device, Cp)
        nv.call("bn_apply_stats", nv.ptr(x), nv.ptr(y), nv.ptr(stat), nv.ptr(gamma), nv.ptr(beta),
                nv.ptr(mod.running_mean) if track else 0, nv.ptr(mod.running_var) if track else 0,
                nv.ptr(mod.num_batches_tracked) if (track and training) else 0,
                nv.ptr(save[0]), nv.ptr(save[1]), nv.ptr(residual), Cres_p, P, C, Cp, mod.eps, mom,
                1 if relu else 0, 1 if training else 0,
                stat[2 * Cp:].data_ptr() if training else 0, 2 * Cp, out_sums, act_of(x), st)
        ctx.mod, ctx.relu, ctx.Cres_p, ctx.training = mod, relu, Cres_p, training
        ctx.save_for_backward(x, y if relu else None, gamma, beta, save)
        if fork_input:
            # second output = the input itself (the identity shortcut of a residual block): its
            # gradient comes back to this node and is summed into dx by the backward kernel
            return y, x.view_as(x)
        return y

    @staticmethod
    def backward(ctx, dy, dalias=None):
        mod = ctx.mod
        x, y, gamma, beta, save = ctx.saved_tensors
        if dy is None:      # only the shortcut output was used
            return dalias, None, None, None, None, None, None, None
        dy = dy.contiguous()
        add_in = dalias.contiguous() if dalias is not None else None
        N, H, W, Cp = x.shape
        P = N * H * W
        st = nv.stream_ptr()
        stat = mod._stat_buffer(x.device, Cp)
        if not ctx.training:
            raise NotImplementedError("backward through eval-mode BatchNorm is not supported")
        gg, acc_g = grad_buffer(gamma)
        gb, acc_b = grad_buffer(beta)
        dx = torch.empty_like(x)
        dres = None
        if ctx.Cres_p and ctx.needs_input_grad[3]:
            dres = torch.empty((N, H, W, ctx.Cres_p), dtype=x.dtype, device=x.device)
        act = act_of(x)
        fits = mod._fused_bwd_fits.get((P, Cp, act))
        if fits is None:
            fits = mod._fused_bwd_fits[(P, Cp, act)] = bool(nv.lib().dtdl_bn_bwd_fused_fits(P, Cp, act))
        if fits:
            # one cooperative kernel: dy / x / y cross L2 -> SM once, not twice (csrc/ops/nn_ops.cu)
            nv.call("bn_bwd_fused", nv.ptr(dy), nv.ptr(x), nv.ptr(y), nv.ptr(save[0]), nv.ptr(save[1]),
                    nv.ptr(gamma), stat[2 * Cp:].data_ptr(), nv.ptr(dx), nv.ptr(dres), ctx.Cres_p,
                    nv.ptr(gg), nv.ptr(gb), 1 if (acc_g or acc_b) else 0, P, mod.num_features, Cp,
                    1 if ctx.relu else 0, nv.ptr(stat), 2 * Cp, nv.ptr(add_in),
                    nv.ptr(mod._grid_barrier(x.device)), act, st)
        else:
            nv.call("bn_bwd_add", nv.ptr(dy), nv.ptr(x), nv.ptr(y), nv.ptr(save[0]), nv.ptr(save[1]),
                    nv.ptr(gamma), stat[2 * Cp:].data_ptr(), nv.ptr(dx), nv.ptr(dres), ctx.Cres_p,
                    nv.ptr(gg), nv.ptr(gb), 1 if (acc_g or acc_b) else 0, P, mod.num_features, Cp,
                    1 if ctx.relu else 0, nv.ptr(stat), 2 * Cp, nv.ptr(add_in), act, st)
        mod._fwd_dirty = False
        _grad_ready(gamma)
        _grad_ready(beta)
        return dx, None, None, dres, None, None, None, None


def batch_norm(x, mod, residual=None, relu=False, stats_for=None, fork_input=False):
    """``stats_for``: the training-mode BatchNorm2d that reads this layer's output next; its batch
    statistics are then accumulated while the output is written.  ``fork_input``: returns
    ``(y, x)`` where the second item stands for the input on the identity-shortcut path."""
    return _BatchNorm.apply(x, mod.weight, mod.bias, residual, mod, relu, stats_for, fork_input)


# --------------------------------------------------------------------------------------
# pooling
# --------------------------------------------------------------------------------------
class _Pool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad, mode, ceil_mode):
        x = x.contiguous()
        N, H, W, Cp = x.shape

        def osz(n):
            num = n + 2 * pad - k
            o = (-(-num // stride) if ceil_mode else num // stride) + 1
            if ceil_mode and (o - 1) * stride >= n + pad:
                o -= 1
            return o

        Ho, Wo = osz(H), osz(W)
        y = torch.empty((N, Ho, Wo, Cp), dtype=x.dtype, device=x.device)
        nv.call("pool_fwd", nv.ptr(x), nv.ptr(y), N, H, W, Ho, Wo, Cp, k, stride, pad, mode, act_of(x),
                nv.stream_ptr())
        ctx.meta = (k, stride, pad, mode)
        ctx.save_for_backward(x, y)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y = ctx.saved_tensors
        k, stride, pad, mode = ctx.meta
        dy = dy.contiguous()
        N, H, W, Cp = x.shape
        dx = torch.empty_like(x)
        nv.call("pool_bwd", nv.ptr(dy), nv.ptr(x), nv.ptr(y), nv.ptr(dx), N, H, W, y.shape[1],
                y.shape[2], Cp, k, stride, pad, mode, act_of(x), nv.stream_ptr())
        return dx, None, None, None, None, None


def avg_pool(x, k, stride=None, pad=0, ceil_mode=False):
    return _Pool.apply(x, k, stride or k, pad, 0, ceil_mode)


def max_pool(x, k, stride=None, pad=0, ceil_mode=False):
    return _Pool.apply(x, k, stride or k, pad, 1, ceil_mode)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        N, H, W, Cp = x.shape
        y = torch.empty((N, 1, 1, Cp), dtype=x.dtype, device=x.device)
        nv.call("gap_fwd", nv.ptr(x), nv.ptr(y), N, H * W, Cp, act_of(x), nv.stream_ptr())
        ctx.shape = (N, H, W, Cp)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, H, W, Cp = ctx.shape
        dy = dy.contiguous()
        dx = torch.empty((N, H, W, Cp), dtype=dy.dtype, device=dy.device)
        nv.call("gap_bwd", nv.ptr(dy), nv.ptr(dx), N, H * W, Cp, act_of(dy), nv.stream_ptr())
        return dx


def global_avg_pool(x):
    return _GlobalAvgPool.apply(x)


class _Flatten(torch.autograd.Function):
    """NHWC [N,H,W,Cp] -> [N,1,1,pad8(H*W*C)] in NHWC feature order (Keras ``Flatten``)."""

    @staticmethod
    def forward(ctx, x, C):
        N, H, W, Cp = x.shape
        F = H * W * C
        Fp = pad8(F)
        ctx.meta = (N, H, W, C, Cp, F, Fp)
        y = x.new_zeros((N, 1, 1, Fp))
        y.view(N, Fp)[:, :F] = x[..., :C].reshape(N, F)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, H, W, C, Cp, F, Fp = ctx.meta
        dx = dy.new_zeros((N, H, W, Cp))
        dx[..., :C] = dy.view(N, Fp)[:, :F].reshape(N, H, W, C)
        return dx, None


def flatten_nhwc(x, C):
    return _Flatten.apply(x, C)


# --------------------------------------------------------------------------------------
# softmax cross entropy (+ accuracy count), fp32 logits
# --------------------------------------------------------------------------------------
class _SoftmaxCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, stats_out):
        logits = logits.contiguous().float()
        N, C = logits.shape
        out = stats_out if stats_out is not None else torch.zeros(2, device=logits.device)
        dl = torch.empty_like(logits)
        nv.call("softmax_ce", nv.ptr(logits), nv.ptr(targets.contiguous()), nv.ptr(out), 0, nv.ptr(dl),
                N, C, C, 1.0 / N, 1.0 / N, nv.stream_ptr())
        ctx.save_for_backward(dl)
        ctx.mark_non_differentiable(out)
        return out[0].clone(), out

    @staticmethod
    def backward(ctx, dloss, _):
        (dl,) = ctx.saved_tensors
        return dl * dloss, None, None


def softmax_cross_entropy(logits, targets, stats_out=None):
    """Mean cross entropy; also returns a 2-vector [loss, #correct] accumulated on device."""
    return _SoftmaxCE.apply(logits, targets, stats_out)
