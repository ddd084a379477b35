"""ncu target: ONE launch of forward, data-gradient and weight-gradient convolution per layer shape and
precision, plus the one-kernel BatchNorm backward, for the per-stage captures under profiles/ncu/.

    python tools/ncu_shapes.py && ncu --set full --clock-control none --import-source on \
        -k regex:"conv_gemm|conv_wgrad_kernel|bn_bwd_fused" -c 60 -o gpurun_out/stages python tools/ncu_shapes.py

Shapes: one PyramidNet layer per stage (batch 64), ResNet-50's 56x56x64 3x3 and 14x14 1x1 1024->256.
"""
import sys

import torch

sys.path.insert(0, ".")
from distributed_training_dl_b200 import _native as nv  # noqa: E402
from distributed_training_dl_b200.ops import BatchNorm2d, Conv2d, functional as fn  # noqa: E402

SHAPES = [  # cin, cout, hw, N, k
    (56, 61, 32, 64, 3), (146, 151, 16, 64, 3), (266, 271, 8, 64, 3),
    (64, 64, 56, 64, 3), (1024, 256, 14, 64, 1)]
precs = sys.argv[1:] or ["bf16", "tf32"]
for prec in precs:
    dt = fn.act_dtype(prec)
    act = fn.act_of(dt)
    for cin, cout, hw, N, k in SHAPES:
        pad = k // 2
        conv = Conv2d(cin, cout, k, padding=pad, bias=False).cuda()
        conv._ensure_packed(dt)
        x = torch.randn(N, hw, hw, conv.Cin_p, device="cuda").to(dt)
        dy = torch.randn(N, hw, hw, conv.Cout_p, device="cuda").to(dt)
        gw = torch.zeros_like(conv.weight)
        need = nv.lib().dtdl_conv_wgrad_workspace_bytes_ex(N, hw, hw, conv.Cin_p, conv.Cout_p, k, k, 1, act)
        ws = fn.scratch(x.device, need)
        stat = torch.zeros(2 * conv.Cout_p, device="cuda")
        torch.cuda.synchronize()
        fn.conv_fprop_raw(x, conv.wf, conv.Cout_p, k, k, pad, stats=stat.data_ptr())        # forward + BN statistics
        fn.conv_fprop_raw(dy, conv.wd, conv.Cin_p, k, k, k - 1 - pad)                        # data gradient
        nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, hw, hw, cin, conv.Cin_p,
                hw, hw, cout, conv.Cout_p, k, k, pad, pad, 0, 1, act, nv.stream_ptr())
        bn = BatchNorm2d(cout).cuda()
        xb = dy.clone().requires_grad_(True)
        yb = bn(xb, relu=True)
        yb.backward(dy)                                                                      # BatchNorm backward
        torch.cuda.synchronize()
        print("ok", prec, (cin, cout, hw, N, k), flush=True)
