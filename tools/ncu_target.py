"""Small single-GPU program for ncu captures: a few launches of each hot kernel on one
PyramidNet stage-1 shape (batch 64, 32x32, 56->61 channels).

    python tools/ncu_target.py [cin cout hw N [precision [stride]]]      precision: bf16 | tf32"""
import sys
import torch
sys.path.insert(0, ".")
from distributed_training_dl_b200 import _native as nv
from distributed_training_dl_b200.ops import Conv2d, functional as fn

cin, cout, hw, N = (int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (56, 61, 32, 64)))
prec = sys.argv[5] if len(sys.argv) > 5 else "bf16"
stride = int(sys.argv[6]) if len(sys.argv) > 6 else 1
dt = fn.act_dtype(prec)
act = fn.act_of(dt)
conv = Conv2d(cin, cout, 3, padding=1, stride=stride, bias=False).cuda()
conv._ensure_packed(dt)
ho = (hw + 2 - 3) // stride + 1
x = torch.randn(N, hw, hw, conv.Cin_p, device="cuda").to(dt)
dy = torch.randn(N, ho, ho, conv.Cout_p, device="cuda").to(dt)
gw = torch.zeros_like(conv.weight)
need = nv.lib().dtdl_conv_wgrad_workspace_bytes_ex(N, ho, ho, conv.Cin_p, conv.Cout_p, 3, 3, stride, act)
ws = fn.scratch(x.device, need)
for _ in range(4):
    y = fn.conv_fprop_raw(x, conv.wf, conv.Cout_p, 3, 3, 1, stride=stride)
    nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, hw, hw, cin, conv.Cin_p,
            ho, ho, cout, conv.Cout_p, 3, 3, 1, 1, 0, stride, act, nv.stream_ptr())
# the same convolution with the BatchNorm statistics epilogue (what training runs), twice
stat = torch.zeros(2 * conv.Cout_p, device="cuda")
for _ in range(2):
    y = fn.conv_fprop_raw(x, conv.wf, conv.Cout_p, 3, 3, 1, stats=stat.data_ptr(), stride=stride)
# bandwidth-bound kernels on the same shape: BN forward (stats + apply) and backward, fused SGD
from distributed_training_dl_b200.ops import BatchNorm2d
bn = BatchNorm2d(cout).cuda()
xb = dy.clone().requires_grad_(True)
for _ in range(2):
    yb = bn(xb, relu=True)
    yb.backward(dy)
nxt = BatchNorm2d(cout).cuda()
with torch.no_grad():
    for _ in range(2):
        nxt(bn(xb, relu=True, stats_for=nxt))       # bn_apply_stats_kernel
# weight pack of a whole PyramidNet (one launch)
from distributed_training_dl_b200.models import pyramidnet
from distributed_training_dl_b200.ops.modules import pack_weights, packed_layers
net = pyramidnet().cuda()
cache = {}
for _ in range(2):
    pack_weights(packed_layers(net), cache)
p = torch.randn(1 << 24, device="cuda"); g = torch.randn_like(p); m = torch.zeros_like(p)
hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0] + [0.0] * 12, device="cuda")
for _ in range(2):
    nv.call("local_update", nv.ptr(p), nv.ptr(g), nv.ptr(m), 0, nv.ptr(hyper), p.numel(), 1, nv.stream_ptr())
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
