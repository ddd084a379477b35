"""Development aid: run the op tests one by one and print error patterns (GPU box)."""
import sys, traceback
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from distributed_training_dl_b200.ops import Conv2d, functional as fn

def bf(x): return x.to(torch.bfloat16).float()

def conv_case(N, Cin, Cout, H, W, k, pad, stride):
    torch.manual_seed(0)
    conv = Conv2d(Cin, Cout, k, stride=stride, padding=pad, bias=False).cuda()
    with torch.no_grad(): conv.weight.copy_(bf(conv.weight))
    x = bf(torch.randn(N, Cin, H, W, device="cuda"))
    xn = fn.to_nhwc(x).requires_grad_(True)
    y = conv(xn); torch.cuda.synchronize()
    rx = x.clone().requires_grad_(True); rw = conv.weight.detach().clone().requires_grad_(True)
    ry = F.conv2d(rx, rw, stride=stride, padding=pad)
    yy = y[..., :Cout].permute(0, 3, 1, 2).float()
    err = (yy - ry).abs()
    print(f"case {(N,Cin,Cout,H,W,k,pad,stride)} fwd max_err {err.max().item():.4g} ref_max {ry.abs().max().item():.4g} "
          f"bad_frac {(err > 0.05 * ry.abs().max()).float().mean().item():.4f}")
    if err.max() > 0.05 * ry.abs().max():
        bad = (err > 0.05 * ry.abs().max()).nonzero()[:8]
        for b in bad: print("   bad idx (n,c,h,w)", b.tolist(), "got", yy[tuple(b)].item(), "ref", ry[tuple(b)].item())
        eb = (err > 0.05 * ry.abs().max())
        print("   bad per-channel:", eb.float().mean(dim=(0, 2, 3))[:32].tolist())
        print("   bad per-row(h):", eb.float().mean(dim=(0, 1, 3))[:32].tolist())
        print("   bad per-n:", eb.float().mean(dim=(1, 2, 3))[:32].tolist())
    dy = bf(torch.randn_like(ry)); ry.backward(dy)
    y.backward(fn.to_nhwc(dy)); torch.cuda.synchronize()
    gx = xn.grad[..., :Cin].permute(0, 3, 1, 2).float()
    e1 = (gx - rx.grad).abs().max().item() / (rx.grad.abs().max().item() + 1e-9)
    e2 = (conv.weight.grad - rw.grad).abs().max().item() / (rw.grad.abs().max().item() + 1e-9)
    print(f"     dgrad rel_err {e1:.4g}  wgrad rel_err {e2:.4g}")
    if e2 > 0.02:
        eb = ((conv.weight.grad - rw.grad).abs() > 0.02 * rw.grad.abs().max())
        print("   wgrad bad per-co:", eb.float().mean(dim=(1, 2, 3))[:40].tolist())
        print("   wgrad bad per-ci:", eb.float().mean(dim=(0, 2, 3))[:40].tolist())
        print("   wgrad bad per-tap:", eb.float().mean(dim=(0, 1)).tolist())
        print("   sample got/ref:", conv.weight.grad.flatten()[:6].tolist(), rw.grad.flatten()[:6].tolist())

if __name__ == "__main__":
    cases = [(2, 16, 16, 8, 8, 1, 0, 1), (4, 64, 64, 8, 8, 1, 0, 1), (4, 21, 26, 32, 32, 3, 1, 1), (4, 101, 106, 32, 32, 3, 1, 2),
             (8, 266, 271, 8, 8, 3, 1, 1), (4, 1, 32, 28, 28, 3, 0, 1), (3, 300, 520, 7, 7, 1, 0, 1)]
    for c in cases:
        try:
            conv_case(*c)
        except Exception:
            traceback.print_exc()
            break
