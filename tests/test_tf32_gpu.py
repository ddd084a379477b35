"""The fp32-activation ("tf32") mode: every op against plain PyTorch fp32, and whole models against the
CPU fp32 module.

This is the precision the reference trains at: fp32 tensors everywhere, TF32 tensor-core math inside
the convolutions only (torch's cuDNN default; pytorch/single_gpu.py:65-72).  Conv operands are
truncated to TF32 (10 mantissa bits) before both paths so the oracle multiplies the same numbers;
what remains is fp32 summation order, hence the tight tolerances.  The stride-2 cases also cover
the native strided kernels in bf16 mode (TMA element strides forward / weight gradient, gather-form
data gradient).
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _tf(x):
    """Truncate to TF32: what tcgen05 kind::tf32 reads from an fp32 operand."""
    return (x.contiguous().view(torch.int32) & -8192).view(torch.float32)


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def _nchw(y, C):
    return y[..., :C].permute(0, 3, 1, 2).float()


def _rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()


CASES = [
    # N, Cin, Cout, H, W, k, pad, stride
    (4, 3, 16, 32, 32, 3, 1, 1),
    (4, 21, 26, 32, 32, 3, 1, 1),
    (2, 64, 72, 32, 32, 3, 1, 1),
    (4, 101, 106, 32, 32, 3, 1, 2),
    (4, 146, 151, 16, 16, 3, 1, 1),
    (4, 186, 191, 16, 16, 3, 1, 2),
    (8, 266, 271, 8, 8, 3, 1, 1),
    (4, 32, 64, 13, 13, 3, 0, 1),
    (2, 64, 256, 14, 14, 1, 0, 1),
    (2, 64, 128, 14, 14, 1, 0, 2),
    (2, 128, 128, 28, 28, 3, 1, 2),
    (2, 64, 64, 56, 56, 3, 1, 1),
    (2, 24, 40, 15, 15, 3, 1, 2),       # odd size, stride 2
]


@pytest.mark.parametrize("precision", ["tf32", "bf16"])
@pytest.mark.parametrize("case", CASES)
def test_conv_forward_backward(case, precision):
    from distributed_training_dl_b200.ops import Conv2d, functional as fn, set_precision
    N, Cin, Cout, H, W, k, pad, stride = case
    rnd, tol = (_tf, 2e-4) if precision == "tf32" else (_bf, 1e-2)
    torch.manual_seed(0)
    conv = set_precision(Conv2d(Cin, Cout, k, stride=stride, padding=pad, bias=False).cuda(), precision)
    with torch.no_grad():
        conv.weight.copy_(rnd(conv.weight))
    x = rnd(torch.randn(N, Cin, H, W, device="cuda"))
    xn = fn.to_nhwc(x, dtype=fn.act_dtype(precision)).requires_grad_(True)
    y = conv(xn)
    assert y.dtype == xn.dtype
    ref_x = x.double().requires_grad_(True)
    ref_w = conv.weight.detach().double().requires_grad_(True)
    ref_y = F.conv2d(ref_x, ref_w, stride=stride, padding=pad)
    assert y.shape[1:3] == ref_y.shape[2:]
    assert _rel(_nchw(y, Cout), ref_y.float()) < tol
    if y.shape[-1] > Cout:
        assert y[..., Cout:].abs().max().item() == 0

    dy = rnd(torch.randn(ref_y.shape, device="cuda"))
    ref_y.backward(dy.double())
    y.backward(fn.to_nhwc(dy, dtype=xn.dtype))
    assert _rel(_nchw(xn.grad, Cin), ref_x.grad.float()) < tol
    if xn.grad.shape[-1] > Cin:
        assert xn.grad[..., Cin:].abs().max().item() == 0
    assert _rel(conv.weight.grad, ref_w.grad.float()) < tol


@pytest.mark.parametrize("case", [(4, 21, 26, 32, 32, 3, 1, 1), (8, 266, 271, 8, 8, 3, 1, 1),
                                  (4, 101, 106, 32, 32, 3, 1, 2), (2, 64, 256, 14, 14, 1, 0, 1)])
def test_conv_epilogue_bn_statistics_tf32(case):
    from distributed_training_dl_b200.ops import BatchNorm2d, Conv2d, functional as fn, set_precision
    N, Cin, Cout, H, W, k, pad, stride = case
    torch.manual_seed(1)
    conv = set_precision(Conv2d(Cin, Cout, k, stride=stride, padding=pad, bias=False).cuda(), "tf32")
    with torch.no_grad():
        conv.weight.copy_(_tf(conv.weight))
    bn_a, bn_b, ref = BatchNorm2d(Cout).cuda(), BatchNorm2d(Cout).cuda(), torch.nn.BatchNorm2d(Cout).cuda()
    for step in range(2):
        x = _tf(torch.randn(N, Cin, H, W, device="cuda") + 0.3 * step)
        xn = fn.to_nhwc(x, dtype=torch.float32)
        ya = bn_a(conv(xn, stats_for=bn_a), relu=True)
        assert not bn_a._stats_from_producer
        yb = bn_b(conv(xn), relu=True)
        ry = F.relu(ref(F.conv2d(x.double(), conv.weight.double(), stride=stride, padding=pad).float()))
        assert _rel(ya, yb) < 1e-4
        assert _rel(_nchw(ya, Cout), ry) < 1e-3
        assert torch.allclose(bn_a.running_var, ref.running_var, rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("N,Fin,Fout", [(64, 784, 1000), (100, 1000, 10), (7, 271, 10), (256, 576, 64)])
def test_linear_tf32_mode(N, Fin, Fout):
    """Small layers run in full fp32 on CUDA cores, large ones through the TF32 tensor-core path."""
    from distributed_training_dl_b200.ops import Linear, functional as fn, set_precision
    torch.manual_seed(0)
    lin = set_precision(Linear(Fin, Fout, activation="relu").cuda(), "tf32")
    small = Fin * Fout <= fn.LINEAR_F32_MAX_WEIGHTS
    rnd = (lambda t: t) if small else _tf
    with torch.no_grad():
        lin.weight.copy_(rnd(lin.weight))
    x = rnd(torch.randn(N, Fin, device="cuda"))
    xn = fn.to_nhwc(x, dtype=torch.float32).requires_grad_(True)
    y = lin(xn)
    rx = x.double().requires_grad_(True)
    rw = lin.weight.detach().double().requires_grad_(True)
    rb = lin.bias.detach().double().requires_grad_(True)
    ry = F.relu(F.linear(rx, rw, rb))
    assert _rel(y.view(N, -1)[:, :Fout], ry.float()) < 2e-4
    dy = rnd(torch.randn(ry.shape, device="cuda"))
    ry.backward(dy.double())
    y.backward(fn.to_nhwc(dy, dtype=torch.float32))
    # the ReLU-masked dy is not TF32-exact on the big path: allow the truncation error there
    tol = 2e-4 if small else 2e-3
    assert _rel(xn.grad.view(N, -1)[:, :Fin], rx.grad.float()) < tol
    assert _rel(lin.weight.grad, rw.grad.float()) < tol
    assert _rel(lin.bias.grad, rb.grad.float()) < 2e-4


@pytest.mark.parametrize("C,Cres,relu,HW,N", [(21, 16, False, 32, 8), (26, 0, True, 16, 8), (271, 271, True, 8, 16)])
def test_batchnorm_tf32_mode(C, Cres, relu, HW, N):
    from distributed_training_dl_b200.ops import BatchNorm2d, functional as fn
    torch.manual_seed(0)
    bn = BatchNorm2d(C).cuda()
    ref = torch.nn.BatchNorm2d(C).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
        ref.weight.copy_(bn.weight)
        ref.bias.copy_(bn.bias)
    x = torch.randn(N, C, HW, HW, device="cuda") * 2 + 0.5
    res = torch.randn(N, Cres, HW, HW, device="cuda") if Cres else None
    xn = fn.to_nhwc(x, dtype=torch.float32).requires_grad_(True)
    rn = fn.to_nhwc(res, dtype=torch.float32).requires_grad_(True) if Cres else None
    nxt = BatchNorm2d(C).cuda()
    y, xs = bn(xn, residual=rn, relu=relu, stats_for=nxt, fork_input=True)
    assert y.dtype == torch.float32
    rx = x.clone().requires_grad_(True)
    rr = res.clone().requires_grad_(True) if Cres else None
    ry = ref(rx)
    if Cres:
        ry = ry + F.pad(rr, (0, 0, 0, 0, 0, C - Cres))
    if relu:
        ry = F.relu(ry)
    assert _rel(_nchw(y, C), ry) < 1e-4
    assert torch.allclose(bn.running_var, ref.running_var, rtol=1e-4, atol=1e-5)
    # statistics handed to the next BatchNorm == its own pass
    with torch.no_grad():
        za = nxt(y.detach())
        zb = BatchNorm2d(C).cuda()(y.detach())
    assert _rel(za, zb) < 1e-4
    dy, dz = torch.randn_like(ry), torch.randn_like(x)
    ry.backward(dy)
    torch.autograd.backward([y, xs], [fn.to_nhwc(dy, dtype=torch.float32), fn.to_nhwc(dz, dtype=torch.float32)])
    assert _rel(_nchw(xn.grad, C), rx.grad + dz) < 1e-3
    assert _rel(bn.weight.grad, ref.weight.grad) < 1e-3
    assert _rel(bn.bias.grad, ref.bias.grad) < 1e-3
    if Cres:
        assert _rel(_nchw(rn.grad, Cres), rr.grad) < 1e-5


@pytest.mark.parametrize("mode,k,s,p,H", [("avg", 2, 2, 0, 32), ("max", 3, 2, 1, 16), ("avg", 8, 1, 0, 8)])
def test_pooling_tf32_mode(mode, k, s, p, H):
    from distributed_training_dl_b200.ops import AvgPool2d, MaxPool2d, functional as fn
    torch.manual_seed(0)
    C, N = 21, 4
    x = torch.randn(N, C, H, H, device="cuda")
    xn = fn.to_nhwc(x, dtype=torch.float32).requires_grad_(True)
    y = (AvgPool2d if mode == "avg" else MaxPool2d)(k, stride=s, padding=p)(xn)
    rx = x.clone().requires_grad_(True)
    ry = (F.avg_pool2d if mode == "avg" else F.max_pool2d)(rx, k, s, p)
    assert _rel(_nchw(y, C), ry) < 1e-5
    dy = torch.randn_like(ry)
    ry.backward(dy)
    y.backward(fn.to_nhwc(dy, dtype=torch.float32).view_as(y))
    assert _rel(_nchw(xn.grad, C), rx.grad) < 1e-5


def _grad_cosines(cpu, gpu):
    out = {}
    for (n, a), (_, b) in zip(cpu.named_parameters(), gpu.named_parameters()):
        assert b.grad is not None, n
        ga, gb = a.grad.flatten().double(), b.grad.cpu().flatten().double()
        if ga.norm() < 1e-5:
            continue
        out[n] = ((torch.dot(ga, gb) / (ga.norm() * gb.norm() + 1e-30)).item(), (gb.norm() / ga.norm()).item())
    return out


def test_models_match_cpu_oracle_tf32():
    """Whole models in tf32 mode vs the same module on CPU fp32 -- every parameter gradient,
    BatchNorm scales and shifts included, for the full 51-block PyramidNet too (the bf16 mode can
    only promise 0.7 there; fp32 storage removes the activation-rounding noise)."""
    import copy
    from distributed_training_dl_b200.models import MLP, MnistCNN, PyramidNet, pyramidnet
    from distributed_training_dl_b200.ops import set_precision
    torch.manual_seed(0)
    shallow = lambda: PyramidNet(num_layers=3, alpha=45.0)  # noqa: E731
    for make, shape, tol, min_cos in [(lambda: MLP(128), (32, 784), 2e-3, 0.999),
                                      (MnistCNN, (16, 1, 28, 28), 2e-3, 0.99),
                                      (shallow, (8, 3, 32, 32), 5e-3, 0.99),
                                      (pyramidnet, (8, 3, 32, 32), 2e-2, 0.95)]:
        cpu = make()
        gpu = set_precision(copy.deepcopy(cpu).cuda(), "tf32")
        x = torch.randn(*shape)
        t = torch.randint(0, 10, (shape[0],))
        lc = F.cross_entropy(cpu(x), t)
        lg = F.cross_entropy(gpu(x.cuda()), t.cuda())
        assert abs(lc.item() - lg.item()) < tol * max(1.0, abs(lc.item())), (lc.item(), lg.item())
        lc.backward()
        lg.backward()
        cos = _grad_cosines(cpu, gpu)
        worst = min(cos.items(), key=lambda kv: kv[1][0])
        print(f"tf32 oracle {type(cpu).__name__}: loss {lc.item():.5f} vs {lg.item():.5f}; "
              f"worst gradient cosine {worst[1][0]:.5f} ({worst[0]})")
        for n, (c, ratio) in cos.items():
            assert c > min_cos, (n, c)
            assert 0.9 < ratio < 1.1, (n, ratio)


def test_train_step_tf32_graph_matches_eager():
    """wrap(dtype='tf32') + FusedSGD + TrainStep: the captured graph reproduces eager steps."""
    import copy
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import PyramidNet
    torch.manual_seed(0)
    dtdl.init(backend="b200")
    try:
        base = PyramidNet(num_layers=3, alpha=45.0)
        xs = [torch.randn(16, 3, 32, 32).pin_memory() for _ in range(6)]
        ts = [torch.randint(0, 10, (16,)).pin_memory() for _ in range(6)]
        losses = {}
        for graph in (False, True):
            net = dtdl.wrap(copy.deepcopy(base).cuda(), dtype="tf32")
            opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
            step = dtdl.TrainStep(net, opt, graph=graph, warmup=2)
            losses[graph] = [step(x, t).item() for x, t in zip(xs, ts)]
        for a, b in zip(losses[False], losses[True]):
            assert abs(a - b) < 2e-2 * max(1.0, abs(a)), (losses[False], losses[True])
        assert losses[True][-1] == losses[True][-1]
    finally:
        dtdl.shutdown()


def test_precision_switch_between_steps_rebuilds_packs():
    """set_precision on a live, wrapped model: the operand packs are re-allocated in the new type and the
    cached pack tables (keyed by the addresses they describe) are rebuilt, so eager training goes on."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import PyramidNet
    torch.manual_seed(0)
    dtdl.init(backend="b200")
    try:
        net = dtdl.wrap(PyramidNet(num_layers=3, alpha=45.0).cuda())
        opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9)
        x = torch.randn(16, 3, 32, 32, device="cuda")
        t = torch.randint(0, 10, (16,), device="cuda")
        losses = []
        for prec in ("bf16", "bf16", "tf32", "tf32", "bf16", "bf16"):
            dtdl.set_precision(net, prec)
            loss = F.cross_entropy(net(x), t)
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.item())
        torch.cuda.synchronize()
        assert all(l == l for l in losses), losses
        assert losses[-1] < losses[0], losses          # same batch every step: the loss must fall
    finally:
        dtdl.shutdown()
