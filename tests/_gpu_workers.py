"""Worker bodies for the multi-GPU tests (one process per GPU, peer-memory backend)."""
import copy
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _ok(outdir, i):
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")


def _init_nvls(dtdl, outdir, i):
    """init() with the multicast collectives forced on; None (and a skip marker) where the box has
    no NVSwitch multicast support -- the forced mode refuses to start there."""
    os.environ["DTDL_NVLS"] = "1"
    try:
        comm = dtdl.init(backend="b200")
    except Exception as e:  # noqa: BLE001 - only the "not available" refusal is a skip
        if "cannot join an NVLS multicast team" not in str(e):
            raise
        open(os.path.join(outdir, f"skip_{i}"), "w").write(str(e))
        return None
    assert comm.peer.use_nvls
    return comm


def collectives(i, world, outdir, nvls=False):
    import distributed_training_dl_b200 as dtdl
    comm = _init_nvls(dtdl, outdir, i) if nvls else dtdl.init(backend="b200")
    if comm is None:
        return
    peer = comm.peer
    dev = comm.device
    assert comm.rank == i and peer.world == world
    n = 1 << 20
    buf = peer.alloc(n * 4)
    out = peer.alloc(n * 4)
    x = buf.view(torch.float32, 0, n)
    y = out.view(torch.float32, 0, n)
    cases = [(256, 0), (4096, 0), (4096, 1), (n, 0), (n, 1), (n - 8, 1)]
    if nvls:    # 2 = multicast variant; the others then run the peer-pointer kernels on cuMem memory
        assert buf.mc_ptr and out.mc_ptr
        cases = [(256, 2), (4096, 2), (n, 2), (n - 8, 2), (4096, 0), (n, 1)]
    for numel, two_shot in cases:
        for rep in range(3):   # repeated iterations: cross-GPU races show up as flaky mismatches
            g = torch.Generator(device="cpu").manual_seed(100 * rep + numel)
            full = torch.randn(world, numel, generator=g)
            x[:numel].copy_(full[i])
            y.zero_()
            peer.allreduce(buf, 0, numel, scale=1.0 / world, two_shot=two_shot, dst=out)
            torch.cuda.synchronize()
            want = full.sum(0) / world
            assert torch.allclose(y[:numel].cpu(), want, atol=1e-5), (numel, two_shot, rep)
            # in place (never the one-shot schedule: one-shot may not overwrite its source)
            peer.allreduce(buf, 0, numel, scale=1.0, two_shot=two_shot)
            torch.cuda.synchronize()
            assert torch.allclose(x[:numel].cpu(), full.sum(0), atol=1e-5), ("inplace", numel, two_shot)
            comm.barrier()
    # bit-identical results on every rank (fixed reduction order)
    x.copy_(torch.randn(n, generator=torch.Generator().manual_seed(i)))
    peer.allreduce(buf, 0, n, two_shot=2 if nvls else 1)
    torch.cuda.synchronize()
    sums = comm.allgather_object(x[:4096].cpu())
    assert all(torch.equal(sums[0], s) for s in sums)
    # broadcast + all-gather
    x.fill_(float(i + 1))
    peer.broadcast(buf, 0, n * 4, root=world - 1)
    torch.cuda.synchronize()
    assert float(x.min()) == float(x.max()) == float(world)
    per = n // world
    x.zero_()
    x[i * per:(i + 1) * per] = i + 1
    peer.allgather(buf, 0, n * 4)
    torch.cuda.synchronize()
    for r in range(world):
        assert float(x[r * per]) == r + 1 and float(x[(r + 1) * per - 1]) == r + 1
    # fused all-reduce + SGD / Adam against the unfused torch arithmetic
    for opt_code, two_shot in ([(1, 2), (2, 2), (3, 2), (1, 0)] if nvls else [(1, 0), (1, 1), (2, 1), (3, 0)]):
        numel = 1 << 16
        g = torch.Generator().manual_seed(7 + opt_code)
        grads = torch.randn(world, numel, generator=g)
        p0 = torch.randn(numel, generator=g)
        params = peer.alloc(numel * 4)
        p = params.view(torch.float32, 0, numel)
        p.copy_(p0)
        x[:numel].copy_(grads[i])
        m = torch.zeros(numel, device=dev)
        v = torch.zeros(numel, device=dev)
        lr, mu, wd, b1, b2, eps = 0.1, 0.9, 1e-2, 0.9, 0.999, 1e-8
        hyper = torch.tensor([lr, mu, wd, 1.0 / world, b1, b2, eps, 1 - b1, (1 - b2) ** 0.5, 0.0] + [0.0] * 6, device=dev)
        if two_shot:
            ms, vs = peer.alloc(numel * 4), peer.alloc(numel * 4)
            m, v = ms.view(torch.float32, 0, numel), vs.view(torch.float32, 0, numel)
        peer.reduce_update(buf, params, m, v, hyper, 0, numel, opt_code, two_shot=two_shot)
        torch.cuda.synchronize()
        gavg = grads.mean(0) + wd * p0
        if opt_code == 1:
            want = p0 - lr * gavg
        else:
            m1, v1 = (1 - b1) * gavg, (1 - b2) * gavg * gavg
            if opt_code == 2:
                want = p0 - lr / (1 - b1) * m1 / (v1.sqrt() / (1 - b2) ** 0.5 + eps)
            else:
                want = p0 - lr * (1 - b2) ** 0.5 / (1 - b1) * m1 / (v1.sqrt() + eps)
        assert torch.allclose(p.cpu(), want, atol=2e-5), (opt_code, two_shot, (p.cpu() - want).abs().max())
        allp = comm.allgather_object(p[:2048].cpu())
        assert all(torch.equal(allp[0], q) for q in allp)
    _ok(outdir, i)
    dtdl.shutdown()


def collectives_sweep(i, world, outdir, nvls=False):
    """every all-reduce variant x message sizes from 1 KiB up x >= 10 repetitions,
    result == the exact sum on every rank.  Inputs are small integers stored as fp32, so every partial
    sum is exact whatever the order of additions: equality is BITWISE, which also shows that all ranks
    hold identical bits, and a cross-GPU race (a rank reading a peer's buffer too early / late) shows
    up as a wrong integer instead of hiding inside a tolerance.  DTDL_TEST_MAX_BYTES caps the sweep
    (default 64 MiB; the 8-GPU evidence run sets 1 GiB)."""
    import distributed_training_dl_b200 as dtdl
    comm = _init_nvls(dtdl, outdir, i) if nvls else dtdl.init(backend="b200")
    if comm is None:
        return
    peer, dev = comm.peer, comm.device
    max_bytes = int(os.environ.get("DTDL_TEST_MAX_BYTES", 64 << 20))
    sizes = [s for s in (1 << 10, 64 << 10, 4 << 20, 64 << 20, 1 << 30) if s <= max_bytes]
    reps = int(os.environ.get("DTDL_TEST_REPS", "10"))
    nmax = max(sizes) // 4
    buf, out = peer.alloc(nmax * 4), peer.alloc(nmax * 4)
    x, y = buf.view(torch.float32, 0, nmax), out.view(torch.float32, 0, nmax)
    idx = torch.arange(nmax, device=dev, dtype=torch.int64)

    def pattern(rank, rep, n):
        return (((idx[:n] * 7 + rank * 13 + rep * 5) % 1024) - 512).float()

    variants = [("nvls", 2)] if nvls else [("one-shot", 0), ("two-shot", 1)]
    for name, two_shot in variants:
        for nbytes in sizes:
            n = nbytes // 4
            for rep in range(reps):
                want = sum(pattern(r, rep, n) for r in range(world))
                x[:n].copy_(pattern(i, rep, n))
                y[:n].fill_(-1.0)
                peer.allreduce(buf, 0, n, scale=1.0, two_shot=two_shot, dst=out)
                torch.cuda.synchronize()
                assert torch.equal(y[:n], want), (name, nbytes, rep, "out of place")
                peer.allreduce(buf, 0, n, scale=1.0, two_shot=two_shot)      # in place
                torch.cuda.synchronize()
                assert torch.equal(x[:n], want), (name, nbytes, rep, "in place")
                comm.barrier()
            if i == 0:
                print(f"[collectives_sweep] world {world} {name:8s} {nbytes:>11d} B x {reps} reps: exact on every rank",
                      flush=True)
    # fused all-reduce + SGD update at every size: equal to the unfused arithmetic, identical bits on all ranks
    for name, two_shot in ([("nvls", 2)] if nvls else [("one-shot", 0), ("two-shot", 1)]):
        for nbytes in sizes:
            n = nbytes // 4
            params = peer.alloc(n * 4)
            ms = peer.alloc(n * 4)
            p, m = params.view(torch.float32, 0, n), ms.view(torch.float32, 0, n)
            lr, mu, wd = 0.125, 0.5, 0.0
            hyper = torch.tensor([lr, mu, wd, 1.0 / world, 0, 0, 0, 1, 1, 0.0] + [0.0] * 6, device=dev)
            p0 = pattern(3, 1, n) / 64
            p.copy_(p0)
            m.zero_()
            x[:n].copy_(pattern(i, 0, n) * world)             # mean over ranks stays an exact integer
            peer.reduce_update(buf, params, m, None, hyper, 0, n, 1, two_shot=two_shot)
            torch.cuda.synchronize()
            gavg = sum(pattern(r, 0, n) for r in range(world))
            want = p0 - lr * gavg
            assert torch.equal(p, want), (name, nbytes, "fused sgd", (p - want).abs().max().item())
            comm.barrier()
            params.free()
            ms.free()
            if i == 0:
                print(f"[collectives_sweep] world {world} fused SGD {name:8s} {nbytes:>11d} B: exact on every rank", flush=True)
    _ok(outdir, i)
    dtdl.shutdown()


def bf16_gradient_transport(i, world, outdir, nvls=False):
    """wrap(..., grad_dtype="bf16"): the kernel-level arithmetic (cast, half-width reduce with fp32
    accumulation, fused SGD, fp32 parameters stored to every rank) and a short training run."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200 import _native as nv
    from distributed_training_dl_b200.models import MLP
    comm = _init_nvls(dtdl, outdir, i) if nvls else dtdl.init(backend="b200")
    if comm is None:
        return
    peer, dev = comm.peer, comm.device
    for numel in (1 << 10, (1 << 20) + 8):
        g = torch.Generator().manual_seed(17)
        grads = torch.randn(world, numel, generator=g)
        p0 = torch.randn(numel, generator=g)
        gf, g16, params, ms = (peer.alloc(numel * 4), peer.alloc(numel * 2), peer.alloc(numel * 4),
                               peer.alloc(numel * 4))
        gf.view(torch.float32, 0, numel).copy_(grads[i])
        p, m = params.view(torch.float32, 0, numel), ms.view(torch.float32, 0, numel)
        p.copy_(p0)
        lr, mu, wd = 0.1, 0.9, 1e-2
        hyper = torch.tensor([lr, mu, wd, 1.0 / world, 0, 0, 0, 1, 1, 0.0] + [0.0] * 6, device=dev)
        nv.call("cast_f32_bf16", gf.local_ptr, g16.local_ptr, numel, nv.stream_ptr())
        assert torch.equal(g16.view(torch.bfloat16, 0, numel).cpu(), grads[i].to(torch.bfloat16))
        peer.reduce_update16(g16, params, m, None, hyper, 0, numel, 1)
        torch.cuda.synchronize()
        gavg = grads.to(torch.bfloat16).float().mean(0)
        want = p0 - lr * (gavg + wd * p0)
        # peer pointers: fp32 sum of the bf16 values, exact up to summation order.  Multicast: the switch
        # returns the fp32-accumulated sum rounded to bf16 (8 mantissa bits) before the 1/W scale.
        in_switch = peer.use_nvls and bool(g16.mc_ptr)      # also when NVLS switched itself on (world >= 4)
        bound = 1e-5 + (lr * 2.0 ** -8 * grads.abs().sum(0).max().item() / world if in_switch else 0.0)
        err = (p.cpu() - want).abs().max().item()
        assert err < bound, (numel, err, bound)
        allp = comm.allgather_object(p[:4096].cpu())
        assert all(torch.equal(allp[0], q) for q in allp), "ranks hold different parameters"
        comm.barrier()
    # training through wrap(): tracks single-process fp32 SGD (bf16 rounding of the gradients only)
    torch.manual_seed(3)
    mlp = MLP(n_units=64)
    ref = copy.deepcopy(mlp)
    net = dtdl.wrap(mlp.to(dev), grad_dtype="bf16")
    assert net.reducer.grad_bf16
    opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    g = torch.Generator().manual_seed(11)
    for it in range(3):
        xs = torch.randn(world, 32, 784, generator=g)
        ts = torch.randint(0, 10, (world, 32), generator=g)
        loss = F.cross_entropy(net(xs[i].to(dev)), ts[i].to(dev))
        opt.zero_grad()
        loss.backward()
        opt.step()
        rl = F.cross_entropy(ref(xs.reshape(-1, 784)), ts.reshape(-1))
        ropt.zero_grad()
        rl.backward()
        ropt.step()
    torch.cuda.synchronize()
    for (n_, q), r in zip(net.module.named_parameters(), ref.parameters()):
        err = (q.detach().cpu() - r).abs().max().item()
        assert err < 1e-2, (n_, err)
    chk = torch.cat([q.detach().flatten()[:64] for q in net.parameters()]).cpu()
    alls = comm.allgather_object(chk)
    assert all(torch.equal(alls[0], a) for a in alls), "replicas diverged"
    sd = opt.state_dict()
    mb = comm.allgather_object(sd["state"][0]["momentum_buffer"].flatten()[:16].cpu())
    assert all(torch.equal(mb[0], q) for q in mb)
    _ok(outdir, i)
    dtdl.shutdown()


def ddp_training(i, world, outdir, use_graph, nvls=False):
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    from distributed_training_dl_b200.models import MLP, PyramidNet
    comm = _init_nvls(dtdl, outdir, i) if nvls else dtdl.init(backend="b200")
    if comm is None:
        return
    dev = comm.device
    torch.manual_seed(1 + i)
    # (a) replicas identical after construction although seeded differently
    net = dtdl.wrap(PyramidNet(num_layers=3, alpha=45.0).to(dev))
    chk = torch.cat([p.detach().flatten()[:32] for p in net.parameters()]).cpu()
    alls = comm.allgather_object(chk)
    assert all(torch.equal(alls[0], a) for a in alls)
    opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=use_graph, warmup=2)
    g = torch.Generator().manual_seed(5 + i)
    losses = []
    for it in range(8):
        x = torch.randn(16, 3, 32, 32, generator=g).pin_memory()
        t = torch.randint(0, 10, (16,), generator=g).pin_memory()
        losses.append(step(x, t).item())
    assert all(l == l for l in losses), losses
    torch.cuda.synchronize()
    # replicas stay bit-identical; running stats follow rank 0
    chk = torch.cat([p.detach().flatten()[:64] for p in net.parameters()]).cpu()
    alls = comm.allgather_object(chk)
    assert all(torch.equal(alls[0], a) for a in alls), "replicas diverged"
    # K-C2 on the peer path: after a step every rank holds rank 0's BatchNorm running statistics
    rs = comm.allgather_object(torch.cat([b.detach().float().flatten() for b in net.module.buffers()]).cpu())
    assert all(torch.equal(rs[0], r) for r in rs), "BatchNorm running statistics differ between ranks"
    # optimizer state in torch format after a sharded (two-shot) update
    sd = opt.state_dict()
    mb = comm.allgather_object(sd["state"][0]["momentum_buffer"].flatten()[:16].cpu())
    assert all(torch.equal(mb[0], q) for q in mb)

    # (c) MLP (no BN): k steps == single-process fp32 run on the concatenated batch
    torch.manual_seed(3)
    mlp = MLP(n_units=64)
    ref = copy.deepcopy(mlp)
    net2 = dtdl.wrap(mlp.to(dev))
    opt2 = dtdl.FusedSGD(net2.parameters(), lr=0.05, momentum=0.9)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    g = torch.Generator().manual_seed(11)
    for it in range(3):
        xs = torch.randn(world, 32, 784, generator=g)
        ts = torch.randint(0, 10, (world, 32), generator=g)
        loss = F.cross_entropy(net2(xs[i].to(dev)), ts[i].to(dev))
        opt2.zero_grad()
        loss.backward()
        opt2.step()
        rl = F.cross_entropy(ref(xs.reshape(-1, 784)), ts.reshape(-1))
        ropt.zero_grad()
        rl.backward()
        ropt.step()
    torch.cuda.synchronize()
    for (n_, p), q in zip(net2.module.named_parameters(), ref.parameters()):
        err = (p.detach().cpu() - q).abs().max().item()
        assert err < 5e-3, (n_, err)
    _ok(outdir, i)
    dtdl.shutdown()


def data_parallel_engine(outdir):
    """Single process driving all visible GPUs (nn.DataParallel contract + graph fast path)."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import PyramidNet
    ndev = torch.cuda.device_count()
    torch.manual_seed(0)
    base = PyramidNet(num_layers=3, alpha=45.0)
    single = copy.deepcopy(base).cuda(0)
    dp = dtdl.DataParallel(base, device_ids=list(range(ndev)))
    assert list(dp.state_dict().keys()) == ["module." + k for k in single.state_dict().keys()]
    x = torch.randn(8 * ndev, 3, 32, 32, device="cuda:0")
    t = torch.randint(0, 10, (8 * ndev,), device="cuda:0")
    # inference contract: logits on GPU 0, equal to the single-GPU model (eval-mode BN)
    dp.eval()
    single.eval()
    with torch.no_grad():
        a, b = dp(x), single(x)
    assert a.device == torch.device("cuda", 0) and a.shape == b.shape
    assert (a - b).abs().max().item() < 2e-2 * max(1.0, b.abs().max().item())
    # eager training through the DataParallel call contract
    dp.train()
    opt = dtdl.FusedSGD(dp.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    for _ in range(2):
        loss = F.cross_entropy(dp(x), t)
        opt.zero_grad()
        loss.backward()
        opt.step()
    for d in range(ndev):
        torch.cuda.synchronize(d)
    assert loss.item() == loss.item()

    def replicas_identical():
        ref = dp.arenas[0].param_flat
        return all(torch.equal(ref.cpu(), a.param_flat.cpu()) for a in dp.arenas[1:])

    assert replicas_identical(), "replicas diverged (eager)"
    del loss  # an autograd graph of the eager path must not be alive while graphs are captured
    # eager call contract on a BN-free model == single-process torch SGD on the whole batch (the
    # replicas' gradients must add up to the gradient of the GLOBAL mean loss, weight decay unscaled);
    # an uneven batch (one sample more than a multiple of the GPU count) exercises the balanced split
    from distributed_training_dl_b200.models import MLP
    torch.manual_seed(4)
    mlp_e = MLP(n_units=64)
    ref_e = copy.deepcopy(mlp_e)
    dpe = dtdl.DataParallel(mlp_e, device_ids=list(range(ndev)))
    opte = dtdl.FusedSGD(dpe.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-2)
    ropte = torch.optim.SGD(ref_e.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-2)
    ge = torch.Generator().manual_seed(12)
    for it in range(3):
        nb = 16 * ndev + (1 if it == 1 else 0)
        xb = torch.randn(nb, 784, generator=ge)
        tb = torch.randint(0, 10, (nb,), generator=ge)
        le = F.cross_entropy(dpe(xb.cuda(0)), tb.cuda(0))
        opte.zero_grad()
        le.backward()
        opte.step()
        rl = F.cross_entropy(ref_e(xb), tb)
        ropte.zero_grad()
        rl.backward()
        ropte.step()
        assert abs(le.item() - rl.item()) < 3e-2, (le.item(), rl.item())
    for d in range(ndev):
        torch.cuda.synchronize(d)
    for (n_, p), q in zip(dpe.module.named_parameters(), ref_e.parameters()):
        err = (p.detach().cpu() - q).abs().max().item()
        assert err < 5e-3, ("eager DataParallel vs torch SGD", n_, err)
    del le
    if ndev > 1:
        try:
            dpe(torch.randn(ndev - 1, 784, device="cuda:0"))
            raise AssertionError("a batch smaller than the GPU count must be rejected")
        except ValueError:
            pass
    # an eager call's autograd graph that is still alive when the step wants to capture is reported by
    # the step itself instead of surfacing as a CUDA capture error
    held = F.cross_entropy(dp(x), t)
    refused = dtdl.DataParallelStep(dp, opt, graph=True, warmup=0)
    try:
        refused(x.cpu().pin_memory(), t.cpu().pin_memory())
        raise AssertionError("capture with a live eager autograd graph must be refused")
    except RuntimeError as e:
        assert "still alive" in str(e), e
    del held, refused
    assert not dp.eager_graph_alive()
    # graph fast path
    step = dtdl.DataParallelStep(dp, opt, graph=True, warmup=2)
    xs, ts = x.cpu().pin_memory(), t.cpu().pin_memory()
    vals = []
    for _ in range(6):
        vals.append(step(xs, ts).cpu())
    for d in range(ndev):
        torch.cuda.synchronize(d)
    assert all(v[0] == v[0] for v in vals), vals
    assert vals[-1][0] < vals[0][0], [float(v[0]) for v in vals]   # same batch: loss goes down
    assert replicas_identical(), "replicas diverged (graph)"
    sd = opt.state_dict()
    assert "momentum_buffer" in sd["state"][0]
    # BN-free model: k graph steps == single-process fp32 SGD on the whole batch
    torch.manual_seed(3)
    mlp = MLP(n_units=64)
    ref = copy.deepcopy(mlp)
    dp2 = dtdl.DataParallel(mlp, device_ids=list(range(ndev)))
    opt2 = dtdl.FusedSGD(dp2.parameters(), lr=0.05, momentum=0.9)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    step2 = dtdl.DataParallelStep(dp2, opt2, graph=True, warmup=1)
    g = torch.Generator().manual_seed(11)
    for it in range(4):
        xb = torch.randn(16 * ndev, 784, generator=g)
        tb = torch.randint(0, 10, (16 * ndev,), generator=g)
        stats = step2(xb.pin_memory(), tb.pin_memory())
        rl = F.cross_entropy(ref(xb), tb)
        ropt.zero_grad()
        rl.backward()
        ropt.step()
        assert abs(float(stats[0]) - rl.item()) < 3e-2, (float(stats[0]), rl.item())
    for d in range(ndev):
        torch.cuda.synchronize(d)
    for (n_, p), q in zip(dp2.module.named_parameters(), ref.parameters()):
        err = (p.detach().cpu() - q).abs().max().item()
        assert err < 5e-3, (n_, err)
    open(os.path.join(outdir, "ok_dp"), "w").write("ok")


def ddp_variants(i, world, outdir, backend):
    """Same bucket logic over other back ends / optimizers: NCCL collectives, a stock torch
    optimizer behind wrap(), and gradient accumulation with no_sync()."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import MLP
    comm = dtdl.init(backend=backend)
    dev = comm.device
    g = torch.Generator().manual_seed(21)
    for opt_kind in ("fused", "torch"):
        torch.manual_seed(5)
        mlp = MLP(n_units=48)
        ref = copy.deepcopy(mlp)
        net = dtdl.wrap(mlp.to(dev))
        if opt_kind == "fused":
            opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9)
        else:
            opt = torch.optim.SGD(net.parameters(), lr=0.05, momentum=0.9)
        ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
        for it in range(3):
            xs = torch.randn(world, 2, 16, 784, generator=g)      # two micro-batches per rank
            ts = torch.randint(0, 10, (world, 2, 16), generator=g)
            opt.zero_grad()
            with net.no_sync():                                    # accumulate locally, no collective
                (F.cross_entropy(net(xs[i, 0].to(dev)), ts[i, 0].to(dev)) / 2).backward()
            (F.cross_entropy(net(xs[i, 1].to(dev)), ts[i, 1].to(dev)) / 2).backward()
            opt.step()
            ropt.zero_grad()
            F.cross_entropy(ref(xs.reshape(-1, 784)), ts.reshape(-1)).backward()
            ropt.step()
        torch.cuda.synchronize()
        for (n_, p), q in zip(net.module.named_parameters(), ref.parameters()):
            err = (p.detach().cpu() - q).abs().max().item()
            assert err < 5e-3, (backend, opt_kind, n_, err)
        chk = torch.cat([p.detach().flatten()[:32] for p in net.parameters()]).cpu()
        alls = comm.allgather_object(chk)
        assert all(torch.equal(alls[0], a) for a in alls), (backend, opt_kind, max((alls[0] - a).abs().max().item() for a in alls))
    _ok(outdir, i)
    dtdl.shutdown()


def frontends_multi_gpu(outdir):
    """Single process, all visible GPUs, through the Keras- and Chainer-shaped front ends:
    MirroredStrategy (Keras-Adam fused into DataParallelStep), load_weights re-syncing every replica,
    and ParallelUpdater inside a Trainer (SURVEY.md DP-c, DP-e)."""
    from distributed_training_dl_b200 import strategy as ks
    from distributed_training_dl_b200 import trainer as ch
    from distributed_training_dl_b200.models import MLP, MnistCNN
    ndev = torch.cuda.device_count()
    g = torch.Generator().manual_seed(0)
    y = torch.randint(0, 10, (512,), generator=g)
    x = torch.randn((512, 1, 28, 28), generator=g) * 0.1 + (y.float().view(-1, 1, 1, 1) - 4.5) / 3
    # --- MirroredStrategy.fit on all GPUs: global batch split over replicas, Adam (Keras epsilon) fused
    strategy = ks.MirroredStrategy()
    assert strategy.num_replicas_in_sync == ndev
    torch.manual_seed(1)
    with strategy.scope():
        model = ks.Model(MnistCNN())
        model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    ckpt = os.path.join(outdir, "ckpt_{epoch}")
    hist = model.fit(x, y, batch_size=64, epochs=3, verbose=0, validation_data=(x[:128], y[:128]),
                     callbacks=[ks.ModelCheckpoint(ckpt)])
    assert hist.history["loss"][-1] < 0.7 * hist.history["loss"][0], hist.history["loss"]
    def replicas_of(m):
        # one visible GPU: MirroredStrategy degenerates to the one-device engine (no replica list)
        return getattr(m.module, "arenas", None) or []

    assert len(replicas_of(model)) == (ndev if ndev > 1 else 0)
    assert all(torch.equal(replicas_of(model)[0].param_flat.cpu(), a.param_flat.cpu())
               for a in replicas_of(model)[1:]), "replicas diverged under fit()"
    # the fused Adam matches torch.optim.Adam(eps=1e-7 inside the sqrt's sum, Keras placement) closely enough
    # that a 3-epoch run reaches a low loss; exact arithmetic is covered by the collectives test
    before = model.evaluate(x, y)
    # --- load_weights into the live model: every replica (and its operand packs) must follow
    with strategy.scope():
        fresh = ks.Model(MnistCNN()).compile()
    fresh.fit(x[:64], y[:64], batch_size=64, epochs=1, verbose=0)      # graph captured, packs exist
    fresh.load_weights(ks.latest_checkpoint(outdir))
    assert all(torch.equal(replicas_of(fresh)[0].param_flat.cpu(), a.param_flat.cpu())
               for a in replicas_of(fresh)[1:]), "load_weights left replicas behind"
    after = fresh.evaluate(x, y)
    assert abs(before[0] - after[0]) < 2e-2 * max(1.0, before[0]) and abs(before[1] - after[1]) < 0.02, (before, after)
    # one more training step from the loaded weights must start from the loaded loss, not the old packs
    h2 = fresh.fit(x, y, batch_size=128, epochs=1, verbose=0, shuffle=False)
    assert h2.history["loss"][0] < 2.0 * max(before[0], 0.05) + 0.2, (h2.history["loss"], before)
    # --- Chainer-shaped ParallelUpdater (train_mnist_gpu.py:52-62,87-93)
    xs = x.reshape(512, 784)
    data = list(zip(xs, y))
    torch.manual_seed(2)
    clf = ch.Classifier(MLP(64, 10))
    opt = ch.Adam().setup(clf)
    devices = {"main": 0, "second": 1} if ndev >= 2 else {"main": 0}
    updater = ch.ParallelUpdater(ch.SerialIterator(data, 64), opt, devices=devices)
    tr = ch.Trainer(updater, (2, "epoch"), out=os.path.join(outdir, "chainer"))
    tr.extend(ch.LogReport())
    tr.run()
    assert updater.iteration == 16
    assert tr.observation["main/loss"] == tr.observation["main/loss"]
    wrapped = updater.wrapped
    ref = wrapped.arenas[0].param_flat.cpu()
    assert all(torch.equal(ref, a.param_flat.cpu()) for a in wrapped.arenas[1:]), "ParallelUpdater replicas diverged"
    log = tr.observation
    assert log["main/loss"] < 2.1 and log["main/accuracy"] > 0.15, log     # 16 Adam steps from ln(10) = 2.30
    open(os.path.join(outdir, "ok_fe"), "w").write("ok")
