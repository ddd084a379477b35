"""Worker bodies for the multi-process (gloo) tests; must be importable for mp spawn."""
import copy
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _gather(comm, t):
    return comm.allgather_object(t.detach().cpu())


def ddp_equivalence(i, world, init_kind, port, outdir):
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import MLP
    torch.set_num_threads(2)
    if init_kind == "tcp":
        comm = dtdl.init(backend="gloo", init_method=f"tcp://127.0.0.1:{port}", rank=i, world_size=world)
    else:  # env:// variables were exported by dtdl.spawn
        comm = dtdl.init(backend="gloo")
    assert comm.rank == i and comm.size == world
    torch.manual_seed(100 + i)           # different init on every rank on purpose
    net = dtdl.wrap(MLP(n_units=24))
    # (a) identical parameters after construction == rank 0's
    for p in net.parameters():
        allp = _gather(comm, p)
        assert all(torch.equal(allp[0], q) for q in allp)
    # (d) state_dict spellings
    bare_keys = list(MLP(n_units=24).state_dict().keys())
    assert list(net.module.state_dict().keys()) == bare_keys
    assert list(net.state_dict().keys()) == ["module." + k for k in bare_keys]

    ref = copy.deepcopy(net.module)      # single-process model fed the concatenated batch
    for p in ref.parameters():
        p.grad = None
    opt = dtdl.FusedSGD(net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    net.reducer.bucket_timing = True      # per-bucket time / bus bandwidth report
    g = torch.Generator().manual_seed(7)
    for step in range(3):
        xs = torch.randn(world, 8, 784, generator=g)
        ts = torch.randint(0, 10, (world, 8), generator=g)
        out = net(xs[i])
        loss = F.cross_entropy(out, ts[i])
        opt.zero_grad()
        loss.backward()
        if step == 0:
            # (b) the reduced gradient is the mean of the per-rank local gradients
            local = copy.deepcopy(ref)
            F.cross_entropy(local(xs[i]), ts[i]).backward()
            net.reducer.flush()
            for work, bucket, _t0 in net.reducer._works:
                work.wait()
            for (n_, p), q in zip(net.module.named_parameters(), local.parameters()):
                mean = sum(_gather(comm, q.grad)) / world
                got = p._dtdl_grad / world   # all_reduce summed; division happens in finish_step
                assert torch.allclose(got, mean, atol=1e-5), n_
        opt.step()
        rl = F.cross_entropy(ref(xs.reshape(-1, 784)), ts.reshape(-1))
        ropt.zero_grad()
        rl.backward()
        ropt.step()
        # (c) k steps == single process on the concatenated batch
        for p, q in zip(net.module.parameters(), ref.parameters()):
            assert torch.allclose(p, q, atol=2e-5)
    for p in net.parameters():
        allp = _gather(comm, p)
        assert all(torch.equal(allp[0], q) for q in allp)
    rows = net.reducer.bucket_report()
    assert len(rows) == 3 * len(net.reducer.arena.buckets)          # every bucket of every step
    assert all(r["ms"] > 0 and r["bus_GBps"] > 0 and r["kind"] == "gloo" for r in rows)
    assert sum(r["bytes"] for r in rows[:len(net.reducer.arena.buckets)]) == 4 * net.reducer.arena.numel
    assert "bus" in dtdl.utils.format_bucket_report(rows)
    assert net.reducer.bucket_report() == []
    # optimizer state in torch format
    sd = opt.state_dict()
    assert "momentum_buffer" in sd["state"][0]
    dtdl.save_model(net.module, os.path.join(outdir, "ddp.pth"), per_rank=True)
    comm.barrier()
    assert os.path.exists(os.path.join(outdir, f"ddp_{i}.pth"))
    # (e) weights loaded on rank 0 alone reach everyone through net.sync_parameters()
    if i == 0:
        with torch.no_grad():
            for p in net.module.parameters():
                p.add_(1.0)
    net.sync_parameters()
    for p in net.parameters():
        allp = _gather(comm, p)
        assert all(torch.equal(allp[0], q) for q in allp)
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")
    dtdl.shutdown()


def ddp_pyramidnet(i, world, port, outdir):
    """BASELINE.json config 1: the reference CNN under the DDP-style wrapper on CPU/gloo."""
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.models import pyramidnet
    torch.set_num_threads(4)
    comm = dtdl.init(backend="gloo", init_method=f"tcp://127.0.0.1:{port}", rank=i, world_size=world)
    torch.manual_seed(i)
    net = dtdl.wrap(pyramidnet())
    opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.1)
    ds = dtdl.SyntheticDataset("cifar", length=8, seed=1)
    sampler = dtdl.DistributedSampler(ds)
    loader = torch.utils.data.DataLoader(ds, batch_size=2, sampler=sampler)
    net.train()
    losses = []
    for x, t in loader:
        loss = F.cross_entropy(net(x), t)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    sched.step()
    assert all(l == l for l in losses)
    # replicas stay bit-identical; BN running stats follow rank 0 (broadcast at every forward)
    chk = torch.cat([p.detach().flatten()[:64] for p in net.parameters()])
    allc = comm.allgather_object(chk)
    assert all(torch.equal(allc[0], c) for c in allc)
    # K-C2: whatever a rank's running statistics hold, the per-forward sync makes them rank 0's
    with torch.no_grad():
        net.module.bn1.running_mean.fill_(float(i + 1))
        net.module.bn1.running_var.fill_(float(10 * (i + 1)))
    net.reducer.sync_buffers()
    rm = comm.allgather_object((net.module.bn1.running_mean.clone(), net.module.bn1.running_var.clone()))
    for mean, var in rm:
        assert torch.equal(mean, torch.full_like(mean, 1.0)) and torch.equal(var, torch.full_like(var, 10.0))
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")
    dtdl.shutdown()


def multi_worker_fit(i, world, port, outdir):
    """MultiWorkerMirroredStrategy.fit on 2 workers (tensorflow2/mnist_multi_worker_strategy.py:14-29,
    85-96): host-list rendezvous, every worker trains on its strided shard with the global batch split
    over the workers, replicas stay identical, the loss falls, rank 0 alone writes checkpoints and
    TensorBoard records."""
    from distributed_training_dl_b200 import strategy as ks
    from distributed_training_dl_b200.models import MnistCNN
    torch.set_num_threads(2)
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK"):
        os.environ.pop(k, None)
    hosts = ",".join(f"127.0.0.1:{port + r}" for r in range(world))
    strategy = ks.MultiWorkerMirroredStrategy(worker_hosts=hosts, task_index=i, backend="gloo")
    assert strategy.num_replicas_in_sync == world
    g = torch.Generator().manual_seed(0)
    y = torch.randint(0, 10, (256,), generator=g)
    x = torch.randn((256, 1, 28, 28), generator=g) * 0.1 + (y.float().view(-1, 1, 1, 1) - 4.5) / 3
    torch.manual_seed(100 + i)      # different initial weights per worker: wrap() must broadcast rank 0's
    with strategy.scope():
        model = ks.Model(MnistCNN())
        model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    logdir = os.path.join(outdir, "tb")
    hist = model.fit(x, y, batch_size=64, epochs=3, verbose=0, validation_data=(x[:64], y[:64]),
                     callbacks=[ks.TensorBoard(log_dir=logdir, histogram_freq=1),
                                ks.ModelCheckpoint(os.path.join(outdir, "ckpt_{epoch}"))])
    assert hist.history["loss"][-1] < hist.history["loss"][0], hist.history["loss"]
    comm = strategy.comm
    chk = torch.cat([p.detach().flatten()[:64] for p in model.module.parameters()])
    allc = comm.allgather_object(chk)
    assert all(torch.equal(allc[0], c) for c in allc), "workers diverged"
    comm.barrier()
    assert os.path.exists(os.path.join(outdir, "ckpt_3")) and os.listdir(logdir)
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")
    import distributed_training_dl_b200 as dtdl
    dtdl.shutdown()


def fd_channel(i, world, port, outdir):
    """FdChannel (parallel/peer.py): descriptors cross processes and still name the same file."""
    import tempfile
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.parallel.peer import FdChannel
    comm = dtdl.init(backend="gloo", init_method=f"tcp://127.0.0.1:{port}", rank=i, world_size=world)
    store = dist.distributed_c10d._get_default_store()
    chan = FdChannel(i, world, store, lambda: dist.barrier())
    for rnd in range(3):          # repeated exchanges must not mix
        files = []
        for k in range(1 + (i == 0)):     # rank 0 sends two descriptors, the others one
            f = tempfile.TemporaryFile()
            f.write(f"rank{i}-round{rnd}-fd{k}".encode())
            f.flush()
            files.append(f)
        got = chan.allgather([f.fileno() for f in files])
        assert len(got) == world
        for r in range(world):
            assert len(got[r]) == 1 + (r == 0)
            for k, fd in enumerate(got[r]):
                assert os.pread(fd, 64, 0).decode() == f"rank{r}-round{rnd}-fd{k}"
                if r != i:
                    os.close(fd)
    chan.close()
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")
    dtdl.shutdown()


def slurm_and_hosts_init(i, world, port, outdir, kind):
    """init() from srun's environment, or from a TF_CONFIG-style host list."""
    import distributed_training_dl_b200 as dtdl
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK"):
        os.environ.pop(k, None)
    if kind == "slurm":
        os.environ.update(SLURM_PROCID=str(i), SLURM_NTASKS=str(world), SLURM_LOCALID=str(i),
                          SLURM_JOB_NODELIST="127.0.0.1", MASTER_PORT=str(port))
        comm = dtdl.init(backend="gloo")
    else:
        hosts = ",".join(f"127.0.0.1:{port + r}" for r in range(world))
        comm = dtdl.init(backend="gloo", worker_hosts=hosts, task_index=i)
    assert (comm.rank, comm.size, comm.intra_rank) == (i, world, i)
    got = comm.allgather_object(("rank", i))
    assert got == [("rank", r) for r in range(world)]
    assert abs(comm.allreduce_mean(float(i)) - (world - 1) / 2.0) < 1e-6
    comm.barrier()
    open(os.path.join(outdir, f"ok_{i}"), "w").write("ok")
    dtdl.shutdown()
