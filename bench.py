#!/usr/bin/env python
"""Headline benchmark: training throughput (images/sec) of the reference's CNN (PyramidNet,
pytorch/model.py) under one-process-per-GPU data parallelism, per-GPU batch 64, synthetic
CIFAR-shaped data, random-init weights (BASELINE.json config 3; config 2 at N=1).

    python bench.py [--gpus N --steps K --warmup W] [--impl ours|reference] [--model pyramidnet]

For N>1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...``.  Rank 0 prints ONE JSON line.

ours       wrap() + FusedSGD + TrainStep: hand-written sm_100a kernels (tcgen05 implicit-GEMM
           convs, fused BN/ReLU/residual, softmax-CE), peer-memory fused all-reduce+SGD over
           NVLink, whole step replayed from a CUDA graph.  bf16 tensor-core math, fp32 master
           weights / accumulation / statistics.
reference  the UNMODIFIED reference model definition (baseline/_ref/pytorch/model.py) driven by
           the reference's own calls -- DistributedDataParallel + NCCL, nn.CrossEntropyLoss,
           optim.SGD(lr .1, momentum .9, wd 1e-4), the train() loop of
           pytorch/distributed_data_parallel.py:118-152 -- on the same synthetic batches.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PER_GPU_BATCH = 64
BASELINE_PUBLISHED = None  # BASELINE.json "published": {} -- no comparable published number


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--model", default="pyramidnet", choices=["pyramidnet", "resnet50", "mnist_cnn", "mlp"])
    p.add_argument("--batch-size", type=int, default=PER_GPU_BATCH, help="per-GPU batch")
    p.add_argument("--no-graph", action="store_true", help="ours: eager launches instead of a CUDA graph")
    p.add_argument("--no-e2e", action="store_true")
    return p.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    return rank, world, local


def synthetic_batches(model, batch, n_sets, pin=True):
    import torch
    shape = {"pyramidnet": (3, 32, 32), "resnet50": (3, 224, 224), "mnist_cnn": (1, 28, 28), "mlp": (784,)}[model]
    classes = 1000 if model == "resnet50" else 10
    g = torch.Generator().manual_seed(1234 + dist_env()[0])
    xs = [torch.randn((batch,) + shape, generator=g) for _ in range(n_sets)]
    ts = [torch.randint(0, classes, (batch,), generator=g) for _ in range(n_sets)]
    if pin and torch.cuda.is_available():
        xs, ts = [x.pin_memory() for x in xs], [t.pin_memory() for t in ts]
    return xs, ts


def timed_region(run_step, steps, flush, barrier, device_index):
    """K steps, barrier + synchronize on both sides, CUDA events, clocks sampled meanwhile."""
    import torch
    from distributed_training_dl_b200.utils.timing import ClockSampler
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(device_index) as clocks:
        e0.record()
        for i in range(steps):
            flush()
            run_step(i)
        e1.record()
        torch.cuda.synchronize()
        barrier()
    return e0.elapsed_time(e1), clocks.summary()


# --------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200 import _native as nv
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    from distributed_training_dl_b200.utils.timing import flush_l2, max_over_ranks

    rank, world, local = dist_env()
    comm = dtdl.init(backend="b200")  # env:// from torchrun, or single process
    dev = comm.device
    make = {"pyramidnet": dtdl.models.pyramidnet, "resnet50": dtdl.models.resnet50,
            "mnist_cnn": dtdl.models.MnistCNN, "mlp": dtdl.models.MLP}[args.model]
    torch.manual_seed(0)
    net = dtdl.wrap(make().to(dev))
    if args.model in ("pyramidnet", "resnet50"):
        opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    else:
        opt = dtdl.FusedAdam(net.parameters(), keras=(args.model == "mnist_cnn"))
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=not args.no_graph, warmup=2)
    net.train()
    B = args.batch_size
    xs, ts = synthetic_batches(args.model, B, 8)
    dxs, dts = [x.to(dev) for x in xs], [t.to(dev) for t in ts]

    def barrier():
        comm.barrier()

    for i in range(max(args.warmup, 3) + 3):  # eager warm-up + capture + replays
        step(dxs[i % 8], dts[i % 8])
    torch.cuda.synchronize()

    # ---- device-timed: inputs already resident, no host read-back ----
    nv.reset_launch_count()
    ms, clocks = timed_region(lambda i: step(dxs[i % 8], dts[i % 8]), args.steps,
                              lambda: flush_l2(dev), barrier, dev.index or 0)
    if step.launches_per_step is not None:
        launches = step.launches_per_step * args.steps
    else:
        launches = nv.launch_count()
    ms = max_over_ranks(ms, comm.control)
    value = world * B * args.steps / (ms / 1e3)

    # ---- end to end: pinned host batch -> H2D -> step -> loss D2H, every step ----
    e2e = None
    if not args.no_e2e:
        host_loss = torch.zeros(1).pin_memory()

        def e2e_step(i):
            loss = step(xs[i % 8], ts[i % 8])           # public API: copies from pinned memory
            host_loss.copy_(loss.detach().reshape(1), non_blocking=False)  # D2H read of the result

        for i in range(2):
            e2e_step(i)
        ms2, _ = timed_region(e2e_step, args.steps, lambda: flush_l2(dev), barrier, dev.index or 0)
        ms2 = max_over_ranks(ms2, comm.control)
        e2e = {"value": world * B * args.steps / (ms2 / 1e3), "unit": "images/sec",
               "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": world * (xs[0].numel() * 4 + ts[0].numel() * 8),
               "d2h_bytes_per_step": world * 4, "last_loss": float(host_loss.item())}
    peer = getattr(comm, "peer", None)
    if world == 1:
        comm_kind = "none (fused SGD update kernel)"
    elif peer is not None and peer.use_nvls:
        comm_kind = "NVLS multimem.ld_reduce/st fused allreduce+SGD (NVSwitch multicast)"
    else:
        comm_kind = "peer-memory two-shot fused allreduce+SGD (NVLink P2P)"
    res = result_line(args, "ours", world, value, ms, clocks, e2e, launches, dtype="bf16",
                      extra={"graph": not args.no_graph, "kernels_per_step": step.launches_per_step,
                             "parallelism": f"dp{world}", "comm": comm_kind})
    if rank == 0:
        print(json.dumps(res), flush=True)
    torch.cuda.synchronize()
    dtdl.shutdown()


# --------------------------------------------------------------------------------------
def load_reference_model():
    for cand in (os.path.join(ROOT, "baseline", "_ref", "pytorch", "model.py"),
                 "/root/reference/pytorch/model.py"):
        if os.path.exists(cand):
            spec = importlib.util.spec_from_file_location("ref_model", cand)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod, cand
    return None, None


def run_reference(args):
    if args.model not in ("pyramidnet", "resnet50"):
        print(json.dumps({"impl": "reference", "unavailable":
                          "TensorFlow 2 / Chainer are not installed on this image (BASELINE.md section 4)"}))
        return
    mod, path = load_reference_model()
    if mod is None:
        print(json.dumps({"impl": "reference", "unavailable":
                          "baseline/_ref/pytorch/model.py missing (run __graft_entry__.build() "
                          "where /root/reference is mounted)"}))
        return
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.optim as optim
    from distributed_training_dl_b200.utils.timing import flush_l2, max_over_ranks

    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # pytorch/distributed_data_parallel.py:61-62 (tcp:// rendezvous, NCCL)
        dist.init_process_group(backend="nccl", init_method="tcp://{}:{}".format(
            os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500")),
            world_size=world, rank=rank)
    torch.manual_seed(0)
    if args.model == "resnet50":   # not in the reference repo: torchvision's definition, bf16 autocast
        import torchvision
        net = torchvision.models.resnet50()
    else:
        net = mod.pyramidnet()
    net.cuda(local)
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local], output_device=local)
    criterion = nn.CrossEntropyLoss().cuda(local)
    optimizer = optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    net.train()
    B = args.batch_size
    xs, ts = synthetic_batches(args.model, B, 8)
    dxs, dts = [x.to(dev) for x in xs], [t.to(dev) for t in ts]
    amp = args.model == "resnet50"

    def barrier():
        if world > 1:
            dist.barrier()

    def core(inputs, targets):
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            outputs = net(inputs)
            loss = criterion(outputs, targets)
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        return outputs, loss

    for i in range(max(args.warmup, 3)):
        core(dxs[i % 8], dts[i % 8])
    ms, clocks = timed_region(lambda i: core(dxs[i % 8], dts[i % 8]), args.steps,
                              lambda: flush_l2(dev), barrier, local)
    ms = max_over_ranks(ms) if world > 1 else ms
    value = world * B * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        state = {"loss": 0.0, "correct": 0, "total": 0}

        def e2e_step(i):  # the reference's train() body, pytorch/distributed_data_parallel.py:123-138
            inputs = xs[i % 8].cuda()
            targets = ts[i % 8].cuda()
            outputs, loss = core(inputs, targets)
            state["loss"] += loss.item()
            _, predicted = outputs.max(1)
            state["total"] += targets.size(0)
            state["correct"] += predicted.eq(targets).sum().item()

        for i in range(2):
            e2e_step(i)
        ms2, _ = timed_region(e2e_step, args.steps, lambda: flush_l2(dev), barrier, local)
        ms2 = max_over_ranks(ms2) if world > 1 else ms2
        e2e = {"value": world * B * args.steps / (ms2 / 1e3), "unit": "images/sec",
               "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": world * (xs[0].numel() * 4 + ts[0].numel() * 8),
               "d2h_bytes_per_step": world * 12}
    res = result_line(args, "reference", world, value, ms, clocks, e2e, 0,
                      dtype="bf16 autocast" if amp else "fp32 (TF32 conv allowed: torch default)",
                      extra={"parallelism": f"dp{world}", "comm": "NCCL" if world > 1 else "none",
                             "model_file": os.path.relpath(path, ROOT) if path.startswith(ROOT) else path})
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.destroy_process_group()


def result_line(args, impl, world, value, ms, clocks, e2e, launches, dtype, extra):
    cfg = {"model": {"pyramidnet": "PyramidNet-110 a270 (pytorch/model.py), CIFAR-shaped 3x32x32",
                     "resnet50": "ResNet-50, ImageNet-shaped 3x224x224, 1000 classes",
                     "mnist_cnn": "Keras MNIST CNN (tensorflow2/mnist_single.py)",
                     "mlp": "Chainer MLP 784-1000-1000-10"}[args.model],
           "global_batch": world * args.batch_size, "per_gpu_batch": args.batch_size, "seq_len": None,
           "optimizer": "SGD lr 0.1 momentum 0.9 wd 1e-4" if args.model in ("pyramidnet", "resnet50") else "Adam",
           "l2": "256 MiB memset between steps, inside the timed region"}
    cfg.update(extra)
    out = {"metric": "images/sec (train step: fwd+loss+bwd+allreduce+SGD update), device-timed, max over ranks",
           "impl": impl, "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": (value / BASELINE_PUBLISHED) if BASELINE_PUBLISHED else None,
           "dtype": dtype, "data": "synthetic (random images/labels of the named shape, random-init weights)",
           "config": cfg, "clocks": clocks, "gpu_launches": launches}
    if e2e is not None:
        out["e2e"] = e2e
    return out


def main():
    args = parse()
    if args.impl == "reference":
        try:
            import torch
            if not torch.cuda.is_available():
                raise RuntimeError("no CUDA device")
        except Exception as exc:  # noqa: BLE001
            print(json.dumps({"impl": "reference", "unavailable": str(exc)}))
            return
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
