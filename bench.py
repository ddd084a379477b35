#!/usr/bin/env python
"""Headline benchmark: training throughput (images/sec) of the reference's CNN (PyramidNet,
pytorch/model.py) under one-process-per-GPU data parallelism, per-GPU batch 64, synthetic
CIFAR-shaped data, random-init weights (BASELINE.json config 3; config 2 at N=1).

    python bench.py [--gpus N --steps K --warmup W] [--impl ours|reference]
                    [--model M --precision P] [--mode ddp|dp] [--no-sub]

For N>1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...``.  Rank 0 prints ONE JSON line.

What the line holds (both arms print the same structure):

  value / e2e      PyramidNet at the REFERENCE'S precision: fp32 tensors, TF32 tensor-core
                   convolutions (ours: ``wrap(..., dtype="tf32")``; reference: the stock fp32 model,
                   where torch's cuDNN default allows TF32 for convolutions).  Same dtype on both
                   arms, so the driver's ratio compares like with like.
  records          named sub-records, each a complete measurement with its own clock sample:
                   ``pyramidnet_bf16``  the same model in bf16 (ours: native bf16 mode; reference:
                                        the fp32 model under ``torch.autocast(bfloat16)``) -- splits
                                        "precision" from "engineering" in the speed-up;
                   ``resnet50_bf16``    ResNet-50, ImageNet-shaped 3x224x224, bf16 (BASELINE.json
                                        config 4; reference: torchvision's definition + autocast).
                   ``--no-sub`` skips them; ``--model`` / ``--precision`` measure one configuration.

ours       wrap() + FusedSGD + TrainStep: hand-written sm_100a kernels (tcgen05 implicit-GEMM
           convs, fused BN/ReLU/residual, softmax-CE), peer-memory fused all-reduce+SGD over
           NVLink, whole step replayed from a CUDA graph.
reference  the UNMODIFIED reference model definition (baseline/_ref/pytorch/model.py) driven by
           the reference's own calls -- DistributedDataParallel + NCCL, nn.CrossEntropyLoss,
           optim.SGD(lr .1, momentum .9, wd 1e-4), the train() loop of
           pytorch/distributed_data_parallel.py:118-152 -- on the same synthetic batches.
--mode dp  single-process replication instead (BASELINE.json config 2): ours = DataParallel +
           DataParallelStep, reference = nn.DataParallel exactly as pytorch/data_parallel.py:69-72,
           102-109; launched as ONE process for any N (global batch 64*N split over the GPUs).
"""
from __future__ import annotations

import argparse
import gc
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PER_GPU_BATCH = 64
BASELINE_PUBLISHED = None  # BASELINE.json "published": {} -- no comparable published number

MODEL_DESC = {"pyramidnet": "PyramidNet-110 a270 (pytorch/model.py), CIFAR-shaped 3x32x32",
              "resnet50": "ResNet-50, ImageNet-shaped 3x224x224, 1000 classes",
              "mnist_cnn": "Keras MNIST CNN (tensorflow2/mnist_single.py)",
              "mlp": "Chainer MLP 784-1000-1000-10"}
METRIC = "images/sec (train step: fwd+loss+bwd+allreduce+SGD update), device-timed, max over ranks"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--model", default=None, choices=["pyramidnet", "resnet50", "mnist_cnn", "mlp"],
                   help="measure this model only (default: PyramidNet tf32 + the sub-records)")
    p.add_argument("--precision", default=None, choices=["tf32", "bf16"],
                   help="with --model: tf32 (fp32 tensors, TF32 convs) or bf16")
    p.add_argument("--mode", default="ddp", choices=["ddp", "dp"],
                   help="ddp: one process per GPU; dp: one process drives all GPUs")
    p.add_argument("--batch-size", type=int, default=PER_GPU_BATCH, help="per-GPU batch")
    p.add_argument("--no-graph", action="store_true", help="ours: eager launches instead of a CUDA graph")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-sub", action="store_true", help="skip the bf16 / ResNet-50 sub-records")
    return p.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    return rank, world, local


def synthetic_batches(model, batch, n_sets, pin=True, seed_offset=None):
    import torch
    shape = {"pyramidnet": (3, 32, 32), "resnet50": (3, 224, 224), "mnist_cnn": (1, 28, 28), "mlp": (784,)}[model]
    classes = 1000 if model == "resnet50" else 10
    g = torch.Generator().manual_seed(1234 + (dist_env()[0] if seed_offset is None else seed_offset))
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
            if i == steps // 2:
                clocks.sample_now()      # the GPU is busy with queued steps right now
        e1.record()
        clocks.sample_now()
        torch.cuda.synchronize()
        barrier()
    return e0.elapsed_time(e1), clocks.summary()


def record(args, impl, model, world, value, ms, clocks, e2e, launches, dtype, extra):
    cfg = {"model": MODEL_DESC[model], "global_batch": world * args.batch_size,
           "per_gpu_batch": args.batch_size, "seq_len": None,
           "optimizer": "SGD lr 0.1 momentum 0.9 wd 1e-4" if model in ("pyramidnet", "resnet50") else "Adam",
           "l2": "256 MiB memset between steps, inside the timed region"}
    cfg.update(extra)
    out = {"metric": METRIC, "impl": impl, "value": value, "unit": "images/sec", "n_gpus": world,
           "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
           "higher_is_better": True, "scaling": "weak",
           "vs_baseline": (value / BASELINE_PUBLISHED) if BASELINE_PUBLISHED else None,
           "dtype": dtype, "data": "synthetic (random images/labels of the named shape, random-init weights)",
           "config": cfg, "clocks": clocks, "gpu_launches": launches}
    if e2e is not None:
        out["e2e"] = e2e
    return out


def plan(args):
    """[(key, model, precision)]: the first entry is the headline, the rest are sub-records."""
    if args.model is not None:
        return [(None, args.model, args.precision or ("tf32" if args.model == "pyramidnet" else "bf16"))]
    jobs = [(None, "pyramidnet", args.precision or "tf32")]
    if not args.no_sub:
        jobs += [("pyramidnet_bf16", "pyramidnet", "bf16"), ("resnet50_bf16", "resnet50", "bf16")]
    return jobs


def assemble(results):
    """Headline record + {"records": {...}} of the others."""
    head = results[0][1]
    subs = {k: r for k, r in results[1:]}
    if subs:
        head["records"] = subs
    return head


# --------------------------------------------------------------------------------------
def _make_ours(model, precision):
    import distributed_training_dl_b200 as dtdl
    make = {"pyramidnet": dtdl.models.pyramidnet, "resnet50": dtdl.models.resnet50,
            "mnist_cnn": dtdl.models.MnistCNN, "mlp": dtdl.models.MLP}[model]
    return dtdl.set_precision(make(), precision)


DTYPE = {"tf32": "fp32 tensors, TF32 tensor-core convolutions", "bf16": "bf16"}


def _ours_dtype(precision):
    return DTYPE[precision]


def bench_ours(args, comm, model, precision):
    """One configuration through wrap() + FusedSGD + TrainStep (one process per GPU)."""
    import torch
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200 import _native as nv
    from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep
    from distributed_training_dl_b200.utils.timing import flush_l2, max_over_ranks

    rank, world, local = dist_env()
    dev = comm.device
    torch.manual_seed(0)
    net = dtdl.wrap(_make_ours(model, precision).to(dev))
    if model in ("pyramidnet", "resnet50"):
        opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    else:
        opt = dtdl.FusedAdam(net.parameters(), keras=(model == "mnist_cnn"))
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=not args.no_graph, warmup=2)
    net.train()
    B = args.batch_size
    xs, ts = synthetic_batches(model, B, 8)
    dxs, dts = [x.to(dev) for x in xs], [t.to(dev) for t in ts]

    def barrier():
        comm.barrier()

    for i in range(max(args.warmup, 3) + 3):  # eager warm-up + capture + replays
        flush_l2(dev)     # also allocates the flush buffer here, not inside the timed region (a cudaMalloc
        step(dxs[i % 8], dts[i % 8])     # with peer mappings is a device-wide stall that grows with the GPU count)
    torch.cuda.synchronize()
    if os.environ.get("DTDL_BENCH_FLUSH") == "0":      # diagnosis only: a line measured this way is not a bench value
        flush_l2 = lambda dev: None  # noqa: E731

    # ---- device-timed: inputs already resident, no host read-back ----
    nv.reset_launch_count()
    ms, clocks = timed_region(lambda i: step(dxs[i % 8], dts[i % 8]), args.steps,
                              lambda: flush_l2(dev), barrier, dev.index or 0)
    launches = step.launches_per_step * args.steps if step.launches_per_step is not None else nv.launch_count()
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
        ms2, clocks2 = timed_region(e2e_step, args.steps, lambda: flush_l2(dev), barrier, dev.index or 0)
        ms2 = max_over_ranks(ms2, comm.control)
        e2e = {"value": world * B * args.steps / (ms2 / 1e3), "unit": "images/sec",
               "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": world * (xs[0].numel() * 4 + ts[0].numel() * 8),
               "d2h_bytes_per_step": world * 4, "last_loss": float(host_loss.item()), "clocks": clocks2}
    peer = getattr(comm, "peer", None)
    if world == 1:
        comm_kind = "none (fused SGD update kernel)"
    elif peer is not None and peer.use_nvls:
        comm_kind = "NVLS multimem.ld_reduce/st fused allreduce+SGD (NVSwitch multicast)"
    else:
        comm_kind = "peer-memory two-shot fused allreduce+SGD (NVLink P2P)"
    res = record(args, "ours", model, world, value, ms, clocks, e2e, launches, _ours_dtype(precision),
                 extra={"precision": precision, "graph": not args.no_graph,
                        "kernels_per_step": step.launches_per_step, "parallelism": f"dp{world}",
                        "comm": comm_kind})
    # release this configuration before the next one is built
    net.reducer.close()
    del step, opt, net, dxs, dts
    gc.collect()
    torch.cuda.empty_cache()
    return res


def bench_ours_dp(args, model, precision):
    """Single process, N GPUs: DataParallel + DataParallelStep (global batch = N x per-GPU batch)."""
    import torch
    import distributed_training_dl_b200 as dtdl
    from distributed_training_dl_b200.utils.timing import flush_l2

    n = args.gpus
    d0 = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    dp = dtdl.DataParallel(_make_ours(model, precision), device_ids=list(range(n)))
    opt = dtdl.FusedSGD(dp.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    step = dtdl.DataParallelStep(dp, opt, graph=not args.no_graph, warmup=2)
    dp.train()
    GB = args.batch_size * n
    xs, ts = synthetic_batches(model, GB, 8, seed_offset=0)
    dxs, dts = [x.to(d0) for x in xs], [t.to(d0) for t in ts]

    def sync_all():
        for d in dp.devices:
            torch.cuda.synchronize(d)

    def flush():
        for d in dp.devices:
            with torch.cuda.device(d):
                flush_l2(d)

    for i in range(max(args.warmup, 3) + 3):
        flush()           # warm the flush buffers too (allocated on first use)
        step(dxs[i % 8], dts[i % 8])
    sync_all()
    ms, clocks = timed_region(lambda i: step(dxs[i % 8], dts[i % 8]), args.steps, flush, sync_all, 0)
    launches = (step.launches_per_step or 0) * args.steps
    value = GB * args.steps / (ms / 1e3)
    e2e = None
    if not args.no_e2e:
        host = torch.zeros(2).pin_memory()

        def e2e_step(i):
            stats = step(xs[i % 8], ts[i % 8])
            host.copy_(stats, non_blocking=False)

        for i in range(2):
            e2e_step(i)
        ms2, clocks2 = timed_region(e2e_step, args.steps, flush, sync_all, 0)
        e2e = {"value": GB * args.steps / (ms2 / 1e3), "unit": "images/sec", "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": xs[0].numel() * 4 + ts[0].numel() * 8, "d2h_bytes_per_step": 8,
               "last_loss": float(host[0].item()), "clocks": clocks2}
    res = record(args, "ours", model, n, value, ms, clocks, e2e, launches, _ours_dtype(precision),
                 extra={"precision": precision, "graph": not args.no_graph, "kernels_per_step": step.launches_per_step,
                        "parallelism": f"single-process replication over {n} GPU(s) (DataParallel + DataParallelStep)",
                        "comm": "in-process peer-memory fused allreduce+SGD", "global_batch": GB})
    res["metric"] = METRIC.replace("max over ranks", "one process, all GPUs")
    sync_all()
    dp.close()
    del step, opt, dp
    gc.collect()
    torch.cuda.empty_cache()
    return res


def run_ours(args):
    import torch
    import distributed_training_dl_b200 as dtdl
    rank, world, local = dist_env()
    results = []
    if args.mode == "dp":
        for key, model, precision in plan(args):
            results.append((key, bench_ours_dp(args, model, precision)))
        print(json.dumps(assemble(results)), flush=True)
        return
    comm = dtdl.init(backend="b200")  # env:// from torchrun, or single process
    for key, model, precision in plan(args):
        results.append((key, bench_ours(args, comm, model, precision)))
    if rank == 0:
        print(json.dumps(assemble(results)), flush=True)
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


def bench_reference(args, mod, path, model, precision):
    """The reference's own calls on synthetic batches.  precision "tf32" = the stock fp32 model
    (torch's defaults: cuDNN convolutions may use TF32, matmuls stay fp32) -- what the reference's
    scripts run; "bf16" = the same model under torch.autocast(bfloat16), an explanatory extra for
    PyramidNet and the configuration BASELINE.json names for ResNet-50."""
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.optim as optim
    from distributed_training_dl_b200.utils.timing import flush_l2, max_over_ranks

    rank, world, local = dist_env()
    dp_mode = args.mode == "dp"
    n = args.gpus if dp_mode else world
    dev = torch.device("cuda", 0 if dp_mode else local)
    torch.manual_seed(0)
    if model == "resnet50":   # not in the reference repo: torchvision's definition
        import torchvision
        net = torchvision.models.resnet50()
    else:
        net = mod.pyramidnet()
    if dp_mode:
        # pytorch/data_parallel.py:69-72: nn.DataParallel(net).to('cuda'); inputs land on GPU 0
        net = nn.DataParallel(net, device_ids=list(range(n))).to("cuda")
    else:
        net.cuda(local)
        if world > 1:
            net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local], output_device=local)
    criterion = nn.CrossEntropyLoss().to(dev)
    optimizer = optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    net.train()
    B = args.batch_size * (n if dp_mode else 1)
    xs, ts = synthetic_batches(model, B, 8, seed_offset=0 if dp_mode else None)
    dxs, dts = [x.to(dev) for x in xs], [t.to(dev) for t in ts]
    amp = precision == "bf16"

    def barrier():
        if dp_mode:
            for d in range(n):
                torch.cuda.synchronize(d)
        elif world > 1:
            dist.barrier()

    def flush():
        if dp_mode:
            for d in range(n):
                with torch.cuda.device(d):
                    flush_l2(torch.device("cuda", d))
        else:
            flush_l2(dev)

    def core(inputs, targets):
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            outputs = net(inputs)
            loss = criterion(outputs, targets)
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        return outputs, loss

    for i in range(max(args.warmup, 3)):
        flush()           # warm the flush buffer too (allocated on first use)
        core(dxs[i % 8], dts[i % 8])
    ms, clocks = timed_region(lambda i: core(dxs[i % 8], dts[i % 8]), args.steps, flush, barrier, dev.index)
    ms = max_over_ranks(ms) if (world > 1 and not dp_mode) else ms
    total = B * (1 if dp_mode else world)
    value = total * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        state = {"loss": 0.0, "correct": 0, "total": 0}

        def e2e_step(i):  # the reference's train() body, pytorch/distributed_data_parallel.py:123-138
            inputs = xs[i % 8].to(dev)
            targets = ts[i % 8].to(dev)
            outputs, loss = core(inputs, targets)
            state["loss"] += loss.item()
            _, predicted = outputs.max(1)
            state["total"] += targets.size(0)
            state["correct"] += predicted.eq(targets).sum().item()

        for i in range(2):
            e2e_step(i)
        ms2, clocks2 = timed_region(e2e_step, args.steps, flush, barrier, dev.index)
        ms2 = max_over_ranks(ms2) if (world > 1 and not dp_mode) else ms2
        e2e = {"value": total * args.steps / (ms2 / 1e3), "unit": "images/sec",
               "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": (1 if dp_mode else world) * (xs[0].numel() * 4 + ts[0].numel() * 8),
               "d2h_bytes_per_step": (1 if dp_mode else world) * 12, "clocks": clocks2}
    how = ("torch defaults: cuDNN convolutions may use TF32, nn.Linear fp32" if not amp
           else "torch.autocast(bfloat16) around forward + loss, fp32 master weights")
    res = record(args, "reference", model, n, value, ms, clocks, e2e, 0, DTYPE[precision],
                 extra={"precision": precision, "precision_how": how,
                        "parallelism": (f"nn.DataParallel over {n} GPU(s), one process" if dp_mode else f"dp{world}"),
                        "comm": ("torch.nn.parallel scatter/replicate/gather" if dp_mode else
                                 ("NCCL" if world > 1 else "none")),
                        "global_batch": total,
                        "model_file": (os.path.relpath(path, ROOT) if path.startswith(ROOT) else path)
                        if model == "pyramidnet" else "torchvision.models.resnet50"})
    if dp_mode:
        res["metric"] = METRIC.replace("max over ranks", "one process, all GPUs")
    del net, optimizer, dxs, dts
    gc.collect()
    torch.cuda.empty_cache()
    return res


def run_reference(args):
    jobs = plan(args)
    if any(m not in ("pyramidnet", "resnet50") for _, m, _ in jobs):
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
    rank, world, local = dist_env()
    if args.mode != "dp":
        torch.cuda.set_device(local)
        if world > 1:
            # pytorch/distributed_data_parallel.py:61-62 (tcp:// rendezvous, NCCL)
            dist.init_process_group(backend="nccl", init_method="tcp://{}:{}".format(
                os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500")),
                world_size=world, rank=rank)
    results = [(key, bench_reference(args, mod, path, model, precision)) for key, model, precision in jobs]
    if rank == 0:
        print(json.dumps(assemble(results)), flush=True)
    if args.mode != "dp" and world > 1:
        dist.destroy_process_group()


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
