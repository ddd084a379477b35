#!/usr/bin/env python
"""BASELINE.json config 2: single-process data parallelism on G GPUs of one process.

    python bench/dp_bench.py --gpus G [--impl ours|reference] [--steps K --warmup W]

ours       DataParallel + FusedSGD + DataParallelStep (persistent replicas, scatter / gather
           folded into the first / last kernels, fused peer-memory all-reduce + SGD, one CUDA
           graph per GPU per step).
reference  the reference's pytorch/data_parallel.py calls: nn.DataParallel(pyramidnet()),
           nn.CrossEntropyLoss, optim.SGD, its train() loop body (input on GPU 0).
Global batch = 64 (the reference's --batch-size is the global batch under DataParallel:
pytorch/data_parallel.py:19-20,63-64) unless --per-gpu-batch is given.  Device-timed on GPU 0
with all devices synchronised on both sides.  Prints one JSON line.
"""
import argparse
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def sync_all(n):
    for d in range(n):
        torch.cuda.synchronize(d)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--per-gpu-batch", type=int, default=0)
    args = ap.parse_args()
    G = args.gpus
    B = args.per_gpu_batch * G if args.per_gpu_batch else 64
    g = torch.Generator().manual_seed(0)
    xs = [torch.randn(B, 3, 32, 32, generator=g).pin_memory() for _ in range(4)]
    ts = [torch.randint(0, 10, (B,), generator=g).pin_memory() for _ in range(4)]
    if args.impl == "ours":
        import distributed_training_dl_b200 as dtdl
        model = dtdl.DataParallel(dtdl.models.pyramidnet(), device_ids=list(range(G)))
        opt = dtdl.FusedSGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
        step = dtdl.DataParallelStep(model, opt)
        model.train()

        def run(i):
            return step(xs[i % 4], ts[i % 4])
    else:
        spec = importlib.util.spec_from_file_location("ref_model", os.path.join(ROOT, "baseline/_ref/pytorch/model.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        net = mod.pyramidnet()
        model = torch.nn.DataParallel(net, device_ids=list(range(G))).to("cuda") if G > 1 else net.to("cuda")
        crit = torch.nn.CrossEntropyLoss().to("cuda")
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
        model.train()

        def run(i):
            inputs, targets = xs[i % 4].to("cuda"), ts[i % 4].to("cuda")
            loss = crit(model(inputs), targets)
            opt.zero_grad()
            loss.backward()
            opt.step()
            return loss

    for i in range(max(args.warmup, 3) + 3):
        run(i)
    sync_all(G)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.device(0):
        e0.record()
        for i in range(args.steps):
            out = run(i)
        sync_all(G)
        e1.record()
        torch.cuda.synchronize(0)
    ms = e0.elapsed_time(e1)
    print(json.dumps({"metric": "images/sec, single-process data parallel (input staged from pinned host memory each step)",
                      "impl": args.impl, "n_gpus": G, "global_batch": B, "steps": args.steps,
                      "ms_per_step": ms / args.steps, "value": B * args.steps / (ms / 1e3),
                      "last": float(out.flatten()[0])}))


if __name__ == "__main__":
    main()
