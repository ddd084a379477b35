#!/usr/bin/env python
"""Gradient all-reduce bus-bandwidth sweep, 1 KB - 1 GB (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py

Compares the peer-memory one-shot / two-shot kernels (out-of-place, fp32 sum) and the fused
all-reduce+SGD kernel with NCCL's all_reduce -- the backend the reference selects
(pytorch/distributed_data_parallel.py:46).  Device-timed with CUDA events, max over ranks;
bus bandwidth = 2(W-1)/W * bytes / time.  Writes a markdown table (rank 0) to --out.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_training_dl_b200 as dtdl  # noqa: E402
from distributed_training_dl_b200.utils.timing import max_over_ranks  # noqa: E402


def time_op(fn, comm, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    comm.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return max_over_ranks(e0.elapsed_time(e1) / iters, comm.control)  # ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/allreduce_sweep.md")
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--blocks", type=int, default=64)
    args = ap.parse_args()
    comm = dtdl.init(backend="b200")
    peer, W = comm.peer, comm.size
    nccl = dist.new_group(backend="nccl") if W > 1 else None
    max_n = args.max_mb * (1 << 20) // 4
    src, dst = peer.alloc(max_n * 4), peer.alloc(max_n * 4)
    params = peer.alloc(max_n * 4)
    mom = peer.alloc(max_n * 4)
    m = mom.view(torch.float32, 0, max_n)
    t = src.view(torch.float32, 0, max_n)
    t.normal_()
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0 / W] + [0.0] * 12, device=comm.device)
    rows = []
    size = 1024
    while size <= args.max_mb * (1 << 20):
        n = size // 4
        iters = 50 if size <= (1 << 22) else (10 if size <= (1 << 27) else 4)
        res = {}
        res["one-shot"] = (time_op(lambda: peer.allreduce(src, 0, n, two_shot=0, dst=dst, max_blocks=args.blocks),
                                   comm, iters) if size <= (1 << 26) else None)
        res["two-shot"] = time_op(lambda: peer.allreduce(src, 0, n, two_shot=1, dst=dst, max_blocks=args.blocks),
                                  comm, iters)
        res["fused-sgd"] = time_op(lambda: peer.reduce_update(src, params, m, None, hyper, 0, n, 1, two_shot=1,
                                                              max_blocks=args.blocks), comm, iters)
        if peer.use_nvls and src.mc_ptr:     # DTDL_NVLS=1: in-switch reduction through the multicast address
            res["nvls"] = time_op(lambda: peer.allreduce(src, 0, n, two_shot=2, dst=dst, max_blocks=args.blocks),
                                  comm, iters)
            res["nvls-fused-sgd"] = time_op(lambda: peer.reduce_update(src, params, m, None, hyper, 0, n, 1,
                                                                       two_shot=2, max_blocks=args.blocks),
                                            comm, iters)
        if nccl is not None:
            x = t[:n]
            res["nccl"] = time_op(lambda: dist.all_reduce(x, group=nccl), comm, iters)
        rows.append((size, res))
        size *= 4
    if comm.rank == 0:
        f = 2.0 * (W - 1) / W
        lines = [f"# all-reduce sweep, {W} x B200, fp32 sum, device-timed max over ranks, "
                 f"comm kernels capped at {args.blocks} CTAs", "",
                 "| bytes | " + " | ".join(f"{k} us | {k} busGB/s" for k in rows[0][1]) + " |",
                 "|---|" + "---|" * (2 * len(rows[0][1]))]
        for size, res in rows:
            cells = []
            for k, ms in res.items():
                cells += ["-", "-"] if ms is None else [f"{ms * 1e3:.1f}", f"{f * size / (ms * 1e-3) / 1e9:.1f}"]
            lines.append(f"| {size} | " + " | ".join(cells) + " |")
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        open(args.out, "w").write("\n".join(lines) + "\n")
        print("\n".join(lines))
    comm.barrier()
    torch.cuda.synchronize()
    dtdl.shutdown()


if __name__ == "__main__":
    main()
