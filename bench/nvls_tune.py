#!/usr/bin/env python
"""Fused all-reduce + SGD (NVLS variant) against its NVLink roofline, over CTA count.

    DTDL_NVLS=1 [DTDL_NVLS_UNROLL=4|8 DTDL_NVLS_THREADS=512|256] python -m torch.distributed.run \
        --nproc-per-node 8 --master-addr 127.0.0.1 bench/nvls_tune.py [--out FILE.md]

Roofline (B200_PROFILING.md): a GPU must move (1 + 1/W) n bytes per direction for the multicast
schedule (its whole bucket out once to the switch + its reduced slice back, then its slice out and
everyone's slices in); target time = that / 770 GB/s (measured peer copy).  Reported: microseconds,
bus bandwidth 2(W-1)/W n / t (NCCL's figure) and the fraction of the roofline, device-timed, max
over ranks, with the closing flag exchange deferred to the end of each timed batch of launches
(what a training step does) and, in the last columns, stand-alone (every launch self-contained).
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_training_dl_b200 as dtdl  # noqa: E402
from distributed_training_dl_b200.utils.timing import max_over_ranks  # noqa: E402

LINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--sizes-mb", default="4,16,64,256")
    ap.add_argument("--blocks", default="16,32,64,128")
    args = ap.parse_args()
    comm = dtdl.init(backend="b200")
    peer, W = comm.peer, comm.size
    sizes = [int(float(s) * (1 << 20)) for s in args.sizes_mb.split(",")]
    max_n = max(sizes) // 4
    grads, params, mom = peer.alloc(max_n * 4), peer.alloc(max_n * 4), peer.alloc(max_n * 4)
    grads.view(torch.float32, 0, max_n).normal_()
    m = mom.view(torch.float32, 0, max_n)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0 / W] + [0.0] * 12, device=comm.device)
    variant = 2 if (peer.use_nvls and grads.mc_ptr) else 1
    rows = []
    for size in sizes:
        n = size // 4
        iters = 20 if size <= (64 << 20) else 8
        for blocks in [int(b) for b in args.blocks.split(",")]:
            def op():
                peer.reduce_update(grads, params, m, None, hyper, 0, n, 1, two_shot=variant, max_blocks=blocks)
            res = {}
            for mode in ("deferred", "standalone"):
                for _ in range(3):
                    op()
                peer.barrier()
                torch.cuda.synchronize()
                comm.barrier()
                if mode == "deferred":
                    peer.defer_done(True)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(iters):
                    op()
                if mode == "deferred":
                    peer.barrier()
                    peer.defer_done(False)
                e1.record()
                torch.cuda.synchronize()
                res[mode] = max_over_ranks(e0.elapsed_time(e1) / iters, comm.control)
            rows.append((size, blocks, res))
    if comm.rank == 0:
        kind = "NVLS multimem" if variant == 2 else "peer-pointer two-shot"
        per_dir = (1 + 1 / W) if variant == 2 else 2 * (W - 1) / W
        lines = [f"# Fused all-reduce + SGD, {kind}, {W} x B200; kernel variant unroll "
                 f"{os.environ.get('DTDL_NVLS_UNROLL', '4')} x {os.environ.get('DTDL_NVLS_THREADS', '512')} threads", "",
                 f"Roofline: {per_dir:.3f} n bytes per direction per GPU at {LINK_GBS:.0f} GB/s (measured peer copy).", "",
                 "| bytes | CTAs | us (deferred) | bus GB/s | of roofline | us (stand-alone) | bus GB/s | of roofline |",
                 "|---|---|---|---|---|---|---|---|"]
        for size, blocks, res in rows:
            cells = []
            for mode in ("deferred", "standalone"):
                t = res[mode] * 1e-3
                cells += [f"{t * 1e6:.1f}", f"{2 * (W - 1) / W * size / t / 1e9:.0f}",
                          f"{(per_dir * size / (LINK_GBS * 1e9)) / t:.2f}"]
            lines.append(f"| {size} | {blocks} | " + " | ".join(cells) + " |")
        text = "\n".join(lines)
        print(text, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            open(args.out, "w").write(text + "\n")
    comm.barrier()
    torch.cuda.synchronize()
    dtdl.shutdown()


if __name__ == "__main__":
    main()
