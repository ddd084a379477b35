"""Device-side timeline of one training step's collectives, taken INSIDE the replayed CUDA graph.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/comm_timeline.py [--model pyramidnet] [--precision bf16] [--out FILE.md]

Every flag-exchange and data kernel of the communication stream stamps %globaltimer at its first
block's start and last block's end (csrc/comm/collectives.cu, TraceSlot); a mark kernel on the main
stream stamps the end of backward.  Rank 0 prints the table: what ran when, and how much
communication was left after backward had finished (the exposed tail).
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_training_dl_b200 as dtdl  # noqa: E402
from distributed_training_dl_b200.engine import CrossEntropyLoss, TrainStep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="pyramidnet")
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    comm = dtdl.init(backend="b200")
    dev = comm.device
    make = {"pyramidnet": dtdl.models.pyramidnet, "resnet50": dtdl.models.resnet50}[args.model]
    shape = {"pyramidnet": (3, 32, 32), "resnet50": (3, 224, 224)}[args.model]
    torch.manual_seed(0)
    net = dtdl.wrap(make().to(dev), dtype=args.precision)
    opt = dtdl.FusedSGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    step = TrainStep(net, opt, CrossEntropyLoss(), graph=True, warmup=2)
    x = torch.randn(args.batch, *shape, device=dev)
    t = torch.randint(0, 10, (args.batch,), device=dev)
    for _ in range(2):
        step(x, t)                       # eager warm-up
    peer = comm.peer
    if comm.size > 1:
        peer.trace_begin(128)            # the launches captured next keep these slots
    for _ in range(6):
        step(x, t)                       # capture, then replays: the buffer holds the last replay
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        step(x, t)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    rows = peer.trace_rows() if comm.size > 1 else []
    all_rows = comm.allgather_object((comm.rank, ms, rows))
    if comm.rank == 0:
        lines = [f"# Collective timeline inside the replayed step graph: {args.model} {args.precision}, "
                 f"{comm.size} GPUs, per-GPU batch {args.batch}", "",
                 f"Step time (CUDA events, 10 replays, max over ranks): {max(r[1] for r in all_rows):.3f} ms.  "
                 "Times below are microseconds on each rank's own %globaltimer, relative to the first traced "
                 "kernel of that rank.", ""]
        for rank, ms_r, rws in all_rows:
            if not rws:
                continue
            mark = [a for k, a, b, n in rws if k == "mark"]
            t_bwd = mark[0] if mark else None
            end = max(b for k, a, b, n in rws)
            busy = sum(b - a for k, a, b, n in rws if k not in ("mark",))
            wait = sum(b - a for k, a, b, n in rws if k.startswith("wait"))
            lines.append(f"## rank {rank}: step {ms_r:.3f} ms; collectives span {end:.0f} us, kernels busy {busy:.0f} us "
                         f"(of which flag waits {wait:.0f} us)"
                         + (f"; backward ends at {t_bwd:.0f} us -> exposed tail {max(0.0, end - t_bwd):.0f} us" if t_bwd is not None else ""))
            if rank in (0, comm.size - 1):
                lines += ["", "| kernel | start us | end us | dur us | bytes |", "|---|---|---|---|---|"]
                for k, a, b, n in sorted(rws, key=lambda r: r[1]):
                    lines.append(f"| {k} | {a:.1f} | {b:.1f} | {b - a:.1f} | {n} |")
                lines.append("")
        text = "\n".join(lines)
        print(text, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            open(args.out, "w").write(text + "\n")
    dtdl.shutdown()


if __name__ == "__main__":
    main()
