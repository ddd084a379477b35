"""Shared pieces of the example scripts (argument groups, data, the training loop)."""
import argparse
import datetime
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_training_dl_b200 as dtdl  # noqa: E402


def pytorch_flags(description, distributed=False):
    """The flag set of the reference's pytorch scripts (pytorch/single_gpu.py:14-39,
    pytorch/distributed_data_parallel.py:45-48); both of its spellings are accepted."""
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--train-dir", "-td", default="./train_dir")
    p.add_argument("--dataset-dir", "-dd", default="./data")
    p.add_argument("--batch-size", "-b", type=int, default=64)
    p.add_argument("--num-workers", type=int, default=4)
    p.add_argument("--test-batch-size", "--test-batchsize", "-tb", type=int, default=1000)
    p.add_argument("--epochs", "-e", type=int, default=10)
    p.add_argument("--gpu-nums", "-g", type=int, default=0)
    p.add_argument("--learning-rate", "--lr", "-lr", type=float, default=0.1)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--seed", type=int, default=1)
    p.add_argument("--log-interval", "-li", type=int, default=20)
    p.add_argument("--save-model", "-sm", action="store_true")
    p.add_argument("--weight-decay", "--wd", "-wd", type=float, default=1e-4)
    p.add_argument("--synthetic", type=int, default=0, metavar="N",
                   help="train on N random CIFAR-shaped images instead of CIFAR-10")
    p.add_argument("--no-graph", action="store_true", help="eager launches instead of a CUDA graph")
    p.add_argument("--precision", default="bf16", choices=["bf16", "tf32"],
                   help="bf16 (fast) or tf32 = fp32 tensors with TF32 convolutions, the reference's precision")
    if distributed:
        p.add_argument("--init-method", default=None,
                       help="tcp://host:port (with --rank/--world-size), env://, slurm; default: auto")
        p.add_argument("--dist-backend", default=None, help="b200 (default on GPU), nccl, gloo")
        p.add_argument("--rank", type=int, default=None)
        p.add_argument("--world-size", type=int, default=None)
    return p


def cifar_dataset(args):
    """CIFAR-10 from --dataset-dir when torchvision finds it there (never downloads: the build
    machines have no network, pytorch/single_gpu.py:57 would fail); otherwise synthetic data."""
    if not args.synthetic:
        try:
            import torchvision.transforms as T
            from torchvision.datasets import CIFAR10
            tf = T.Compose([T.RandomCrop(32, padding=4), T.RandomHorizontalFlip(), T.ToTensor(),
                            T.Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))])
            return CIFAR10(root=args.dataset_dir, train=True, download=False, transform=tf)
        except Exception as exc:  # noqa: BLE001
            print(f"CIFAR-10 not available under {args.dataset_dir} ({exc}); using synthetic data")
    return dtdl.SyntheticDataset("cifar", length=args.synthetic or 5000, seed=args.seed)


def train_epoch(epoch, step, loader, device, log_interval, logger, rank_prefix=True):
    """The reference's train() (pytorch/single_gpu.py:88-120) without its per-step host syncs:
    loss / accuracy accumulate on the device and are read at the log interval."""
    meter = dtdl.utils.TrainMeter(device)
    epoch_start = time.time()
    n = len(loader)
    # batches reach the step through a ring of reusable page-locked buffers filled by a background
    # thread, so the host-to-device copy is asynchronous and nothing is pinned per step
    feed = dtdl.utils.PinnedPrefetcher(loader, device=device) if device.type == "cuda" else loader
    for batch_idx, (inputs, targets) in enumerate(feed):
        start = time.time()
        out = step(inputs, targets)
        stats = out if out.numel() == 2 else step.last_stats
        meter.update(stats[0], targets=targets, correct=stats[1])
        if batch_idx % log_interval == 0:
            line = meter.line(epoch, batch_idx, n, time.time() - start)
            logger.info(line) if rank_prefix else print(line, flush=True)
    msg = "Training time {}".format(datetime.timedelta(seconds=time.time() - epoch_start))
    logger.info(msg) if rank_prefix else print(msg, flush=True)
