#!/usr/bin/env python
"""One process per GPU -- the job of the reference's pytorch/distributed_data_parallel.py.

Launch any of these ways:
  torchrun --nproc-per-node 8 distributed_data_parallel.py                  (env://)
  srun -n 16 python distributed_data_parallel.py                            (SLURM)
  python distributed_data_parallel.py --init-method tcp://c1:20201 --rank r --world-size W
  python distributed_data_parallel.py --spawn 8                             (mp.spawn, one node)

Batch size.  ``--batch-size`` is the batch of ONE process, i.e. of one GPU: the global
batch is ``batch_size x world_size`` and every epoch has ``len(dataset) / (batch_size x world_size)``
steps.  The reference divides the flag by the GPUs visible to the process instead
(pytorch/distributed_data_parallel.py:71, ``args.batch_size = int(args.batch_size / ngpus_per_node)``);
its README launches one process per shell with a single visible GPU, where that division is by one,
so the effective rule is the same: 64 per process, 64 x W globally.  Launchers that make all GPUs
visible to every process (torchrun, ``--spawn``) therefore do NOT shrink the per-process batch here --
pass ``--batch-size 8`` yourself if you want a global batch of 64 on 8 GPUs.

``--precision tf32`` trains at the reference's own precision (fp32 tensors, TF32 convolutions),
``bf16`` (default) is the fast mode.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _common import cifar_dataset, dtdl, pytorch_flags, train_epoch  # noqa: E402


def main_worker(local_rank, args):
    comm = dtdl.init(backend=args.dist_backend, init_method=args.init_method, rank=args.rank,
                     world_size=args.world_size)
    logger = dtdl.utils.RankLogger()
    logger.info(f"Use device {comm.device} for training")
    torch.manual_seed(args.seed)
    net = dtdl.wrap(dtdl.models.pyramidnet().to(comm.device), comm=comm, dtype=args.precision)
    logger.info(f"per-process batch {args.batch_size}, global batch {args.batch_size * comm.size} "
                f"({comm.size} process(es)); precision {args.precision}")
    logger.info(f"The number of parameters of model is {sum(p.numel() for p in net.parameters())}")
    dataset = cifar_dataset(args)
    sampler = dtdl.DistributedSampler(dataset)
    loader = torch.utils.data.DataLoader(dataset, batch_size=args.batch_size, sampler=sampler,
                                         num_workers=args.num_workers, drop_last=True)
    optimizer = dtdl.FusedSGD(net.parameters(), lr=args.learning_rate, momentum=args.momentum,
                              weight_decay=args.weight_decay)
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=2, gamma=0.1)
    step = dtdl.TrainStep(net, optimizer, dtdl.CrossEntropyLoss(), graph=not args.no_graph)
    net.reducer.bucket_timing = args.bucket_timing    # eager (warm-up / --no-graph) steps are timed
    net.train()
    for epoch in range(1, args.epochs + 1):
        sampler.set_epoch(epoch)   # the reference never reshuffles between epochs
        train_epoch(epoch, step, loader, comm.device, args.log_interval, logger)
        scheduler.step()
        if args.bucket_timing:
            rows = net.reducer.bucket_report()
            if rows:
                logger.info("gradient buckets (collective time, achieved bus bandwidth):\n"
                            + dtdl.utils.format_bucket_report(rows))
    if args.save_model:
        path = dtdl.save_model(net.module, os.path.join(args.train_dir, "distributed_data_parallel.pth"),
                               per_rank=True)
        logger.info(f"model saved to {path}")
    dtdl.shutdown()


def main():
    parser = pytorch_flags("B200-native CIFAR-10 training, one process per GPU", distributed=True)
    parser.add_argument("--spawn", type=int, default=0, help="start N local worker processes")
    parser.add_argument("--bucket-timing", action="store_true",
                        help="report per-bucket all-reduce time and bus bandwidth after each epoch")
    args = parser.parse_args()
    if args.spawn:
        dtdl.spawn(main_worker, nprocs=args.spawn, args=(args,))
    else:
        main_worker(int(os.environ.get("LOCAL_RANK", 0)), args)


if __name__ == "__main__":
    main()
