#!/usr/bin/env python
"""One process, several GPUs -- the job of the reference's pytorch/data_parallel.py
(nn.DataParallel), on this framework's single-process engine.
`python data_parallel.py -g 2 --synthetic 4096`"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _common import cifar_dataset, dtdl, pytorch_flags, train_epoch  # noqa: E402


def main():
    args = pytorch_flags("B200-native CIFAR-10 training, single-process data parallel").parse_args()
    if args.gpu_nums < 1 or torch.cuda.device_count() < args.gpu_nums:
        raise SystemExit(f"data_parallel.py needs -g N with 1 <= N <= visible GPUs (have {torch.cuda.device_count()}); "
                         "one GPU runs a single replica, as nn.DataParallel does")
    torch.manual_seed(args.seed)
    device = torch.device("cuda:0")
    loader = torch.utils.data.DataLoader(cifar_dataset(args), batch_size=args.batch_size, shuffle=True,
                                         num_workers=args.num_workers, drop_last=True)
    print("==> Making model..")
    model = dtdl.DataParallel(dtdl.set_precision(dtdl.models.pyramidnet(), args.precision),
                              device_ids=list(range(args.gpu_nums)))
    print("The number of parameters of model is", sum(p.numel() for p in model.parameters() if p.requires_grad))
    optimizer = dtdl.FusedSGD(model.parameters(), lr=args.learning_rate, momentum=args.momentum,
                              weight_decay=args.weight_decay)
    step = dtdl.DataParallelStep(model, optimizer, graph=not args.no_graph)
    model.train()
    logger = dtdl.utils.RankLogger()
    for epoch in range(1, args.epochs + 1):
        train_epoch(epoch, step, loader, device, args.log_interval, logger, rank_prefix=False)
    if args.save_model:  # wrapper spelling ("module." keys), as the reference saves it
        path = dtdl.save_model(model, os.path.join(args.train_dir, "data_parallel_model.pth"))
        print("model saved to", path)


if __name__ == "__main__":
    main()
