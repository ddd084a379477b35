#!/usr/bin/env python
"""Single-device training of PyramidNet on CIFAR-10 -- the job of the reference's
pytorch/single_gpu.py, on this framework's kernels.  `python single_gpu.py -g 1 --synthetic 2048`"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _common import cifar_dataset, dtdl, pytorch_flags, train_epoch  # noqa: E402


def main():
    args = pytorch_flags("B200-native CIFAR-10 training, one device").parse_args()
    if args.gpu_nums > 1:
        raise SystemExit("single_gpu.py drives one device; use data_parallel.py or distributed_data_parallel.py")
    torch.manual_seed(args.seed)  # the reference parses --seed but never applies it
    device = torch.device("cuda:0" if torch.cuda.is_available() and args.gpu_nums != 0 or
                          (args.gpu_nums == 0 and torch.cuda.is_available()) else "cpu")
    comm = dtdl.init(backend="b200" if device.type == "cuda" else "gloo", device=device)
    loader = torch.utils.data.DataLoader(cifar_dataset(args), batch_size=args.batch_size, shuffle=True,
                                         num_workers=args.num_workers, drop_last=True)
    print("==> Making model..")
    net = dtdl.wrap(dtdl.models.pyramidnet().to(device), comm=comm, dtype=args.precision if device.type == "cuda" else None)
    print("The number of parameters of model is", sum(p.numel() for p in net.parameters() if p.requires_grad))
    optimizer = dtdl.FusedSGD(net.parameters(), lr=args.learning_rate, momentum=args.momentum,
                              weight_decay=args.weight_decay)
    step = dtdl.TrainStep(net, optimizer, dtdl.CrossEntropyLoss(), graph=not args.no_graph)
    net.train()
    logger = dtdl.utils.RankLogger()
    for epoch in range(1, args.epochs + 1):
        train_epoch(epoch, step, loader, device, args.log_interval, logger, rank_prefix=False)
    if args.save_model:
        path = dtdl.save_model(net.module, os.path.join(args.train_dir, "single_gpu_model.pth"))
        print("model saved to", path)


if __name__ == "__main__":
    main()
