#!/usr/bin/env python
"""MLP on MNIST on 1..N GPUs of one process (the job of chainer/train_mnist_gpu.py, which is
limited to two GPUs; this one is not)."""
from _chainer_common import MLP, add_reporting, ch, flags, get_mnist


def main():
    p = flags("MLP on MNIST, ParallelUpdater")
    p.add_argument("--gpu-num", "-g", type=int, default=1)
    args = p.parse_args()
    model = ch.Classifier(MLP(args.unit, 10))
    optimizer = ch.Adam().setup(model)
    train, test = get_mnist(args.dataset_dir)
    it = ch.SerialIterator(train, args.batchsize)
    if args.gpu_num <= 1:
        updater = ch.StandardUpdater(it, optimizer, device=0)
    else:
        devices = {"main": 0, **{f"gpu{i}": i for i in range(1, args.gpu_num)}}
        updater = ch.ParallelUpdater(it, optimizer, devices=devices)
    trainer = ch.Trainer(updater, (args.epoch, "epoch"), out=args.out)
    trainer.extend(ch.Evaluator(ch.SerialIterator(test, args.batchsize, repeat=False, shuffle=False), model))
    add_reporting(trainer, args)
    trainer.run()


if __name__ == "__main__":
    main()
