#!/usr/bin/env python
"""MLP on MNIST, CPU or one GPU (the job of chainer/train_mnist.py)."""
from _chainer_common import MLP, add_reporting, ch, flags, get_mnist


def main():
    p = flags("MLP on MNIST")
    p.add_argument("--gpu", "-g", type=int, default=-1, help="GPU id (negative: CPU)")
    args = p.parse_args()
    print(f"# gpu = {args.gpu}, unit = {args.unit}, minibatch = {args.batchsize}, epoch = {args.epoch}")
    model = ch.Classifier(MLP(args.unit, 10))
    optimizer = ch.Adam().setup(model)
    train, test = get_mnist(args.dataset_dir)
    updater = ch.StandardUpdater(ch.SerialIterator(train, args.batchsize), optimizer, device=args.gpu)
    trainer = ch.Trainer(updater, (args.epoch, "epoch"), out=args.out)
    trainer.extend(ch.Evaluator(ch.SerialIterator(test, args.batchsize, repeat=False, shuffle=False), model))
    add_reporting(trainer, args)
    trainer.run()


if __name__ == "__main__":
    main()
