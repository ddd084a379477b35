#!/usr/bin/env python
"""MLP on MNIST, one process per GPU across nodes (the job of chainer/train_mnist_multi.py).
Launch with torchrun, srun, or --init-method/--rank/--world-size instead of mpiexec."""
from _chainer_common import MLP, add_reporting, ch, flags, get_mnist


def main():
    p = flags("MLP on MNIST, multi node")
    p.add_argument("--communicator", default="pure_nccl", help="pure_nccl (GPU peer memory) or naive (CPU)")
    p.add_argument("--gpu", "-g", action="store_true")
    args = p.parse_args()
    comm = ch.create_communicator(args.communicator if args.gpu else "naive")
    if comm.rank == 0:
        print(f"Num process: {comm.size}  communicator: {args.communicator}  unit: {args.unit}  "
              f"minibatch: {args.batchsize}  epoch: {args.epoch}")
    model = ch.Classifier(MLP(args.unit, 10))
    optimizer = ch.create_multi_node_optimizer(ch.Adam(), comm).setup(model)
    train, test = get_mnist(args.dataset_dir) if comm.rank == 0 else (None, None)
    train = ch.scatter_dataset(train, comm, shuffle=True)
    test = ch.scatter_dataset(test, comm, shuffle=True)
    updater = ch.StandardUpdater(ch.SerialIterator(train, args.batchsize), optimizer, device=comm.device)
    trainer = ch.Trainer(updater, (args.epoch, "epoch"), out=args.out)
    evaluator = ch.Evaluator(ch.SerialIterator(test, args.batchsize, repeat=False, shuffle=False), model)
    trainer.extend(ch.create_multi_node_evaluator(evaluator, comm))
    add_reporting(trainer, args, rank0=comm.rank == 0)
    trainer.run()


if __name__ == "__main__":
    main()
