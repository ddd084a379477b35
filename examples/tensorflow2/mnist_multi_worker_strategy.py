#!/usr/bin/env python
"""Keras-style MNIST CNN, one process per worker (the job of
tensorflow2/mnist_multi_worker_strategy.py):
  python mnist_multi_worker_strategy.py --worker_hosts a:2222,b:2222 --task_index 0   (on a)
  python mnist_multi_worker_strategy.py --worker_hosts a:2222,b:2222 --task_index 1   (on b)
Without --worker_hosts the torchrun / SLURM environment is used."""
from _tf2_common import flags, ks, train_and_eval

if __name__ == "__main__":
    args = flags("MNIST CNN, MultiWorkerMirroredStrategy", multi_worker=True).parse_args()
    strategy = ks.MultiWorkerMirroredStrategy(worker_hosts=args.worker_hosts, task_index=args.task_index,
                                              job_name=args.job_name)
    train_and_eval(args, strategy)
