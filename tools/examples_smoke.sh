#!/usr/bin/env bash
# Runs every example script for one short epoch on synthetic data on ONE GPU and prints PASS / FAIL
# per script (logs under gpurun_out/examples/).  The scripts that open a rendezvous run one after the
# other (they would share the default port); the rest run side by side.
# usage: tools/examples_smoke.sh
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/examples
mkdir -p $OUT
export MASTER_ADDR=127.0.0.1
run() {  # name, timeout, command...
  local name=$1 limit=$2; shift 2
  if timeout "$limit" "$@" > $OUT/$name.log 2>&1; then
    echo "PASS $name: $(grep -v '^\s*$' $OUT/$name.log | tail -1 | cut -c1-140)"
  else
    echo "FAIL $name (rc $?)"; tail -12 $OUT/$name.log | cut -c1-240
  fi
}
P=examples/pytorch T=examples/tensorflow2 C=examples/chainer
run single_gpu 150 python $P/single_gpu.py -g 1 --synthetic 512 -e 1 -td $OUT/td_single &
run data_parallel 150 python $P/data_parallel.py -g 1 --synthetic 512 -e 1 -td $OUT/td_dp &
run mnist_single 150 python $T/mnist_single.py -e 1 -td $OUT/td_tf_single &
run mnist_mirror 150 python $T/mnist_mirror_strategy.py -e 1 -g 1 -td $OUT/td_tf_mirror &
run train_mnist 150 python $C/train_mnist.py -e 1 -g 0 -u 128 -o $OUT/ch_single &
run train_mnist_gpu 150 python $C/train_mnist_gpu.py -e 1 -g 1 -u 128 -o $OUT/ch_gpu &
(
  run ddp_spawn 150 python $P/distributed_data_parallel.py --spawn 1 --synthetic 512 -e 1 -td $OUT/td_ddp
  MASTER_PORT=29611 run mnist_multi_worker 150 python $T/mnist_multi_worker_strategy.py -e 1 \
    --worker_hosts 127.0.0.1:29612 --task_index 0 -td $OUT/td_tf_multi
  MASTER_PORT=29613 RANK=0 WORLD_SIZE=1 run train_mnist_multi 150 python $C/train_mnist_multi.py -e 1 -g -u 128 -o $OUT/ch_multi
) &
wait
