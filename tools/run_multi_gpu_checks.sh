#!/usr/bin/env bash
# One multi-GPU box visit: NVLS tests at full world size, all-reduce sweep with the multicast
# columns, and the flagship bench with the peer-pointer and the multicast collectives.
# usage: tools/run_multi_gpu_checks.sh N   (writes gpurun_out/*_${N}gpu*)
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
DTDL_TEST_WORLD=$N timeout 240 python -m pytest tests/test_comm_gpu.py -x -q -k "nvls" 2>&1 | tail -4
DTDL_NVLS=1 timeout 200 $TR --master-port 29541 bench/allreduce_sweep.py --max-mb 256 \
  --out gpurun_out/allreduce_sweep_nvls_${N}gpu.md > gpurun_out/sweep_${N}.log 2>&1 || tail -20 gpurun_out/sweep_${N}.log
for nv in 0 1; do
  DTDL_NVLS=$nv timeout 200 $TR --master-port $((29542 + nv)) bench.py --gpus $N --steps 30 --warmup 5 \
    > gpurun_out/bench${N}_nvls${nv}.log 2>&1
  tail -1 gpurun_out/bench${N}_nvls${nv}.log | tee gpurun_out/bench${N}_ours_nvls${nv}.json | cut -c1-330
done
cat gpurun_out/allreduce_sweep_nvls_${N}gpu.md 2>/dev/null | cut -c1-260
