#!/usr/bin/env bash
# 2-GPU bench under a few communication settings (one line each).
set -u
run() { env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29550 bench.py --gpus 2 --steps 30 --warmup 5 --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['value'])"; }
run DTDL_X=default
run DTDL_BROADCAST_BUFFERS=0
run DTDL_COMM_BLOCKS=8
run DTDL_COMM_BLOCKS=64
run DTDL_BUCKET_MB=8
run DTDL_WGRAD_FORK=0
