#!/usr/bin/env bash
# Full single-GPU check: test suite, smoke(), both bench arms, ResNet-50 line.  Writes gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py --impl reference --steps 30 --warmup 5 > gpurun_out/bench1_ref.log 2>&1; tail -1 gpurun_out/bench1_ref.log | tee gpurun_out/bench1_ref.json | cut -c1-250
timeout 300 python bench.py --steps 50 --warmup 5 > gpurun_out/bench1_ours.log 2>&1; tail -1 gpurun_out/bench1_ours.log | tee gpurun_out/bench1_ours.json
timeout 300 python bench.py --model resnet50 --steps 30 --warmup 5 > gpurun_out/bench1_rn.log 2>&1; tail -1 gpurun_out/bench1_rn.log | tee gpurun_out/bench1_rn_ours.json | cut -c1-250
