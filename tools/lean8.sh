#!/usr/bin/env bash
# Lean multi-GPU evidence visit (GPU-minutes are scarce): see tools/run_multi_gpu_checks.sh for the full one.
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
step() { echo "=== $* ($(date +%T))"; }
step tests
DTDL_TEST_WORLD=$N DTDL_TEST_MAX_BYTES=$((1 << 30)) timeout 420 python -m pytest tests/test_comm_gpu.py -x -q -s -m gpu -k "sweep or ddp_training or bf16" \
  > gpurun_out/comm_tests_w${N}.log 2>&1
grep -c "exact on every rank" gpurun_out/comm_tests_w${N}.log; tail -2 gpurun_out/comm_tests_w${N}.log
step timeline
timeout 150 $TR --master-port 29551 tools/comm_timeline.py --precision bf16 --out gpurun_out/timeline_w${N}_bf16.md 2>&1 | grep "^## rank" | head -8
step bench
timeout 300 $TR --master-port 29553 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_ours_w${N}.json 2> gpurun_out/bench_ours_w${N}.err || tail -5 gpurun_out/bench_ours_w${N}.err
python - <<PY
import json
try:
    o = json.loads([l for l in open("gpurun_out/bench_ours_w${N}.json").read().splitlines() if l.startswith("{")][-1])
    print("tf32", round(o["value"]), round(o["ms_per_step"], 2), o["clocks"])
    for k, v in o.get("records", {}).items():
        print(k, round(v["value"]), round(v["ms_per_step"], 2), v["clocks"])
except Exception as e:
    print("bench parse failed", e)
PY
step dp
timeout 240 python bench.py --mode dp --gpus $N --steps 20 --no-sub > gpurun_out/dp_ours_w${N}.json 2> gpurun_out/dp_ours_w${N}.err || tail -5 gpurun_out/dp_ours_w${N}.err
timeout 240 python bench.py --impl reference --mode dp --gpus $N --steps 8 --warmup 3 --no-sub > gpurun_out/dp_ref_w${N}.json 2> gpurun_out/dp_ref_w${N}.err || tail -5 gpurun_out/dp_ref_w${N}.err
grep -h -o '"impl": "[a-z]*", "value": [0-9.]*' gpurun_out/dp_ours_w${N}.json gpurun_out/dp_ref_w${N}.json
step sweep
DTDL_NVLS=$([ "$N" -ge 4 ] && echo 1 || echo 0) timeout 240 $TR --master-port 29555 bench/allreduce_sweep.py --max-mb 1024 \
  --out gpurun_out/allreduce_sweep_w${N}.md > gpurun_out/sweep_w${N}.log 2>&1 || tail -20 gpurun_out/sweep_w${N}.log
cut -c1-220 gpurun_out/allreduce_sweep_w${N}.md | tail -13
if [ "$N" -ge 4 ]; then
  step nvls_tune
  for v in "4 512" "8 256"; do
    set -- $v
    DTDL_NVLS=1 DTDL_NVLS_UNROLL=$1 DTDL_NVLS_THREADS=$2 timeout 150 $TR --master-port 29556 bench/nvls_tune.py \
      --sizes-mb 4,64,256 --blocks 32,64,128 --out gpurun_out/nvls_tune_w${N}_u$1_t$2.md 2>&1 | grep "^|" | tail -9
  done
fi
step done
