#!/usr/bin/env bash
# One multi-GPU box visit that leaves evidence under gpurun_out/ (copy what should be judged to
# profiles/): collective / DDP / front-end tests at full world size with the exact-sum sweep up to
# 1 GiB, the in-graph collective timeline, the benchmark on both arms (DDP style and
# single-process DataParallel style), the all-reduce sweep against NCCL and the NVLS kernel-variant
# table against the NVLink roofline.
# usage: tools/run_multi_gpu_checks.sh N [quick]   (writes gpurun_out/*_w${N}*)
set -u
N=${1:-8}
QUICK=${2:-}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
step() { echo "=== $* ($(date +%T))"; }

step "pytest tests/test_comm_gpu.py at world $N, exact sums up to 1 GiB"
DTDL_TEST_WORLD=$N DTDL_TEST_MAX_BYTES=$((1 << 30)) timeout 900 python -m pytest tests/test_comm_gpu.py -x -q -s -m gpu \
  > gpurun_out/comm_tests_w${N}.log 2>&1
grep -c "exact on every rank" gpurun_out/comm_tests_w${N}.log; tail -3 gpurun_out/comm_tests_w${N}.log

step "collective timeline inside the step graph"
timeout 200 $TR --master-port 29551 tools/comm_timeline.py --precision bf16 --out gpurun_out/timeline_w${N}_bf16.md 2>&1 | grep "^## rank" | head -8
timeout 200 $TR --master-port 29552 tools/comm_timeline.py --precision tf32 --out gpurun_out/timeline_w${N}_tf32.md 2>&1 | grep "^## rank" | head -2

step "benchmark, DDP style, both arms"
timeout 400 $TR --master-port 29553 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_ours_w${N}.json 2> gpurun_out/bench_ours_w${N}.err || tail -5 gpurun_out/bench_ours_w${N}.err
timeout 600 $TR --master-port 29554 bench.py --impl reference --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_ref_w${N}.json 2> gpurun_out/bench_ref_w${N}.err || tail -5 gpurun_out/bench_ref_w${N}.err
python - <<EOF
import json
def last(p):
    try:
        return json.loads([l for l in open(p).read().splitlines() if l.startswith("{")][-1])
    except Exception:
        return None
o, r = last("gpurun_out/bench_ours_w${N}.json"), last("gpurun_out/bench_ref_w${N}.json")
for name, get in [("pyramidnet tf32", lambda d: d), ("pyramidnet bf16", lambda d: d["records"]["pyramidnet_bf16"]),
                  ("resnet50 bf16", lambda d: d["records"]["resnet50_bf16"])]:
    try:
        a = get(o); b = get(r) if r else None
        print(f"{name}: ours {a['value']:.0f} img/s ({a['ms_per_step']:.2f} ms, clocks {a['clocks']})"
              + (f" | reference {b['value']:.0f} ({b['ms_per_step']:.2f} ms) ratio {a['value']/b['value']:.2f}" if b else ""))
    except Exception as e:
        print(name, "missing", e)
EOF

if [ -z "$QUICK" ]; then
  step "benchmark, single-process DataParallel style, both arms"
  timeout 300 python bench.py --mode dp --gpus $N --steps 20 --no-sub > gpurun_out/dp_ours_w${N}.json 2> gpurun_out/dp_ours_w${N}.err || tail -5 gpurun_out/dp_ours_w${N}.err
  timeout 300 python bench.py --impl reference --mode dp --gpus $N --steps 10 --warmup 3 --no-sub > gpurun_out/dp_ref_w${N}.json 2> gpurun_out/dp_ref_w${N}.err || tail -5 gpurun_out/dp_ref_w${N}.err
  grep -h -o '"impl": "[a-z]*", "value": [0-9.]*' gpurun_out/dp_ours_w${N}.json gpurun_out/dp_ref_w${N}.json

  step "all-reduce sweep 1 KiB - 1 GiB vs NCCL"
  DTDL_NVLS=$([ "$N" -ge 4 ] && echo 1 || echo 0) timeout 300 $TR --master-port 29555 bench/allreduce_sweep.py --max-mb 1024 \
    --out gpurun_out/allreduce_sweep_w${N}.md > gpurun_out/sweep_w${N}.log 2>&1 || tail -20 gpurun_out/sweep_w${N}.log
  cut -c1-200 gpurun_out/allreduce_sweep_w${N}.md | tail -14

  if [ "$N" -ge 4 ]; then
    step "NVLS fused reduce+SGD vs roofline, kernel variants"
    for v in "4 512" "8 512" "8 256"; do
      set -- $v
      DTDL_NVLS=1 DTDL_NVLS_UNROLL=$1 DTDL_NVLS_THREADS=$2 timeout 200 $TR --master-port 29556 bench/nvls_tune.py \
        --sizes-mb 4,64,256 --blocks 32,64,128 --out gpurun_out/nvls_tune_w${N}_u$1_t$2.md 2>&1 | grep "^|" | tail -9
    done
  fi
fi
step done
