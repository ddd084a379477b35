#!/usr/bin/env bash
# Builds libdtdl_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
OUT=distributed_training_dl_b200/lib
mkdir -p "$OUT" build
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC
       -Xptxas -v --expt-relaxed-constexpr -Iinclude)
SRCS=(csrc/common/host.cpp csrc/common/api.cpp csrc/ops/conv_gemm.cu csrc/ops/conv_gemm_v2.cu csrc/ops/conv_wgrad.cu
      csrc/ops/nn_ops.cu csrc/ops/pack.cu csrc/comm/collectives.cu csrc/comm/vmm.cpp)
OBJS=()
pids=()
for s in "${SRCS[@]}"; do
  o=build/$(echo "$s" | tr '/' '_').o
  OBJS+=("$o")
  if [[ ! -f "$o" || "$s" -nt "$o" || csrc/common/ptx.cuh -nt "$o" || csrc/common/host.h -nt "$o" || csrc/ops/epilogue.cuh -nt "$o" || include/dtdl_b200.h -nt "$o" ]]; then
    ( "$NVCC" "${FLAGS[@]}" -x cu -c "$s" -o "$o" > "build/$(basename "$o").log" 2>&1 || { cat "build/$(basename "$o").log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [[ -n "$p" ]] && wait "$p"; done
"$NVCC" -shared -o "$OUT/libdtdl_b200.so" "${OBJS[@]}" -lcudart
echo "built $OUT/libdtdl_b200.so"
