// Epilogue helpers shared by the implicit-GEMM kernels: per-channel sum / sum-of-squares of the
// tile that is being written (the statistics the following training-mode BatchNorm needs),
// taken from the fp32 accumulators after rounding to the stored bf16 value.  This is the
// "statistics reduction fused into the producer conv epilogue" of SURVEY.md 2.5 K-O5; the
// reference runs a separate BatchNorm kernel after every convolution (pytorch/model.py:41-46).
//
// Inside the CTA nothing is atomic: every epilogue warp owns one row part[2*n_tile] of a scratch
// area (the kernels lend their pipeline stage memory, idle once the accumulator is complete), and
// lane L stores the warp total of column/quantity L.  (Shared-memory float atomics are
// compare-and-swap loops; four warps hitting the same 32 words every chunk made them the
// dominant cost of an earlier version.)  The flush adds the four rows and issues one global
// atomic per channel and quantity per CTA.
#pragma once
#include "../common/ptx.cuh"

namespace dtdl {

// Butterfly reduce-scatter across a warp: on return lane L holds sum over all lanes of v[L].
// 31 shuffles for 32 values (a plain all-reduce per value would need 160).
__device__ __forceinline__ float warp_reduce_scatter32(float (&v)[32], int lane) {
#pragma unroll
  for (int half = 16; half >= 1; half >>= 1) {
    const bool upper = (lane & half) != 0;
#pragma unroll
    for (int j = 0; j < half; ++j) {
      const float send = upper ? v[j] : v[j + half];
      const float keep = upper ? v[j + half] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, half);
    }
  }
  return v[0];
}

// One 16-column chunk of one warp's 32 rows -> this warp's row of partial sums.  `first` is false
// when the warp visits the same columns again (second sub-tile of the halo kernel).
__device__ __forceinline__ void epilogue_col_stats_part(const float (&f)[16], bool valid, int lane,
                                                        float* part, int n_tile, int jcol, bool first) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float x = valid ? __bfloat162float(__float2bfloat16(f[i])) : 0.f;
    v[i] = x;
    v[16 + i] = x * x;
  }
  const float tot = warp_reduce_scatter32(v, lane);
  float* dst = part + (lane < 16 ? 0 : n_tile) + jcol + (lane & 15);
  *dst = first ? tot : *dst + tot;
}

// named barrier over the 128 epilogue threads (id 1; id 0 is __syncthreads)
__device__ __forceinline__ void epilogue_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// parts = [4 warps][2*n_tile]; stats = [2][Cout_p] in global memory (pre-zeroed by the caller).
__device__ __forceinline__ void epilogue_stats_flush_parts(const float* parts, float* stats, int n_tile,
                                                           int col_base, int Cout_p, int etid) {
  epilogue_bar();
  const int w = 2 * n_tile;
  for (int t = etid; t < w; t += 128) {
    const int col = col_base + (t % n_tile);
    if (col < Cout_p)
      atomicAdd(stats + (t < n_tile ? 0 : Cout_p) + col, parts[t] + parts[w + t] + parts[2 * w + t] + parts[3 * w + t]);
  }
}

}  // namespace dtdl
