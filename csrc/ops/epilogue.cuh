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
// when the warp visits the same columns again (second sub-tile of the halo kernel).  ROUND_BF16:
// the statistics are those of the stored bf16 values (fp32 outputs are summed as they are).
template <bool ROUND_BF16>
__device__ __forceinline__ void epilogue_col_stats_part(const float (&f)[16], bool valid, int lane,
                                                        float* part, int n_tile, int jcol, bool first) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float x = !valid ? 0.f : (ROUND_BF16 ? __bfloat162float(__float2bfloat16(f[i])) : f[i]);
    v[i] = x;
    v[16 + i] = x * x;
  }
  const float tot = warp_reduce_scatter32(v, lane);
  float* dst = part + (lane < 16 ? 0 : n_tile) + jcol + (lane & 15);
  *dst = first ? tot : *dst + tot;
}

// 16 consecutive output channels of one pixel; `remain` = channels left in the row from f[0] on
// (a multiple of 8: rows are padded to 8 channels, tiles to 16).
__device__ __forceinline__ void epilogue_store_f32(float* out, const float (&f)[16], int remain) {
#pragma unroll
  for (int h = 0; h < 4; ++h)
    if (h * 4 < remain)
      *reinterpret_cast<float4*>(out + h * 4) = make_float4(f[h * 4], f[h * 4 + 1], f[h * 4 + 2], f[h * 4 + 3]);
}
__device__ __forceinline__ void epilogue_store_bf16(__nv_bfloat16* out, const float (&f)[16], int remain) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    if (h * 8 < remain) {
      uint4 pk;
      __nv_bfloat162 p0 = __floats2bfloat162_rn(f[h * 8 + 0], f[h * 8 + 1]);
      __nv_bfloat162 p1 = __floats2bfloat162_rn(f[h * 8 + 2], f[h * 8 + 3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(f[h * 8 + 4], f[h * 8 + 5]);
      __nv_bfloat162 p3 = __floats2bfloat162_rn(f[h * 8 + 6], f[h * 8 + 7]);
      pk.x = *reinterpret_cast<uint32_t*>(&p0);
      pk.y = *reinterpret_cast<uint32_t*>(&p1);
      pk.z = *reinterpret_cast<uint32_t*>(&p2);
      pk.w = *reinterpret_cast<uint32_t*>(&p3);
      *reinterpret_cast<uint4*>(out + h * 8) = pk;
    }
  }
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
