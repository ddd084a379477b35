// Implicit-GEMM convolution forward / data-gradient, second generation: shared-memory reuse
// across filter rows and across two pixel sub-tiles.  Same capability as conv_gemm.cu (K-O1 / K-O3
// of SURVEY.md 2.5: the 3x3 convolutions of pytorch/model.py:28-32,62-63 and their data gradients,
// which the reference gets from cuDNN), for the 32- and 16-pixel-wide stages.
//
// The first-generation kernel (conv_gemm.cu) loads one shifted activation box per filter tap,
// i.e. the same pixels travel L2 -> shared memory KH*KW times, and every 128-pixel tile
// re-loads the full weight slice.  Measured on B200 those kernels sit at the L2 -> SM
// bandwidth limit (~10 TB/s aggregate), not at the tensor-core limit.  This kernel cuts that
// traffic ~2.2x for the 32x32 and 16x16 stages of PyramidNet:
//
//   * For each filter COLUMN s and 64-channel chunk, ONE activation box is loaded that is
//     KH-1 image rows taller than the output tile.  Because the tile spans full image rows
//     whose width is a multiple of 8 pixels, the operand for filter ROW r is the same box
//     viewed r*BW rows further down -- a multiple of the 8-row / 1024-byte swizzle atom, so
//     it is just a different start address in the tcgen05 shared-memory descriptor.
//   * A CTA owns up to two vertically adjacent 128-pixel sub-tiles (two TMEM accumulators);
//     both use the same weight tiles, halving weight traffic per pixel.
//
// Pipeline stage = {activation halo box, KH weight boxes}; per stage the MMA thread issues
// sub-tiles x KH x (<=4) tcgen05.mma instructions.  Roles as in conv_gemm.cu.
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"
#include "epilogue.cuh"

#include <stdlib.h>

namespace dtdl {

struct ConvV2Args {
  int KH, KW, pad_h, pad_w;
  int chunks, last_slices, Cin_p;
  int CB;              // channels per 128-byte row: 64 (bf16) or 32 (fp32 read as TF32)
  int BW, BH, MS;      // sub-tile = BW x BH = 128 pixels, MS sub-tiles per CTA
  int halo_rows;       // (MS*BH + KH - 1) * BW
  int tiles_h;
  int Ho, Wo, N;
  int Cout_p, n_tile, tmem_cols;
  int stages, a_stage_bytes, b_box_bytes, stage_bytes;
  int flags;
  void* y;
  const float* bias;
  float* stats;
};

template <int ACT>
__global__ void __launch_bounds__(256, 2)
conv_gemm_v2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const ConvV2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.stages * a.stage_bytes);
  uint64_t* empty = full + a.stages;
  uint64_t* tmem_full = empty + a.stages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int th_i = blockIdx.x % a.tiles_h;
  const int n_img = blockIdx.x / a.tiles_h;
  const int h0 = th_i * a.MS * a.BH;
  const int nt = blockIdx.y;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < a.stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, (uint32_t)a.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int n_groups = a.KW * a.chunks;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx = (uint32_t)(a.halo_rows * 128 + a.KH * a.n_tile * 128);
      int stage = 0, phase = 0;
      for (int s = 0; s < a.KW; ++s) {
        for (int c = 0; c < a.chunks; ++c) {
          uint8_t* st = smem + stage * a.stage_bytes;
          mbar_wait(&empty[stage], phase ^ 1, 21);
          mbar_arrive_expect_tx(&full[stage], tx);
          tma_load_4d(st, &tmA, &full[stage], c * a.CB, s - a.pad_w, h0 - a.pad_h, n_img);
          for (int r = 0; r < a.KH; ++r)
            tma_load_2d(st + a.a_stage_bytes + r * a.b_box_bytes, &tmB, &full[stage],
                        (r * a.KW + s) * a.Cin_p + c * a.CB, nt * a.n_tile);
          if (++stage == a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc<ACT>(128, a.n_tile, 0, 0);
      const uint64_t desc0 = umma_smem_desc(smem_u32(smem), 16, 1024);
      const uint32_t row_bytes = (uint32_t)(a.BW * 128);          // one image row of the halo box
      const uint32_t sub_bytes = (uint32_t)(a.BH * a.BW * 128);   // one 128-pixel sub-tile
      int stage = 0, phase = 0, c = 0;
      for (int g = 0; g < n_groups; ++g) {
        mbar_wait(&full[stage], phase, 22);
        tc_fence_after();
        const int nsl = (c == a.chunks - 1) ? a.last_slices : 4;
        if (++c == a.chunks) c = 0;
        const uint64_t sa = umma_desc_advance(desc0, (uint32_t)(stage * a.stage_bytes));
        const uint64_t sb = umma_desc_advance(sa, (uint32_t)a.a_stage_bytes);
        for (int m = 0; m < a.MS; ++m) {
          uint64_t ad = umma_desc_advance(sa, (uint32_t)(m * sub_bytes));
          uint64_t bd = sb;
          for (int r = 0; r < a.KH; ++r) {
            umma<ACT>(tmem_base + m * a.n_tile, ad, bd, idesc, (g > 0 || r > 0) ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < 4; ++k)
              if (k < nsl)
                umma<ACT>(tmem_base + m * a.n_tile, umma_desc_advance(ad, k * 32), umma_desc_advance(bd, k * 32), idesc, 1u);
            ad = umma_desc_advance(ad, row_bytes);
            bd = umma_desc_advance(bd, (uint32_t)a.b_box_bytes);
          }
        }
        umma_commit(&empty[stage]);
        if (++stage == a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int row = q * 32 + lane;
    const int wo = row % a.BW;
    const int hi = row / a.BW;
    const bool relu = a.flags & 1;
    const bool f32 = ACT == 1 || (a.flags & 2);
    // statistics scratch = the pipeline's stage memory (free once the accumulators are complete)
    float* s_parts = reinterpret_cast<float*>(smem);
    const int etid = threadIdx.x - 128;
    mbar_wait(tmem_full, 0, 23);
    tc_fence_after();
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int m = 0; m < a.MS; ++m) {
      const int ho = h0 + m * a.BH + hi;
      const bool valid = ho < a.Ho && wo < a.Wo;
      const size_t pix = ((size_t)n_img * a.Ho + ho) * a.Wo + wo;
      for (int j = 0; j < a.n_tile / 16; ++j) {
        uint32_t v[16];
        tmem_ld16(lane_base + m * a.n_tile + j * 16, v);
        tmem_ld_wait();
        const int col0 = nt * a.n_tile + j * 16;
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          f[i] = __uint_as_float(v[i]);
          if (a.bias != nullptr && col0 + i < a.Cout_p) f[i] += a.bias[col0 + i];
          if (relu) f[i] = fmaxf(f[i], 0.f);
        }
        if (a.stats != nullptr)
          epilogue_col_stats_part<ACT == 0>(f, valid, lane, s_parts + q * 2 * a.n_tile, a.n_tile, j * 16, m == 0);
        if (!valid) continue;
        if (f32) {
          epilogue_store_f32(reinterpret_cast<float*>(a.y) + pix * a.Cout_p + col0, f, a.Cout_p - col0);
        } else {
          epilogue_store_bf16(reinterpret_cast<__nv_bfloat16*>(a.y) + pix * a.Cout_p + col0, f, a.Cout_p - col0);
        }
      }
    }
    if (a.stats != nullptr) epilogue_stats_flush_parts(s_parts, a.stats, a.n_tile, nt * a.n_tile, a.Cout_p, etid);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, (uint32_t)a.tmem_cols);
}

// Returns 0 when launched, -1 when the shape does not fit this kernel (caller falls back to
// the per-tap kernel), >0 on error.
int conv_fprop_v2(const void* x, const void* wpack, void* y, const float* bias, float* stats, int N,
                  int H, int W, int Cin_p, int Ho, int Wo, int Cout_p, int KH, int KW, int pad_h,
                  int pad_w, int flags, int act, cudaStream_t stream) {
  const int esz = act_esize(act);
  if (Wo % 8 != 0 || Wo > 128 || 128 % Wo != 0) return -1;
  const int BW = Wo, BH = 128 / Wo;
  if (Ho < BH) return -1;
  ConvV2Args a;
  a.KH = KH;
  a.KW = KW;
  a.pad_h = pad_h;
  a.pad_w = pad_w;
  a.CB = 128 / esz;
  a.chunks = (Cin_p + a.CB - 1) / a.CB;
  a.last_slices = ((Cin_p - a.CB * (a.chunks - 1)) * esz + 31) / 32;
  a.Cin_p = Cin_p;
  a.BW = BW;
  a.BH = BH;
  const int Np16 = (Cout_p + 15) / 16 * 16;
  // N tiling: a single tile when it fits 192 columns (two accumulators <= 384 TMEM columns)
  // tuning overrides for measurements (tools/bench_kernels.py): widest N tile, sub-tiles per CTA, stages
  static const int env_ntile = getenv("DTDL_V2_NTILE") ? atoi(getenv("DTDL_V2_NTILE")) : 192;
  static const int env_ms = getenv("DTDL_V2_MS") ? atoi(getenv("DTDL_V2_MS")) : 2;
  static const int env_stages = getenv("DTDL_V2_STAGES") ? atoi(getenv("DTDL_V2_STAGES")) : 6;
  int n_tiles = (Np16 + env_ntile - 1) / env_ntile;
  int n_tile = ((Np16 + n_tiles - 1) / n_tiles + 15) / 16 * 16;
  int MS = (Ho >= 2 * BH && env_ms >= 2) ? 2 : 1;
  a.b_box_bytes = (n_tile * 128 + 1023) / 1024 * 1024;
  for (;;) {
    a.halo_rows = (MS * BH + KH - 1) * BW;
    a.a_stage_bytes = (a.halo_rows * 128 + 1023) / 1024 * 1024;
    a.stage_bytes = a.a_stage_bytes + KH * a.b_box_bytes;
    if (2 * a.stage_bytes <= 220 * 1024 && MS * BH + KH - 1 <= 256) break;
    if (MS == 2) {
      MS = 1;
      continue;
    }
    return -1;
  }
  // keep enough CTAs in flight: prefer one sub-tile per CTA when the grid would not fill the GPU
  const long long ctas2 = (long long)N * ((Ho + 2 * BH - 1) / (2 * BH)) * n_tiles;
  if (MS == 2 && ctas2 < sm_count()) {
    MS = 1;
    a.halo_rows = (BH + KH - 1) * BW;
    a.a_stage_bytes = (a.halo_rows * 128 + 1023) / 1024 * 1024;
    a.stage_bytes = a.a_stage_bytes + KH * a.b_box_bytes;
  }
  a.MS = MS;
  a.tiles_h = (Ho + MS * BH - 1) / (MS * BH);
  a.Ho = Ho;
  a.Wo = Wo;
  a.N = N;
  a.Cout_p = Cout_p;
  a.n_tile = n_tile;
  const int cols = MS * n_tile;
  a.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  // Two CTAs per SM whenever two stages fit in half of the shared memory and the grid is deep
  // enough: one CTA's prologue / epilogue then overlaps the other's main loop.
  const long long ctas = (long long)N * a.tiles_h * n_tiles;
  const bool two_per_sm = 2 * a.stage_bytes <= 108 * 1024 && a.tmem_cols <= 256 && ctas > sm_count();
  int stages = ((two_per_sm ? 108 : 220) * 1024) / a.stage_bytes;
  if (stages > env_stages) stages = env_stages;
  const int n_groups = KW * a.chunks;
  if (stages > n_groups) stages = n_groups < 2 ? 2 : n_groups;
  a.stages = stages;
  a.flags = flags;
  a.y = y;
  a.bias = bias;
  a.stats = stats;

  CUtensorMap tmA, tmB;
  {
    uint64_t dims[4] = {(uint64_t)Cin_p, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)Cin_p * esz, (uint64_t)W * Cin_p * esz, (uint64_t)H * W * Cin_p * esz};
    uint32_t box[4] = {(uint32_t)a.CB, (uint32_t)BW, (uint32_t)(MS * BH + KH - 1), 1};
    if (encode_tmap_act(&tmA, x, 4, dims, str, box, act, false)) return 1;
  }
  {
    const uint64_t Ktot = (uint64_t)KH * KW * Cin_p;
    uint64_t dims[2] = {Ktot, (uint64_t)Np16};
    uint64_t str[1] = {Ktot * esz};
    uint32_t box[2] = {(uint32_t)a.CB, (uint32_t)n_tile};
    if (encode_tmap_act(&tmB, wpack, 2, dims, str, box, act, false)) return 1;
  }
  const size_t smem = (size_t)stages * a.stage_bytes + 1024 + (2 * stages + 1) * 8 + 16;
  static bool attr_set[64] = {false};
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    // 227 KB less the static epilogue accumulators
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_v2_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_v2_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set[dev & 63] = true;
  }
  dim3 grid(N * a.tiles_h, n_tiles);
  if (act == kActF32)
    conv_gemm_v2_kernel<1><<<grid, 256, smem, stream>>>(tmA, tmB, a);
  else
    conv_gemm_v2_kernel<0><<<grid, 256, smem, stream>>>(tmA, tmB, a);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dtdl
