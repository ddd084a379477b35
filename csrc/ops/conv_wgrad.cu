// Weight-gradient of a stride-1 convolution / linear layer for sm_100a.
//
//   dw[co, ci, r, s] = sum_{n,ho,wo} dy[n,ho,wo,co] * x[n, ho+r-pad_h, wo+s-pad_w, ci]
//
// GEMM view: M = output channels (128 per CTA), N = input channels (64 per
// CTA), K = pixels.  Both operands are "MN-major" for tcgen05 (channels are the
// contiguous dimension of an NHWC tensor, the reduction runs over pixel rows),
// so the TMA boxes [pixels][64 channels] that land in 128-byte-swizzled shared
// memory are consumed as-is with the a_major/b_major bits set.  As in the
// forward kernel the filter tap is a coordinate shift of the x box with the
// zero padding supplied by the TMA unit.
//
// Work split: grid.x = pixel-range split (split-K), grid.y = (filter row,
// group of <=3 filter columns), grid.z = (Cout tile, Cin chunk).  A CTA keeps
// one 128x64 fp32 accumulator per filter column of its group in TMEM.
// Partial sums go to a scratch workspace [split][tap][Cout_p][Cin_w] with plain
// vector stores; dtdl_wgrad_reduce sums the splits in a fixed order (the
// result is bitwise reproducible, unlike an atomic split-K) and writes the
// fp32 gradient in the OIHW layout of the torch parameter.
//
// Two instantiations (see conv_gemm.cu): ACT = 0 bf16 operands, boxes of 64 channels, 16 pixels per
// MMA; ACT = 1 fp32 operands read as TF32, boxes of 32 channels (128 bytes), 8 pixels per MMA.
// MN-major 32-bit operands must use tcgen05's SWIZZLE_128B_BASE32B layout (32-byte swizzle atoms,
// 4-row groups), which the TMA produces with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
// A stride-2 convolution samples x through TMA element strides (row mode only).
//
// Capability covered: K-O4 / K-O12 of SURVEY.md section 2.5 (the reference
// reaches cuDNN/cuBLAS through autograd of nn.Conv2d / nn.Linear:
// pytorch/model.py:28-32,77; pytorch/single_gpu.py:102).
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"

namespace dtdl {

void choose_pixel_tile(int N, int Ho, int Wo, int* BW, int* BH, int* BNi);

struct WgradArgs {
  int KW, pad_h, pad_w;
  int stride;          // 1 or 2 (2: row mode)
  int CB;              // channels per box: 64 (bf16) / 32 (fp32)
  int n_boxes;         // x boxes per filter column: 64 / CB
  int cols_per_cta;    // row mode: filter columns per CTA (3 bf16, 1 fp32)
  int x_box_bytes;     // bytes between the x boxes of one stage
  int groups_per_row;  // ceil(KW / cols_per_cta)
  int BW, BH, BNi, rows, kslices;
  int tiles_w, tiles_h, tiles_n, n_ptiles;
  int tiles_per_split;
  int ci_chunks;
  int m_boxes;  // dy boxes per stage: ceil(min(Cout_p, 128) / CB) (the others are never loaded)
  int Cout_p, Cin_w, taps;
  int stages, stage_bytes;
  int zero_fill;
  int col_mode;    // 1: CTA owns one filter COLUMN; its KH filter rows are views of one x halo box
  int KH, x_off, halo_rows;
  float* ws;
};

static constexpr int kBox = 128 * 128;  // [128 pixel rows][64 bf16]

template <int ACT>
__global__ void __launch_bounds__(256, 2)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDy, const __grid_constant__ CUtensorMap tmX,
                  const WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + a.stages * a.stage_bytes);
  uint64_t* empty = full + a.stages;
  uint64_t* tmem_full = empty + a.stages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int split = blockIdx.x;
  // row mode : CTA = (filter row r, up to 3 filter columns s0..), one shifted x box per column
  // col mode : CTA = filter column s0, all KH filter rows, ONE x halo box; the operand of filter
  //            row r is that box viewed r image rows (r*BW*128 bytes, 1 KB aligned) further down
  const int r = a.col_mode ? 0 : blockIdx.y / a.groups_per_row;
  const int s0 = a.col_mode ? blockIdx.y : (blockIdx.y % a.groups_per_row) * a.cols_per_cta;
  const int ns = a.col_mode ? a.KH : ((a.KW - s0) < a.cols_per_cta ? (a.KW - s0) : a.cols_per_cta);
  const int co_tile = blockIdx.z / a.ci_chunks;
  const int ci_chunk = blockIdx.z % a.ci_chunks;
  const int co0 = co_tile * 128, ci0 = ci_chunk * 64;

  const int pt_begin = split * a.tiles_per_split;
  int pt_end = pt_begin + a.tiles_per_split;
  if (pt_end > a.n_ptiles) pt_end = a.n_ptiles;
  const int n_steps = pt_end - pt_begin;

  if (a.zero_fill) {
    // Pixel tiles with a row count that is not a multiple of 16 leave tail
    // rows the TMA never writes; they are K padding and must read as zero.
    uint4* p = reinterpret_cast<uint4*>(smem);
    const int n16 = a.stages * a.stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) p[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDy);
    tma_prefetch_desc(&tmX);
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
    tmem_alloc(tmem_ptr, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx = a.col_mode ? (uint32_t)(a.rows * 128 * a.m_boxes + a.halo_rows * 128 * a.n_boxes)
                                     : (uint32_t)(a.rows * 128 * (a.m_boxes + ns * a.n_boxes));
      int stage = 0, phase = 0;
      for (int pt = pt_begin; pt < pt_end; ++pt) {
        int t = pt;
        const int w0 = (t % a.tiles_w) * a.BW;
        t /= a.tiles_w;
        const int h0 = (t % a.tiles_h) * a.BH;
        const int n0 = (t / a.tiles_h) * a.BNi;
        uint8_t* st = smem + stage * a.stage_bytes;
        mbar_wait(&empty[stage], phase ^ 1, 11);
        mbar_arrive_expect_tx(&full[stage], tx);
        for (int b = 0; b < a.m_boxes; ++b)
          tma_load_4d(st + b * kBox, &tmDy, &full[stage], co0 + b * a.CB, w0, h0, n0);
        if (a.col_mode) {
          for (int nb = 0; nb < a.n_boxes; ++nb)
            tma_load_4d(st + a.x_off + nb * a.x_box_bytes, &tmX, &full[stage], ci0 + nb * a.CB,
                        w0 + s0 - a.pad_w, h0 - a.pad_h, n0);
        } else {
          for (int s = 0; s < ns; ++s)
            for (int nb = 0; nb < a.n_boxes; ++nb)
              tma_load_4d(st + a.x_off + (s * a.n_boxes + nb) * kBox, &tmX, &full[stage], ci0 + nb * a.CB,
                          a.stride * w0 + s0 + s - a.pad_w, a.stride * h0 + r - a.pad_h, n0);
        }
        if (++stage == a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr int kPix = ACT ? 8 : 16;          // pixels (K) per MMA
      constexpr uint32_t kSlice = kPix * 128;     // bytes of one K slice
      constexpr int kMaxSlices = 128 / kPix;
      const uint32_t idesc = umma_idesc<ACT>(128, 64, 1, 1);
      // bf16: SWIZZLE_128B, 8-row K groups; fp32: SWIZZLE_128B_BASE32B, 4-row K groups
      const uint32_t layout = ACT ? 1u : 2u, sbo = ACT ? 512u : 1024u;
      const uint64_t adesc0 = umma_smem_desc(smem_u32(smem), kBox, sbo, layout);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem), (uint32_t)a.x_box_bytes, sbo, layout);
      uint32_t koff[kMaxSlices];  // byte offset of K slice k inside the B operand (computed once)
#pragma unroll
      for (int k = 0; k < kMaxSlices; ++k) {
        const int kpi = (a.BW * a.BH) / kPix;
        koff[k] = a.col_mode ? (uint32_t)(((k / kpi) * (a.BH + a.KH - 1) * a.BW) * 128 + (k % kpi) * kSlice)
                             : (uint32_t)(k * kSlice);
      }
      int stage = 0, phase = 0;
      for (int i = 0; i < n_steps; ++i) {
        mbar_wait(&full[stage], phase, 12);
        tc_fence_after();
        const uint64_t ad0 = umma_desc_advance(adesc0, (uint32_t)(stage * a.stage_bytes));
        const uint64_t bs0 = umma_desc_advance(bdesc0, (uint32_t)(stage * a.stage_bytes + a.x_off));
        for (int s = 0; s < ns; ++s) {
          // B operand of accumulator s: row mode = its own shifted box(es); col mode = the halo box
          // viewed s image rows further down (koff[] already holds the per-slice image jumps)
          const uint64_t bd0 = umma_desc_advance(bs0, a.col_mode ? (uint32_t)(s * a.BW * 128)
                                                                : (uint32_t)(s * a.n_boxes * kBox));
#pragma unroll
          for (int k = 0; k < kMaxSlices; ++k) {
            if (k < a.kslices)
              umma<ACT>(tmem_base + s * 64, umma_desc_advance(ad0, k * kSlice), umma_desc_advance(bd0, koff[k]),
                        idesc, (i > 0 || k > 0) ? 1u : 0u);
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
    const int co = co0 + q * 32 + lane;
    const bool valid = co < a.Cout_p;
    mbar_wait(tmem_full, 0, 13);
    tc_fence_after();
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int s = 0; s < ns; ++s) {
      const int t = a.col_mode ? s * a.KW + s0 : r * a.KW + s0 + s;
      float* dst =
          a.ws + (((size_t)split * a.taps + t) * a.Cout_p + (valid ? co : 0)) * a.Cin_w + ci0;
      for (int j = 0; j < 4; ++j) {
        uint32_t v[16];
        tmem_ld16(lane_base + s * 64 + j * 16, v);
        tmem_ld_wait();
        if (valid) {
          if (n_steps > 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<uint4*>(dst + j * 16)[i] =
                  make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<uint4*>(dst + j * 16)[i] = make_uint4(0, 0, 0, 0);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 256);
}

// dw[co][ci][t] (OIHW, fp32) = (accumulate ? dw : 0) + sum_split ws[split][t][co][ci]
// Block = 64 (tap, 4-channel vector) items x 4 split lanes; every lane sums its
// splits (s = lane, lane+4, ...) with float4 loads, the four lanes are combined
// through shared memory in a fixed order, so the result does not depend on timing.
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int splits, int taps,
                    int Cout, int Cin, int Cout_p, int Cin_w, int accumulate) {
  __shared__ float4 part[4][64];
  const int item = blockIdx.x * 64 + (threadIdx.x & 63);
  const int lane = threadIdx.x >> 6;
  const int co = blockIdx.y;
  const int vec_per_tap = Cin_w >> 2;
  const int n_items = taps * vec_per_tap;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const int t = item / vec_per_tap, c4 = item % vec_per_tap;
  if (item < n_items) {
    const size_t stride4 = ((size_t)taps * Cout_p * Cin_w) >> 2;
    const float4* p = reinterpret_cast<const float4*>(ws + ((size_t)t * Cout_p + co) * Cin_w) + c4;
    int s = lane;
    for (; s + 12 < splits; s += 16) {
      const float4 v0 = p[(size_t)s * stride4];
      const float4 v1 = p[(size_t)(s + 4) * stride4];
      const float4 v2 = p[(size_t)(s + 8) * stride4];
      const float4 v3 = p[(size_t)(s + 12) * stride4];
      acc.x += (v0.x + v1.x) + (v2.x + v3.x);
      acc.y += (v0.y + v1.y) + (v2.y + v3.y);
      acc.z += (v0.z + v1.z) + (v2.z + v3.z);
      acc.w += (v0.w + v1.w) + (v2.w + v3.w);
    }
    for (; s < splits; s += 4) {
      const float4 v = p[(size_t)s * stride4];
      acc.x += v.x;
      acc.y += v.y;
      acc.z += v.z;
      acc.w += v.w;
    }
  }
  part[lane][threadIdx.x & 63] = acc;
  __syncthreads();
  if (lane == 0 && item < n_items) {
    const float4 a0 = part[0][threadIdx.x], a1 = part[1][threadIdx.x], a2 = part[2][threadIdx.x],
                 a3 = part[3][threadIdx.x];
    const float r[4] = {(a0.x + a1.x) + (a2.x + a3.x), (a0.y + a1.y) + (a2.y + a3.y),
                        (a0.z + a1.z) + (a2.z + a3.z), (a0.w + a1.w) + (a2.w + a3.w)};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ci = c4 * 4 + j;
      if (ci < Cin) {
        float* o = dw + ((size_t)co * Cin + ci) * taps + t;
        *o = accumulate ? (*o + r[j]) : r[j];
      }
    }
  }
}

}  // namespace dtdl

using namespace dtdl;

#include <stdlib.h>
#include <atomic>
static int wgrad_waves() {
  static int w = getenv("DTDL_WGRAD_WAVES") ? atoi(getenv("DTDL_WGRAD_WAVES")) : 2;
  return w < 1 ? 1 : w;
}

// Plans the split-K decomposition; shared by the size query and the launch.
static void wgrad_plan(int N, int Ho, int Wo, int Cin_p, int Cout_p, int KH, int KW, int stride, int act,
                       WgradArgs* a, int* splits) {
  const int kpix = act == kActF32 ? 8 : 16;
  choose_pixel_tile(N, Ho, Wo, &a->BW, &a->BH, &a->BNi);
  a->rows = a->BW * a->BH * a->BNi;
  a->kslices = (a->rows + kpix - 1) / kpix;
  a->zero_fill = (a->rows % kpix) != 0;
  a->stride = stride;
  a->CB = act == kActF32 ? 32 : 64;
  a->n_boxes = 64 / a->CB;
  a->cols_per_cta = act == kActF32 ? 1 : 3;
  a->tiles_w = (Wo + a->BW - 1) / a->BW;
  a->tiles_h = (Ho + a->BH - 1) / a->BH;
  a->tiles_n = (N + a->BNi - 1) / a->BNi;
  a->n_ptiles = a->tiles_w * a->tiles_h * a->tiles_n;
  a->KW = KW;
  a->groups_per_row = (KW + a->cols_per_cta - 1) / a->cols_per_cta;
  a->ci_chunks = (Cin_p + 63) / 64;
  a->Cin_w = a->ci_chunks * 64;
  a->Cout_p = Cout_p;
  a->m_boxes = ((Cout_p < 128 ? Cout_p : 128) + a->CB - 1) / a->CB;
  a->taps = KH * KW;
  a->KH = KH;
  static const bool no_col = getenv("DTDL_WGRAD_ROWMODE") != nullptr;
  a->col_mode = (!no_col && stride == 1 && a->BW % 8 == 0 && (a->BW * a->BH) % 16 == 0 && a->rows == 128 &&
                 KH <= 4 && KH > 1) ? 1 : 0;
  a->halo_rows = a->BNi * (a->BH + KH - 1) * a->BW;
  a->x_off = a->m_boxes * kBox;
  a->x_box_bytes = a->col_mode ? (a->halo_rows * 128 + 1023) / 1024 * 1024 : kBox;
  const int co_tiles = (Cout_p + 127) / 128;
  const int groups = (a->col_mode ? KW : KH * a->groups_per_row) * co_tiles * a->ci_chunks;
  // few output tiles (narrow layers): one wave of fat CTAs amortises set-up and halves the
  // partials the reduction has to read; wider layers do better with two waves
  const int waves = getenv("DTDL_WGRAD_WAVES") ? wgrad_waves() : (groups <= 3 ? 1 : 2);
  int sp = (waves * sm_count() + groups - 1) / groups;
  if (sp > a->n_ptiles) sp = a->n_ptiles;
  if (sp < 1) sp = 1;
  a->tiles_per_split = (a->n_ptiles + sp - 1) / sp;
  sp = (a->n_ptiles + a->tiles_per_split - 1) / a->tiles_per_split;
  *splits = sp;
}

// The scratch need depends on the split count only, which does not depend on stride or act for a
// given OUTPUT size -- except through the operating mode (stride 2 and fp32 use more CTA groups).
// Callers size for the mode they run; this entry point keeps the bf16 stride-1 figure and
// dtdl_conv_wgrad_workspace_bytes_ex answers for any mode.
extern "C" long long dtdl_conv_wgrad_workspace_bytes_ex(int N, int Ho, int Wo, int Cin_p, int Cout_p,
                                                        int KH, int KW, int stride, int act) {
  WgradArgs a;
  int splits;
  wgrad_plan(N, Ho, Wo, Cin_p, Cout_p, KH, KW, stride, act, &a, &splits);
  return (long long)splits * a.taps * Cout_p * a.Cin_w * 4;
}

extern "C" long long dtdl_conv_wgrad_workspace_bytes(int N, int Ho, int Wo, int Cin_p, int Cout_p,
                                                     int KH, int KW) {
  return dtdl_conv_wgrad_workspace_bytes_ex(N, Ho, Wo, Cin_p, Cout_p, KH, KW, 1, kActBf16);
}

static int conv_wgrad_impl(const void* x, const void* dy, float* dw, void* workspace,
                           long long workspace_bytes, int N, int H, int W, int Cin, int Cin_p, int Ho,
                           int Wo, int Cout, int Cout_p, int KH, int KW, int pad_h, int pad_w,
                           int accumulate, int stride, int act, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTDL_REQUIRE(act == kActBf16 || act == kActF32, "act must be 0 (bf16) or 1 (fp32 / TF32)");
  DTDL_REQUIRE(Cin_p % 8 == 0 && Cout_p % 8 == 0, "channel counts must be padded to 8");
  DTDL_REQUIRE(stride == 1 || stride == 2, "stride 1 or 2");
  DTDL_REQUIRE(Ho >= 1 && Wo >= 1 && Ho <= (H + 2 * pad_h - KH) / stride + 1 &&
                   Wo <= (W + 2 * pad_w - KW) / stride + 1,
               "output size does not match the convolution");
  const int esz = act_esize(act);
  WgradArgs a;
  int splits;
  wgrad_plan(N, Ho, Wo, Cin_p, Cout_p, KH, KW, stride, act, &a, &splits);
  a.pad_h = pad_h;
  a.pad_w = pad_w;
  const long long need = (long long)splits * a.taps * Cout_p * a.Cin_w * 4;
  DTDL_REQUIRE(workspace != nullptr && workspace_bytes >= need,
               "wgrad workspace too small (see dtdl_conv_wgrad_workspace_bytes_ex)");
  a.ws = reinterpret_cast<float*>(workspace);
  if (a.col_mode) {
    a.stage_bytes = a.x_off + a.n_boxes * a.x_box_bytes;  // dy box(es) + x halo box(es)
    a.stages = a.stage_bytes <= 40 * 1024 ? 3 : 2;
  } else {
    // dy boxes + the shifted x boxes of this CTA's filter columns (bf16: <= 2 + 3, fp32: <= 4 + 2)
    a.stage_bytes = a.x_off + a.cols_per_cta * a.n_boxes * kBox;
    a.stages = 2;
  }
  // The MMA always reads M = 128 output channels = 128 / CB box slots from the start of a stage, loaded
  // or not (rows of boxes that were not loaded are never stored): the slots must exist.
  if (a.stage_bytes < (128 / a.CB) * kBox) a.stage_bytes = (128 / a.CB) * kBox;
  const int steps = a.tiles_per_split;
  if (steps < a.stages) a.stages = steps < 1 ? 1 : steps;

  CUtensorMap tmDy, tmX;
  {
    uint64_t dims[4] = {(uint64_t)Cout_p, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)Cout_p * esz, (uint64_t)Wo * Cout_p * esz,
                       (uint64_t)Ho * Wo * Cout_p * esz};
    uint32_t box[4] = {(uint32_t)a.CB, (uint32_t)a.BW, (uint32_t)a.BH, (uint32_t)a.BNi};
    if (encode_tmap_act(&tmDy, dy, 4, dims, str, box, act, true)) return 1;
  }
  {
    uint64_t dims[4] = {(uint64_t)Cin_p, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)Cin_p * esz, (uint64_t)W * Cin_p * esz, (uint64_t)H * W * Cin_p * esz};
    uint32_t box[4] = {(uint32_t)a.CB, (uint32_t)(a.BW * stride),
                       (uint32_t)((a.col_mode ? a.BH + KH - 1 : a.BH) * stride), (uint32_t)a.BNi};
    uint32_t est[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
    if (encode_tmap_act(&tmX, x, 4, dims, str, box, act, true, est)) return 1;
  }
  const size_t smem = (size_t)a.stages * a.stage_bytes + 1024 + (2 * a.stages + 1) * 8 + 16;
  DTDL_REQUIRE(smem <= 227 * 1024, "wgrad pipeline does not fit in shared memory");
  static bool attr_set[64] = {false};
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int co_tiles = (Cout_p + 127) / 128;
  dim3 grid(splits, a.col_mode ? KW : KH * a.groups_per_row, co_tiles * a.ci_chunks);
  if (act == kActF32)
    conv_wgrad_kernel<1><<<grid, 256, smem, stream>>>(tmDy, tmX, a);
  else
    conv_wgrad_kernel<0><<<grid, 256, smem, stream>>>(tmDy, tmX, a);
  DTDL_CHECK_CUDA(cudaGetLastError());

  dim3 rgrid((a.taps * (a.Cin_w / 4) + 63) / 64, Cout);
  wgrad_reduce_kernel<<<rgrid, 256, 0, stream>>>(a.ws, dw, splits, a.taps, Cout, Cin, Cout_p,
                                                 a.Cin_w, accumulate);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dtdl_conv_wgrad(const void* x, const void* dy, float* dw, void* workspace,
                               long long workspace_bytes, int N, int H, int W, int Cin, int Cin_p,
                               int Ho, int Wo, int Cout, int Cout_p, int KH, int KW, int pad_h,
                               int pad_w, int accumulate, int stride, int act, void* stream_) {
  return conv_wgrad_impl(x, dy, dw, workspace, workspace_bytes, N, H, W, Cin, Cin_p, Ho, Wo, Cout, Cout_p,
                         KH, KW, pad_h, pad_w, accumulate, stride, act, stream_);
}

// ---------------------------------------------------------------------------
// Weight gradient for callers that cannot allocate device memory.
//
// dtdl_conv_wgrad needs a split-K scratch buffer, and normally the caller
// allocates it.  Some callers have no way to do that: a host program whose own
// memory pool has already claimed the whole GPU, a plug-in that is handed
// tensors but no allocator, or code that is entered while a stream is being
// captured, where cudaMalloc is not permitted.  For them the scratch lives
// inside this library, in the fixed-size array below.  The price is that every
// GPU that loads the library carries these 32 MiB until the process exits, and
// that problems needing more scratch than that are refused.
//
// Inside host code `g_wgrad_static_scratch` is only the host-side handle of the
// symbol, not an address a kernel can use, so the device address is looked up
// with cudaGetSymbolAddress -- once per device, with that device current -- and
// that pointer is what the kernels receive.
//
// What a caller needs to know:
//   * The first call on each device performs that look-up.  Make it before any
//     stream capture begins on that device (a plain warm-up call is enough).
//   * The call itself allocates nothing: no cudaMalloc / cudaFree, no library
//     routine that allocates.  Inputs and the output gradient are the caller's.
//   * All calls on a device share the one array and it keeps whatever the last
//     call left in it.  Nothing in it has to start at zero: every partial sum
//     the reduction pass reads was written earlier in the same call.  Because
//     of the sharing, two calls that may run on the GPU at the same time
//     (different streams, no dependency between them) race; keep them ordered.
//   * Shapes whose scratch need (dtdl_conv_wgrad_workspace_bytes) exceeds
//     dtdl_conv_wgrad_static_scratch_bytes() = 32 MiB are rejected with an error
//     and nothing is launched.
static constexpr long long kStaticScratchBytes = 32ll << 20;
static __device__ __align__(256) unsigned char g_wgrad_static_scratch[kStaticScratchBytes];

extern "C" long long dtdl_conv_wgrad_static_scratch_bytes(void) { return kStaticScratchBytes; }

extern "C" int dtdl_conv_wgrad_static(const void* x, const void* dy, float* dw, int N, int H, int W,
                                      int Cin, int Cin_p, int Ho, int Wo, int Cout, int Cout_p,
                                      int KH, int KW, int pad_h, int pad_w, int accumulate,
                                      void* stream_) {
  // one slot per device; atomic because host threads driving different streams may race here
  // (both would store the same address, so a plain exchange is enough)
  static std::atomic<void*> scratch_of_device[64];
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  DTDL_REQUIRE(dev >= 0 && dev < 64, "device index out of range for the static scratch table");
  void* scratch = scratch_of_device[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    // first call on this device: resolve the device address of the array
    DTDL_CHECK_CUDA(cudaGetSymbolAddress(&scratch, g_wgrad_static_scratch));
    DTDL_REQUIRE(scratch != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    scratch_of_device[dev].store(scratch, std::memory_order_release);
  }
  const long long need = dtdl_conv_wgrad_workspace_bytes(N, Ho, Wo, Cin_p, Cout_p, KH, KW);
  if (need > kStaticScratchBytes) {
    set_last_error("dtdl_conv_wgrad_static: this shape needs " + std::to_string(need) +
                   " bytes of scratch but the built-in buffer holds " +
                   std::to_string(kStaticScratchBytes) +
                   "; use dtdl_conv_wgrad with a caller-provided workspace");
    return 3;
  }
  return conv_wgrad_impl(x, dy, dw, scratch, kStaticScratchBytes, N, H, W, Cin, Cin_p,
                         Ho, Wo, Cout, Cout_p, KH, KW, pad_h, pad_w, accumulate, 1, kActBf16, stream_);
}
