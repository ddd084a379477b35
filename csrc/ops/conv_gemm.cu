// Implicit-GEMM convolution / linear forward (and data-gradient) for sm_100a.
//
//   y[n,ho,wo,co] = sum_{taps,ci} x[n, S*ho+dh(tap), S*wo+dw(tap), ci] * w[co, k(tap), ci]
//
// GEMM view: M = output pixels (128 per CTA), N = output channels, K = taps x
// input channels.  Nothing is im2col'ed: for each filter tap the A operand is
// the *same* NHWC activation tensor loaded by a 4-D tiled TMA box shifted by
// the tap offset; out-of-image elements (the zero padding) and the channel
// tail come back as zeros from the TMA unit.  A stride-2 convolution uses the
// same box with TMA element strides of 2 in W and H, so only the pixels that
// are needed leave L2 (no full-resolution pass followed by a sub-sample).
// B is the packed weight matrix [Cout_p16][taps*Cin_p] (K-major).  Both land
// in shared memory in the 128-byte-swizzled K-major layout tcgen05.mma
// consumes directly; the fp32 accumulator lives in TMEM and is drained by four
// epilogue warps.
//
// Two instantiations: ACT = 0, bf16 activations / weights (tcgen05 kind::f16,
// 64 channels per 128-byte row) and ACT = 1, fp32 activations / weights read
// by the tensor core as TF32 (kind::tf32, 32 channels per row) -- the
// precision the reference trains at (fp32 tensors, cuDNN's default TF32
// convolutions: pytorch/single_gpu.py:65-72).
//
// Roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected
// lane), warp 2 = TMEM allocator, warps 4-7 = epilogue (TMEM -> regs -> global,
// with optional bias / ReLU and bf16 or fp32 output).
//
// The taps are a table (offset, index into the weight pack), which is what lets
// the same kernel compute
//   * the data gradient of a stride-1 convolution: dy against the tap-flipped,
//     channel-transposed pack (pack.cu);
//   * the data gradient of a stride-2 convolution in gather form: the four
//     output phases (h%2, w%2) are four launches, each a stride-1 convolution
//     over dy with the sub-filter of taps that reach that phase, written to
//     every other pixel of dx (1x the MACs; no zero-inserted dy);
//   * a Linear layer: the 1x1 case with H = W = 1.
//
// Capability covered: K-O1/K-O2/K-O3/K-O12/K-O20/K-O21 of SURVEY.md section 2.5
// (the reference reaches cuDNN/cuBLAS through nn.Conv2d / nn.Linear:
// pytorch/model.py:28-32,62-63,69,71,77).
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"
#include "epilogue.cuh"

#include <stdlib.h>

namespace dtdl {

static constexpr int kMaxTaps = 64;

struct ConvTap {
  short dh, dw;  // input pixel = stride * output pixel + (dh, dw)
  short k;       // tap index inside the weight pack
  short pad_;
};

struct ConvGemmArgs {
  int ntaps;
  int chunks;       // ceil(Cin_p / CB): K blocks of one 128-byte row per tap
  int last_slices;  // 32-byte MMA K slices that carry data in the last chunk
  int Cin_p;        // K extent per tap inside the weight pack
  int CB;           // channels per 128-byte row: 64 (bf16) or 32 (fp32)
  int stride;
  int BW, BH, BNi, rows;
  int tiles_w, tiles_h, tiles_n;
  int Ho, Wo, N;  // extent of the pixel grid this launch computes
  // where grid pixel (ho, wo) is stored: row oh0 + ho*osh, column ow0 + wo*osw of [N, oH, oW]
  int oH, oW, osh, osw, oh0, ow0;
  int Cout_p;  // channels per output pixel row
  int n_tile;  // N per CTA, multiple of 16, <= 256
  int tmem_cols;
  int stages;
  int b_stage_bytes;
  int flags;  // 1 = ReLU, 2 = fp32 output
  void* y;
  const float* bias;
  float* stats;  // optional [2*Cout_p]: per-channel sum / sum of squares of y (pre-zeroed)
  // optional second, fp32 copy of the first y2_cols channels, row stride y2_ld floats: the single-process
  // engine points it at GPU 0's logits buffer, so "gather" is a store of the last layer's epilogue (K-C6)
  float* y2;
  int y2_ld, y2_cols;
  ConvTap tap[kMaxTaps];
};

static constexpr int kAStage = 128 * 128;  // 128 pixel rows x 128 bytes

template <int ACT>
__global__ void __launch_bounds__(256, 2)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ ConvGemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  uint8_t* sA = smem;
  uint8_t* sB = smem + a.stages * kAStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + a.stages * a.b_stage_bytes);
  uint64_t* empty = full + a.stages;
  uint64_t* tmem_full = empty + a.stages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // tile coordinates
  int mt = blockIdx.x;
  const int tw_i = mt % a.tiles_w;
  mt /= a.tiles_w;
  const int th_i = mt % a.tiles_h;
  const int tn_i = mt / a.tiles_h;
  const int w0 = tw_i * a.BW, h0 = th_i * a.BH, n0 = tn_i * a.BNi;
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

  const int n_kblocks = a.ntaps * a.chunks;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx = (uint32_t)(a.rows * 128 + a.n_tile * 128);
      int stage = 0, phase = 0;
      for (int t = 0; t < a.ntaps; ++t) {
        const int dh = a.tap[t].dh, dw = a.tap[t].dw, kt = a.tap[t].k;
        for (int c = 0; c < a.chunks; ++c) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          mbar_arrive_expect_tx(&full[stage], tx);
          tma_load_4d(sA + stage * kAStage, &tmA, &full[stage], c * a.CB, a.stride * w0 + dw,
                      a.stride * h0 + dh, n0);
          tma_load_2d(sB + stage * a.b_stage_bytes, &tmB, &full[stage], kt * a.Cin_p + c * a.CB,
                      nt * a.n_tile);
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
      const uint64_t adesc0 = umma_smem_desc(smem_u32(sA), 16, 1024);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(sB), 16, 1024);
      int stage = 0, phase = 0, c = 0;
      for (int kb = 0; kb < n_kblocks; ++kb) {
        mbar_wait(&full[stage], phase, 2);
        tc_fence_after();
        const int nsl = (c == a.chunks - 1) ? a.last_slices : 4;
        if (++c == a.chunks) c = 0;
        const uint64_t ad = umma_desc_advance(adesc0, (uint32_t)(stage * kAStage));
        const uint64_t bd = umma_desc_advance(bdesc0, (uint32_t)(stage * a.b_stage_bytes));
        umma<ACT>(tmem_base, ad, bd, idesc, kb > 0 ? 1u : 0u);
#pragma unroll
        for (int k = 1; k < 4; ++k)
          if (k < nsl) umma<ACT>(tmem_base, umma_desc_advance(ad, k * 32), umma_desc_advance(bd, k * 32), idesc, 1u);
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
    const int wi = row % a.BW;
    const int hi = (row / a.BW) % a.BH;
    const int ni = row / (a.BW * a.BH);
    const int wo = w0 + wi, ho = h0 + hi, n = n0 + ni;
    const bool valid = row < a.rows && wo < a.Wo && ho < a.Ho && n < a.N;
    const size_t pix = ((size_t)n * a.oH + (a.oh0 + ho * a.osh)) * a.oW + (a.ow0 + wo * a.osw);
    const bool relu = a.flags & 1;
    const bool f32 = ACT == 1 || (a.flags & 2);

    // statistics scratch = the pipeline's stage memory (free once the accumulator is complete)
    float* s_parts = reinterpret_cast<float*>(sA);
    const int etid = threadIdx.x - 128;
    mbar_wait(tmem_full, 0, 3);
    tc_fence_after();
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int j = 0; j < a.n_tile / 16; ++j) {
      uint32_t v[16];
      tmem_ld16(lane_base + j * 16, v);
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
        epilogue_col_stats_part<ACT == 0>(f, valid, lane, s_parts + q * 2 * a.n_tile, a.n_tile, j * 16, true);
      if (!valid) continue;
      if (f32) {
        epilogue_store_f32(reinterpret_cast<float*>(a.y) + pix * a.Cout_p + col0, f, a.Cout_p - col0);
      } else {
        epilogue_store_bf16(reinterpret_cast<__nv_bfloat16*>(a.y) + pix * a.Cout_p + col0, f, a.Cout_p - col0);
      }
      if (a.y2 != nullptr) {
        float* o2 = a.y2 + pix * a.y2_ld + col0;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (col0 + i < a.y2_cols) o2[i] = f[i];
      }
    }
    if (a.stats != nullptr) epilogue_stats_flush_parts(s_parts, a.stats, a.n_tile, nt * a.n_tile, a.Cout_p, etid);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, (uint32_t)a.tmem_cols);
}

// Pixel tile (BW x BH x BNi <= 128 rows) shared by fprop and wgrad so both see
// the same decomposition of the output image.
void choose_pixel_tile(int N, int Ho, int Wo, int* BW, int* BH, int* BNi) {
  int bw = Wo < 128 ? Wo : 128;
  int bh = 128 / bw;
  if (bh > Ho) bh = Ho;
  if (bh < 1) bh = 1;
  int bn = 1;
  if (bh == Ho && bw == Wo) {
    bn = 128 / (bw * bh);
    if (bn > N) bn = N;
    if (bn < 1) bn = 1;
  }
  *BW = bw;
  *BH = bh;
  *BNi = bn;
}

int conv_fprop_v2(const void* x, const void* wpack, void* y, const float* bias, float* stats, int N,
                  int H, int W, int Cin_p, int Ho, int Wo, int Cout_p, int KH, int KW, int pad_h,
                  int pad_w, int flags, int act, cudaStream_t stream);

// One launch of the tap-table kernel.  (N, H, W, Cin_p) describe the tensor the TMA reads, (gH, gW)
// the pixel grid computed, and the o* fields where a grid pixel lands in the output tensor.
static int launch_conv_taps(const void* x, const void* wpack, void* y, const float* bias, float* stats,
                            int N, int H, int W, int Cin_p, int gH, int gW, int Cout_p,
                            const ConvTap* taps, int ntaps, int pack_taps, int stride, int oH, int oW,
                            int osh, int osw, int oh0, int ow0, int flags, int act,
                            cudaStream_t stream, float* y2 = nullptr, int y2_ld = 0, int y2_cols = 0) {
  DTDL_REQUIRE(ntaps >= 1 && ntaps <= kMaxTaps, "between 1 and 64 filter taps per launch");
  const int esz = act_esize(act);
  ConvGemmArgs a;
  a.ntaps = ntaps;
  for (int t = 0; t < ntaps; ++t) a.tap[t] = taps[t];
  a.CB = 128 / esz;
  a.chunks = (Cin_p + a.CB - 1) / a.CB;
  a.last_slices = ((Cin_p - a.CB * (a.chunks - 1)) * esz + 31) / 32;
  a.Cin_p = Cin_p;
  a.stride = stride;
  choose_pixel_tile(N, gH, gW, &a.BW, &a.BH, &a.BNi);
  a.rows = a.BW * a.BH * a.BNi;
  a.tiles_w = (gW + a.BW - 1) / a.BW;
  a.tiles_h = (gH + a.BH - 1) / a.BH;
  a.tiles_n = (N + a.BNi - 1) / a.BNi;
  a.Ho = gH;
  a.Wo = gW;
  a.N = N;
  a.oH = oH;
  a.oW = oW;
  a.osh = osh;
  a.osw = osw;
  a.oh0 = oh0;
  a.ow0 = ow0;
  a.Cout_p = Cout_p;
  const int Np16 = (Cout_p + 15) / 16 * 16;
  const int m_tiles = a.tiles_w * a.tiles_h * a.tiles_n;
  int n_tile;
  if (Np16 <= 256) {
    int split = 1;
    while (split < 4 && m_tiles * split < 120 && Np16 / (split + 1) >= 32) ++split;
    n_tile = ((Np16 + split - 1) / split + 15) / 16 * 16;
  } else {
    n_tile = (m_tiles * ((Np16 + 255) / 256) >= 148) ? 256 : 128;
  }
  a.n_tile = n_tile;
  const int n_tiles = (Np16 + n_tile - 1) / n_tile;
  a.tmem_cols = n_tile <= 32 ? 32 : n_tile <= 64 ? 64 : n_tile <= 128 ? 128 : 256;
  a.b_stage_bytes = (n_tile * 128 + 1023) / 1024 * 1024;
  const int stage_bytes = kAStage + a.b_stage_bytes;
  // Many small tiles: keep shared memory under half an SM so two CTAs are resident and one
  // tile's epilogue overlaps the other's main loop.  Few big tiles: one deep pipeline.
  const bool two_per_sm = (long long)m_tiles * n_tiles > sm_count() && 3 * stage_bytes <= 108 * 1024 &&
                          a.tmem_cols <= 256;
  int stages = ((two_per_sm ? 108 : 200) * 1024) / stage_bytes;
  if (stages > 8) stages = 8;
  const int n_kblocks = a.ntaps * a.chunks;
  if (stages > n_kblocks) stages = n_kblocks < 2 ? 2 : n_kblocks;
  a.stages = stages;
  a.flags = flags;
  a.y = y;
  a.bias = bias;
  a.stats = stats;
  a.y2 = y2;
  a.y2_ld = y2_ld;
  a.y2_cols = y2_cols;

  CUtensorMap tmA, tmB;
  {
    uint64_t dims[4] = {(uint64_t)Cin_p, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)Cin_p * esz, (uint64_t)W * Cin_p * esz, (uint64_t)H * W * Cin_p * esz};
    uint32_t box[4] = {(uint32_t)a.CB, (uint32_t)(a.BW * stride), (uint32_t)(a.BH * stride), (uint32_t)a.BNi};
    uint32_t est[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
    if (encode_tmap_act(&tmA, x, 4, dims, str, box, act, false, est)) return 1;
  }
  {
    const uint64_t Ktot = (uint64_t)pack_taps * Cin_p;
    uint64_t dims[2] = {Ktot, (uint64_t)Np16};
    uint64_t str[1] = {Ktot * esz};
    uint32_t box[2] = {(uint32_t)a.CB, (uint32_t)n_tile};
    if (encode_tmap_act(&tmB, wpack, 2, dims, str, box, act, false)) return 1;
  }
  const size_t smem = (size_t)stages * stage_bytes + 1024 + (2 * stages + 1) * 8 + 16;
  static bool attr_set[64] = {false};
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    // 227 KB less the static epilogue accumulators
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set[dev & 63] = true;
  }
  dim3 grid(m_tiles, n_tiles);
  if (act == kActF32)
    conv_gemm_kernel<1><<<grid, 256, smem, stream>>>(tmA, tmB, a);
  else
    conv_gemm_kernel<0><<<grid, 256, smem, stream>>>(tmA, tmB, a);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dtdl

using namespace dtdl;

extern "C" int dtdl_conv_fprop(const void* x, const void* wpack, void* y, const float* bias, int N,
                               int H, int W, int Cin_p, int Ho, int Wo, int Cout_p, int KH, int KW,
                               int pad_h, int pad_w, int flags, int stride, int act, void* stream_) {
  return dtdl_conv_fprop_stats(x, wpack, y, bias, nullptr, N, H, W, Cin_p, Ho, Wo, Cout_p, KH, KW, pad_h,
                               pad_w, flags, stride, act, stream_);
}

extern "C" int dtdl_conv_fprop_stats(const void* x, const void* wpack, void* y, const float* bias,
                                     float* stats, int N, int H, int W, int Cin_p, int Ho, int Wo,
                                     int Cout_p, int KH, int KW, int pad_h, int pad_w, int flags,
                                     int stride, int act, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTDL_REQUIRE(act == kActBf16 || act == kActF32, "act must be 0 (bf16) or 1 (fp32 / TF32)");
  DTDL_REQUIRE(Cin_p % 8 == 0 && Cout_p % 8 == 0, "channel counts must be padded to 8");
  DTDL_REQUIRE(stride == 1 || stride == 2, "stride 1 or 2");
  DTDL_REQUIRE(KH * KW <= kMaxTaps, "at most 64 filter taps");
  // Ho / Wo may be smaller than the full output (asymmetric padding: the caller simply asks for
  // fewer rows / columns on the high side).
  DTDL_REQUIRE(Ho >= 1 && Wo >= 1 && Ho <= (H + 2 * pad_h - KH) / stride + 1 &&
                   Wo <= (W + 2 * pad_w - KW) / stride + 1,
               "output size does not match the convolution");
  static const bool force_v1 = getenv("DTDL_CONV_V1") != nullptr;
  if (!force_v1 && stride == 1) {
    const int rc = conv_fprop_v2(x, wpack, y, bias, stats, N, H, W, Cin_p, Ho, Wo, Cout_p, KH, KW, pad_h,
                                 pad_w, flags, act, stream);
    if (rc >= 0) return rc;  // launched (0) or failed (>0); -1 = shape not covered, fall through
  }
  ConvTap taps[kMaxTaps];
  for (int r = 0; r < KH; ++r)
    for (int s = 0; s < KW; ++s) {
      ConvTap& t = taps[r * KW + s];
      t.dh = (short)(r - pad_h);
      t.dw = (short)(s - pad_w);
      t.k = (short)(r * KW + s);
      t.pad_ = 0;
    }
  return launch_conv_taps(x, wpack, y, bias, stats, N, H, W, Cin_p, Ho, Wo, Cout_p, taps, KH * KW, KH * KW,
                          stride, Ho, Wo, 1, 1, 0, 0, flags, act, stream);
}

// dtdl_conv_fprop for the LAST layer of a replica in the single-process engine: besides y, the first
// y2_cols channels are stored as fp32 rows of stride y2_ld floats at y2, which may be (peer-mapped)
// memory of another GPU -- the logits land in GPU 0's output buffer from the GEMM epilogue, there is
// no gather pass (nn.DataParallel's gather: pytorch/data_parallel.py:104).  Stride 1.
extern "C" int dtdl_conv_fprop_gather(const void* x, const void* wpack, void* y, const float* bias,
                                      float* y2, int y2_ld, int y2_cols, int N, int H, int W, int Cin_p,
                                      int Ho, int Wo, int Cout_p, int KH, int KW, int pad_h, int pad_w,
                                      int flags, int act, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTDL_REQUIRE(act == kActBf16 || act == kActF32, "act must be 0 (bf16) or 1 (fp32 / TF32)");
  DTDL_REQUIRE(Cin_p % 8 == 0 && Cout_p % 8 == 0, "channel counts must be padded to 8");
  DTDL_REQUIRE(KH * KW <= kMaxTaps, "at most 64 filter taps");
  DTDL_REQUIRE(y2 != nullptr && y2_ld >= y2_cols && y2_cols <= Cout_p, "bad gather destination");
  DTDL_REQUIRE(Ho >= 1 && Wo >= 1 && Ho <= H + 2 * pad_h - KH + 1 && Wo <= W + 2 * pad_w - KW + 1,
               "output size does not match the convolution");
  ConvTap taps[kMaxTaps];
  for (int r = 0; r < KH; ++r)
    for (int s = 0; s < KW; ++s) {
      ConvTap& t = taps[r * KW + s];
      t.dh = (short)(r - pad_h);
      t.dw = (short)(s - pad_w);
      t.k = (short)(r * KW + s);
      t.pad_ = 0;
    }
  return launch_conv_taps(x, wpack, y, bias, nullptr, N, H, W, Cin_p, Ho, Wo, Cout_p, taps, KH * KW, KH * KW, 1,
                          Ho, Wo, 1, 1, 0, 0, flags, act, stream, y2, y2_ld, y2_cols);
}

// Data gradient of a stride-2 convolution, gather form.  dx[n,h,w,:] receives, for every filter tap
// (r,s) with (h+pad-r) and (w+pad-s) even, dy[n,(h+pad-r)/2,(w+pad-s)/2,:] times w[:,:,r,s]: per
// output phase (h%2, w%2) that is a stride-1 convolution over dy with a sub-filter.  `wdpack` is the
// ordinary dgrad pack wd[ci][t'][co] = w[co][ci][taps-1-t'].
extern "C" int dtdl_conv_dgrad_s2(const void* dy, const void* wdpack, void* dx, int N, int Hy, int Wy,
                                  int Cout_p, int H, int W, int Cin_p, int KH, int KW, int pad_h,
                                  int pad_w, int act, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTDL_REQUIRE(act == kActBf16 || act == kActF32, "act must be 0 (bf16) or 1 (fp32 / TF32)");
  DTDL_REQUIRE(Cin_p % 8 == 0 && Cout_p % 8 == 0, "channel counts must be padded to 8");
  DTDL_REQUIRE(KH >= 2 && KW >= 2 && KH * KW <= kMaxTaps, "filter between 2x2 and 64 taps");
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int gH = (H - ph + 1) / 2, gW = (W - pw + 1) / 2;
      if (gH <= 0 || gW <= 0) continue;
      ConvTap taps[kMaxTaps];
      int nt = 0;
      for (int r = 0; r < KH; ++r) {
        if ((ph + pad_h - r) & 1) continue;
        for (int s = 0; s < KW; ++s) {
          if ((pw + pad_w - s) & 1) continue;
          ConvTap& t = taps[nt++];
          // floor division: (ph + pad - r) is even, so an arithmetic shift is exact for negatives too
          t.dh = (short)((ph + pad_h - r) >> 1);
          t.dw = (short)((pw + pad_w - s) >> 1);
          t.k = (short)(KH * KW - 1 - (r * KW + s));
          t.pad_ = 0;
        }
      }
      DTDL_REQUIRE(nt > 0, "stride-2 data gradient: an output phase without taps (filter too small)");
      const int rc = launch_conv_taps(dy, wdpack, dx, nullptr, nullptr, N, Hy, Wy, Cout_p, gH, gW, Cin_p,
                                      taps, nt, KH * KW, 1, H, W, 2, 2, ph, pw, 0, act, stream);
      if (rc) return rc;
    }
  return 0;
}
