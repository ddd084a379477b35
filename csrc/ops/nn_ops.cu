// Bandwidth-bound layer kernels on NHWC activations whose channel count is padded
// to a multiple of 8.  Pad channels are kept at exactly zero by every kernel here.
// Every kernel is instantiated for both activation storage types: a thread always
// moves one 16-byte vector, i.e. V = 8 channels of bf16 (act 0) or V = 4 channels
// of fp32 (act 1, the reference's own precision: pytorch/single_gpu.py:65-66 keeps
// the model in fp32).  Arithmetic and statistics are fp32 in both.
//
//   batch norm (train/eval) forward  : stats reduction + fused apply(+residual)(+ReLU)
//   batch norm backward              : fused (ReLU mask) reduction + apply
//   average / max pooling, 2x subsample, global average pool
//   softmax cross entropy (loss, dlogits, correct count in one kernel)
//   layout conversion NCHW fp32 -> NHWC bf16 (the source pointer may be peer memory)
//
// Capability covered: K-O5..K-O11, K-O13, K-O17, K-O19 of SURVEY.md section 2.5
// (reference call sites: pytorch/model.py:10-20,27-34,40-50,74-76,107-111;
// pytorch/single_gpu.py:99,105-108; tensorflow2/mnist_single.py:19-26).
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"

#include <stdlib.h>

namespace dtdl {

// V channels of one pixel as floats; V = 8 <-> a uint4 of bf16, V = 4 <-> a uint4 of fp32.
template <int V>
struct FV {
  float v[V];
};
template <int V>
__device__ __forceinline__ FV<V> unpackv(uint4 u) {
  FV<V> r;
  if (V == 8) {
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < V / 2; ++i) {
      float2 f = __bfloat1622float2(p[i]);
      r.v[2 * i] = f.x;
      r.v[2 * i + 1] = f.y;
    }
  } else {
    r.v[0] = __uint_as_float(u.x);
    r.v[1] = __uint_as_float(u.y);
    r.v[2] = __uint_as_float(u.z);
    r.v[3] = __uint_as_float(u.w);
  }
  return r;
}
template <int V>
__device__ __forceinline__ uint4 packv(const FV<V>& f) {
  uint4 u;
  if (V == 8) {
    __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < V / 2; ++i) p[i] = __floats2bfloat162_rn(f.v[2 * i], f.v[2 * i + 1]);
  } else {
    u = make_uint4(__float_as_uint(f.v[0]), __float_as_uint(f.v[1]), __float_as_uint(f.v[2]),
                   __float_as_uint(f.v[3]));
  }
  return u;
}

static inline int block_for_groups(int G) {
  int b = G * (256 / G);
  if (b < G) b = G;  // G > 256: one row of G threads (G <= 1024 is required)
  return b;
}
static inline int grid_for(long long work_items, int per_block, int max_blocks) {
  long long g = (work_items + per_block - 1) / per_block;
  if (g > max_blocks) g = max_blocks;
  if (g < 1) g = 1;
  return (int)g;
}

// ---------------------------------------------------------------------------
// NCHW fp32 -> NHWC bf16 (channel padded).  One thread per (pixel, 8-channel group).
// ---------------------------------------------------------------------------
template <int V>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, uint4* __restrict__ y, int N, int C,
                                    int HW, int G) {
  const long long total = (long long)N * HW * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const long long pix = i / G;
    const int hw = (int)(pix % HW);
    const long long n = pix / HW;
    FV<V> f;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const int c = g * V + k;
      f.v[k] = c < C ? x[(n * C + c) * HW + hw] : 0.f;
    }
    y[i] = packv<V>(f);
  }
}

// NCHW fp32 -> 2x2 space-to-depth NHWC bf16: y[n, ho, wo, (bh*2+bw)*C + c] = x[n, c, 2ho+bh, 2wo+bw].
// Turns a stride-2 KxK convolution into a stride-1 ceil(K/2) x ceil(K/2) one over 4C channels
// (ResNet's 7x7/2 stem becomes 4x4/1 over 12 channels).
template <int V>
__global__ void nchw_to_s2d_kernel(const float* __restrict__ x, uint4* __restrict__ y, int N, int C,
                                   int H, int W, int G) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = (long long)N * Ho * Wo * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    long long t = i / G;
    const int wo = (int)(t % Wo);
    t /= Wo;
    const int ho = (int)(t % Ho);
    const long long n = t / Ho;
    FV<V> f;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const int ch = g * V + k;
      float v = 0.f;
      if (ch < 4 * C) {
        const int phase = ch / C, c = ch % C;
        v = x[((n * C + c) * H + 2 * ho + (phase >> 1)) * W + 2 * wo + (phase & 1)];
      }
      f.v[k] = v;
    }
    y[i] = packv<V>(f);
  }
}

// ---------------------------------------------------------------------------
// Per-channel sum / sum of squares (accumulated with atomics into sums[2*Cp]).
// ---------------------------------------------------------------------------
template <int V>
__global__ void bn_stats_kernel(const uint4* __restrict__ x, float* __restrict__ sums, long long P,
                                int G) {
  extern __shared__ float red[];  // [rows][G][16]
  const int rows = blockDim.x / G;
  const int g = threadIdx.x % G, rrow = threadIdx.x / G;
  float s1[V], s2[V];
#pragma unroll
  for (int k = 0; k < V; ++k) s1[k] = s2[k] = 0.f;
  for (long long p = (long long)blockIdx.x * rows + rrow; p < P; p += (long long)gridDim.x * rows) {
    const FV<V> f = unpackv<V>(x[p * G + g]);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      s1[k] += f.v[k];
      s2[k] += f.v[k] * f.v[k];
    }
  }
  float* mine = red + (size_t)threadIdx.x * (2 * V);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    mine[k] = s1[k];
    mine[V + k] = s2[k];
  }
  __syncthreads();
  // threads [0, G*16) each own one (g, k) pair and sum it over the rows
  for (int o = threadIdx.x; o < G * 2 * V; o += blockDim.x) {
    const int og = o / (2 * V), ok = o % (2 * V);
    float acc = 0.f;
    for (int rr = 0; rr < rows; ++rr) acc += red[((size_t)rr * G + og) * (2 * V) + ok];
    const int Cp = G * V;
    if (ok < V)
      atomicAdd(&sums[og * V + ok], acc);
    else
      atomicAdd(&sums[Cp + og * V + (ok - V)], acc);
  }
}

// ---------------------------------------------------------------------------
// BN apply: y = [relu]( (x-mean)*rstd*gamma + beta [+ residual[..., :Cres_p]] )
// training=1: mean/var from `sums`; block 0 also updates the running stats and
// stores mean/rstd for backward.  training=0: mean/var from the running stats.
// ---------------------------------------------------------------------------
struct BnApplyArgs {
  const uint4* x;
  uint4* y;
  const float* sums;
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;
  float* save_mean;
  float* save_rstd;
  const uint4* residual;
  int Gres;
  long long P;
  int C, G;
  float eps, momentum;
  int relu, training;
  float* zero_buf;  // scratch of another pass that this launch resets (may be null)
  int zero_n;
  float* out_sums;  // optional [2*Cp]: sum / sum of squares of y, for the BatchNorm that reads y next
};

// Per-channel scale/shift into shared memory; block 0 also writes the saved / running statistics
// and resets `zero_buf`.  Ends with __syncthreads().
template <int V>
__device__ __forceinline__ void bn_apply_prologue(const BnApplyArgs& a, float* sc, float* sh) {
  const int Cp = a.G * V;
  for (int c = threadIdx.x; c < Cp; c += blockDim.x) {
    float scale = 0.f, shift = 0.f;
    if (c < a.C) {
      float mean, var;
      if (a.training) {
        mean = a.sums[c] / (float)a.P;
        var = fmaxf(a.sums[Cp + c] / (float)a.P - mean * mean, 0.f);
      } else {
        mean = a.running_mean[c];
        var = a.running_var[c];
      }
      const float rstd = rsqrtf(var + a.eps);
      scale = a.gamma[c] * rstd;
      shift = a.beta[c] - mean * scale;
      if (a.training && blockIdx.x == 0) {
        a.save_mean[c] = mean;
        a.save_rstd[c] = rstd;
        if (a.running_mean != nullptr) {
          const float unbiased = a.P > 1 ? var * (float)a.P / (float)(a.P - 1) : var;
          a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * mean;
          a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * unbiased;
        }
      }
    }
    sc[c] = scale;
    sh[c] = shift;
  }
  if (a.training && blockIdx.x == 0 && threadIdx.x == 0 && a.num_batches_tracked != nullptr)
    *a.num_batches_tracked += 1;
  if (blockIdx.x == 0 && a.zero_buf != nullptr)
    for (int i = threadIdx.x; i < a.zero_n; i += blockDim.x) a.zero_buf[i] = 0.f;
  __syncthreads();
}

template <int V>
__global__ void bn_apply_kernel(const BnApplyArgs a) {
  extern __shared__ float ss[];  // scale[Cp], shift[Cp]
  const int Cp = a.G * V;
  float* sc = ss;
  float* sh = ss + Cp;
  bn_apply_prologue<V>(a, sc, sh);

  const long long total = a.P * a.G;
  // (g, p) = (i % G, i / G) follow i by carry arithmetic: a 64-bit division per element is a
  // subroutine call of several dozen instructions, more than the rest of the loop body
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  int g = (int)(i % a.G);
  long long p = i / a.G;
  const int dg = (int)(stride % a.G);
  const long long dp = stride / a.G;
  for (; i < total; i += stride, g += dg, p += dp) {
    if (g >= a.G) {
      g -= a.G;
      ++p;
    }
    FV<V> f = unpackv<V>(a.x[i]);
#pragma unroll
    for (int k = 0; k < V; ++k) f.v[k] = f.v[k] * sc[g * V + k] + sh[g * V + k];
    if (a.residual != nullptr && g < a.Gres) {
      const FV<V> r = unpackv<V>(a.residual[p * a.Gres + g]);
#pragma unroll
      for (int k = 0; k < V; ++k) f.v[k] += r.v[k];
    }
    if (a.relu) {
#pragma unroll
      for (int k = 0; k < V; ++k) f.v[k] = fmaxf(f.v[k], 0.f);
    }
    a.y[i] = packv<V>(f);
  }
}

// Same, with the thread layout of bn_stats_kernel (a thread keeps its 8-channel group), so the
// sums of the values it stores can ride along: the next BatchNorm needs no statistics pass.
template <int V>
__global__ void bn_apply_stats_kernel(const BnApplyArgs a) {
  extern __shared__ float ss[];  // scale[Cp], shift[Cp], red[blockDim][16]
  const int Cp = a.G * V;
  float* sc = ss;
  float* sh = ss + Cp;
  float* red = ss + 2 * Cp;
  bn_apply_prologue<V>(a, sc, sh);

  const int G = a.G;
  const int rows = blockDim.x / G;
  const int g = threadIdx.x % G, rrow = threadIdx.x / G;
  float s1[V], s2[V], kscale[V], kshift[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    s1[k] = s2[k] = 0.f;
    kscale[k] = sc[g * V + k];
    kshift[k] = sh[g * V + k];
  }
  if (rrow < rows) {
    for (long long p = (long long)blockIdx.x * rows + rrow; p < a.P; p += (long long)gridDim.x * rows) {
      FV<V> f = unpackv<V>(a.x[p * G + g]);
#pragma unroll
      for (int k = 0; k < V; ++k) f.v[k] = f.v[k] * kscale[k] + kshift[k];
      if (a.residual != nullptr && g < a.Gres) {
        const FV<V> r = unpackv<V>(a.residual[p * a.Gres + g]);
#pragma unroll
        for (int k = 0; k < V; ++k) f.v[k] += r.v[k];
      }
      if (a.relu) {
#pragma unroll
        for (int k = 0; k < V; ++k) f.v[k] = fmaxf(f.v[k], 0.f);
      }
      const uint4 packed = packv<V>(f);
      a.y[p * G + g] = packed;
      const FV<V> q = unpackv<V>(packed);  // statistics of what the consumer will read (bf16-rounded)
#pragma unroll
      for (int k = 0; k < V; ++k) {
        s1[k] += q.v[k];
        s2[k] += q.v[k] * q.v[k];
      }
    }
  }
  float* mine = red + (size_t)threadIdx.x * (2 * V);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    mine[k] = s1[k];
    mine[V + k] = s2[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < G * 2 * V; o += blockDim.x) {
    const int og = o / (2 * V), ok = o % (2 * V);
    float acc = 0.f;
    for (int rr = 0; rr < rows; ++rr) acc += red[((size_t)rr * G + og) * (2 * V) + ok];
    if (ok < V)
      atomicAdd(&a.out_sums[og * V + ok], acc);
    else
      atomicAdd(&a.out_sums[Cp + og * V + (ok - V)], acc);
  }
}

// ---------------------------------------------------------------------------
// BN backward, pass 1: sums2[c] = sum g, sums2[Cp+c] = sum g * xhat, g = dy*[y>0]
// ---------------------------------------------------------------------------
template <int V>
__global__ void bn_bwd_reduce_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ x,
                                     const uint4* __restrict__ y, const float* __restrict__ mean,
                                     const float* __restrict__ rstd, float* __restrict__ sums2,
                                     long long P, int C, int G, int relu) {
  extern __shared__ float red[];
  const int rows = blockDim.x / G;
  const int g = threadIdx.x % G, rrow = threadIdx.x / G;
  float m[V], rs[V], s1[V], s2[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int c = g * V + k;
    m[k] = c < C ? mean[c] : 0.f;
    rs[k] = c < C ? rstd[c] : 0.f;
    s1[k] = s2[k] = 0.f;
  }
  for (long long p = (long long)blockIdx.x * rows + rrow; p < P; p += (long long)gridDim.x * rows) {
    FV<V> d = unpackv<V>(dy[p * G + g]);
    const FV<V> xv = unpackv<V>(x[p * G + g]);
    if (relu) {
      const FV<V> yv = unpackv<V>(y[p * G + g]);
#pragma unroll
      for (int k = 0; k < V; ++k)
        if (!(yv.v[k] > 0.f)) d.v[k] = 0.f;
    }
#pragma unroll
    for (int k = 0; k < V; ++k) {
      s1[k] += d.v[k];
      s2[k] += d.v[k] * (xv.v[k] - m[k]) * rs[k];
    }
  }
  float* mine = red + (size_t)threadIdx.x * (2 * V);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    mine[k] = s1[k];
    mine[V + k] = s2[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < G * 2 * V; o += blockDim.x) {
    const int og = o / (2 * V), ok = o % (2 * V);
    float acc = 0.f;
    for (int rr = 0; rr < rows; ++rr) acc += red[((size_t)rr * G + og) * (2 * V) + ok];
    const int Cp = G * V;
    if (ok < V)
      atomicAdd(&sums2[og * V + ok], acc);
    else
      atomicAdd(&sums2[Cp + og * V + (ok - V)], acc);
  }
}

// ---------------------------------------------------------------------------
// BN backward, pass 2:
//   dx = gamma*rstd*(g - sum_g/P - xhat*sum_gx/P);  dres = g[..., :Cres_p]
//   dgamma = sum_gx, dbeta = sum_g  (block 0; written or accumulated)
// ---------------------------------------------------------------------------
struct BnBwdArgs {
  const uint4* dy;
  const uint4* x;
  const uint4* y;
  const float* mean;
  const float* rstd;
  const float* gamma;
  const float* sums2;
  uint4* dx;
  uint4* dres;
  int Gres;
  float* dgamma;
  float* dbeta;
  int accumulate;
  long long P;
  int C, G, relu;
  float* zero_buf;
  int zero_n;
  const uint4* add_in;  // optional [P][Cp]: another gradient of the same input, summed into dx
};

template <int V>
__global__ void bn_bwd_apply_kernel(const BnBwdArgs a) {
  extern __shared__ float ss[];  // k1[Cp] = gamma*rstd, k2[Cp] = mean_g, k3[Cp] = mean_gx*rstd, m, rs
  const int Cp = a.G * V;
  float* k1 = ss;
  float* k2 = ss + Cp;
  float* k3 = ss + 2 * Cp;
  float* mm = ss + 3 * Cp;
  float* rr = ss + 4 * Cp;
  for (int c = threadIdx.x; c < Cp; c += blockDim.x) {
    if (c < a.C) {
      const float sg = a.sums2[c], sgx = a.sums2[Cp + c];
      k1[c] = a.gamma[c] * a.rstd[c];
      k2[c] = sg / (float)a.P;
      k3[c] = sgx / (float)a.P;
      mm[c] = a.mean[c];
      rr[c] = a.rstd[c];
      if (blockIdx.x == 0) {
        if (a.accumulate) {
          a.dgamma[c] += sgx;
          a.dbeta[c] += sg;
        } else {
          a.dgamma[c] = sgx;
          a.dbeta[c] = sg;
        }
      }
    } else {
      k1[c] = k2[c] = k3[c] = mm[c] = rr[c] = 0.f;
    }
  }
  if (blockIdx.x == 0 && a.zero_buf != nullptr)
    for (int i = threadIdx.x; i < a.zero_n; i += blockDim.x) a.zero_buf[i] = 0.f;
  __syncthreads();
  const long long total = a.P * a.G;
  const long long stride = (long long)gridDim.x * blockDim.x;   // carry arithmetic, see bn_apply_kernel
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  int g = (int)(i % a.G);
  long long p = i / a.G;
  const int dg = (int)(stride % a.G);
  const long long dp = stride / a.G;
  for (; i < total; i += stride, g += dg, p += dp) {
    if (g >= a.G) {
      g -= a.G;
      ++p;
    }
    FV<V> d = unpackv<V>(a.dy[i]);
    const FV<V> xv = unpackv<V>(a.x[i]);
    if (a.relu) {
      const FV<V> yv = unpackv<V>(a.y[i]);
#pragma unroll
      for (int k = 0; k < V; ++k)
        if (!(yv.v[k] > 0.f)) d.v[k] = 0.f;
    }
    if (a.dres != nullptr && g < a.Gres) a.dres[p * a.Gres + g] = packv<V>(d);
    FV<V> o;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const int c = g * V + k;
      const float xhat = (xv.v[k] - mm[c]) * rr[c];
      o.v[k] = k1[c] * (d.v[k] - k2[c] - xhat * k3[c]);
    }
    if (a.add_in != nullptr) {
      const FV<V> e = unpackv<V>(a.add_in[i]);
#pragma unroll
      for (int k = 0; k < V; ++k) o.v[k] += e.v[k];
    }
    a.dx[i] = packv<V>(o);
  }
}

// ---------------------------------------------------------------------------
// BN backward in ONE kernel (both passes above): the two per-channel sums need every pixel before
// any dx can be written, so the two-kernel form reads dy, x (and the ReLU mask y) twice.  Here each
// CTA keeps its slice of the masked gradient and of x in shared memory across a grid-wide barrier:
//   phase 1  load dy, x, y once; store g = dy*[y>0] and x into shared memory; per-channel partial
//            sums -> block reduction -> atomics into sums2
//   barrier  all CTAs (cooperative launch: the grid is co-resident)
//   phase 2  coefficients from sums2; dx (+ add_in) and dres from the shared-memory copies
// Per layer that is 3 tensor reads + 1 write instead of 6 + 1, and one launch instead of two.  The
// aggregate shared memory of the GPU (148 x ~200 KB) bounds the layer size; larger layers take the
// two-kernel path.  Covers K-O6 / K-O7 of SURVEY.md 2.5 (autograd of pytorch/model.py:27,30,33,34).
// ---------------------------------------------------------------------------
struct GridBarrier {
  unsigned int count;
  unsigned int generation;
};

__device__ __forceinline__ void grid_barrier(GridBarrier* b) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int gen = *reinterpret_cast<volatile unsigned int*>(&b->generation);
    if (atomicAdd(&b->count, 1u) == gridDim.x - 1) {
      b->count = 0;
      __threadfence();
      atomicAdd(&b->generation, 1u);
    } else {
      long long t0 = clock64();
      while (*reinterpret_cast<volatile unsigned int*>(&b->generation) == gen) {
        __nanosleep(32);
        if (clock64() - t0 > DTDL_SPIN_TIMEOUT_CYCLES) {
          printf("[dtdl] grid barrier timed out: block %d of %d\n", blockIdx.x, gridDim.x);
          __trap();
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

template <int V>
__global__ void __launch_bounds__(512, 1)
bn_bwd_fused_kernel(const BnBwdArgs a, float* sums2, GridBarrier* bar, int rows_per_cta, int keep_x) {
  extern __shared__ uint4 fsm[];
  const int G = a.G, Cp = G * V;
  const int rows = blockDim.x / G;                 // pixel rows handled per pass by this CTA
  const int g = threadIdx.x % G, rrow = threadIdx.x / G;
  uint4* s_g = fsm;                                // [rows_per_cta][G] masked gradient
  uint4* s_x = fsm + (size_t)rows_per_cta * G;     // [rows_per_cta][G] BN input (keep_x only:
                                                   // wide fp32 layers re-read x in phase 2 instead)
  float* s_f = reinterpret_cast<float*>(fsm + (size_t)(keep_x ? 2 : 1) * rows_per_cta * G);   // reduction / coefficients
  const long long p0 = (long long)blockIdx.x * rows_per_cta;
  long long p1 = p0 + rows_per_cta;
  if (p1 > a.P) p1 = a.P;

  float m[V], rs[V], s1[V], s2[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int c = g * V + k;
    m[k] = c < a.C ? a.mean[c] : 0.f;
    rs[k] = c < a.C ? a.rstd[c] : 0.f;
    s1[k] = s2[k] = 0.f;
  }
  // One CTA per SM leaves few warps to hide the load latency, so every thread keeps kU rows' worth
  // of loads in flight before it consumes the first.
  constexpr int kU = 4;
  if (rrow < rows) {
    for (long long pb = p0 + rrow; pb < p1; pb += (long long)rows * kU) {
      uint4 dr[kU], xr[kU], yr[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long p = pb + (long long)u * rows;
        if (p < p1) {
          dr[u] = a.dy[p * G + g];
          xr[u] = a.x[p * G + g];
          if (a.relu) yr[u] = a.y[p * G + g];
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long p = pb + (long long)u * rows;
        if (p < p1) {
          FV<V> d = unpackv<V>(dr[u]);
          const FV<V> xv = unpackv<V>(xr[u]);
          if (a.relu) {
            const FV<V> yv = unpackv<V>(yr[u]);
#pragma unroll
            for (int k = 0; k < V; ++k)
              if (!(yv.v[k] > 0.f)) d.v[k] = 0.f;
          }
          const size_t li = (size_t)(p - p0) * G + g;
          s_g[li] = packv<V>(d);
          if (keep_x) s_x[li] = xr[u];
#pragma unroll
          for (int k = 0; k < V; ++k) {
            s1[k] += d.v[k];
            s2[k] += d.v[k] * (xv.v[k] - m[k]) * rs[k];
          }
        }
      }
    }
  }
  float* mine = s_f + (size_t)threadIdx.x * (2 * V);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    mine[k] = s1[k];
    mine[V + k] = s2[k];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < G * 2 * V; o += blockDim.x) {
    const int og = o / (2 * V), ok = o % (2 * V);
    float acc = 0.f;
    for (int rr = 0; rr < rows; ++rr) acc += s_f[((size_t)rr * G + og) * (2 * V) + ok];
    if (ok < V)
      atomicAdd(&sums2[og * V + ok], acc);
    else
      atomicAdd(&sums2[Cp + og * V + (ok - V)], acc);
  }
  grid_barrier(bar);

  // coefficients (same arithmetic as bn_bwd_apply_kernel)
  float* k1 = s_f;
  float* k2 = s_f + Cp;
  float* k3 = s_f + 2 * Cp;
  for (int c = threadIdx.x; c < Cp; c += blockDim.x) {
    if (c < a.C) {
      const float sg = __ldcg(sums2 + c), sgx = __ldcg(sums2 + Cp + c);
      k1[c] = a.gamma[c] * a.rstd[c];
      k2[c] = sg / (float)a.P;
      k3[c] = sgx / (float)a.P;
      if (blockIdx.x == 0) {
        if (a.accumulate) {
          a.dgamma[c] += sgx;
          a.dbeta[c] += sg;
        } else {
          a.dgamma[c] = sgx;
          a.dbeta[c] = sg;
        }
      }
    } else {
      k1[c] = k2[c] = k3[c] = 0.f;
    }
  }
  if (blockIdx.x == 0 && a.zero_buf != nullptr)
    for (int i = threadIdx.x; i < a.zero_n; i += blockDim.x) a.zero_buf[i] = 0.f;
  __syncthreads();
  if (rrow < rows) {
    float c1[V], c2[V], c3[V];
#pragma unroll
    for (int k = 0; k < V; ++k) {
      c1[k] = k1[g * V + k];
      c2[k] = k2[g * V + k];
      c3[k] = k3[g * V + k];
    }
    for (long long pb = p0 + rrow; pb < p1; pb += (long long)rows * kU) {
      uint4 xr[kU], er[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long p = pb + (long long)u * rows;
        if (p < p1) {
          if (!keep_x) xr[u] = a.x[p * G + g];
          if (a.add_in != nullptr) er[u] = a.add_in[p * G + g];
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long p = pb + (long long)u * rows;
        if (p < p1) {
          const size_t li = (size_t)(p - p0) * G + g;
          const uint4 gr = s_g[li];
          const FV<V> d = unpackv<V>(gr);
          const FV<V> xv = unpackv<V>(keep_x ? s_x[li] : xr[u]);
          if (a.dres != nullptr && g < a.Gres) a.dres[p * a.Gres + g] = gr;
          FV<V> o;
#pragma unroll
          for (int k = 0; k < V; ++k) {
            const float xhat = (xv.v[k] - m[k]) * rs[k];
            o.v[k] = c1[k] * (d.v[k] - c2[k] - xhat * c3[k]);
          }
          if (a.add_in != nullptr) {
            const FV<V> e = unpackv<V>(er[u]);
#pragma unroll
            for (int k = 0; k < V; ++k) o.v[k] += e.v[k];
          }
          a.dx[p * G + g] = packv<V>(o);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// relu backward on a saved output: dx = dy * [y > 0]
// ---------------------------------------------------------------------------
template <int V>
__global__ void relu_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y,
                                uint4* __restrict__ dx, long long n16) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
       i += (long long)gridDim.x * blockDim.x) {
    FV<V> d = unpackv<V>(dy[i]);
    const FV<V> yv = unpackv<V>(y[i]);
#pragma unroll
    for (int k = 0; k < V; ++k)
      if (!(yv.v[k] > 0.f)) d.v[k] = 0.f;
    dx[i] = packv<V>(d);
  }
}

// out = a + b  (bf16 vectors)
template <int V>
__global__ void add_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b,
                           uint4* __restrict__ o, long long n16) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
       i += (long long)gridDim.x * blockDim.x) {
    FV<V> x = unpackv<V>(a[i]);
    const FV<V> y = unpackv<V>(b[i]);
#pragma unroll
    for (int k = 0; k < V; ++k) x.v[k] += y.v[k];
    o[i] = packv<V>(x);
  }
}

// a[..., :Gb] += b  (channel-prefix add: the gradient of PyramidNet's zero-padded shortcut)
template <int V>
__global__ void add_prefix_kernel(uint4* __restrict__ a, const uint4* __restrict__ b, long long P,
                                  int Ga, int Gb) {
  const long long total = P * Gb;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % Gb);
    const long long p = i / Gb;
    FV<V> x = unpackv<V>(a[p * Ga + g]);
    const FV<V> y = unpackv<V>(b[i]);
#pragma unroll
    for (int k = 0; k < V; ++k) x.v[k] += y.v[k];
    a[p * Ga + g] = packv<V>(x);
  }
}

// ---------------------------------------------------------------------------
// Pooling on NHWC.  mode 0 = average (count includes only in-bounds taps when
// ceil-mode windows overhang; the reference's AvgPool2d(2, stride=2,
// ceil_mode=True) on even sizes never overhangs), mode 1 = max.
// ---------------------------------------------------------------------------
template <int V>
__global__ void pool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int N, int H,
                                int W, int Ho, int Wo, int G, int k, int stride, int pad, int mode) {
  const long long total = (long long)N * Ho * Wo * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    long long t = i / G;
    const int wo = (int)(t % Wo);
    t /= Wo;
    const int ho = (int)(t % Ho);
    const long long n = t / Ho;
    FV<V> acc;
#pragma unroll
    for (int q = 0; q < V; ++q) acc.v[q] = mode ? -3.0e38f : 0.f;
    int cnt = 0;
    for (int r = 0; r < k; ++r) {
      const int h = ho * stride - pad + r;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < k; ++s) {
        const int w = wo * stride - pad + s;
        if (w < 0 || w >= W) continue;
        const FV<V> v = unpackv<V>(x[((n * H + h) * W + w) * G + g]);
        ++cnt;
#pragma unroll
        for (int q = 0; q < V; ++q) acc.v[q] = mode ? fmaxf(acc.v[q], v.v[q]) : acc.v[q] + v.v[q];
      }
    }
    if (!mode) {
      const float inv = cnt > 0 ? 1.f / (float)cnt : 0.f;
#pragma unroll
      for (int q = 0; q < V; ++q) acc.v[q] *= inv;
    }
    y[i] = packv<V>(acc);
  }
}

// Gather-form backward (no atomics): each input element sums the windows it belongs to.
template <int V>
__global__ void pool_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ x,
                                const uint4* __restrict__ y, uint4* __restrict__ dx, int N, int H,
                                int W, int Ho, int Wo, int G, int k, int stride, int pad, int mode) {
  const long long total = (long long)N * H * W * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    long long t = i / G;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const long long n = t / H;
    FV<V> acc;
#pragma unroll
    for (int q = 0; q < V; ++q) acc.v[q] = 0.f;
    FV<V> xv;
    if (mode) xv = unpackv<V>(x[i]);
    for (int ho = 0; ho < Ho; ++ho) {
      const int r = h - (ho * stride - pad);
      if (r < 0 || r >= k) continue;
      for (int wo = 0; wo < Wo; ++wo) {
        const int s = w - (wo * stride - pad);
        if (s < 0 || s >= k) continue;
        const long long oi = ((n * Ho + ho) * Wo + wo) * G + g;
        const FV<V> d = unpackv<V>(dy[oi]);
        if (mode) {
          // torch routes the gradient to the FIRST maximum of the window (row-major scan);
          // bf16 data ties often, so check that no earlier tap holds the same value.
          const FV<V> yv = unpackv<V>(y[oi]);
          bool mine[V];
#pragma unroll
          for (int q = 0; q < V; ++q) mine[q] = xv.v[q] == yv.v[q];
          for (int rr = 0; rr <= r; ++rr) {
            const int hh = ho * stride - pad + rr;
            if (hh < 0 || hh >= H) continue;
            const int s_end = rr < r ? k : s;
            for (int sc = 0; sc < s_end; ++sc) {
              const int ww = wo * stride - pad + sc;
              if (ww < 0 || ww >= W) continue;
              const FV<V> ev = unpackv<V>(x[((n * H + hh) * W + ww) * G + g]);
#pragma unroll
              for (int q = 0; q < V; ++q)
                if (ev.v[q] == yv.v[q]) mine[q] = false;
            }
          }
#pragma unroll
          for (int q = 0; q < V; ++q)
            if (mine[q]) acc.v[q] += d.v[q];
        } else {
          int hs = ho * stride - pad, ws = wo * stride - pad;
          int he = hs + k < H ? hs + k : H, we = ws + k < W ? ws + k : W;
          hs = hs < 0 ? 0 : hs;
          ws = ws < 0 ? 0 : ws;
          const float inv = 1.f / (float)((he - hs) * (we - ws));
#pragma unroll
          for (int q = 0; q < V; ++q) acc.v[q] += d.v[q] * inv;
        }
      }
    }
    dx[i] = packv<V>(acc);
  }
}

// y[n,ho,wo] = x[n, 2ho, 2wo]   and its adjoint (zero insertion)
template <int V>
__global__ void subsample2_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int N, int H,
                                  int W, int Ho, int Wo, int G, int adjoint) {
  if (!adjoint) {
    const long long total = (long long)N * Ho * Wo * G;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
      const int g = (int)(i % G);
      long long t = i / G;
      const int wo = (int)(t % Wo);
      t /= Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      y[i] = x[((n * H + 2 * ho) * W + 2 * wo) * G + g];
    }
  } else {
    // here x is the small [N,Ho,Wo] gradient and y the big [N,H,W] one
    const long long total = (long long)N * H * W * G;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
      const int g = (int)(i % G);
      long long t = i / G;
      const int w = (int)(t % W);
      t /= W;
      const int h = (int)(t % H);
      const long long n = t / H;
      uint4 v = make_uint4(0, 0, 0, 0);
      if ((h & 1) == 0 && (w & 1) == 0 && h / 2 < Ho && w / 2 < Wo)
        v = x[((n * Ho + h / 2) * Wo + w / 2) * G + g];
      y[i] = v;
    }
  }
}

// Global average pool [N,HW,Cp] -> [N,Cp] and its backward (broadcast / HW).
template <int V>
__global__ void gap_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int N, int HW,
                               int G) {
  const long long total = (long long)N * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const long long n = i / G;
    FV<V> acc;
#pragma unroll
    for (int q = 0; q < V; ++q) acc.v[q] = 0.f;
    for (int p = 0; p < HW; ++p) {
      const FV<V> v = unpackv<V>(x[(n * HW + p) * G + g]);
#pragma unroll
      for (int q = 0; q < V; ++q) acc.v[q] += v.v[q];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int q = 0; q < V; ++q) acc.v[q] *= inv;
    y[i] = packv<V>(acc);
  }
}
template <int V>
__global__ void gap_bwd_kernel(const uint4* __restrict__ dy, uint4* __restrict__ dx, int N, int HW,
                               int G) {
  const long long total = (long long)N * HW * G;
  const float inv = 1.f / (float)HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const long long n = i / ((long long)HW * G);
    FV<V> d = unpackv<V>(dy[n * G + g]);
#pragma unroll
    for (int q = 0; q < V; ++q) d.v[q] *= inv;
    dx[i] = packv<V>(d);
  }
}

// ---------------------------------------------------------------------------
// Softmax cross entropy (mean over the batch).  One warp per sample.
//   out[0] += loss_n / N_total      out[1] += [argmax == target]
//   dlogits[n, c] = (softmax - onehot) * grad_scale        (bf16, pad channels 0)
// probs_input=1 treats `logits` as probabilities (Keras' softmax-in-model +
// sparse_categorical_crossentropy pairing, tensorflow2/mnist_single.py:26,79).
// ---------------------------------------------------------------------------
__global__ void softmax_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ tgt,
                                  float* __restrict__ out, __nv_bfloat16* __restrict__ dlogits,
                                  float* __restrict__ dlogits_f32, int N, int C, int Cp,
                                  float inv_n, float grad_scale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= N) return;
  const float* row = logits + (size_t)warp * Cp;
  float mx = -3.0e38f;
  int am = 0;
  for (int c = lane; c < C; c += 32) {
    const float v = row[c];
    if (v > mx) {
      mx = v;
      am = c;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float omx = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oam = __shfl_xor_sync(0xffffffffu, am, o);
    if (omx > mx || (omx == mx && oam < am)) {
      mx = omx;
      am = oam;
    }
  }
  float se = 0.f;
  for (int c = lane; c < C; c += 32) se += __expf(row[c] - mx);
  for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
  const int t = (int)tgt[warp];
  const float lse = mx + __logf(se);
  if (lane == 0) {
    atomicAdd(&out[0], (lse - row[t]) * inv_n);
    if (am == t) atomicAdd(&out[1], 1.f);
  }
  for (int c = lane; c < Cp; c += 32) {
    float gval = 0.f;
    if (c < C) gval = (__expf(row[c] - lse) - (c == t ? 1.f : 0.f)) * grad_scale;
    if (dlogits != nullptr) dlogits[(size_t)warp * Cp + c] = __float2bfloat16(gval);
    if (dlogits_f32 != nullptr) dlogits_f32[(size_t)warp * Cp + c] = gval;
  }
}

// Column sum of a bf16 [P][Cp] matrix into fp32 out[C] (bias gradients).
template <int V>
__global__ void colsum_kernel(const uint4* __restrict__ x, float* __restrict__ out, long long P,
                              int C, int G) {
  extern __shared__ float red[];  // [rows][G][V]
  const int rows = blockDim.x / G;
  const int g = threadIdx.x % G, rrow = threadIdx.x / G;
  float s[V];
#pragma unroll
  for (int k = 0; k < V; ++k) s[k] = 0.f;
  for (long long p = (long long)blockIdx.x * rows + rrow; p < P; p += (long long)gridDim.x * rows) {
    const FV<V> f = unpackv<V>(x[p * G + g]);
#pragma unroll
    for (int k = 0; k < V; ++k) s[k] += f.v[k];
  }
#pragma unroll
  for (int k = 0; k < V; ++k) red[(size_t)threadIdx.x * V + k] = s[k];
  __syncthreads();
  for (int o = threadIdx.x; o < G * V; o += blockDim.x) {
    const int og = o / V, ok = o % V;
    float acc = 0.f;
    for (int rr = 0; rr < rows; ++rr) acc += red[((size_t)rr * G + og) * V + ok];
    if (og * V + ok < C) atomicAdd(&out[og * V + ok], acc);
  }
}

// fp32 [rows][cols] -> activation type [rows][cols_p] (zero padded) and back
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16(v);
}
template <>
__device__ __forceinline__ float from_f32<float>(float v) {
  return v;
}
__device__ __forceinline__ float to_f32v(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32v(float v) { return v; }

template <typename T>
__global__ void pad_cast_kernel(const float* __restrict__ x, T* __restrict__ y, long long rows, int cols,
                                int cols_p) {
  const long long total = rows * cols_p;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cols_p);
    const long long r = i / cols_p;
    y[i] = from_f32<T>(c < cols ? x[r * cols + c] : 0.f);
  }
}
template <typename T>
__global__ void unpad_cast_kernel(const T* __restrict__ x, float* __restrict__ y, long long rows, int cols,
                                  int cols_p) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cols);
    const long long r = i / cols;
    y[i] = to_f32v(x[r * cols_p + c]);
  }
}

// ---------------------------------------------------------------------------
// fp32 Linear on CUDA cores (fp32 activation mode only).  The reference's nn.Linear runs in full
// fp32 (torch keeps TF32 off for matmuls by default: pytorch/model.py:77,111), so in the
// matched-precision mode the fully connected layers do not go through the TF32 tensor-core path.
// The layers this serves are small (271 -> 10 in PyramidNet); one thread per output element with
// float4 loads along K is enough.
//   fwd  : y[n][o]  = [relu](sum_k x[n][k] w[o][k] + b[o])         x [N][Kp], y [N][Op] (pads 0)
//   dgrad: dx[n][k] = sum_o dy[n][o] w[o][k]
//   wgrad: dw[o][k] (+)= sum_n dy[n][o] x[n][k]
// ---------------------------------------------------------------------------
__global__ void linear_f32_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                      const float* __restrict__ bias, float* __restrict__ y, int N,
                                      int K, int Kp, int O, int Op, int relu, float* __restrict__ y2,
                                      int y2_ld) {
  // one warp per output element: the lanes stride over K (coalesced reads of the x row and the w row)
  // and combine with shuffles -- a thread per output would walk K serially, one dependent load at a time
  const int lane = threadIdx.x & 31;
  const long long total = (long long)N * Op;
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < total; i += warps) {
    const int o = (int)(i % Op);
    const long long n = i / Op;
    float acc = 0.f;
    if (o < O) {
      const float* xr = x + n * Kp;
      const float* wr = w + (long long)o * K;
      for (int k = lane; k < K; k += 32) acc = fmaf(xr[k], wr[k], acc);
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
      if (bias != nullptr) acc += bias[o];
      if (relu) acc = fmaxf(acc, 0.f);
    }
    if (lane == 0) {
      if (o < O && y2 != nullptr) y2[n * y2_ld + o] = acc;   // second copy, e.g. a row of GPU 0's logits buffer
      y[i] = acc;
    }
  }
}
__global__ void linear_f32_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                        float* __restrict__ dx, int N, int K, int Kp, int O, int Op) {
  const long long total = (long long)N * Kp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % Kp);
    const long long n = i / Kp;
    float acc = 0.f;
    if (k < K) {
      const float* dr = dy + n * Op;
      for (int o = 0; o < O; ++o) acc = fmaf(dr[o], w[(long long)o * K + k], acc);
    }
    dx[i] = acc;
  }
}
__global__ void linear_f32_wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                        float* __restrict__ dw, int N, int K, int Kp, int O, int Op,
                                        int accumulate) {
  const long long total = (long long)O * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % K);
    const int o = (int)(i / K);
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc = fmaf(dy[(long long)n * Op + o], x[(long long)n * Kp + k], acc);
    dw[i] = accumulate ? dw[i] + acc : acc;
  }
}

}  // namespace dtdl

using namespace dtdl;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)
static const int kMaxBlocks = 148 * 8;

// act -> channels per 16-byte vector
#define DTDL_ACT_OK(act) DTDL_REQUIRE((act) == kActBf16 || (act) == kActF32, "act must be 0 (bf16) or 1 (fp32)")
static inline int vec_of(int act) { return act == kActF32 ? 4 : 8; }
// launches KERNEL<8> (bf16) or KERNEL<4> (fp32) with the same configuration and arguments
#define LAUNCH_ACT(act, KERNEL, grid, block, smem, stream, ...)              \
  do {                                                                         \
    if ((act) == kActF32)                                                      \
      KERNEL<4><<<(grid), (block), (smem), STREAM(stream)>>>(__VA_ARGS__);     \
    else                                                                       \
      KERNEL<8><<<(grid), (block), (smem), STREAM(stream)>>>(__VA_ARGS__);     \
    DTDL_CHECK_CUDA(cudaGetLastError());                                       \
  } while (0)

extern "C" int dtdl_nchw_to_nhwc(const float* x, void* y, int N, int C, int H, int W, int Cp, int act,
                                 void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cp >= C, "bad channel padding");
  const int G = Cp / vec_of(act);
  const long long total = (long long)N * H * W * G;
  LAUNCH_ACT(act, nchw_to_nhwc_kernel, grid_for(total, 256, kMaxBlocks), 256, 0, stream, x,
             reinterpret_cast<uint4*>(y), N, C, H * W, G);
  return 0;
}

extern "C" int dtdl_nchw_to_s2d(const float* x, void* y, int N, int C, int H, int W, int Cp, int act,
                                void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cp >= 4 * C && H % 2 == 0 && W % 2 == 0, "bad space-to-depth shape");
  const int G = Cp / vec_of(act);
  const long long total = (long long)N * (H / 2) * (W / 2) * G;
  LAUNCH_ACT(act, nchw_to_s2d_kernel, grid_for(total, 256, kMaxBlocks), 256, 0, stream, x,
             reinterpret_cast<uint4*>(y), N, C, H, W, G);
  return 0;
}

// One-time opt-in to large dynamic shared memory for the reduction kernels of both instantiations.
static int set_reduce_attrs() {
  static bool attr[64] = {false};
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr[dev & 63]) {
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_stats_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_stats_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_reduce_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_reduce_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_apply_stats_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_apply_stats_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
    attr[dev & 63] = true;
  }
  return 0;
}

extern "C" int dtdl_bn_stats(const void* x, float* sums, long long P, int Cp, int act, void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cp / V <= 1024, "bad channel padding");
  const int G = Cp / V;
  const int block = block_for_groups(G);
  const int rows = block / G;
  const size_t smem = (size_t)block * 2 * V * sizeof(float);
  if (set_reduce_attrs()) return 1;
  LAUNCH_ACT(act, bn_stats_kernel, grid_for(P, rows * 8, 148 * 4), block, smem, stream,
             reinterpret_cast<const uint4*>(x), sums, P, G);
  return 0;
}

extern "C" int dtdl_bn_apply(const void* x, void* y, const float* sums, const float* gamma,
                             const float* beta, float* running_mean, float* running_var,
                             long long* num_batches_tracked, float* save_mean, float* save_rstd,
                             const void* residual, int Cres_p, long long P, int C, int Cp, float eps,
                             float momentum, int relu, int training, float* zero_buf, int zero_n,
                             int act, void* stream) {
  return dtdl_bn_apply_stats(x, y, sums, gamma, beta, running_mean, running_var, num_batches_tracked,
                             save_mean, save_rstd, residual, Cres_p, P, C, Cp, eps, momentum, relu,
                             training, zero_buf, zero_n, nullptr, act, stream);
}

extern "C" int dtdl_bn_apply_stats(const void* x, void* y, const float* sums, const float* gamma,
                                   const float* beta, float* running_mean, float* running_var,
                                   long long* num_batches_tracked, float* save_mean, float* save_rstd,
                                   const void* residual, int Cres_p, long long P, int C, int Cp,
                                   float eps, float momentum, int relu, int training, float* zero_buf,
                                   int zero_n, float* out_sums, int act, void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cres_p % 8 == 0 && Cres_p <= Cp, "bad channel padding");
  BnApplyArgs a;
  a.x = reinterpret_cast<const uint4*>(x);
  a.y = reinterpret_cast<uint4*>(y);
  a.sums = sums;
  a.gamma = gamma;
  a.beta = beta;
  a.running_mean = running_mean;
  a.running_var = running_var;
  a.num_batches_tracked = num_batches_tracked;
  a.save_mean = save_mean;
  a.save_rstd = save_rstd;
  a.residual = reinterpret_cast<const uint4*>(residual);
  a.Gres = Cres_p / V;
  a.P = P;
  a.C = C;
  a.G = Cp / V;
  a.eps = eps;
  a.momentum = momentum;
  a.relu = relu;
  a.training = training;
  a.zero_buf = zero_buf;
  a.zero_n = zero_n;
  a.out_sums = out_sums;
  if (out_sums != nullptr) {
    DTDL_REQUIRE(a.G <= 1024, "bad channel padding");
    const int block = block_for_groups(a.G);
    const int rows = block / a.G;
    const size_t smem = (size_t)(2 * Cp + block * 2 * V) * sizeof(float);
    if (set_reduce_attrs()) return 1;
    LAUNCH_ACT(act, bn_apply_stats_kernel, grid_for(P, rows * 4, kMaxBlocks), block, smem, stream, a);
    return 0;
  }
  const long long total = P * a.G;
  LAUNCH_ACT(act, bn_apply_kernel, grid_for(total, 256 * 4, kMaxBlocks), 256, 2 * Cp * sizeof(float), stream, a);
  return 0;
}

extern "C" int dtdl_bn_bwd(const void* dy, const void* x, const void* y, const float* save_mean,
                           const float* save_rstd, const float* gamma, float* sums2, void* dx,
                           void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                           long long P, int C, int Cp, int relu, float* zero_buf, int zero_n, int act,
                           void* stream) {
  return dtdl_bn_bwd_add(dy, x, y, save_mean, save_rstd, gamma, sums2, dx, dres, Cres_p, dgamma, dbeta,
                         accumulate, P, C, Cp, relu, zero_buf, zero_n, nullptr, act, stream);
}

extern "C" int dtdl_bn_bwd_add(const void* dy, const void* x, const void* y, const float* save_mean,
                               const float* save_rstd, const float* gamma, float* sums2, void* dx,
                               void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                               long long P, int C, int Cp, int relu, float* zero_buf, int zero_n,
                               const void* add_in, int act, void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cres_p % 8 == 0 && Cres_p <= Cp && Cp / V <= 1024, "bad channel padding");
  DTDL_REQUIRE(!relu || y != nullptr, "relu backward needs the saved output");
  const int G = Cp / V;
  {
    const int block = block_for_groups(G);
    const int rows = block / G;
    const size_t smem = (size_t)block * 2 * V * sizeof(float);
    if (set_reduce_attrs()) return 1;
    LAUNCH_ACT(act, bn_bwd_reduce_kernel, grid_for(P, rows * 8, 148 * 4), block, smem, stream,
               reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(x),
               reinterpret_cast<const uint4*>(y), save_mean, save_rstd, sums2, P, C, G, relu);
  }
  BnBwdArgs a;
  a.dy = reinterpret_cast<const uint4*>(dy);
  a.x = reinterpret_cast<const uint4*>(x);
  a.y = reinterpret_cast<const uint4*>(y);
  a.mean = save_mean;
  a.rstd = save_rstd;
  a.gamma = gamma;
  a.sums2 = sums2;
  a.dx = reinterpret_cast<uint4*>(dx);
  a.dres = reinterpret_cast<uint4*>(dres);
  a.Gres = Cres_p / V;
  a.dgamma = dgamma;
  a.dbeta = dbeta;
  a.accumulate = accumulate;
  a.P = P;
  a.C = C;
  a.G = G;
  a.relu = relu;
  a.zero_buf = zero_buf;
  a.zero_n = zero_n;
  a.add_in = reinterpret_cast<const uint4*>(add_in);
  const long long total = P * G;
  LAUNCH_ACT(act, bn_bwd_apply_kernel, grid_for(total, 256 * 4, kMaxBlocks), 256, 5 * Cp * sizeof(float),
             stream, a);
  return 0;
}

// Launch plan of the one-kernel BatchNorm backward; false when the layer does not fit.
static bool bn_fused_plan(long long P, int Cp, int act, int* grid, int* block, int* rows_per_cta, size_t* smem,
                          int* keep_x) {
  static const bool off = getenv("DTDL_BN_FUSED_BWD") != nullptr && atoi(getenv("DTDL_BN_FUSED_BWD")) == 0;
  if (off) return false;
  const int V = vec_of(act);
  const int G = Cp / V;
  if (G < 1 || G > 512 || P < 1) return false;
  const int sms = sm_count();
  *block = G * (512 / G);
  *rows_per_cta = (int)((P + sms - 1) / sms);
  *grid = (int)((P + *rows_per_cta - 1) / *rows_per_cta);
  const size_t fl = (size_t)(*block) * 2 * V > (size_t)3 * Cp ? (size_t)(*block) * 2 * V : (size_t)3 * Cp;
  // both tiles resident when they fit; otherwise only the masked gradient (x is read again in phase 2)
  *keep_x = 1;
  *smem = (size_t)2 * (*rows_per_cta) * G * 16 + fl * sizeof(float);
  if (*smem <= 216 * 1024) return true;
  *keep_x = 0;
  *smem = (size_t)(*rows_per_cta) * G * 16 + fl * sizeof(float);
  return *smem <= 216 * 1024;
}

extern "C" int dtdl_bn_bwd_fused_fits(long long P, int Cp, int act) {
  int grid, block, rows, keep_x;
  size_t smem;
  return (act == kActBf16 || act == kActF32) && Cp % 8 == 0 &&
                 bn_fused_plan(P, Cp, act, &grid, &block, &rows, &smem, &keep_x)
             ? 1
             : 0;
}

// Same contract as dtdl_bn_bwd_add; `barrier` = 8 zero-initialised bytes of device memory owned by
// the caller (the grid barrier's counter and generation; one per concurrently running call).
extern "C" int dtdl_bn_bwd_fused(const void* dy, const void* x, const void* y, const float* save_mean,
                                 const float* save_rstd, const float* gamma, float* sums2, void* dx,
                                 void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                                 long long P, int C, int Cp, int relu, float* zero_buf, int zero_n,
                                 const void* add_in, void* barrier, int act, void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cres_p % 8 == 0 && Cres_p <= Cp, "bad channel padding");
  DTDL_REQUIRE(!relu || y != nullptr, "relu backward needs the saved output");
  DTDL_REQUIRE(barrier != nullptr, "the fused BatchNorm backward needs its barrier words");
  int grid, block, rows_per_cta, keep_x;
  size_t smem;
  DTDL_REQUIRE(bn_fused_plan(P, Cp, act, &grid, &block, &rows_per_cta, &smem, &keep_x),
               "layer too large for the one-kernel BatchNorm backward (see dtdl_bn_bwd_fused_fits)");
  static bool attr[64] = {false};
  int dev = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr[dev & 63]) {
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_fused_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DTDL_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_fused_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr[dev & 63] = true;
  }
  BnBwdArgs a;
  a.dy = reinterpret_cast<const uint4*>(dy);
  a.x = reinterpret_cast<const uint4*>(x);
  a.y = reinterpret_cast<const uint4*>(y);
  a.mean = save_mean;
  a.rstd = save_rstd;
  a.gamma = gamma;
  a.sums2 = sums2;
  a.dx = reinterpret_cast<uint4*>(dx);
  a.dres = reinterpret_cast<uint4*>(dres);
  a.Gres = Cres_p / V;
  a.dgamma = dgamma;
  a.dbeta = dbeta;
  a.accumulate = accumulate;
  a.P = P;
  a.C = C;
  a.G = Cp / V;
  a.relu = relu;
  a.zero_buf = zero_buf;
  a.zero_n = zero_n;
  a.add_in = reinterpret_cast<const uint4*>(add_in);
  GridBarrier* bar = reinterpret_cast<GridBarrier*>(barrier);
  void* args[] = {&a, &sums2, &bar, &rows_per_cta, &keep_x};
  const void* fn = act == kActF32 ? reinterpret_cast<const void*>(bn_bwd_fused_kernel<4>)
                                  : reinterpret_cast<const void*>(bn_bwd_fused_kernel<8>);
  // cooperative launch: every CTA of the grid is resident at once, which the grid barrier needs
  DTDL_CHECK_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(block), args, smem, STREAM(stream)));
  return 0;
}

extern "C" int dtdl_relu_bwd(const void* dy, const void* y, void* dx, long long n_elems, int act,
                             void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(n_elems % 8 == 0, "element count must be a multiple of 8");
  const long long n16 = n_elems / vec_of(act);
  LAUNCH_ACT(act, relu_bwd_kernel, grid_for(n16, 256 * 4, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(y),
             reinterpret_cast<uint4*>(dx), n16);
  return 0;
}

extern "C" int dtdl_add(const void* a, const void* b, void* out, long long n_elems, int act,
                        void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(n_elems % 8 == 0, "element count must be a multiple of 8");
  const long long n16 = n_elems / vec_of(act);
  LAUNCH_ACT(act, add_kernel, grid_for(n16, 256 * 4, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b),
             reinterpret_cast<uint4*>(out), n16);
  return 0;
}

extern "C" int dtdl_add_prefix(void* a, const void* b, long long P, int Ca_p, int Cb_p, int act,
                               void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Ca_p % 8 == 0 && Cb_p % 8 == 0 && Cb_p <= Ca_p, "bad channel padding");
  LAUNCH_ACT(act, add_prefix_kernel, grid_for(P * (Cb_p / V), 256 * 4, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<uint4*>(a), reinterpret_cast<const uint4*>(b), P, Ca_p / V, Cb_p / V);
  return 0;
}

extern "C" int dtdl_pool_fwd(const void* x, void* y, int N, int H, int W, int Ho, int Wo, int Cp,
                             int k, int stride, int pad, int mode, int act, void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0, "bad channel padding");
  const int G = Cp / vec_of(act);
  const long long total = (long long)N * Ho * Wo * G;
  LAUNCH_ACT(act, pool_fwd_kernel, grid_for(total, 256, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(y), N, H, W, Ho, Wo, G, k,
             stride, pad, mode);
  return 0;
}

extern "C" int dtdl_pool_bwd(const void* dy, const void* x, const void* y, void* dx, int N, int H,
                             int W, int Ho, int Wo, int Cp, int k, int stride, int pad, int mode,
                             int act, void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0, "bad channel padding");
  const int G = Cp / vec_of(act);
  const long long total = (long long)N * H * W * G;
  LAUNCH_ACT(act, pool_bwd_kernel, grid_for(total, 256, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(x),
             reinterpret_cast<const uint4*>(y), reinterpret_cast<uint4*>(dx), N, H, W, Ho, Wo, G, k,
             stride, pad, mode);
  return 0;
}

extern "C" int dtdl_subsample2(const void* x, void* y, int N, int H, int W, int Cp, int adjoint,
                               int act, void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0, "bad channel padding");
  const int G = Cp / vec_of(act);
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  const long long total = (long long)N * (adjoint ? H * W : Ho * Wo) * G;
  LAUNCH_ACT(act, subsample2_kernel, grid_for(total, 256, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(y), N, H, W, Ho, Wo, G, adjoint);
  return 0;
}

extern "C" int dtdl_gap_fwd(const void* x, void* y, int N, int HW, int Cp, int act, void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0, "bad channel padding");
  const int G = Cp / vec_of(act);
  LAUNCH_ACT(act, gap_fwd_kernel, grid_for((long long)N * G, 64, kMaxBlocks), 64, 0, stream,
             reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(y), N, HW, G);
  return 0;
}
extern "C" int dtdl_gap_bwd(const void* dy, void* dx, int N, int HW, int Cp, int act, void* stream) {
  DTDL_ACT_OK(act);
  DTDL_REQUIRE(Cp % 8 == 0, "bad channel padding");
  const int G = Cp / vec_of(act);
  LAUNCH_ACT(act, gap_bwd_kernel, grid_for((long long)N * HW * G, 256, kMaxBlocks), 256, 0, stream,
             reinterpret_cast<const uint4*>(dy), reinterpret_cast<uint4*>(dx), N, HW, G);
  return 0;
}

extern "C" int dtdl_softmax_ce(const float* logits, const long long* targets, float* out2,
                               void* dlogits_bf16, float* dlogits_f32, int N, int C, int Cp,
                               float inv_n, float grad_scale, void* stream) {
  const int warps_per_block = 4;
  softmax_ce_kernel<<<(N + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0,
                      STREAM(stream)>>>(logits, targets, out2,
                                        reinterpret_cast<__nv_bfloat16*>(dlogits_bf16), dlogits_f32,
                                        N, C, Cp, inv_n, grad_scale);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dtdl_colsum(const void* x, float* out, long long P, int C, int Cp, int act,
                           void* stream) {
  DTDL_ACT_OK(act);
  const int V = vec_of(act);
  DTDL_REQUIRE(Cp % 8 == 0 && Cp / V <= 1024, "bad channel padding");
  const int G = Cp / V;
  const int block = block_for_groups(G);
  const int rows = block / G;
  LAUNCH_ACT(act, colsum_kernel, grid_for(P, rows * 8, 148 * 2), block, (size_t)block * V * sizeof(float),
             stream, reinterpret_cast<const uint4*>(x), out, P, C, G);
  return 0;
}

extern "C" int dtdl_pad_cast(const float* x, void* y, long long rows, int cols, int cols_p, int act,
                             void* stream) {
  DTDL_ACT_OK(act);
  const int grid = grid_for(rows * cols_p, 256, kMaxBlocks);
  if (act == kActF32)
    pad_cast_kernel<float><<<grid, 256, 0, STREAM(stream)>>>(x, reinterpret_cast<float*>(y), rows, cols, cols_p);
  else
    pad_cast_kernel<__nv_bfloat16><<<grid, 256, 0, STREAM(stream)>>>(x, reinterpret_cast<__nv_bfloat16*>(y),
                                                                     rows, cols, cols_p);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int dtdl_unpad_cast(const void* x, float* y, long long rows, int cols, int cols_p, int act,
                               void* stream) {
  DTDL_ACT_OK(act);
  const int grid = grid_for(rows * cols, 256, kMaxBlocks);
  if (act == kActF32)
    unpad_cast_kernel<float><<<grid, 256, 0, STREAM(stream)>>>(reinterpret_cast<const float*>(x), y, rows, cols, cols_p);
  else
    unpad_cast_kernel<__nv_bfloat16><<<grid, 256, 0, STREAM(stream)>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                                       y, rows, cols, cols_p);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dtdl_linear_f32_fwd(const float* x, const float* w, const float* bias, float* y, float* y2,
                                   int y2_ld, int N, int K, int Kp, int O, int Op, int relu, void* stream) {
  DTDL_REQUIRE(y2 == nullptr || y2_ld >= O, "bad gather destination");
  linear_f32_fwd_kernel<<<grid_for((long long)N * Op, 8, kMaxBlocks), 256, 0, STREAM(stream)>>>(
      x, w, bias, y, N, K, Kp, O, Op, relu, y2, y2_ld);   // 8 warps = 8 outputs per block
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int dtdl_linear_f32_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw,
                                   int N, int K, int Kp, int O, int Op, int accumulate, void* stream) {
  if (dx != nullptr) {
    linear_f32_dgrad_kernel<<<grid_for((long long)N * Kp, 128, kMaxBlocks), 128, 0, STREAM(stream)>>>(
        dy, w, dx, N, K, Kp, O, Op);
    DTDL_CHECK_CUDA(cudaGetLastError());
  }
  linear_f32_wgrad_kernel<<<grid_for((long long)O * K, 128, kMaxBlocks), 128, 0, STREAM(stream)>>>(
      dy, x, dw, N, K, Kp, O, Op, accumulate);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
