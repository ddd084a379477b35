// Peer-memory collectives over NVLink 5 / NVSwitch, with the optimizer fused in.
//
// Every rank owns "symmetric" buffers (same size on every rank) and holds a
// table of pointers to each peer's copy (CUDA IPC across processes, plain peer
// access inside one process).  Kernels read and write peer memory directly
// with ordinary ld/st; ordering between ranks uses system-scope release /
// acquire on per-rank flag words with a monotonically increasing epoch that
// lives in device memory, so the same launch can be replayed from a CUDA graph.
// The flag exchanges are one-CTA kernels in front of / behind the data kernels
// (sync_kernel below), so a data kernel never spins on a peer.
//
//   one-shot : every rank reads all W copies of the whole bucket and reduces
//              locally (latency optimal, W x read traffic; small buckets)
//   two-shot : rank r reduces slice r from all W copies, then stores the
//              result into all W destinations (bandwidth optimal on NVSwitch)
//
//   multicast: (NVLS, buffers from the vmm.cpp allocator) rank r reads slice r through the
//              multicast address with multimem.ld_reduce -- NVSwitch sums the W copies and
//              returns one value -- and writes the result with multimem.st, which the switch
//              replicates to all W ranks: 1/W of the NVLink traffic of two-shot per GPU
//
// The fused variants apply cast + 1/W + weight decay + SGD-momentum or Adam to
// the reduced gradient and (two-shot) scatter the *updated parameters* instead
// of gradients, so gradient all-reduce, averaging and optimizer.step() are one
// kernel and one pass over memory.
//
// Reductions always sum ranks in rank order 0..W-1, so every rank computes
// bit-identical results and runs are reproducible.
//
// Capability covered: K-C1, K-C2, K-C3, K-C5, K-C7, K-C8, K-C10, K-C11, K-O14,
// K-O15 of SURVEY.md section 2.5 (the reference reaches NCCL through
// DistributedDataParallel / DataParallel and torch.optim.SGD:
// pytorch/distributed_data_parallel.py:74,94-95,132-133; pytorch/data_parallel.py:71,108-109).
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"

#include <stdlib.h>
#include <string.h>

namespace dtdl {

static constexpr int kMaxWorld = 16;

// Cross-GPU synchronisation is done by one-CTA kernels of its own (sync_kernel), not inside the data
// kernels: a data kernel of up to 64 CTAs x 512 threads that spins on peer flags holds SM slots that
// the backward pass running next to it wants, for as long as the slowest rank is late.  With the
// wait in a 32-thread kernel in front of it, the data kernel starts only when every rank is ready and
// never spins.  Two flag rows with one epoch word each:
//   START : "everything I enqueued before this collective is complete" -- exchanged before the data
//           kernel reads peer buffers (each bucket);
//   DONE  : "my data kernels have finished reading / writing your buffers" -- exchanged before the
//           buffers are touched again.  In a training step only ONE such exchange is needed, after the
//           last bucket (the buckets touch disjoint memory, and bucket k+1's START already implies
//           that every rank has finished bucket k): dtdl_comm_defer_done switches the per-collective
//           DONE off and the caller closes the step with dtdl_comm_barrier.
// Epochs live in device memory and only ever grow, so the same launches can be replayed from a graph.
struct CommDev {
  int rank, world;
  uint32_t* flags_local;           // [2][kMaxWorld]: START row, DONE row
  uint32_t* flags_peer[kMaxWorld]; // peers' flags_local
  uint32_t* epochs;                // device words [2]: last completed epoch of each row
};

// Timeline of the collectives of one step, written by the kernels themselves so that it also works
// inside a replayed graph: slot = {first block's start, last block's end (globaltimer ns), kind,
// bytes}.  Slots are handed out in launch order from dtdl_comm_trace(); a captured launch keeps the
// slot it got at capture time.
struct TraceSlot {
  unsigned long long t0, t1, kind, bytes;
};
enum { TR_START = 1, TR_DONE = 2, TR_REDUCE = 3, TR_NVLS_REDUCE = 4, TR_BROADCAST = 5, TR_ALLGATHER = 6,
       TR_MARK = 7, TR_LOCAL_UPDATE = 8 };

struct CommHost {
  CommDev d;
  int device;
  int defer_done;
  TraceSlot* trace;
  int trace_cap, trace_next;
};

struct PeerPtrs {
  void* p[kMaxWorld];
};

// hyper-parameter block in device memory (written by the host before a step)
//  [0] lr  [1] momentum  [2] weight_decay  [3] inv_world  [4] beta1  [5] beta2
//  [6] eps [7] bias_correction1  [8] bias_correction2_sqrt  [9] dampening
enum { H_LR = 0, H_MOM, H_WD, H_INVW, H_B1, H_B2, H_EPS, H_BC1, H_BC2S, H_DAMP };

enum { OPT_NONE = 0, OPT_SGD = 1, OPT_ADAM = 2, OPT_ADAM_KERAS = 3 };

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void trace_enter(TraceSlot* s, int kind, long long bytes) {
  if (s != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    s->t0 = globaltimer_ns();
    s->kind = (unsigned long long)kind;
    s->bytes = (unsigned long long)bytes;
  }
}
__device__ __forceinline__ void trace_exit(TraceSlot* s) {
  if (s != nullptr && threadIdx.x == 0) atomicMax(&s->t1, globaltimer_ns());
}

__device__ __forceinline__ void spin_until_ge(const uint32_t* p, uint32_t want, int tag) {
  long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(p) - want) < 0) {
    __nanosleep(40);
    if (clock64() - t0 > DTDL_SPIN_TIMEOUT_CYCLES) {
      printf("[dtdl] peer flag wait timed out (row %d): thread %d want %u have %u\n", tag, threadIdx.x,
             want, ld_relaxed_sys(p));
      __trap();
    }
  }
}

// One CTA of 32 threads: thread r announces this rank's epoch to rank r and waits for rank r's.
__global__ void __launch_bounds__(32)
sync_kernel(const CommDev c, int row, TraceSlot* slot) {
  trace_enter(slot, row == 0 ? TR_START : TR_DONE, 0);
  const uint32_t e = c.epochs[row] + 1;
  if (threadIdx.x < c.world) {
    __threadfence_system();
    st_release_sys(c.flags_peer[threadIdx.x] + row * kMaxWorld + c.rank, e);
    spin_until_ge(c.flags_local + row * kMaxWorld + threadIdx.x, e, row);
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    c.epochs[row] = e;
    __threadfence();
  }
  trace_exit(slot);
}

__global__ void trace_mark_kernel(TraceSlot* slot, int kind) {
  const unsigned long long t = globaltimer_ns();
  slot->t0 = t;
  atomicMax(&slot->t1, t);
  slot->kind = (unsigned long long)kind;
  slot->bytes = 0;
}

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) {
  return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 ld_peer_f4(const float* base, long long i4) {
  // peer reads bypass the local L2; keep them out of L1 too (data is read once)
  return ld_nc_f4(reinterpret_cast<const float4*>(base) + i4);
}

__device__ __forceinline__ float sgd_one(float g, float& p, float& m, const float* h) {
  g = g * h[H_INVW] + h[H_WD] * p;
  m = h[H_MOM] * m + (1.f - h[H_DAMP]) * g;
  p -= h[H_LR] * m;
  return p;
}
__device__ __forceinline__ float adam_one(float g, float& p, float& m, float& v, const float* h,
                                          int keras) {
  g = g * h[H_INVW] + h[H_WD] * p;
  m = h[H_B1] * m + (1.f - h[H_B1]) * g;
  v = h[H_B2] * v + (1.f - h[H_B2]) * g * g;
  if (keras) {
    // lr_t = lr * sqrt(1-b2^t)/(1-b1^t);  p -= lr_t * m / (sqrt(v) + eps)
    p -= h[H_LR] * h[H_BC2S] / h[H_BC1] * m / (sqrtf(v) + h[H_EPS]);
  } else {
    p -= (h[H_LR] / h[H_BC1]) * m / (sqrtf(v) / h[H_BC2S] + h[H_EPS]);
  }
  return p;
}

struct ReduceArgs {
  PeerPtrs grads;    // fp32 gradient bucket on every rank (same offset everywhere)
  PeerPtrs dst;      // two-shot plain: reduced output on every rank; fused: params on every rank
  float* param;      // local params   (fused)
  float* mom;        // local momentum / Adam m (fused)
  float* var;        // local Adam v
  const float* hyper;
  long long n;       // elements, multiple of 4
  int opt;           // OPT_*
  int two_shot;
  float scale;       // plain all-reduce: multiply the sum by this (1 or 1/W)
  const float* mc_grads;  // multicast variant: multicast address of the gradient bucket
  float* mc_dst;          //                    multicast address of the destination bucket
};

// Applies the epilogue (scale, or optimizer update) to one reduced 4-vector and stores it.
__device__ __forceinline__ float4 reduce_epilogue(const ReduceArgs& a, long long i, float4 s) {
  float4 out;
  if (a.opt == OPT_NONE) {
    out = make_float4(s.x * a.scale, s.y * a.scale, s.z * a.scale, s.w * a.scale);
  } else {
    float4 p = reinterpret_cast<float4*>(a.param)[i];
    float4 m = reinterpret_cast<float4*>(a.mom)[i];
    if (a.opt == OPT_SGD) {
      sgd_one(s.x, p.x, m.x, a.hyper);
      sgd_one(s.y, p.y, m.y, a.hyper);
      sgd_one(s.z, p.z, m.z, a.hyper);
      sgd_one(s.w, p.w, m.w, a.hyper);
    } else {
      float4 v = reinterpret_cast<float4*>(a.var)[i];
      const int keras = a.opt == OPT_ADAM_KERAS;
      adam_one(s.x, p.x, m.x, v.x, a.hyper, keras);
      adam_one(s.y, p.y, m.y, v.y, a.hyper, keras);
      adam_one(s.z, p.z, m.z, v.z, a.hyper, keras);
      adam_one(s.w, p.w, m.w, v.w, a.hyper, keras);
      reinterpret_cast<float4*>(a.var)[i] = v;
    }
    reinterpret_cast<float4*>(a.mom)[i] = m;
    out = p;
  }
  return out;
}

template <int WT>
__device__ __forceinline__ void reduce_finish(const CommDev& c, const ReduceArgs& a, int W,
                                              long long i, float4 s) {
  const float4 out = reduce_epilogue(a, i, s);
  if (a.two_shot && W > 1) {
#pragma unroll
    for (int r = 0; r < (WT ? WT : kMaxWorld); ++r)
      if (r < W) reinterpret_cast<float4*>(a.dst.p[r])[i] = out;
  } else {
    reinterpret_cast<float4*>(a.dst.p[c.rank])[i] = out;
  }
}

// One kernel for every (variant, optimizer) pair; branches are grid-uniform.  WT is the world
// size as a template constant (0 = generic) so that a thread's peer loads -- kUnroll vectors
// from each of W ranks -- are all in flight together: a peer load over NVLink takes 2-3 us, so
// bytes in flight per SM, not instruction count, set the achieved bandwidth.
template <int WT>
__global__ void __launch_bounds__(512)
reduce_kernel(const CommDev c, const ReduceArgs a, TraceSlot* slot) {
  constexpr int kUnroll = 4;
  trace_enter(slot, TR_REDUCE, a.n * 4);

  const int W = WT ? WT : c.world;
  long long begin4 = 0, end4 = a.n / 4;
  if (a.two_shot && W > 1) {
    const long long per = (end4 + W - 1) / W;
    begin4 = per * c.rank;
    end4 = begin4 + per < end4 ? begin4 + per : end4;
    if (begin4 > end4) begin4 = end4;
  }
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = begin4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < end4;
       i0 += stride * kUnroll) {
    float4 s[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      s[u] = i < end4 ? ld_peer_f4(reinterpret_cast<const float*>(a.grads.p[0]), i)
                      : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int r = 1; r < (WT ? WT : kMaxWorld); ++r) {
      if (r < W) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const long long i = i0 + u * stride;
          if (i < end4)
            s[u] = f4_add(s[u], ld_peer_f4(reinterpret_cast<const float*>(a.grads.p[r]), i));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < end4) reduce_finish<WT>(c, a, W, i, s[u]);
    }
  }
  trace_exit(slot);
}

static void launch_reduce(const CommDev& d, const ReduceArgs& a, int grid, cudaStream_t stream,
                          TraceSlot* slot) {
  switch (d.world) {
    case 1: reduce_kernel<1><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 2: reduce_kernel<2><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 4: reduce_kernel<4><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 8: reduce_kernel<8><<<grid, 512, 0, stream>>>(d, a, slot); break;
    default: reduce_kernel<0><<<grid, 512, 0, stream>>>(d, a, slot); break;
  }
}

// ---- NVLS (multicast) variants ---------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// Rank r owns slice r: one in-switch reduction per 16 bytes in, one replicated store out.  The
// slice's parameters / optimizer state are read and written locally, as in two-shot.  The order
// in which the switch adds the W values is fixed by the hardware, not by rank; results are
// still identical on every rank because every rank receives the same stored value.
//
// kUnroll x 16 bytes of in-switch reductions are in flight per thread before the first is consumed;
// (kUnroll, kThreads) is picked at run time from the measured sweep (DTDL_NVLS_UNROLL /
// DTDL_NVLS_THREADS override): 512 threads x 4 keeps the register count under the 128 that a
// 512-thread CTA allows, 256 x 8 trades occupancy for deeper per-thread queues.
template <int kUnroll, int kThreads>
__global__ void __launch_bounds__(kThreads)
nvls_reduce_kernel(const CommDev c, const ReduceArgs a, TraceSlot* slot) {
  trace_enter(slot, TR_NVLS_REDUCE, a.n * 4);
  const int W = c.world;
  const long long n4 = a.n / 4;
  const long long per = (n4 + W - 1) / W;
  long long begin4 = per * c.rank;
  const long long end4 = begin4 + per < n4 ? begin4 + per : n4;
  if (begin4 > end4) begin4 = end4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = begin4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < end4;
       i0 += stride * kUnroll) {
    float4 s[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < end4) s[u] = multimem_ld_reduce_f4(a.mc_grads + 4 * i);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < end4) multimem_st_f4(a.mc_dst + 4 * i, reduce_epilogue(a, i, s[u]));
    }
  }
  trace_exit(slot);
}

// ---- bf16 gradient transport -------------------------------------------------------------------
// Option of the fused reducer (wrap(..., grad_dtype="bf16")): the bucket's fp32 gradients are cast to
// bf16 into a second symmetric arena (cast_f32_bf16_kernel, on the communication stream), and the
// reduce kernels below move HALF the bytes over NVLink: every 16-byte peer / multicast load carries 8
// gradients, the sum is accumulated in fp32 (multimem: .acc::f32), and the fp32 epilogue -- 1/W,
// weight decay, SGD / Adam, fp32 parameter store to every rank -- is the same as above.  This is the
// "fp32/bf16 cast fused with the collective and the update" of the north star; the reference's DDP
// moves fp32 gradients (pytorch/distributed_data_parallel.py:132), so fp32 stays the default.
__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, long long n4) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = src[i];
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    dst[i] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
  }
}

struct F8s {
  float4 lo, hi;
};
__device__ __forceinline__ F8s bf16x8_to_f32(uint4 u) {
  F8s r;
  const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
  const float2 a = __bfloat1622float2(p[0]), b = __bfloat1622float2(p[1]), c = __bfloat1622float2(p[2]),
               d = __bfloat1622float2(p[3]);
  r.lo = make_float4(a.x, a.y, b.x, b.y);
  r.hi = make_float4(c.x, c.y, d.x, d.y);
  return r;
}
__device__ __forceinline__ uint4 ld_nc_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}

// Two-shot over peer pointers, bf16 gradients (grads.p[r] = rank r's bf16 bucket).  a.n is a multiple of 8.
template <int WT>
__global__ void __launch_bounds__(512)
reduce16_kernel(const CommDev c, const ReduceArgs a, TraceSlot* slot) {
  trace_enter(slot, TR_REDUCE, a.n * 2);
  const int W = WT ? WT : c.world;
  long long begin8 = 0, end8 = a.n / 8;
  if (W > 1) {
    const long long per = (end8 + W - 1) / W;
    begin8 = per * c.rank;
    end8 = begin8 + per < end8 ? begin8 + per : end8;
    if (begin8 > end8) begin8 = end8;
  }
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = begin8 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < end8; i += stride) {
    uint4 raw[WT ? WT : kMaxWorld];
#pragma unroll
    for (int r = 0; r < (WT ? WT : kMaxWorld); ++r)
      if (r < W) raw[r] = ld_nc_u4(reinterpret_cast<const uint4*>(a.grads.p[r]) + i);
    F8s s = bf16x8_to_f32(raw[0]);
#pragma unroll
    for (int r = 1; r < (WT ? WT : kMaxWorld); ++r) {
      if (r < W) {
        const F8s t = bf16x8_to_f32(raw[r]);
        s.lo = f4_add(s.lo, t.lo);
        s.hi = f4_add(s.hi, t.hi);
      }
    }
    const float4 o0 = reduce_epilogue(a, 2 * i, s.lo), o1 = reduce_epilogue(a, 2 * i + 1, s.hi);
#pragma unroll
    for (int r = 0; r < (WT ? WT : kMaxWorld); ++r) {
      if (r < W) {
        reinterpret_cast<float4*>(a.dst.p[r])[2 * i] = o0;
        reinterpret_cast<float4*>(a.dst.p[r])[2 * i + 1] = o1;
      }
    }
  }
  trace_exit(slot);
}

// Multicast form: one in-switch bf16 reduction (fp32 accumulate) per 16 bytes = 8 gradients.
__global__ void __launch_bounds__(512)
nvls_reduce16_kernel(const CommDev c, const ReduceArgs a, TraceSlot* slot) {
  constexpr int kUnroll = 4;
  trace_enter(slot, TR_NVLS_REDUCE, a.n * 2);
  const int W = c.world;
  const long long n8 = a.n / 8;
  const long long per = (n8 + W - 1) / W;
  long long begin8 = per * c.rank;
  const long long end8 = begin8 + per < n8 ? begin8 + per : n8;
  if (begin8 > end8) begin8 = end8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const char* mc = reinterpret_cast<const char*>(a.mc_grads);
  for (long long i0 = begin8 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < end8;
       i0 += stride * kUnroll) {
    uint4 raw[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < end8) raw[u] = multimem_ld_reduce_bf16x8(mc + 16 * i);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < end8) {
        const F8s s = bf16x8_to_f32(raw[u]);
        multimem_st_f4(a.mc_dst + 8 * i, reduce_epilogue(a, 2 * i, s.lo));
        multimem_st_f4(a.mc_dst + 8 * i + 4, reduce_epilogue(a, 2 * i + 1, s.hi));
      }
    }
  }
  trace_exit(slot);
}

// The root pushes its copy once; the switch delivers it to every rank (the root included).
__global__ void __launch_bounds__(512)
nvls_broadcast_kernel(const CommDev c, const float* local, float* mc, long long n16, int root,
                      TraceSlot* slot) {
  trace_enter(slot, TR_BROADCAST, n16 * 16);
  if (c.rank == root) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride)
      multimem_st_f4(mc + 4 * i, reinterpret_cast<const float4*>(local)[i]);
  }
  trace_exit(slot);
}

// Every non-root rank pulls `n16` 16-byte words from the root's copy.
__global__ void __launch_bounds__(512)
broadcast_kernel(const CommDev c, const PeerPtrs buf, long long n16, int root, TraceSlot* slot) {
  trace_enter(slot, TR_BROADCAST, n16 * 16);
  if (c.rank != root) {
    const uint4* src = reinterpret_cast<const uint4*>(buf.p[root]);
    uint4* dst = reinterpret_cast<uint4*>(buf.p[c.rank]);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride)
      dst[i] = src[i];
  }
  trace_exit(slot);
}

// all-gather: rank r pushes its slice of `buf` into every peer's copy.
__global__ void __launch_bounds__(512)
allgather_kernel(const CommDev c, const PeerPtrs buf, long long n16, TraceSlot* slot) {
  trace_enter(slot, TR_ALLGATHER, n16 * 16);
  const int W = c.world;
  const long long per = (n16 + W - 1) / W;
  const long long b = per * c.rank;
  const long long en = b + per < n16 ? b + per : n16;
  const uint4* src = reinterpret_cast<const uint4*>(buf.p[c.rank]);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = b + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < en; i += stride) {
    const uint4 v = src[i];
    for (int r = 0; r < W; ++r)
      if (r != c.rank) reinterpret_cast<uint4*>(buf.p[r])[i] = v;
  }
  trace_exit(slot);
}

// Single-rank optimizer step on a flat fp32 arena (W = 1, no flags touched).
__global__ void __launch_bounds__(512)
local_update_kernel(float* __restrict__ param, const float* __restrict__ grad,
                    float* __restrict__ mom, float* __restrict__ var,
                    const float* __restrict__ hyper, long long n4, int opt) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 g = reinterpret_cast<const float4*>(grad)[i];
    float4 p = reinterpret_cast<float4*>(param)[i];
    float4 m = reinterpret_cast<float4*>(mom)[i];
    if (opt == OPT_SGD) {
      sgd_one(g.x, p.x, m.x, hyper);
      sgd_one(g.y, p.y, m.y, hyper);
      sgd_one(g.z, p.z, m.z, hyper);
      sgd_one(g.w, p.w, m.w, hyper);
    } else {
      float4 v = reinterpret_cast<float4*>(var)[i];
      const int keras = opt == OPT_ADAM_KERAS;
      adam_one(g.x, p.x, m.x, v.x, hyper, keras);
      adam_one(g.y, p.y, m.y, v.y, hyper, keras);
      adam_one(g.z, p.z, m.z, v.z, hyper, keras);
      adam_one(g.w, p.w, m.w, v.w, hyper, keras);
      reinterpret_cast<float4*>(var)[i] = v;
    }
    reinterpret_cast<float4*>(mom)[i] = m;
    reinterpret_cast<float4*>(param)[i] = p;
  }
}

static void launch_nvls_reduce(const CommDev& d, const ReduceArgs& a, int grid, cudaStream_t stream,
                               TraceSlot* slot) {
  static const int unroll = getenv("DTDL_NVLS_UNROLL") ? atoi(getenv("DTDL_NVLS_UNROLL")) : 4;
  static const int threads = getenv("DTDL_NVLS_THREADS") ? atoi(getenv("DTDL_NVLS_THREADS")) : 512;
  if (threads == 256) {
    if (unroll >= 8)
      nvls_reduce_kernel<8, 256><<<grid * 2, 256, 0, stream>>>(d, a, slot);
    else
      nvls_reduce_kernel<4, 256><<<grid * 2, 256, 0, stream>>>(d, a, slot);
  } else {
    if (unroll >= 8)
      nvls_reduce_kernel<8, 512><<<grid, 512, 0, stream>>>(d, a, slot);
    else
      nvls_reduce_kernel<4, 512><<<grid, 512, 0, stream>>>(d, a, slot);
  }
}

static int comm_grid(long long n4, int world, int two_shot, int max_blocks) {
  long long per_rank = (two_shot && world > 1) ? (n4 + world - 1) / world : n4;
  long long g = (per_rank + 512 * 8 - 1) / (512 * 8);
  if (g > max_blocks) g = max_blocks;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace dtdl

using namespace dtdl;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int dtdl_comm_create(int rank, int world, void** handle, void** flags_ptr) {
  DTDL_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "bad rank/world");
  CommHost* h = new CommHost();
  memset(h, 0, sizeof(*h));
  h->d.rank = rank;
  h->d.world = world;
  DTDL_CHECK_CUDA(cudaGetDevice(&h->device));
  uint32_t* mem = nullptr;
  // flags + the two epoch words in one allocation: [2*kMaxWorld flags][START epoch][DONE epoch]
  DTDL_CHECK_CUDA(cudaMalloc(&mem, (2 * kMaxWorld + 2) * sizeof(uint32_t)));
  DTDL_CHECK_CUDA(cudaMemset(mem, 0, (2 * kMaxWorld + 2) * sizeof(uint32_t)));
  h->d.flags_local = mem;
  h->d.epochs = mem + 2 * kMaxWorld;
  h->d.flags_peer[rank] = mem;
  *handle = h;
  *flags_ptr = mem;
  return 0;
}

extern "C" int dtdl_comm_set_peer_flags(void* handle, int peer, void* flags_ptr) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(peer >= 0 && peer < h->d.world, "bad peer");
  h->d.flags_peer[peer] = reinterpret_cast<uint32_t*>(flags_ptr);
  return 0;
}

extern "C" int dtdl_comm_destroy(void* handle) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  if (h) {
    cudaFree(h->d.flags_local);
    delete h;
  }
  return 0;
}

extern "C" int dtdl_symm_alloc(long long bytes, void** ptr) {
  DTDL_REQUIRE(bytes > 0, "bad size");
  const long long padded = (bytes + 255) / 256 * 256;
  DTDL_CHECK_CUDA(cudaMalloc(ptr, (size_t)padded));
  DTDL_CHECK_CUDA(cudaMemset(*ptr, 0, (size_t)padded));
  return 0;
}
extern "C" int dtdl_symm_free(void* ptr) {
  DTDL_CHECK_CUDA(cudaFree(ptr));
  return 0;
}
extern "C" int dtdl_ipc_get_handle(void* ptr, void* out64) {
  cudaIpcMemHandle_t h;
  DTDL_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is expected to be 64 bytes");
  memcpy(out64, &h, 64);
  return 0;
}
extern "C" int dtdl_ipc_open_handle(const void* in64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, in64, 64);
  DTDL_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int dtdl_ipc_close_handle(void* ptr) {
  DTDL_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
extern "C" int dtdl_enable_peer_access(int dev, int peer) {
  int can = 0;
  DTDL_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer));
  DTDL_REQUIRE(can, "devices cannot access each other's memory");
  int cur = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&cur));
  DTDL_CHECK_CUDA(cudaSetDevice(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  cudaSetDevice(cur);
  DTDL_CHECK_CUDA(e);
  return 0;
}

static TraceSlot* next_slot(CommHost* h) {
  if (h->trace == nullptr || h->trace_next >= h->trace_cap) return nullptr;
  return h->trace + h->trace_next++;
}

// START exchange in front of a data kernel / DONE exchange behind it (see CommDev).
static int sync_row(CommHost* h, int row, cudaStream_t stream) {
  if (h->d.world == 1) return 0;
  sync_kernel<<<1, 32, 0, stream>>>(h->d, row, next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
static int sync_before(CommHost* h, cudaStream_t stream) { return sync_row(h, 0, stream); }
static int sync_after(CommHost* h, cudaStream_t stream) {
  return h->defer_done ? 0 : sync_row(h, 1, stream);
}

// defer = 1: the collectives that follow skip their trailing DONE exchange; the caller must issue
// dtdl_comm_barrier on the same stream before any rank touches the buffers again.  defer = 0 (the
// default): every collective is self-contained.
extern "C" int dtdl_comm_defer_done(void* handle, int defer) {
  reinterpret_cast<CommHost*>(handle)->defer_done = defer ? 1 : 0;
  return 0;
}

// Device-side timeline: `base` = device buffer of `capacity` slots of 4 x u64 {t0 ns, t1 ns, kind,
// bytes} (zero it first); every sync / data kernel launched through this handle from now on records
// into the next slot.  base = NULL switches tracing off.  Kinds: 1 START wait, 2 DONE wait, 3 reduce,
// 4 NVLS reduce, 5 broadcast, 6 all-gather, 7 user mark.
extern "C" int dtdl_comm_trace(void* handle, void* base, int capacity) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  h->trace = reinterpret_cast<TraceSlot*>(base);
  h->trace_cap = base ? capacity : 0;
  h->trace_next = 0;
  return 0;
}
extern "C" int dtdl_comm_trace_mark(void* handle, int kind, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  TraceSlot* s = next_slot(h);
  if (s == nullptr) return 0;
  trace_mark_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(s, kind);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int dtdl_comm_trace_count(void* handle) { return reinterpret_cast<CommHost*>(handle)->trace_next; }

static int fill_peers(PeerPtrs* pp, void* const* ptrs, int world, long long offset_bytes) {
  memset(pp, 0, sizeof(*pp));
  for (int r = 0; r < world; ++r) {
    if (ptrs[r] == nullptr) return 1;
    pp->p[r] = reinterpret_cast<char*>(ptrs[r]) + offset_bytes;
  }
  return 0;
}

extern "C" int dtdl_comm_barrier(void* handle, void* stream) {
  // one DONE-row exchange: every rank has finished everything it enqueued before this call
  return sync_row(reinterpret_cast<CommHost*>(handle), 1, STREAM(stream));
}

// Plain all-reduce (sum * scale) of n fp32 elements at byte offset `offset` of
// the symmetric buffer whose per-rank base pointers are grads[0..W).  The
// result lands in dst[rank] (one-shot) or in every dst[r] (two-shot).
extern "C" int dtdl_comm_allreduce(void* handle, void* const* grads, void* const* dst,
                                   long long offset_bytes, long long n, float scale, int two_shot,
                                   int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 4 == 0 && offset_bytes % 16 == 0, "buckets must be 16-byte granular");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  DTDL_REQUIRE(!fill_peers(&a.grads, grads, h->d.world, offset_bytes), "missing peer pointer");
  DTDL_REQUIRE(!fill_peers(&a.dst, dst, h->d.world, offset_bytes), "missing peer pointer");
  a.n = n;
  a.opt = OPT_NONE;
  // In place (grads == dst) the one-shot schedule would let a rank overwrite its bucket while peers
  // are still reading it: in-place calls always run the two-shot schedule, where rank r reduces and
  // rewrites only the slice it owns.
  if (h->d.world > 1 && grads[h->d.rank] == dst[h->d.rank]) two_shot = 1;
  a.two_shot = two_shot;
  a.scale = scale;
  const int grid = comm_grid(n / 4, h->d.world, two_shot, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  launch_reduce(h->d, a, grid, STREAM(stream), next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

// Fused gradient all-reduce + 1/W + weight decay + SGD-momentum / Adam update.
// params[r] are the per-rank base pointers of the symmetric parameter arena,
// mom / var the local optimizer-state arenas, all addressed at `offset_bytes`.
extern "C" int dtdl_comm_reduce_update(void* handle, void* const* grads, void* const* params,
                                       float* mom, float* var, const float* hyper,
                                       long long offset_bytes, long long n, int opt, int two_shot,
                                       int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 4 == 0 && offset_bytes % 16 == 0, "buckets must be 16-byte granular");
  DTDL_REQUIRE(opt == OPT_SGD || opt == OPT_ADAM || opt == OPT_ADAM_KERAS, "unknown optimizer");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  DTDL_REQUIRE(!fill_peers(&a.grads, grads, h->d.world, offset_bytes), "missing peer pointer");
  DTDL_REQUIRE(!fill_peers(&a.dst, params, h->d.world, offset_bytes), "missing peer pointer");
  a.param = reinterpret_cast<float*>(a.dst.p[h->d.rank]);
  a.mom = reinterpret_cast<float*>(reinterpret_cast<char*>(mom) + offset_bytes);
  a.var = var ? reinterpret_cast<float*>(reinterpret_cast<char*>(var) + offset_bytes) : nullptr;
  DTDL_REQUIRE(opt == OPT_SGD || a.var != nullptr, "Adam needs the second-moment arena");
  a.hyper = hyper;
  a.n = n;
  a.opt = opt;
  a.two_shot = two_shot;
  a.scale = 1.f;
  const int grid = comm_grid(n / 4, h->d.world, two_shot, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  launch_reduce(h->d, a, grid, STREAM(stream), next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_comm_broadcast(void* handle, void* const* bufs, long long offset_bytes,
                                   long long bytes, int root, int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(bytes % 16 == 0 && offset_bytes % 16 == 0, "buffers must be 16-byte granular");
  PeerPtrs pp;
  DTDL_REQUIRE(!fill_peers(&pp, bufs, h->d.world, offset_bytes), "missing peer pointer");
  const long long n16 = bytes / 16;
  long long g = (n16 + 2047) / 2048;
  const int cap = max_blocks > 0 ? max_blocks : 64;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  if (sync_before(h, STREAM(stream))) return 1;
  broadcast_kernel<<<(int)g, 512, 0, STREAM(stream)>>>(h->d, pp, n16, root, next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_comm_allgather(void* handle, void* const* bufs, long long offset_bytes,
                                   long long bytes, int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(bytes % 16 == 0 && offset_bytes % 16 == 0, "buffers must be 16-byte granular");
  PeerPtrs pp;
  DTDL_REQUIRE(!fill_peers(&pp, bufs, h->d.world, offset_bytes), "missing peer pointer");
  const long long n16 = bytes / 16;
  long long g = (n16 / h->d.world + 2047) / 2048;
  const int cap = max_blocks > 0 ? max_blocks : 64;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  if (sync_before(h, STREAM(stream))) return 1;
  allgather_kernel<<<(int)g, 512, 0, STREAM(stream)>>>(h->d, pp, n16, next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_local_update(float* param, const float* grad, float* mom, float* var,
                                 const float* hyper, long long n, int opt, void* stream) {
  DTDL_REQUIRE(n % 4 == 0, "arena must be a multiple of 4 elements");
  DTDL_REQUIRE(opt == OPT_SGD || var != nullptr, "Adam needs the second-moment arena");
  long long g = (n / 4 + 2047) / 2048;
  if (g > 148 * 4) g = 148 * 4;
  if (g < 1) g = 1;
  local_update_kernel<<<(int)g, 512, 0, STREAM(stream)>>>(param, grad, mom, var, hyper, n / 4, opt);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- NVLS entry points: `mc_*` are multicast addresses of buffers from dtdl_vmm_* / dtdl_mc_* ----
extern "C" int dtdl_comm_nvls_allreduce(void* handle, void* mc_src, void* mc_dst, long long offset_bytes,
                                        long long n, float scale, int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 4 == 0 && offset_bytes % 16 == 0, "buckets must be 16-byte granular");
  DTDL_REQUIRE(mc_src != nullptr && mc_dst != nullptr, "buffer has no multicast mapping");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  a.mc_grads = reinterpret_cast<const float*>(reinterpret_cast<char*>(mc_src) + offset_bytes);
  a.mc_dst = reinterpret_cast<float*>(reinterpret_cast<char*>(mc_dst) + offset_bytes);
  a.n = n;
  a.opt = OPT_NONE;
  a.two_shot = 1;
  a.scale = scale;
  const int grid = comm_grid(n / 4, h->d.world, 1, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  launch_nvls_reduce(h->d, a, grid, STREAM(stream), next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_comm_nvls_reduce_update(void* handle, void* mc_grads, void* mc_params, float* param,
                                            float* mom, float* var, const float* hyper,
                                            long long offset_bytes, long long n, int opt, int max_blocks,
                                            void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 4 == 0 && offset_bytes % 16 == 0, "buckets must be 16-byte granular");
  DTDL_REQUIRE(opt == OPT_SGD || opt == OPT_ADAM || opt == OPT_ADAM_KERAS, "unknown optimizer");
  DTDL_REQUIRE(mc_grads != nullptr && mc_params != nullptr, "buffer has no multicast mapping");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  a.mc_grads = reinterpret_cast<const float*>(reinterpret_cast<char*>(mc_grads) + offset_bytes);
  a.mc_dst = reinterpret_cast<float*>(reinterpret_cast<char*>(mc_params) + offset_bytes);
  a.param = reinterpret_cast<float*>(reinterpret_cast<char*>(param) + offset_bytes);
  a.mom = reinterpret_cast<float*>(reinterpret_cast<char*>(mom) + offset_bytes);
  a.var = var ? reinterpret_cast<float*>(reinterpret_cast<char*>(var) + offset_bytes) : nullptr;
  DTDL_REQUIRE(opt == OPT_SGD || a.var != nullptr, "Adam needs the second-moment arena");
  a.hyper = hyper;
  a.n = n;
  a.opt = opt;
  a.two_shot = 1;
  a.scale = 1.f;
  const int grid = comm_grid(n / 4, h->d.world, 1, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  launch_nvls_reduce(h->d, a, grid, STREAM(stream), next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_comm_nvls_broadcast(void* handle, void* local, void* mc, long long offset_bytes,
                                        long long bytes, int root, int max_blocks, void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(bytes % 16 == 0 && offset_bytes % 16 == 0, "buffers must be 16-byte granular");
  DTDL_REQUIRE(local != nullptr && mc != nullptr, "buffer has no multicast mapping");
  const long long n16 = bytes / 16;
  long long g = (n16 + 2047) / 2048;
  const int cap = max_blocks > 0 ? max_blocks : 64;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  if (sync_before(h, STREAM(stream))) return 1;
  nvls_broadcast_kernel<<<(int)g, 512, 0, STREAM(stream)>>>(
      h->d, reinterpret_cast<const float*>(reinterpret_cast<char*>(local) + offset_bytes),
      reinterpret_cast<float*>(reinterpret_cast<char*>(mc) + offset_bytes), n16, root, next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

// ---- bf16 gradient transport: entry points ------------------------------------------------------
extern "C" int dtdl_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  DTDL_REQUIRE(n % 4 == 0, "element count must be a multiple of 4");
  long long g = (n / 4 + 1023) / 1024;
  if (g > 148 * 4) g = 148 * 4;
  if (g < 1) g = 1;
  cast_f32_bf16_kernel<<<(int)g, 256, 0, STREAM(stream)>>>(reinterpret_cast<const float4*>(src),
                                                           reinterpret_cast<uint2*>(dst), n / 4);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static void launch_reduce16(const CommDev& d, const ReduceArgs& a, int grid, cudaStream_t stream,
                            TraceSlot* slot) {
  switch (d.world) {
    case 1: reduce16_kernel<1><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 2: reduce16_kernel<2><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 4: reduce16_kernel<4><<<grid, 512, 0, stream>>>(d, a, slot); break;
    case 8: reduce16_kernel<8><<<grid, 512, 0, stream>>>(d, a, slot); break;
    default: reduce16_kernel<0><<<grid, 512, 0, stream>>>(d, a, slot); break;
  }
}

// dtdl_comm_reduce_update with bf16 gradient buckets: grads16[r] = rank r's bf16 gradient arena;
// `offset_bytes` is the bucket's offset in the fp32 arenas (the bf16 arena is addressed at half of it).
extern "C" int dtdl_comm_reduce_update16(void* handle, void* const* grads16, void* const* params,
                                         float* mom, float* var, const float* hyper,
                                         long long offset_bytes, long long n, int opt, int max_blocks,
                                         void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 8 == 0 && offset_bytes % 32 == 0, "bf16 buckets must be 8-element granular");
  DTDL_REQUIRE(opt == OPT_SGD || opt == OPT_ADAM || opt == OPT_ADAM_KERAS, "unknown optimizer");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  DTDL_REQUIRE(!fill_peers(&a.grads, grads16, h->d.world, offset_bytes / 2), "missing peer pointer");
  DTDL_REQUIRE(!fill_peers(&a.dst, params, h->d.world, offset_bytes), "missing peer pointer");
  a.param = reinterpret_cast<float*>(a.dst.p[h->d.rank]);
  a.mom = reinterpret_cast<float*>(reinterpret_cast<char*>(mom) + offset_bytes);
  a.var = var ? reinterpret_cast<float*>(reinterpret_cast<char*>(var) + offset_bytes) : nullptr;
  DTDL_REQUIRE(opt == OPT_SGD || a.var != nullptr, "Adam needs the second-moment arena");
  a.hyper = hyper;
  a.n = n;
  a.opt = opt;
  a.two_shot = 1;
  a.scale = 1.f;
  const int grid = comm_grid(n / 8, h->d.world, 1, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  launch_reduce16(h->d, a, grid, STREAM(stream), next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}

extern "C" int dtdl_comm_nvls_reduce_update16(void* handle, void* mc_grads16, void* mc_params, float* param,
                                              float* mom, float* var, const float* hyper,
                                              long long offset_bytes, long long n, int opt, int max_blocks,
                                              void* stream) {
  CommHost* h = reinterpret_cast<CommHost*>(handle);
  DTDL_REQUIRE(n % 8 == 0 && offset_bytes % 32 == 0, "bf16 buckets must be 8-element granular");
  DTDL_REQUIRE(opt == OPT_SGD || opt == OPT_ADAM || opt == OPT_ADAM_KERAS, "unknown optimizer");
  DTDL_REQUIRE(mc_grads16 != nullptr && mc_params != nullptr, "buffer has no multicast mapping");
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  a.mc_grads = reinterpret_cast<const float*>(reinterpret_cast<char*>(mc_grads16) + offset_bytes / 2);
  a.mc_dst = reinterpret_cast<float*>(reinterpret_cast<char*>(mc_params) + offset_bytes);
  a.param = reinterpret_cast<float*>(reinterpret_cast<char*>(param) + offset_bytes);
  a.mom = reinterpret_cast<float*>(reinterpret_cast<char*>(mom) + offset_bytes);
  a.var = var ? reinterpret_cast<float*>(reinterpret_cast<char*>(var) + offset_bytes) : nullptr;
  DTDL_REQUIRE(opt == OPT_SGD || a.var != nullptr, "Adam needs the second-moment arena");
  a.hyper = hyper;
  a.n = n;
  a.opt = opt;
  a.two_shot = 1;
  a.scale = 1.f;
  const int grid = comm_grid(n / 8, h->d.world, 1, max_blocks > 0 ? max_blocks : 64);
  if (sync_before(h, STREAM(stream))) return 1;
  nvls_reduce16_kernel<<<grid, 512, 0, STREAM(stream)>>>(h->d, a, next_slot(h));
  DTDL_CHECK_CUDA(cudaGetLastError());
  return sync_after(h, STREAM(stream));
}
