// fp32 master weights (torch OIHW / [out,in]) -> bf16 (or, in fp32 activation mode, fp32/TF32) operand packs for the
// tcgen05 kernels, for a whole model in ONE launch (a device-side table lists
// the layers).  Runs once per optimizer step, right after the update.  (No counterpart in the
// reference: cuDNN reads torch's OIHW fp32 weights directly, pytorch/model.py:28-32.  The packs
// are what lets the TMA-fed tcgen05 kernels take K-major bf16 operands without a transform.)
//
//   fprop pack  wf[co][t][ci]   co < Cout_p16, ci < Cin_p   : w[co][ci][t]
//   dgrad pack  wd[ci][t'][co]  ci < Cin_p16, co < Cout_p   : w[co][ci][taps-1-t']
//
// Rows / columns beyond the real channel counts are written as zeros; the
// kernels rely on that to keep the activation pad channels at zero.
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../../include/dtdl_b200.h"

namespace dtdl {

struct PackEntry {
  const float* w;
  void* wf;
  void* wd;  // may be null (first layer needs no data gradient)
  int act;   // element type of the packs: 0 = bf16, 1 = fp32 (read by the tensor core as TF32)
  int Cout, Cin, taps;
  int Cin_p, Cout_p, Cin_p16, Cout_p16;
  long long f_elems, d_elems;  // elements in each pack
  long long first_block;       // prefix sum of blocks
};

__device__ __forceinline__ void pack_store(void* base, long long idx, float v, int act) {
  if (act)
    reinterpret_cast<float*>(base)[idx] = v;
  else
    reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16(v);
}

static constexpr int kPackBlock = 256;
static constexpr int kPackPerBlock = 256 * 8;
// Tiled path (taps <= kTileMaxTaps): one block stages a 32 (co) x 32 (ci) x taps cube of the fp32
// weights in shared memory with row-contiguous loads, then writes its piece of both packs in
// 64-byte runs.  The element-wise path below reads the source with a stride of `taps` floats for
// the fprop pack and of a whole filter row for the dgrad pack and divides per element (200 us for
// PyramidNet's 28.5 M weights; the tiled path: 112 us).
static constexpr int kTile = 32;
static constexpr int kTileMaxTaps = 9;

// TAPS > 0: compile-time filter size (index arithmetic becomes multiplies and shifts -- with run-time
// divisors the kernel is bound by integer division, not by memory); TAPS = 0: any size.
template <int TAPS>
__device__ __forceinline__ void pack_tile(const PackEntry& e, long long tile, float* s) {
  const int taps = TAPS ? TAPS : e.taps;
  const int tiles_ci = (e.Cin_p16 + kTile - 1) / kTile;
  const int co0 = (int)(tile / tiles_ci) * kTile, ci0 = (int)(tile % tiles_ci) * kTile;
  const int row = kTile * taps;   // floats of one co row of the cube
  const int rstride = row + 1;    // +1: the dgrad pass walks down a column
  // fixed roles, no per-element index division: a warp owns rows (or columns) w, w+8, ...
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int kvalid = (e.Cin - ci0 < kTile ? (e.Cin - ci0 > 0 ? e.Cin - ci0 : 0) : kTile) * taps;
  for (int r = w; r < kTile; r += kPackBlock / 32) {
    const int co = co0 + r;
    const float* src = e.w + ((long long)co * e.Cin + ci0) * taps;
#pragma unroll 3
    for (int k = lane; k < row; k += 32)
      s[r * rstride + k] = (co < e.Cout && k < kvalid) ? src[k] : 0.f;
  }
  __syncthreads();
  // fprop pack: wf[co][t][ci], lanes run along ci
  {
    const int ci = ci0 + lane;
    if (ci < e.Cin_p) {
      for (int r = w; r < kTile; r += kPackBlock / 32) {
        const int co = co0 + r;
        if (co >= e.Cout_p16) break;
        const long long dst = (long long)co * taps * e.Cin_p + ci;
        const float* sp = s + r * rstride + lane * taps;
#pragma unroll
        for (int t = 0; t < (TAPS ? TAPS : 1); ++t) pack_store(e.wf, dst + (long long)t * e.Cin_p, sp[t], e.act);
        if (!TAPS)
          for (int t = 1; t < taps; ++t) pack_store(e.wf, dst + (long long)t * e.Cin_p, sp[t], e.act);
      }
    }
  }
  // dgrad pack: wd[ci][t'][co] = w[co][ci][taps-1-t'], lanes run along co
  if (e.wd != nullptr) {
    const int co = co0 + lane;
    if (co < e.Cout_p) {
      for (int cl = w; cl < kTile; cl += kPackBlock / 32) {
        const int ci = ci0 + cl;
        if (ci >= e.Cin_p16) break;
        const long long dst = (long long)ci * taps * e.Cout_p + co;
        const float* sp = s + lane * rstride + cl * taps + (taps - 1);
#pragma unroll
        for (int t = 0; t < (TAPS ? TAPS : 1); ++t) pack_store(e.wd, dst + (long long)t * e.Cout_p, sp[-t], e.act);
        if (!TAPS)
          for (int t = 1; t < taps; ++t) pack_store(e.wd, dst + (long long)t * e.Cout_p, sp[-t], e.act);
      }
    }
  }
}

static constexpr int kMaxSmemEntries = 1024;

// Persistent blocks walk the global list of work items (tiles, or 2048-element runs for large
// filters).  Which layer an item belongs to is a binary search over the layers' first-item prefix
// sums, staged once per block in shared memory: searching the table in global memory cost eight
// dependent L2 round trips per item and was, at 28 K items, what the whole kernel waited for.
__global__ void __launch_bounds__(kPackBlock)
pack_weights_kernel(const PackEntry* __restrict__ table, int n_entries, long long total_blocks) {
  __shared__ float s_tile[kTile * (kTile * kTileMaxTaps + 1)];
  __shared__ long long s_first[kMaxSmemEntries];
  const bool staged = n_entries <= kMaxSmemEntries;
  if (staged)
    for (int i = threadIdx.x; i < n_entries; i += kPackBlock) s_first[i] = table[i].first_block;
  __syncthreads();
  for (long long item = blockIdx.x; item < total_blocks; item += gridDim.x) {
    int lo = 0, hi = n_entries - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      const long long fb = staged ? s_first[mid] : table[mid].first_block;
      if (fb <= item)
        lo = mid;
      else
        hi = mid - 1;
    }
    const PackEntry e = table[lo];
    if (e.taps <= kTileMaxTaps) {
      const long long tile = item - e.first_block;
      if (e.taps == 9)
        pack_tile<9>(e, tile, s_tile);
      else if (e.taps == 1)
        pack_tile<1>(e, tile, s_tile);
      else
        pack_tile<0>(e, tile, s_tile);
      __syncthreads();  // the next item overwrites the tile
      continue;
    }
    const long long base = (item - e.first_block) * kPackPerBlock;
    const long long total = e.f_elems + (e.wd ? e.d_elems : 0);
    for (int k = 0; k < 8; ++k) {
      const long long i = base + k * kPackBlock + threadIdx.x;
      if (i >= total) break;
      if (i < e.f_elems) {
        const int ci = (int)(i % e.Cin_p);
        const long long r = i / e.Cin_p;
        const int t = (int)(r % e.taps);
        const int co = (int)(r / e.taps);
        float v = 0.f;
        if (co < e.Cout && ci < e.Cin) v = e.w[((long long)co * e.Cin + ci) * e.taps + t];
        pack_store(e.wf, i, v, e.act);
      } else {
        const long long j = i - e.f_elems;
        const int co = (int)(j % e.Cout_p);
        const long long r = j / e.Cout_p;
        const int t = (int)(r % e.taps);
        const int ci = (int)(r / e.taps);
        float v = 0.f;
        if (co < e.Cout && ci < e.Cin)
          v = e.w[((long long)co * e.Cin + ci) * e.taps + (e.taps - 1 - t)];
        pack_store(e.wd, j, v, e.act);
      }
    }
  }
}

}  // namespace dtdl

using namespace dtdl;

extern "C" int dtdl_pack_entry_size() { return (int)sizeof(PackEntry); }

// Fills one host-side table entry; returns the number of blocks it needs.
extern "C" long long dtdl_pack_entry_fill(void* entry, const float* w, void* wf, void* wd, int Cout,
                                          int Cin, int taps, int Cin_p, int Cout_p,
                                          long long first_block, int act) {
  PackEntry* e = reinterpret_cast<PackEntry*>(entry);
  e->w = w;
  e->wf = wf;
  e->wd = wd;
  e->act = act ? 1 : 0;
  e->Cout = Cout;
  e->Cin = Cin;
  e->taps = taps;
  e->Cin_p = Cin_p;
  e->Cout_p = Cout_p;
  e->Cin_p16 = (Cin_p + 15) / 16 * 16;
  e->Cout_p16 = (Cout_p + 15) / 16 * 16;
  e->f_elems = (long long)e->Cout_p16 * taps * Cin_p;
  e->d_elems = (long long)e->Cin_p16 * taps * Cout_p;
  e->first_block = first_block;
  if (taps <= kTileMaxTaps)
    return (long long)((e->Cout_p16 + kTile - 1) / kTile) * ((e->Cin_p16 + kTile - 1) / kTile);
  const long long total = e->f_elems + (wd ? e->d_elems : 0);
  return (total + kPackPerBlock - 1) / kPackPerBlock;
}

extern "C" int dtdl_pack_weights(const void* device_table, int n_entries, long long total_blocks,
                                 void* stream) {
  DTDL_REQUIRE(n_entries > 0 && total_blocks > 0 && total_blocks < (1ll << 31), "bad pack table");
  const long long resident = (long long)sm_count() * 5;  // 45 KB of shared memory per block
  const unsigned grid = (unsigned)(total_blocks < resident ? total_blocks : resident);
  pack_weights_kernel<<<grid, kPackBlock, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const PackEntry*>(device_table), n_entries, total_blocks);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
