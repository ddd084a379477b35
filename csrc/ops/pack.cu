// fp32 master weights (torch OIHW / [out,in]) -> bf16 operand packs for the
// tcgen05 kernels, for a whole model in ONE launch (a device-side table lists
// the layers).  Runs once per optimizer step, right after the update.
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
  __nv_bfloat16* wf;
  __nv_bfloat16* wd;  // may be null (first layer needs no data gradient)
  int Cout, Cin, taps;
  int Cin_p, Cout_p, Cin_p16, Cout_p16;
  long long f_elems, d_elems;  // elements in each pack
  long long first_block;       // prefix sum of blocks
};

static constexpr int kPackBlock = 256;
static constexpr int kPackPerBlock = 256 * 8;

__global__ void pack_weights_kernel(const PackEntry* __restrict__ table, int n_entries) {
  // locate the entry owning this block (tables are small: linear scan from a guess)
  int lo = 0, hi = n_entries - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (table[mid].first_block <= (long long)blockIdx.x)
      lo = mid;
    else
      hi = mid - 1;
  }
  const PackEntry e = table[lo];
  const long long base = ((long long)blockIdx.x - e.first_block) * kPackPerBlock;
  const long long total = e.f_elems + (e.wd ? e.d_elems : 0);
  for (int k = 0; k < 8; ++k) {
    const long long i = base + k * kPackBlock + threadIdx.x;
    if (i >= total) return;
    if (i < e.f_elems) {
      const int ci = (int)(i % e.Cin_p);
      const long long r = i / e.Cin_p;
      const int t = (int)(r % e.taps);
      const int co = (int)(r / e.taps);
      float v = 0.f;
      if (co < e.Cout && ci < e.Cin) v = e.w[((long long)co * e.Cin + ci) * e.taps + t];
      e.wf[i] = __float2bfloat16(v);
    } else {
      const long long j = i - e.f_elems;
      const int co = (int)(j % e.Cout_p);
      const long long r = j / e.Cout_p;
      const int t = (int)(r % e.taps);
      const int ci = (int)(r / e.taps);
      float v = 0.f;
      if (co < e.Cout && ci < e.Cin)
        v = e.w[((long long)co * e.Cin + ci) * e.taps + (e.taps - 1 - t)];
      e.wd[j] = __float2bfloat16(v);
    }
  }
}

}  // namespace dtdl

using namespace dtdl;

extern "C" int dtdl_pack_entry_size() { return (int)sizeof(PackEntry); }

// Fills one host-side table entry; returns the number of blocks it needs.
extern "C" long long dtdl_pack_entry_fill(void* entry, const float* w, void* wf, void* wd, int Cout,
                                          int Cin, int taps, int Cin_p, int Cout_p,
                                          long long first_block) {
  PackEntry* e = reinterpret_cast<PackEntry*>(entry);
  e->w = w;
  e->wf = reinterpret_cast<__nv_bfloat16*>(wf);
  e->wd = reinterpret_cast<__nv_bfloat16*>(wd);
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
  const long long total = e->f_elems + (wd ? e->d_elems : 0);
  return (total + kPackPerBlock - 1) / kPackPerBlock;
}

extern "C" int dtdl_pack_weights(const void* device_table, int n_entries, long long total_blocks,
                                 void* stream) {
  DTDL_REQUIRE(n_entries > 0 && total_blocks > 0 && total_blocks < (1ll << 31), "bad pack table");
  pack_weights_kernel<<<(unsigned)total_blocks, kPackBlock, 0,
                        reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const PackEntry*>(device_table), n_entries);
  DTDL_CHECK_CUDA(cudaGetLastError());
  return 0;
}
