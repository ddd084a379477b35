// Host-side helpers shared by every translation unit of libdtdl_b200.so:
// error handling for the C ABI, TMA tensor-map encoding (through the driver
// entry point so the library does not link libcuda), and device properties.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace dtdl {

// Last error message of the calling thread; every extern "C" function returns
// 0 on success and non-zero after storing a message here.
void set_last_error(const std::string& msg);
const char* get_last_error();

#define DTDL_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::dtdl::set_last_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +   \
                             " (" __FILE__ ":" + std::to_string(__LINE__) + ")");          \
      return 1;                                                                            \
    }                                                                                      \
  } while (0)

#define DTDL_REQUIRE(cond, msg)                                                            \
  do {                                                                                     \
    if (!(cond)) {                                                                         \
      ::dtdl::set_last_error(std::string(msg) + " [" #cond "] (" __FILE__ ":" +            \
                             std::to_string(__LINE__) + ")");                              \
      return 2;                                                                            \
    }                                                                                      \
  } while (0)

// Encode a bf16 tiled tensor map with 128-byte swizzle.  `dims`/`box` are in
// elements, fastest dimension first; `strides_bytes[i]` is the byte stride of
// dimension i+1 (rank-1 entries).  Out-of-bounds box elements read as zero.
// Returns 0 on success.
int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box);

// Activation storage types of the op kernels ("act" in the C ABI).
enum { kActBf16 = 0, kActF32 = 1 };

// Same for either activation type: act = kActBf16 -> bf16 elements, kActF32 -> fp32 elements (the
// tensor core reads them as TF32).  `mn_major32` selects the 128-byte swizzle with 32-byte atoms
// that tcgen05 requires for MN-major 32-bit operands (weight-gradient GEMM in fp32 mode); it is
// ignored for bf16.  `elem_strides` (rank entries, may be null = all 1) lets a box sample every
// other pixel (stride-2 convolutions).
int encode_tmap_act(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, int act, bool mn_major32,
                    const uint32_t* elem_strides = nullptr);

static inline int act_esize(int act) { return act == kActF32 ? 4 : 2; }

int sm_count();

// 1 / 0 = the device can / cannot join an NVLS multicast object, -1 = query failed.
int multicast_supported(int dev);

}  // namespace dtdl
