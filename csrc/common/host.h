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

int sm_count();

// 1 / 0 = the device can / cannot join an NVLS multicast object, -1 = query failed.
int multicast_supported(int dev);

}  // namespace dtdl
