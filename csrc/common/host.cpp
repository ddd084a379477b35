#include "host.h"

#include <mutex>

namespace dtdl {

static thread_local std::string g_last_error;

void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* get_last_error() { return g_last_error.c_str(); }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

int encode_tmap_act(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, int act, bool mn_major32,
                    const uint32_t* elem_strides) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled is not available from this driver");
    return 1;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bdim[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = elem_strides ? elem_strides[i] : 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  const CUtensorMapDataType dt =
      act == kActF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const CUtensorMapSwizzle sw = (act == kActF32 && mn_major32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                                               : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bdim, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    std::string m = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r) +
                    " act=" + std::to_string(act) + " rank=" + std::to_string(rank) + " dims=";
    for (int i = 0; i < rank; ++i) m += std::to_string(dims[i]) + ",";
    m += " box=";
    for (int i = 0; i < rank; ++i) m += std::to_string(box[i]) + ",";
    m += " strides=";
    for (int i = 0; i + 1 < rank; ++i) m += std::to_string(strides_bytes[i]) + ",";
    set_last_error(m);
    return 1;
  }
  return 0;
}

int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box) {
  return encode_tmap_act(out, base, rank, dims, strides_bytes, box, kActBf16, false, nullptr);
}

int multicast_supported(int dev) {
  // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED through the runtime's driver entry point (no -lcuda)
  typedef CUresult (*GetAttrFn)(int*, CUdevice_attribute, CUdevice);
  typedef CUresult (*GetDevFn)(CUdevice*, int);
  void *pa = nullptr, *pd = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &pa, cudaEnableDefault, &q) != cudaSuccess || !pa) return -1;
  if (cudaGetDriverEntryPoint("cuDeviceGet", &pd, cudaEnableDefault, &q) != cudaSuccess || !pd) return -1;
  CUdevice d;
  if (reinterpret_cast<GetDevFn>(pd)(&d, dev) != CUDA_SUCCESS) return -1;
  int v = 0;
  if (reinterpret_cast<GetAttrFn>(pa)(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) != CUDA_SUCCESS) return -1;
  return v;
}

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace dtdl
