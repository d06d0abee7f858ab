// cnx_util.cu — host utilities shared by every launcher: error string, device check,
// TMA descriptor encoding through the driver entry point (no link-time libcuda dependency).
#include "cnx_common.cuh"
#include "cnx_kernels.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <stdlib.h>

namespace cnx {

namespace {
thread_local char g_err[512] = "";
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
std::once_flag g_encode_once;
int g_sm_count = 0;
int g_cc_major = -1, g_cc_minor = -1;
std::once_flag g_dev_once;

void query_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return;
  g_sm_count = prop.multiProcessorCount;
  g_cc_major = prop.major;
  g_cc_minor = prop.minor;
}
}  // namespace

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

namespace {
int* g_flag_host = nullptr;
int* g_flag_dev = nullptr;
std::once_flag g_flag_once;
}  // namespace

int* timeout_flag_ptr() {
  std::call_once(g_flag_once, [] {
    void* h = nullptr;
    if (cudaHostAlloc(&h, sizeof(int), cudaHostAllocMapped) != cudaSuccess) return;
    *static_cast<int*>(h) = 0;
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) return;
    g_flag_host = static_cast<int*>(h);
    g_flag_dev = static_cast<int*>(d);
  });
  return g_flag_dev;
}

int read_and_clear_timeout_flag() {
  if (timeout_flag_ptr() == nullptr) return -1;
  const int v = *reinterpret_cast<volatile int*>(g_flag_host);
  *reinterpret_cast<volatile int*>(g_flag_host) = 0;
  return v;
}

void debug_set_timeout_flag(int v) {
  if (timeout_flag_ptr() != nullptr) *reinterpret_cast<volatile int*>(g_flag_host) = v;
}

int check_sticky_timeout() {
  // the flag page exists only once a TMA / tcgen05 kernel has been launched
  if (g_flag_host != nullptr && *reinterpret_cast<volatile int*>(g_flag_host) != 0) {
    set_last_error("an earlier libcnx_b200 kernel timed out waiting on an mbarrier: its results "
                   "and those of every launch since are invalid (cnx_debug_timeout_flag() reads "
                   "and clears the condition)");
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int zero_async(void* p, size_t bytes, cudaStream_t stream, const char* who) {
  cudaError_t e = cudaMemsetAsync(p, 0, bytes, stream);
  if (e != cudaSuccess) {
    set_last_error("%s: cudaMemsetAsync(%zu B): %s", who, bytes, cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int ensure_dyn_smem_impl(const void* kernel, size_t bytes, const char* who) {
  static std::mutex mu;
  static std::unordered_map<const void*, unsigned long long> done;   // kernel -> device bit mask
  const int dev = current_device_slot();
  std::lock_guard<std::mutex> lock(mu);
  unsigned long long& mask = done[kernel];
  if (mask & (1ull << dev)) return CNX_OK;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(bytes));
  if (e != cudaSuccess) {
    set_last_error("%s: cudaFuncSetAttribute(dynamic shared memory %zu B): %s", who, bytes,
                   cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  mask |= 1ull << dev;
  return CNX_OK;
}

int current_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
  return dev & 63;
}

int check_device_sm100() {
  std::call_once(g_dev_once, query_device);
  if (g_cc_major != 10) {
    set_last_error("libcnx_b200 requires an sm_100a device (found compute capability %d.%d); "
                   "there is no fallback path", g_cc_major, g_cc_minor);
    return CNX_ERR_NOT_SM100;
  }
  return CNX_OK;
}

bool pdl_enabled() {
  static const bool on = getenv("CNX_NO_PDL") == nullptr;
  return on;
}

int sm_count() {
  std::call_once(g_dev_once, query_device);
  return g_sm_count > 0 ? g_sm_count : 148;
}

int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank,
                    const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                    int swizzle) {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
  });
  if (g_encode == nullptr) {
    set_last_error("cuTensorMapEncodeTiled driver entry point not available");
    return CNX_ERR_DRIVER;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) {
    set_last_error("TMA base address %p is not 16-byte aligned", base);
    return CNX_ERR_BAD_ARG;
  }
  CUtensorMapDataType dt;
  switch (elem_bytes) {
    case 2: dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; break;
    case 4: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; break;
    default:
      set_last_error("make_tensor_map: unsupported element size %d", elem_bytes);
      return CNX_ERR_BAD_ARG;
  }
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (box[i] == 0 || box[i] > 256) {
      set_last_error("make_tensor_map: box[%d]=%u out of range", i, box[i]);
      return CNX_ERR_BAD_SHAPE;
    }
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    if ((gstr[i] & 15) != 0) {
      set_last_error("make_tensor_map: stride[%d]=%llu bytes is not a multiple of 16", i,
                     (unsigned long long)gstr[i]);
      return CNX_ERR_BAD_SHAPE;
    }
  }
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle == 1) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  if (swizzle == 2) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  if (swizzle == 3) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = g_encode(map, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims,
                        gstr, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, box0 %u)",
                   static_cast<int>(r), rank, box[0]);
    return CNX_ERR_DRIVER;
  }
  return CNX_OK;
}

}  // namespace cnx
