// Host-side plumbing of libb200_paged.so: error strings, device probe, TMA descriptor encoding.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <unordered_set>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
static std::atomic<int> g_pdl{1};
bool pdl_enabled() { return g_pdl.load(std::memory_order_relaxed) != 0; }

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_last_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  // libcuda is reached through the runtime so the .so has no link-time dependency on the driver
  // (it must load, and export its symbols, on the GPU-less build box).
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, CUtensorMapDataType dt, uint32_t elem_bytes, uint64_t cols,
                 uint64_t rows, uint64_t row_pitch_bytes, uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no driver?)");
    return B200_ERR_DRIVER;
  }
  if (((uintptr_t)base & 15) || (row_pitch_bytes & 15) || box_cols * elem_bytes != 128 || box_rows > 256) {
    set_last_error("tensor map: base %p pitch %llu box %ux%u elem %u violates TMA alignment rules", base,
                   (unsigned long long)row_pitch_bytes, box_cols, box_rows, elem_bytes);
    return B200_ERR_INVALID_ARG;
  }
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed: CUresult %d (cols=%llu rows=%llu pitch=%llu box=%ux%u)", (int)r,
                   (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)row_pitch_bytes, box_cols,
                   box_rows);
    return B200_ERR_DRIVER;
  }
  return 0;
}

int make_tmap_2d_plain(CUtensorMap* out, const void* base, CUtensorMapDataType dt, uint32_t elem_bytes, uint64_t cols,
                       uint64_t rows, uint64_t row_pitch_bytes, uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no driver?)");
    return B200_ERR_DRIVER;
  }
  if (((uintptr_t)base & 15) || (row_pitch_bytes & 15) || (box_cols * elem_bytes) % 16 != 0 || box_rows > 256) {
    set_last_error("tensor map (plain): base %p pitch %llu box %ux%u elem %u violates TMA alignment rules", base,
                   (unsigned long long)row_pitch_bytes, box_cols, box_rows, elem_bytes);
    return B200_ERR_INVALID_ARG;
  }
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (plain) failed: CUresult %d (cols=%llu rows=%llu pitch=%llu box=%ux%u)", (int)r,
                   (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)row_pitch_bytes, box_cols,
                   box_rows);
    return B200_ERR_DRIVER;
  }
  return 0;
}

int make_tmap_2d_sw64(CUtensorMap* out, const void* base, CUtensorMapDataType dt, uint32_t elem_bytes, uint64_t cols,
                      uint64_t rows, uint64_t row_pitch_bytes, uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled entry point not available (no driver?)");
    return B200_ERR_DRIVER;
  }
  if (((uintptr_t)base & 15) || (row_pitch_bytes & 15) || box_cols * elem_bytes != 64 || box_rows > 256) {
    set_last_error("tensor map (64B swizzle): base %p pitch %llu box %ux%u elem %u violates TMA alignment rules", base,
                   (unsigned long long)row_pitch_bytes, box_cols, box_rows, elem_bytes);
    return B200_ERR_INVALID_ARG;
  }
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (64B swizzle) failed: CUresult %d (cols=%llu rows=%llu pitch=%llu box=%ux%u)",
                   (int)r, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)row_pitch_bytes, box_cols,
                   box_rows);
    return B200_ERR_DRIVER;
  }
  return 0;
}

void unify_carveout(const void* kern) {
  static const int on = getenv("B200_CARVEOUT") ? atoi(getenv("B200_CARVEOUT")) : 0;
  if (!on) return;
  static std::mutex mu;
  static std::unordered_set<const void*> done;
  std::lock_guard<std::mutex> g(mu);
  if (done.insert(kern).second) {
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaGetLastError();
  }
}

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaDeviceProp p;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaGetDeviceProperties(&p, dev) == cudaSuccess) n = p.multiProcessorCount;
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace b200

extern "C" {

int b200_abi_version(void) { return B200_ABI_VERSION; }

const char* b200_last_error(void) { return b200::g_err; }

uint64_t b200_launch_count(void) { return b200::g_launches.load(); }

int b200_set_pdl(int enabled) {
  int old = b200::g_pdl.exchange(enabled ? 1 : 0);
  return old;
}

int b200_device_check(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device >= n) {
    b200::set_last_error("no CUDA device %d (count %d)", device, n);
    cudaGetLastError();
    return B200_ERR_NO_DEVICE;
  }
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device);
  if (major != 10) {
    b200::set_last_error("device %d is sm_%d%d; libb200_paged.so holds sm_100a code only", device, major, minor);
    return B200_ERR_NO_DEVICE;
  }
  return B200_OK;
}

}  // extern "C"
