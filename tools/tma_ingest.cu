// Micro-benchmark: how fast can one SM ingest bytes through TMA, by request shape?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I vllm_neuron_b200/csrc tools/tma_ingest.cu -lcuda -o gpurun_out/tma_ingest
// Modes: 0 = tensor 2D box [128 rows x 128 B] out of a row-major matrix with pitch P (128B swizzle)   (decode GEMM W tile)
//        1 = tensor 2D box [128 rows x  64 B], 64B swizzle                                             (weight-only raw tile)
//        2 = tensor 2D box [256 rows x 128 B] (one 32 KB request)
//        3 = 1D bulk copy of 16 KB contiguous (tile-major packed weights)
//        4 = 1D bulk copy of 32 KB contiguous
//        5 = tensor 2D box [128 x 128 B] out of a tile-major tensor (pitch 128 B: the box is one contiguous 16 KB)
//        6 = tensor 2D box [ 64 rows x 128 B] (activation tile, M = 64) re-read from a small L2-resident matrix
// Each CTA streams its own slice with a ring of `stages` buffers; consumer = one thread that waits and releases.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (++spins > (1u << 26)) { printf("timeout\n"); __trap(); }
  }
}
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
               ::"r"(smem_u32(dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// units: one unit = one request of `req_bytes`; CTA c handles units [c*per, (c+1)*per)
__global__ void __launch_bounds__(160) ingest(const __grid_constant__ CUtensorMap map, const uint8_t* base, int mode, int per,
                                             int stages, int req_bytes, int kb_per_row, int box_rows, int box_bytes, int nprod) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + (size_t)stages * req_bytes);
  uint64_t* empty = full + stages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  const long long u0 = (long long)blockIdx.x * per;
  // producers: lane 0 of warps 0..nprod-1, request i belongs to producer i % nprod (does a CTA's TMA stream speed up when
  // its requests come from several warps?); consumer: lane 0 of warp 4
  if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < nprod) {
    for (int i = threadIdx.x >> 5; i < per; i += nprod) {
      const int s = i % stages;
      if (i >= stages) mbar_wait(&empty[s], ((i / stages) - 1) & 1);
      mbar_expect_tx(&full[s], req_bytes);
      const long long u = u0 + i;
      if (mode == 3 || mode == 4) {
        bulk_1d(smem + (size_t)s * req_bytes, base + u * req_bytes, req_bytes, &full[s]);
      } else if (mode == 6) {
        tma_2d(smem + (size_t)s * req_bytes, &map, &full[s], (int)(u % kb_per_row) * box_bytes, 0);
      } else {
        // unit u -> (row tile, k block): consecutive units walk along K like a GEMM CTA does
        const int kb = (int)(u % kb_per_row);
        const long long rt = u / kb_per_row;
        tma_2d(smem + (size_t)s * req_bytes, &map, &full[s], kb * box_bytes, (int)(rt * box_rows));
      }
    }
  } else if (threadIdx.x == 128) {
    for (int i = 0; i < per; ++i) {
      const int s = i % stages;
      mbar_wait(&full[s], (i / stages) & 1);
      mbar_arrive(&empty[s]);
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  CK(cudaSetDevice(0));
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qr));
  const size_t BYTES = (size_t)1 << 30;  // 1 GiB streamed (>> L2)
  uint8_t* buf;
  CK(cudaMalloc(&buf, BYTES));
  CK(cudaMemset(buf, 1, BYTES));
  const int P = 4096;  // row pitch of the "weight matrix" in bytes (K = 4096 fp8)
  const long long rows = BYTES / P;
  CK(cudaFuncSetAttribute(ingest, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  struct Cfg { int mode, box_bytes, box_rows; CUtensorMapSwizzle sw; const char* name; };
  Cfg cfgs[] = {{0, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B, "tensor 128x128B pitch 4096"},
                {1, 64, 128, CU_TENSOR_MAP_SWIZZLE_64B, "tensor 128x64B  pitch 4096"},
                {2, 128, 256, CU_TENSOR_MAP_SWIZZLE_128B, "tensor 256x128B pitch 4096"},
                {3, 0, 0, CU_TENSOR_MAP_SWIZZLE_NONE, "bulk 1D 16 KB"},
                {4, 0, 0, CU_TENSOR_MAP_SWIZZLE_NONE, "bulk 1D 32 KB"},
                {5, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B, "tensor 128x128B tile-major (pitch 128)"},
                {6, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B, "tensor 64x128B L2-resident activations"}};
  for (auto& c : cfgs) {
    CUtensorMap map;
    memset(&map, 0, sizeof(map));
    int req_bytes = c.mode == 3 ? 16384 : c.mode == 4 ? 32768 : c.box_bytes * c.box_rows;
    int kb_per_row = 1;
    if (c.mode != 3 && c.mode != 4) {
      cuuint64_t pitch = c.mode == 5 ? 128 : P;
      cuuint64_t nrows = c.mode == 6 ? 64 : (c.mode == 5 ? BYTES / 128 : rows);
      cuuint64_t gdim[2] = {pitch, nrows};
      cuuint64_t gstr[1] = {pitch};
      cuuint32_t box[2] = {(cuuint32_t)c.box_bytes, (cuuint32_t)c.box_rows};
      cuuint32_t es[2] = {1, 1};
      CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, c.sw,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("encode failed %d for %s\n", (int)r, c.name); continue; }
      kb_per_row = (int)(pitch / c.box_bytes);
    }
    for (int per_sm = 1; per_sm <= 2; ++per_sm) {
      for (int nprod : {1, 2, 4})
      for (int stages : {4, 8}) {
        if (nprod > 1 && (stages != 8 || per_sm != 1)) continue;
        const size_t smem = (size_t)stages * req_bytes + 2 * stages * 8 + 1024;
        if (smem * per_sm > 226 * 1024) continue;
        const int grid = 148 * per_sm;
        const size_t total_bytes = c.mode == 6 ? (size_t)256 << 20 : BYTES / 2;
        const int per = (int)(total_bytes / req_bytes / grid);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        for (int it = 0; it < 2; ++it) {
          CK(cudaEventRecord(e0));
          ingest<<<grid, 160, smem>>>(map, buf, c.mode, per, stages, req_bytes, kb_per_row, c.box_rows, c.box_bytes, nprod);
          CK(cudaEventRecord(e1));
          CK(cudaDeviceSynchronize());
        }
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double gb = (double)per * grid * req_bytes / 1e9;
        printf("%-42s ctas/SM %d producers %d stages %d: %8.1f GB/s total, %6.1f GB/s per SM, %.2f cyc@1.9GHz per 128B\n", c.name, per_sm,
               nprod, stages, gb / (ms * 1e-3), gb / (ms * 1e-3) / 148, 1.9e9 * (ms * 1e-3) / ((double)per * per_sm * req_bytes / 128));
      }
    }
  }
  return 0;
}
