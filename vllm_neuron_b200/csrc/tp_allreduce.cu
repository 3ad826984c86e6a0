// Tensor-parallel exchange step of the decode path: all-reduce of the row-parallel GEMM outputs (o_proj, down_proj)
// over NVLink peer memory, fused with the RMSNorm (+ dynamic per-token quantisation) that consumes the sum.
//
// One-shot, pull based: every rank's GEMM writes its partial [T][H] bf16 into a buffer of its own SYMMETRIC region (the
// same virtual layout on every GPU, mapped into every peer).  One CTA per token row then
//   1. signals "my partial of row t is written" into every peer's flag area (one 4-byte store over NVLink) and waits for
//      the same signal from every peer (spinning on its own, local, flags),
//   2. loads row t from all ranks (peer loads), adds them in RANK ORDER in fp32 -- so every rank computes bit-identical
//      sums -- rounds to bf16 = the new residual stream, and
//   3. applies RMSNorm and the per-token quantisation exactly like rmsnorm_kernel (same reductions, same rounding).
// No NCCL call, no separate norm kernel, no exit barrier: the two exchange sites of a layer alternate between two buffers,
// and a rank can only overwrite buffer A after it has passed the entry barrier of the exchange on buffer B, by which time
// every peer has finished reading A (same-stream order on the peer).
// Flags carry a per-row epoch that only grows, so nothing is ever reset and a CUDA graph can replay the step.
//
// NVLS variant (b200_allreduce_rmsnorm_nvls): when the symmetric region also has a MULTICAST mapping (NVSwitch), step 2
// is ONE multimem.ld_reduce per 16 bytes on the multicast address: the switch fetches the row from every rank, adds the
// bf16 pairs with an fp32 accumulator and returns the rounded sum.  Each rank then receives T*H*2 bytes per exchange
// instead of (world - 1) x that (3.5 MB -> 0.5 MB at hidden 4096, TP=8; 7 -> 1 MB at hidden 8192), which is what bounded
// the pull at TP=8.  The order of the fp32 additions is the switch's, not rank order: sums may differ from the pull
// variant in the last bf16 bit of a few elements; every rank still receives identical bits (checked in tools/tp_check.py).
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int AR_MAX_TP = 8;
constexpr int AR_THREADS = 256;
constexpr int AR_MAXV = 4;  // 4 x 8 bf16 per thread -> H <= 8192

struct ArPeers {
  const uint4* buf[AR_MAX_TP];  // rank p's partial [T][H] bf16
  uint32_t* flags[AR_MAX_TP];   // rank p's flag area: [rows][AR_MAX_TP] epochs, written by the peers
  const uint4* mc_buf;          // NVLS: multicast address of the partial (all ranks' replicas), else null
};

// Flag traffic is RELAXED at system scope on purpose.  Measured on 2 x B200 (profiles/r01_tp_exchange_*.md): a
// system-scope fence costs ~3.2 us and st.release.sys as much again, more than the NVLink hop itself.  They are not needed:
// the partial this flag announces was written by the PREVIOUS kernel of the stream, whose completion (griddepcontrol.wait)
// already made it visible at its home L2, which is where a peer's NVLink read is served; and the reader issues its data
// loads only after the flag value it spins on has arrived (control dependency, then a CTA barrier).
__device__ __forceinline__ void st_flag_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer memory must not be served from this SM's L1
__device__ __forceinline__ uint4 ld_peer(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// in-switch sum over all replicas of 8 bf16 (fp32 accumulate, rounded back to bf16)
__device__ __forceinline__ uint4 ld_reduce_mc(const uint4* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void unpack8f(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __low2float(h[i]);
    f[2 * i + 1] = __high2float(h[i]);
  }
}
__device__ __forceinline__ float ar_block_sum(float v, float* sh) {
  v = warp_sum(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = (l < nw) ? sh[l] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float ar_block_max(float v, float* sh) {
  v = warp_max(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = (l < nw) ? sh[l] : 0.f;
  return warp_max(r);
}

template <int QD, bool NVLS>
__global__ void __launch_bounds__(AR_THREADS)
    allreduce_rmsnorm_kernel(ArPeers peers, uint32_t* __restrict__ epoch, int rank, int world, uint4* __restrict__ x_out,
                             const uint4* __restrict__ w, uint4* __restrict__ out_bf16, uint2* __restrict__ out_q,
                             float* __restrict__ out_scale, int H, float eps, int dbg) {
  griddep_launch();
  griddep_wait();  // this rank's GEMM has written its partial
  __shared__ float sh[32];
  const int t = blockIdx.x, nvec = H / 8;

  // ---- entry barrier for row t
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    const uint32_t e = epoch[t] + 1;
    __syncwarp((1u << world) - 1u);  // everyone has read the epoch before lane 0 may advance it
    st_flag_sys(peers.flags[peer] + (size_t)t * AR_MAX_TP + rank, e);
    const uint32_t* mine = peers.flags[rank] + (size_t)t * AR_MAX_TP + peer;
    const long long t0 = clock64();
    while (!(dbg & 2) && (int32_t)(ld_flag_sys(mine) - e) < 0) {
      if (clock64() - t0 > 6000000000LL) {  // ~3 s: a missing peer must not hang the box
        printf("b200: all-reduce barrier timeout rank %d row %d peer %d epoch %u\n", rank, t, peer, e);
        __trap();
      }
    }
    if (peer == 0) epoch[t] = e;
  }
  __syncthreads();

  // ---- sum over ranks in rank order (all loads in flight before the first add)
  float xv[AR_MAXV][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < AR_MAXV; ++i) {
    const int vi = threadIdx.x + i * AR_THREADS;
    if (vi < nvec) {
      uint32_t pk[4];
      if constexpr (NVLS) {
        const uint4 r = ld_reduce_mc(peers.mc_buf + (size_t)t * nvec + vi);
        pk[0] = r.x, pk[1] = r.y, pk[2] = r.z, pk[3] = r.w;
      } else {
        uint4 part[AR_MAX_TP];
#pragma unroll
        for (int p = 0; p < AR_MAX_TP; ++p)
          if (p < world) part[p] = ld_peer(peers.buf[(dbg & 4) ? rank : p] + (size_t)t * nvec + vi);
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int p = 0; p < AR_MAX_TP; ++p) {
          if (p < world) {
            float f[8];
            unpack8f(part[p], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = __fadd_rn(acc[e], f[e]);
          }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) pk[e] = pack_bf16(acc[2 * e], acc[2 * e + 1]);
      }
      const uint4 xr = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      x_out[(size_t)t * nvec + vi] = xr;
      unpack8f(xr, xv[i]);  // the norm sees the bf16 residual stream, like rmsnorm_kernel does
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += xv[i][e] * xv[i][e];
    }
  }
  if (w == nullptr) return;  // plain all-reduce

  // ---- RMSNorm (+ quantisation): same arithmetic as rmsnorm_kernel in kv_ops.cu
  ss = ar_block_sum(ss, sh);
  const float rstd = rsqrtf(ss / (float)H + eps);
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < AR_MAXV; ++i) {
    const int vi = threadIdx.x + i * AR_THREADS;
    if (vi < nvec) {
      float wv[8];
      unpack8f(w[vi], wv);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        xv[i][e] = __bfloat162float(__float2bfloat16_rn(xv[i][e] * rstd * wv[e]));
        amax = fmaxf(amax, fabsf(xv[i][e]));
      }
      if (out_bf16) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) pk[e] = pack_bf16(xv[i][2 * e], xv[i][2 * e + 1]);
        out_bf16[(size_t)t * nvec + vi] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
    }
  }
  if constexpr (QD != B200_BF16) {
    amax = ar_block_max(amax, sh);
    const float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
    const float scale = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
    if (threadIdx.x == 0) out_scale[t] = scale;
#pragma unroll
    for (int i = 0; i < AR_MAXV; ++i) {
      const int vi = threadIdx.x + i * AR_THREADS;
      if (vi < nvec) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) lo |= (uint32_t)quant_byte<QD>(xv[i][e], scale) << (8 * e);
#pragma unroll
        for (int e = 0; e < 4; ++e) hi |= (uint32_t)quant_byte<QD>(xv[i][4 + e], scale) << (8 * e);
        out_q[(size_t)t * nvec + vi] = make_uint2(lo, hi);
      }
    }
  }
}

// Greedy merge across the vocabulary shards of a tensor-parallel group (replaces two NCCL all-gathers plus a handful of
// eager element-wise kernels per step): every rank's argmax kernel has written its local (max logit, GLOBAL index) pairs
// into its symmetric region (vals [B] fp32, idxs [B] int64); this kernel signals "my pairs of this step are written",
// waits for every peer's signal, pulls the `world` pairs of each row and keeps the largest value, ties to the lowest
// index -- the single-GPU argmax rule, so every rank computes the same token.  Same flag discipline as the exchange above
// (relaxed system-scope epoch flags; the pairs were written by the PREVIOUS kernel of the stream).  One buffer suffices:
// a rank rewrites its pairs one full decode step later, with dozens of all-rank exchange barriers in between.
struct MergePeers {
  const float* vals[AR_MAX_TP];
  const long long* idxs[AR_MAX_TP];
  uint32_t* flags[AR_MAX_TP];  // [AR_MAX_TP] epochs per rank, written by the peers
};
__global__ void __launch_bounds__(128)
    tp_argmax_merge_kernel(MergePeers peers, uint32_t* __restrict__ epoch, int rank, int world, long long* __restrict__ out_ids,
                           int B) {
  griddep_launch();
  griddep_wait();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    const uint32_t e = epoch[0] + 1;
    __syncwarp((1u << world) - 1u);
    st_flag_sys(peers.flags[peer] + rank, e);
    const uint32_t* mine = peers.flags[rank] + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_flag_sys(mine) - e) < 0) {
      if (clock64() - t0 > 6000000000LL) {
        printf("b200: argmax merge barrier timeout rank %d peer %d epoch %u\n", rank, peer, e);
        __trap();
      }
    }
    if (peer == 0) epoch[0] = e;
  }
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    // Start from rank 0's pair, so that out_ids is always an index some rank's argmax wrote (b200_argmax never writes one
    // outside its shard) -- also for a row whose maxima are all NaN, where no comparison below is true.  A NaN maximum
    // loses to any number; among numbers the largest wins, ties to the lowest index.
    float best = 0.f;
    long long bi = 0;
    for (int p = 0; p < world; ++p) {
      float v;
      long long i;
      asm volatile("ld.volatile.global.f32 %0, [%1];\n" : "=f"(v) : "l"(peers.vals[p] + b) : "memory");
      asm volatile("ld.volatile.global.s64 %0, [%1];\n" : "=l"(i) : "l"(peers.idxs[p] + b) : "memory");
      if (p == 0 || v > best || (v == best && i < bi) || (best != best && v == v)) {
        best = v;
        bi = i;
      }
    }
    out_ids[b] = bi;
  }
}

}  // namespace b200

using namespace b200;

namespace {
int allreduce_common(const void* const* host_peer_bufs, void* const* host_peer_flags, const void* mc_buf, uint32_t* epoch,
                     int rank, int world, void* x_out, const void* norm_w, void* out_bf16, void* out_q, float* out_scale,
                     int T, int H, float eps, int q_dtype, void* stream) {
  B200_REQUIRE(host_peer_flags && epoch && x_out && (host_peer_bufs || mc_buf), "allreduce_rmsnorm: null pointer");
  B200_REQUIRE(world >= 1 && world <= AR_MAX_TP && rank >= 0 && rank < world, "allreduce_rmsnorm: bad rank %d / world %d",
               rank, world);
  B200_REQUIRE(H > 0 && H % 8 == 0 && H <= 8 * AR_THREADS * AR_MAXV, "allreduce_rmsnorm: H=%d unsupported", H);
  B200_REQUIRE(q_dtype == B200_BF16 || q_dtype == B200_FP8_E4M3 || q_dtype == B200_INT8, "allreduce_rmsnorm: bad q_dtype");
  B200_REQUIRE(!norm_w || q_dtype == B200_BF16 ? true : (out_q && out_scale), "allreduce_rmsnorm: out_q/out_scale missing");
  B200_REQUIRE(!norm_w || q_dtype != B200_BF16 || out_bf16, "allreduce_rmsnorm: out_bf16 missing");
  if (T <= 0) return B200_OK;
  ArPeers peers{};
  for (int p = 0; p < world; ++p) {
    B200_REQUIRE(host_peer_flags[p] && (mc_buf || host_peer_bufs[p]), "allreduce_rmsnorm: null peer pointer %d", p);
    peers.buf[p] = mc_buf ? nullptr : (const uint4*)host_peer_bufs[p];
    peers.flags[p] = (uint32_t*)host_peer_flags[p];
  }
  peers.mc_buf = (const uint4*)mc_buf;
  cudaStream_t s = (cudaStream_t)stream;
  static const int dbg = getenv("B200_AR_DEBUG") ? atoi(getenv("B200_AR_DEBUG")) : 0;  // timing experiments only
#define B200_AR_LAUNCH(QD, NV, OQ, OS)                                                                                     \
  B200_CUDA_OK(launch_kernel(allreduce_rmsnorm_kernel<QD, NV>, dim3(T), dim3(AR_THREADS), 0, s, peers, epoch, rank, world, \
                             (uint4*)x_out, (const uint4*)norm_w, (uint4*)out_bf16, (uint2*)(OQ), (float*)(OS), H, eps, dbg))
  const bool nv = mc_buf != nullptr;
  if (norm_w && q_dtype == B200_FP8_E4M3) {
    if (nv) B200_AR_LAUNCH(B200_FP8_E4M3, true, out_q, out_scale); else B200_AR_LAUNCH(B200_FP8_E4M3, false, out_q, out_scale);
  } else if (norm_w && q_dtype == B200_INT8) {
    if (nv) B200_AR_LAUNCH(B200_INT8, true, out_q, out_scale); else B200_AR_LAUNCH(B200_INT8, false, out_q, out_scale);
  } else {
    if (nv) B200_AR_LAUNCH(B200_BF16, true, nullptr, nullptr); else B200_AR_LAUNCH(B200_BF16, false, nullptr, nullptr);
  }
#undef B200_AR_LAUNCH
  return B200_OK;
}
}  // namespace

extern "C" int b200_allreduce_rmsnorm(const void* const* host_peer_bufs, void* const* host_peer_flags, uint32_t* epoch,
                                      int rank, int world, void* x_out, const void* norm_w, void* out_bf16, void* out_q,
                                      float* out_scale, int T, int H, float eps, int q_dtype, void* stream) {
  B200_REQUIRE(host_peer_bufs, "allreduce_rmsnorm: null pointer");
  return allreduce_common(host_peer_bufs, host_peer_flags, nullptr, epoch, rank, world, x_out, norm_w, out_bf16, out_q,
                          out_scale, T, H, eps, q_dtype, stream);
}

extern "C" int b200_allreduce_rmsnorm_nvls(const void* mc_buf, void* const* host_peer_flags, uint32_t* epoch, int rank,
                                           int world, void* x_out, const void* norm_w, void* out_bf16, void* out_q,
                                           float* out_scale, int T, int H, float eps, int q_dtype, void* stream) {
  B200_REQUIRE(mc_buf, "allreduce_rmsnorm_nvls: null multicast pointer");
  return allreduce_common(nullptr, host_peer_flags, mc_buf, epoch, rank, world, x_out, norm_w, out_bf16, out_q, out_scale, T,
                          H, eps, q_dtype, stream);
}

extern "C" int b200_tp_argmax_merge(const void* const* host_peer_vals, const void* const* host_peer_idxs,
                                    void* const* host_peer_flags, uint32_t* epoch, int rank, int world, int64_t* out_ids,
                                    int B, void* stream) {
  B200_REQUIRE(host_peer_vals && host_peer_idxs && host_peer_flags && epoch && out_ids, "tp_argmax_merge: null pointer");
  B200_REQUIRE(world >= 1 && world <= AR_MAX_TP && rank >= 0 && rank < world, "tp_argmax_merge: bad rank %d / world %d", rank,
               world);
  if (B <= 0) return B200_OK;
  MergePeers peers{};
  for (int p = 0; p < world; ++p) {
    B200_REQUIRE(host_peer_vals[p] && host_peer_idxs[p] && host_peer_flags[p], "tp_argmax_merge: null peer pointer %d", p);
    peers.vals[p] = (const float*)host_peer_vals[p];
    peers.idxs[p] = (const long long*)host_peer_idxs[p];
    peers.flags[p] = (uint32_t*)host_peer_flags[p];
  }
  B200_CUDA_OK(launch_kernel(tp_argmax_merge_kernel, dim3(1), dim3(128), 0, (cudaStream_t)stream, peers, epoch, rank, world,
                             (long long*)out_ids, B));
  return B200_OK;
}
