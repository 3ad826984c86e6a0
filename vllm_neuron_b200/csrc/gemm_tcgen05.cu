// Per-channel-scaled weight GEMM for the decoder linears (QKV, O, gate/up, down, lm_head) on tcgen05.
//
//   out[m][n] = (sum_k a[m][k] w[n][k]) * a_scale[m] * w_scale[n] + bias[n] + residual[m][n]
//
// "swap-AB" tile everywhere: the 128 weight rows of a tile fill the UMMA M slot, the activation rows are the UMMA N.
// Operands are staged by TMA (cp.async.bulk.tensor, 128B swizzle) in a ring of stages; one thread issues tcgen05.mma
// (kind::f8f6f4 / kind::f16 / kind::i8) into a TMEM accumulator; four epilogue warps read it back with tcgen05.ld,
// apply the per-token and per-channel scales and write bf16/fp32.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = epilogue.
//
// Two schedules:
//  * DECODE (M <= 128, one m-tile): bound by streaming W from HBM exactly once, and so short (2.7 .. 19 us of traffic per
//    linear of Llama-3.1-8B) that fixed latencies decide.  One CTA per (weight tile, K split); the K splits of a tile
//    form a THREAD-BLOCK CLUSTER (1/2/4/8 CTAs) and are summed through distributed shared memory: every CTA pushes its
//    fp32 partial, slice by slice, into the peer that owns those weight rows (st.shared::cluster), one cluster barrier,
//    then each CTA adds the slices in fixed rank order (bit-reproducible) and runs the epilogue for its 128/S rows.
//    No global workspace, no atomics, no second kernel.  CTAs are sized (<= 105 KB smem, <= 64 TMEM columns) so that
//    two are resident per SM: the prologue/epilogue of one overlaps the streaming of the other.
//  * PREFILL (M > 128): persistent stream-K, one CTA per SM; the flattened (tile, k-block) space is cut into equal
//    contiguous spans; a tile cut by a span boundary is finished by whichever contributor arrives last (partials parked
//    in the workspace, summed in fixed contributor order).
// Both use programmatic dependent launch: the grid becomes resident while the previous kernel drains and the TMA
// producer fills the smem ring with WEIGHT tiles (which no kernel of the step writes) before griddepcontrol.wait;
// activations are only touched after the wait.
#include "common.cuh"

namespace b200 {

constexpr int GEMM_BN = 128;        // weight rows per tile  (UMMA M)
constexpr int GEMM_BK_BYTES = 128;  // one 128B swizzle row of K per stage
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_CNT_INTS = 1024;  // fix-up counters at the head of the workspace (one per possible first contributor)

template <int KIND>
struct KindTraits;
template <>
struct KindTraits<KIND_F16> {
  static constexpr int ELEM = 2, CFMT = 1, ABFMT = 1;  // bf16 in, f32 accumulate
};
template <>
struct KindTraits<KIND_F8F6F4> {
  static constexpr int ELEM = 1, CFMT = 1, ABFMT = 0;  // e4m3 in, f32 accumulate
};
template <>
struct KindTraits<KIND_I8> {
  static constexpr int ELEM = 1, CFMT = 2, ABFMT = 1;  // s8 in, s32 accumulate
};

enum { EPI_LINEAR = 0, EPI_SILU_MUL = 1 };

struct GemmEpilogue {
  const float* a_scale;
  const float* w_scale;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  void* out;
  int out_f32;
  float* group_amax;  // EPI_SILU_MUL: [M][N/32] max |act| over the 16 channels of each interleaved row group
};

__device__ __forceinline__ float linear_value(const GemmEpilogue& ep, float acc, int m, int n, int N) {
  float y = acc;
  if (ep.a_scale) y *= ep.a_scale[m];
  if (ep.w_scale) y *= ep.w_scale[n];
  if (ep.bias) y += __bfloat162float(ep.bias[n]);
  if (ep.residual) y += __bfloat162float(ep.residual[(int64_t)m * N + n]);
  return y;
}
__device__ __forceinline__ void store_out(const GemmEpilogue& ep, float acc, int m, int n, int N) {
  float y = linear_value(ep, acc, m, n, N);
  if (ep.out_f32)
    reinterpret_cast<float*>(ep.out)[(int64_t)m * N + n] = y;
  else
    reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * N + n] = __float2bfloat16_rn(y);
}

// span of CTA c: units [U*c/P, U*(c+1)/P); owner(x) = the CTA whose span holds unit x
__device__ __forceinline__ long long span_begin(long long U, int P, int c) { return U * c / P; }
__device__ __forceinline__ int span_owner(long long U, int P, long long x) { return (int)(((x + 1) * P - 1) / U); }

// SiLU(gate) * up for one channel; both inputs already rounded to bf16 like the unfused path (b200_gemm + b200_silu_mul)
__device__ __forceinline__ float silu_mul_bf16(float g, float u) {
  const float sg = __fdiv_rn(1.f, 1.f + expf(-g));
  return __bfloat162float(__float2bfloat16_rn(g * sg * u));
}

// Epilogue of one 16-token column chunk held in v[] by the lane that owns weight row n (EPI_LINEAR), or by the lane pair
// (l, l^16) that owns the gate and the up row of one channel (EPI_SILU_MUL; W rows interleaved 16 gate / 16 up per warp).
template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmEpilogue& ep, const float (&v)[16], int m_base, int M, int n,
                                               bool n_ok, int N, int lane) {
  if constexpr (EPI == EPI_LINEAR) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int m = m_base + j;
      if (m < M && n_ok) store_out(ep, v[j], m, n, N);
    }
  } else {
    // this lane's row: gate (lane < 16) or up (lane >= 16) of channel ch; both rounded to bf16 like the unfused path
    const int I = N >> 1, n_groups = N >> 5;
    const int ch = (n >> 5) * 16 + (lane & 15);
    float mine[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int m = m_base + j;
      float y = 0.f;
      if (m < M && n_ok) y = __bfloat162float(__float2bfloat16_rn(linear_value(ep, v[j], m, n, N)));
      mine[j] = y;
    }
    // lanes < 16 finish tokens 0..7 of the chunk, lanes >= 16 finish tokens 8..15
    const bool hi = lane >= 16;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float send = hi ? mine[j] : mine[j + 8];  // what the partner needs from me
      const float got = __shfl_xor_sync(0xffffffffu, send, 16);
      const float g = hi ? got : mine[j];
      const float u = hi ? mine[j + 8] : got;
      const int m = m_base + (hi ? 8 : 0) + j;
      const bool ok = m < M && n_ok;
      const float a = silu_mul_bf16(g, u);
      if (ok) reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * I + ch] = __float2bfloat16_rn(a);
      float x = ok ? fabsf(a) : 0.f;  // max over the 16 channels of the group (one half-warp)
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
      if ((lane & 15) == 0 && ok) ep.group_amax[(int64_t)m * n_groups + (n >> 5)] = x;
    }
  }
}

template <int KIND, int BM, int NSTAGES, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_streamk_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap a_map,
                        GemmEpilogue ep, float* __restrict__ partials, int* __restrict__ counters, int M, int N, int m_tiles,
                        int num_kb, int tiles, int w_stream_once) {
  using T = KindTraits<KIND>;
  constexpr int BK_ELEMS = GEMM_BK_BYTES / T::ELEM;
  constexpr int W_TILE = GEMM_BN * GEMM_BK_BYTES;
  constexpr int A_TILE = BM * GEMM_BK_BYTES;
  constexpr int STAGE = W_TILE + A_TILE;
  constexpr uint32_t TMEM_COLS = 2 * BM < 32 ? 32 : 2 * BM;  // two accumulator stages
  constexpr uint32_t IDESC = umma_idesc(T::CFMT, T::ABFMT, GEMM_BN, BM);
  static_assert(2 * BM <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM columns");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NSTAGES * STAGE);
  uint64_t* empty = full + NSTAGES;
  uint64_t* tfull = empty + NSTAGES;   // [2] accumulator stage ready for the epilogue
  uint64_t* tempty = tfull + 2;        // [2] accumulator stage drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* fix_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = gridDim.x, c = blockIdx.x;
  const long long U = (long long)tiles * num_kb;
  const long long u0 = span_begin(U, P, c), u1 = span_begin(U, P, c + 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&a_map);
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(&tfull[0], 1);
    mbar_init(&tfull[1], 1);
    mbar_init(&tempty[0], 4);
    mbar_init(&tempty[1], 4);
    fence_mbar_init();
    fence_proxy_async();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_launch();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && u1 > u0) {
      const uint64_t pol_w = w_stream_once ? policy_evict_first() : policy_evict_last();
      const uint64_t pol_a = policy_evict_last();  // activations are re-read by every CTA
      const int total = (int)(u1 - u0);
      const int npre = total < NSTAGES ? total : NSTAGES;
      // weights do not depend on any kernel of the step: fill the ring with W tiles before the dependency resolves
      for (int i = 0; i < npre; ++i) {
        const long long u = u0 + i;
        const int tile = (int)(u / num_kb), kb = (int)(u % num_kb);
        mbar_expect_tx(&full[i], STAGE);
        tma_load_2d_hint(smem + (size_t)i * STAGE, &w_map, &full[i], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN, pol_w);
      }
      griddep_wait();
      for (int i = 0; i < npre; ++i) {
        const long long u = u0 + i;
        const int tile = (int)(u / num_kb), kb = (int)(u % num_kb);
        tma_load_2d_hint(smem + (size_t)i * STAGE + W_TILE, &a_map, &full[i], kb * BK_ELEMS, (tile % m_tiles) * BM, pol_a);
      }
      for (int i = npre; i < total; ++i) {
        const int s = i % NSTAGES;
        const long long u = u0 + i;
        const int tile = (int)(u / num_kb), kb = (int)(u % num_kb);
        mbar_wait(&empty[s], ((i / NSTAGES) - 1) & 1);
        uint8_t* ws = smem + (size_t)s * STAGE;
        mbar_expect_tx(&full[s], STAGE);
        tma_load_2d_hint(ws, &w_map, &full[s], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN, pol_w);
        tma_load_2d_hint(ws + W_TILE, &a_map, &full[s], kb * BK_ELEMS, (tile % m_tiles) * BM, pol_a);
      }
    }
  } else if (warp == 1) {
    int i = 0, seg = 0;
    for (long long u = u0; u < u1; ++seg) {
      const int kb0 = (int)(u % num_kb);
      const long long left = u1 - u;
      const int nk = left < (long long)(num_kb - kb0) ? (int)left : num_kb - kb0;
      const int as = seg & 1;
      if (seg >= 2) mbar_wait(&tempty[as], ((seg >> 1) - 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + (uint32_t)(as * BM);
      for (int k = 0; k < nk; ++k, ++i) {
        const int s = i % NSTAGES;
        mbar_wait(&full[s], (i / NSTAGES) & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t wa = smem_u32(smem + (size_t)s * STAGE);
          const uint32_t aa = wa + W_TILE;
#pragma unroll
          for (int kk = 0; kk < GEMM_BK_BYTES / 32; ++kk)  // UMMA_K = 32 bytes of K
            umma_ss<KIND>(acc, umma_desc_sw128(wa + kk * 32), umma_desc_sw128(aa + kk * 32), IDESC, (uint32_t)((k | kk) != 0));
          tc_commit(&empty[s]);  // smem slot is free once these MMAs retire
        }
        __syncwarp();
      }
      if (lane == 0) tc_commit(&tfull[as]);
      __syncwarp();
      u += nk;
    }
  } else {
    // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31; lane = weight row of the tile
    const int lg = warp & 3;
    const int et = threadIdx.x - 64;  // 0..127
    griddep_wait();
    int seg = 0;
    for (long long u = u0; u < u1; ++seg) {
      const int tile = (int)(u / num_kb), kb0 = (int)(u % num_kb);
      const long long left = u1 - u;
      const int nk = left < (long long)(num_kb - kb0) ? (int)left : num_kb - kb0;
      const bool complete = kb0 == 0 && nk == num_kb;
      const int as = seg & 1;
      const int n_tile = tile / m_tiles, m0 = (tile % m_tiles) * BM;
      const int nl = lg * 32 + lane;
      const int n = n_tile * GEMM_BN + nl;
      const bool n_ok = n < N;
      mbar_wait(&tfull[as], (seg >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * BM);
      float* slot = partials + ((size_t)c * 2 + (seg == 0 ? 0 : 1)) * (size_t)(BM * GEMM_BN);
#pragma unroll 1
      for (int cc = 0; cc < BM / 16; ++cc) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
        if (complete) {
          epilogue_chunk<EPI>(ep, f, m0 + cc * 16, M, n, n_ok, N, lane);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) slot[(size_t)(cc * 16 + j) * GEMM_BN + nl] = f[j];
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
      if (!complete) {
        // park-and-count: the contributor that arrives last reduces all partials of the tile in contributor order
        const long long t0 = (long long)tile * num_kb;
        const int c_first = span_owner(U, P, t0), c_last = span_owner(U, P, t0 + num_kb - 1);
        __threadfence();
        named_bar_sync(1, 128);
        if (et == 0) {
          const int old = atomicAdd(&counters[c_first], 1);
          *fix_flag = old == c_last - c_first;
        }
        named_bar_sync(1, 128);
        const bool last = *fix_flag != 0;
        named_bar_sync(1, 128);  // everyone has read the flag before a later segment may rewrite it
        if (last) {
          __threadfence();
          if (et == 0) counters[c_first] = 0;  // leave the workspace ready for the next launch
#pragma unroll 1
          for (int mb = 0; mb < BM; mb += 16) {
            float f[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = 0.f;
            for (int cb = c_first; cb <= c_last; ++cb) {
              const bool first_seg = (int)(span_begin(U, P, cb) / num_kb) == tile;
              const float* ps = partials + ((size_t)cb * 2 + (first_seg ? 0 : 1)) * (size_t)(BM * GEMM_BN) + nl;
#pragma unroll
              for (int j = 0; j < 16; ++j) f[j] += __ldcg(ps + (size_t)(mb + j) * GEMM_BN);
            }
            epilogue_chunk<EPI>(ep, f, m0 + mb, M, n, n_ok, N, lane);
          }
        }
      }
      u += nk;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}


// ============================================================ decode schedule (M <= 128): cluster split-K
// grid = n_tiles * S CTAs, cluster = the S CTAs (S = 1, 2, 4, 8) that share weight tile blockIdx.x / S; cluster rank r
// accumulates k-blocks [num_kb*r/S, num_kb*(r+1)/S).  Reduction buffer `red` (S > 1) in every CTA:
//   red[src rank][token m][row] fp32, row = weight row inside this CTA's slice of 128/S rows,
// written by the peers with st.shared::cluster (32 lanes -> 32 consecutive floats), read locally after one cluster
// barrier.  EPI_SILU_MUL needs whole 32-row groups per slice: S <= 4.
template <int KIND, int BM, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, (BM <= 64 ? 2 : 1))
    gemm_decode_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap a_map,
                       GemmEpilogue ep, int M, int N, int num_kb, int nstages) {
  using T = KindTraits<KIND>;
  constexpr int BK_ELEMS = GEMM_BK_BYTES / T::ELEM;
  constexpr int W_TILE = GEMM_BN * GEMM_BK_BYTES;
  constexpr int A_TILE = BM * GEMM_BK_BYTES;
  constexpr int STAGE = W_TILE + A_TILE;
  constexpr uint32_t TMEM_COLS = BM < 32 ? 32 : BM;
  constexpr uint32_t IDESC = umma_idesc(T::CFMT, T::ABFMT, GEMM_BN, BM);
  constexpr int MAX_STAGES = 8;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = (int)cluster_nctarank(), r = (int)cluster_ctarank();
  float* red = reinterpret_cast<float*>(smem + (size_t)nstages * STAGE);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)nstages * STAGE + (S > 1 ? (size_t)GEMM_BN * BM * 4 : 0));
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x / S;
  const int kb0 = (int)((long long)num_kb * r / S), kb1 = (int)((long long)num_kb * (r + 1) / S);
  const int nk = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&a_map);
    for (int s = 0; s < nstages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tfull, 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_launch();
  // phase 1 of the cluster barrier: "this CTA runs" -- peers wait for it before their first DSMEM store
  if (S > 1) cluster_arrive_relaxed();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && nk > 0) {
      const uint64_t pol_w = policy_evict_first();  // every weight byte is read once per step
      const uint64_t pol_a = policy_evict_last();   // activations are re-read by every tile
      const int npre = nk < nstages ? nk : nstages;
      // weights do not depend on any kernel of the step: fill the ring with W tiles before the dependency resolves
      for (int i = 0; i < npre; ++i) {
        mbar_expect_tx(&full[i], STAGE);
        tma_load_2d_hint(smem + (size_t)i * STAGE, &w_map, &full[i], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
      }
      griddep_wait();
      for (int i = 0; i < npre; ++i)
        tma_load_2d_hint(smem + (size_t)i * STAGE + W_TILE, &a_map, &full[i], (kb0 + i) * BK_ELEMS, 0, pol_a);
      int s = 0, ph = 0;  // stage / parity of the empty barrier for refills
      for (int i = npre; i < nk; ++i) {
        mbar_wait(&empty[s], ph);
        uint8_t* ws = smem + (size_t)s * STAGE;
        mbar_expect_tx(&full[s], STAGE);
        tma_load_2d_hint(ws, &w_map, &full[s], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
        tma_load_2d_hint(ws + W_TILE, &a_map, &full[s], (kb0 + i) * BK_ELEMS, 0, pol_a);
        if (++s == nstages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    int s = 0, ph = 0;
    for (int i = 0; i < nk; ++i) {
      mbar_wait(&full[s], ph);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t wa = smem_u32(smem + (size_t)s * STAGE);
        const uint32_t aa = wa + W_TILE;
#pragma unroll
        for (int kk = 0; kk < GEMM_BK_BYTES / 32; ++kk)  // UMMA_K = 32 bytes of K
          umma_ss<KIND>(tmem_base, umma_desc_sw128(wa + kk * 32), umma_desc_sw128(aa + kk * 32), IDESC,
                        (uint32_t)((i | kk) != 0));
        tc_commit(&empty[s]);
      }
      __syncwarp();
      if (++s == nstages) {
        s = 0;
        ph ^= 1;
      }
    }
    if (lane == 0) tc_commit(tfull);
    __syncwarp();
  } else {
    // epilogue warps: warp w may touch TMEM lanes 32*(w%4) .. +31; lane = weight row of the tile
    const int lg = warp & 3;
    const int nl = lg * 32 + lane;
    griddep_wait();
    mbar_wait(tfull, 0);
    tc_fence_after();
    const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16);
    if (S == 1) {
      const int n = n_tile * GEMM_BN + nl;
#pragma unroll 1
      for (int cc = 0; cc < BM / 16; ++cc) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
        epilogue_chunk<EPI>(ep, f, cc * 16, M, n, n < N, N, lane);
      }
    } else {
      // push this CTA's partial of weight row nl to the CTA that owns the row (once every peer is known to run)
      cluster_wait();
      const int rows_per = GEMM_BN / S;
      const uint32_t dst = mapa_shared(smem_u32(red), (uint32_t)(nl / rows_per)) +
                           (uint32_t)(((size_t)r * BM * rows_per + (nl % rows_per)) * 4);
#pragma unroll 1
      for (int cc = 0; cc < BM / 16; ++cc) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float f = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
          st_cluster_f32(dst + (uint32_t)((cc * 16 + j) * rows_per * 4), f);
        }
      }
    }
  }
  tc_fence_before();
  __syncwarp();
  if (S > 1) {
    if (warp < 2) cluster_wait();  // phase 1 for the warps that do not push
    cluster_sync_all();            // phase 2: all partials have landed
    if (warp >= 2) {
      const int rows_per = GEMM_BN / S;
      const int et = threadIdx.x - 64;
      const int n_base = n_tile * GEMM_BN + r * rows_per;
      // element (m, row) of this CTA's slice: consecutive threads take consecutive rows (conflict-free smem reads,
      // contiguous global stores); S partials added in rank order
#pragma unroll 1
      for (int idx = et; idx < BM * rows_per; idx += 128) {
        const int row = idx % rows_per, m = idx / rows_per;
        float acc = 0.f;
        for (int q = 0; q < S; ++q) acc += red[((size_t)q * BM + m) * rows_per + row];
        const int n = n_base + row;
        const bool ok = m < M && n < N;
        if constexpr (EPI == EPI_LINEAR) {
          if (ok) store_out(ep, acc, m, n, N);
        } else {
          // rows_per is 32, 64 or 128 here: this warp holds one whole 32-row group of token m (lane = row & 31)
          const int I = N >> 1, n_groups = N >> 5;
          const float mine = ok ? __bfloat162float(__float2bfloat16_rn(linear_value(ep, acc, m, n, N))) : 0.f;
          const float other = __shfl_xor_sync(0xffffffffu, mine, 16);
          const bool is_gate = lane < 16;
          const float a = silu_mul_bf16(is_gate ? mine : other, is_gate ? other : mine);
          if (ok && is_gate)
            reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * I + (n >> 5) * 16 + lane] = __float2bfloat16_rn(a);
          const float x = warp_max(ok ? fabsf(a) : 0.f);
          if (lane == 0 && ok) ep.group_amax[(int64_t)m * n_groups + (n >> 5)] = x;
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace b200

using namespace b200;

namespace {

struct GemmPlan {
  bool decode;  // M <= 128: cluster split-K kernel; else persistent stream-K
  int bm, m_tiles, n_tiles, tiles, num_kb;
  int split, nstages;  // decode
  int ctas;            // stream-K
  size_t smem, ws_bytes;
};

int elem_bytes(int dt) { return dt == B200_BF16 ? 2 : 1; }
constexpr int stages_for(int bm) { return bm == 32 ? 10 : bm == 64 ? 8 : bm == 128 ? 6 : 4; }
constexpr size_t smem_for(int bm) {
  return (size_t)stages_for(bm) * (GEMM_BN * GEMM_BK_BYTES + bm * GEMM_BK_BYTES) + (2 * stages_for(bm) + 4) * 8 + 8 + 1024;
}
// decode kernel: ring depth chosen so that two CTAs fit on an SM (BM <= 64); the reduction buffer takes one stage's room
int decode_stages(int bm, int split) {
  if (bm == 32) return split > 1 ? 4 : 5;
  if (bm == 64) return split > 1 ? 3 : 4;
  return split > 1 ? 4 : 6;
}
size_t decode_smem(int bm, int split) {
  return 1024 + (size_t)decode_stages(bm, split) * (GEMM_BN + bm) * GEMM_BK_BYTES +
         (split > 1 ? (size_t)GEMM_BN * bm * 4 : 0) + 256;
}

// split_k: decode = cluster size along K (0 = auto; 1, 2, 4, 8); stream-K = cap on the persistent CTAs (0 = one per SM)
int plan_gemm(GemmPlan& p, int M, int N, int K, int ab_dtype, int split_k, int epi) {
  p.decode = M <= 128;
  p.bm = M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  p.m_tiles = (M + p.bm - 1) / p.bm;
  p.n_tiles = (N + GEMM_BN - 1) / GEMM_BN;
  p.tiles = p.m_tiles * p.n_tiles;
  int bk = GEMM_BK_BYTES / elem_bytes(ab_dtype);
  p.num_kb = (K + bk - 1) / bk;
  p.split = 1;
  p.nstages = 0;
  p.ctas = 0;
  if (p.decode) {
    const int max_split = epi == EPI_SILU_MUL ? 4 : 8;
    int s = split_k;
    if (s == 0) {
      // fill the CTA slots of the chip (two per SM for BM <= 64) without dropping below 2 k-blocks per CTA
      const long slots = (long)sm_count() * (p.bm <= 64 ? 2 : 1);
      s = 1;
      while (s * 2 <= max_split && (long)p.n_tiles * s * 2 <= slots && p.num_kb >= s * 4) s *= 2;
    }
    if (!(s == 1 || s == 2 || s == 4 || s == 8) || s > max_split || s > p.num_kb) {
      set_last_error("gemm: split_k %d invalid here (1, 2, 4%s; at most %d k-blocks)", split_k,
                     epi == EPI_SILU_MUL ? "" : ", 8", p.num_kb);
      return B200_ERR_INVALID_ARG;
    }
    p.split = s;
    p.nstages = decode_stages(p.bm, s);
    p.smem = decode_smem(p.bm, s);
    p.ws_bytes = 0;
    return B200_OK;
  }
  long long units = (long long)p.tiles * p.num_kb;
  int ctas = sm_count();
  if (split_k > 0 && split_k < ctas) ctas = split_k;
  if (ctas > GEMM_CNT_INTS) ctas = GEMM_CNT_INTS;
  if ((long long)ctas > units) ctas = (int)units;
  p.ctas = ctas;
  p.smem = smem_for(p.bm);
  p.ws_bytes = (size_t)GEMM_CNT_INTS * sizeof(int) + (size_t)ctas * 2 * p.bm * GEMM_BN * sizeof(float);
  return B200_OK;
}

template <int KIND, int EPI>
int launch_streamk(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, void* workspace, int M, int N,
                   const GemmPlan& p, cudaStream_t stream) {
  constexpr int BM = 256;
  auto kern = gemm_streamk_kernel<KIND, BM, stages_for(BM), EPI>;
  constexpr size_t smem = smem_for(BM);
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  int* counters = (int*)workspace;
  float* partials = (float*)((uint8_t*)workspace + GEMM_CNT_INTS * sizeof(int));
  B200_CUDA_OK(launch_kernel(kern, dim3(p.ctas), dim3(GEMM_THREADS), smem, stream, wm, am, ep, partials, counters, M, N,
                             p.m_tiles, p.num_kb, p.tiles, p.m_tiles == 1 ? 1 : 0));
  return B200_OK;
}

template <int KIND, int BM, int EPI>
int launch_decode(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, int M, int N, const GemmPlan& p,
                  cudaStream_t stream) {
  auto kern = gemm_decode_kernel<KIND, BM, EPI>;
  static bool attr_done = false;
  if (!attr_done) {
    const size_t mx = decode_smem(BM, 1) > decode_smem(BM, 2) ? decode_smem(BM, 1) : decode_smem(BM, 2);
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mx));
    attr_done = true;
  }
  B200_CUDA_OK(launch_kernel_cluster(kern, dim3(p.n_tiles * p.split), dim3(GEMM_THREADS), p.smem, p.split, stream, wm, am,
                                     ep, M, N, p.num_kb, p.nstages));
  return B200_OK;
}

template <int KIND, int EPI>
int dispatch_gemm(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, void* ws, int M, int N,
                  const GemmPlan& p, cudaStream_t stream) {
  if (!p.decode) return launch_streamk<KIND, EPI>(wm, am, ep, ws, M, N, p, stream);
  if (p.bm == 32) return launch_decode<KIND, 32, EPI>(wm, am, ep, M, N, p, stream);
  if (p.bm == 64) return launch_decode<KIND, 64, EPI>(wm, am, ep, M, N, p, stream);
  return launch_decode<KIND, 128, EPI>(wm, am, ep, M, N, p, stream);
}

int gemm_common(const void* a, const void* w, const GemmEpilogue& ep, int epi, int M, int N, int K, int ab_dtype,
                int split_k, void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && w && ep.out, "gemm: null pointer");
  B200_REQUIRE(ab_dtype == B200_BF16 || ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8, "gemm: bad ab_dtype %d", ab_dtype);
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  B200_REQUIRE(split_k >= 0, "gemm: bad split_k %d", split_k);
  int eb = elem_bytes(ab_dtype);
  B200_REQUIRE(((int64_t)K * eb) % 16 == 0, "gemm: K*elem_bytes must be a multiple of 16 (TMA row pitch), K=%d", K);
  GemmPlan p;
  int rc = plan_gemm(p, M, N, K, ab_dtype, split_k, epi);
  if (rc) return rc;
  if (p.ws_bytes && (!workspace || workspace_bytes < p.ws_bytes)) {
    set_last_error("gemm: workspace %zu < %zu bytes", workspace_bytes, p.ws_bytes);
    return B200_ERR_WORKSPACE;
  }
  CUtensorMapDataType dt = ab_dtype == B200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  int bk = GEMM_BK_BYTES / eb;
  CUtensorMap wm, am;
  rc = make_tmap_2d(&wm, w, dt, eb, (uint64_t)K, (uint64_t)N, (uint64_t)K * eb, bk, GEMM_BN);
  if (rc) return rc;
  rc = make_tmap_2d(&am, a, dt, eb, (uint64_t)K, (uint64_t)M, (uint64_t)K * eb, bk, p.bm);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (epi == EPI_LINEAR) {
    if (ab_dtype == B200_BF16) return dispatch_gemm<KIND_F16, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, s);
    if (ab_dtype == B200_FP8_E4M3) return dispatch_gemm<KIND_F8F6F4, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, s);
    return dispatch_gemm<KIND_I8, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, s);
  }
  if (ab_dtype == B200_BF16) return dispatch_gemm<KIND_F16, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, s);
  if (ab_dtype == B200_FP8_E4M3) return dispatch_gemm<KIND_F8F6F4, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, s);
  return dispatch_gemm<KIND_I8, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, s);
}

}  // namespace

extern "C" {

size_t b200_gemm_workspace_bytes(int M, int N, int K, int ab_dtype, int split_k) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  GemmPlan p;
  if (plan_gemm(p, M, N, K, ab_dtype, split_k, EPI_LINEAR)) return 0;
  return p.ws_bytes;
}

int b200_gemm(const void* a, const void* w, void* out, const float* a_scale, const float* w_scale, const void* bias,
              const void* residual, int M, int N, int K, int ab_dtype, int out_dtype, int split_k, void* workspace,
              size_t workspace_bytes, void* stream) {
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F32, "gemm: bad out_dtype %d", out_dtype);
  GemmEpilogue ep{a_scale, w_scale, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)residual, out, out_dtype == B200_F32,
                  nullptr};
  return gemm_common(a, w, ep, EPI_LINEAR, M, N, K, ab_dtype, split_k, workspace, workspace_bytes, stream);
}

int b200_gemm_silu_mul(const void* a, const void* w_gu, void* out_act, const float* a_scale, const float* w_scale,
                       const void* bias, float* group_amax, int M, int N2, int K, int ab_dtype, int split_k,
                       void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(group_amax, "gemm_silu_mul: group_amax required");
  B200_REQUIRE(N2 % 32 == 0, "gemm_silu_mul: 2*I must be a multiple of 32 (16 gate + 16 up rows per group), got %d", N2);
  GemmEpilogue ep{a_scale, w_scale, (const __nv_bfloat16*)bias, nullptr, out_act, 0, group_amax};
  return gemm_common(a, w_gu, ep, EPI_SILU_MUL, M, N2, K, ab_dtype, split_k, workspace, workspace_bytes, stream);
}

}  // extern "C"
