// Per-channel-scaled weight GEMM for the decoder linears (QKV, O, gate/up, down, lm_head) on tcgen05.
//
//   out[m][n] = (sum_k a[m][k] w[n][k]) * a_scale[m] * w_scale[n] + bias[n] + residual[m][n]
//
// Decode has M = batch (<= 128) and is bound by streaming W from HBM, so the tile is "swap-AB": the 128 weight rows of
// a tile fill the UMMA M slot, the (few) activation rows are the UMMA N.  Operands are staged by TMA
// (cp.async.bulk.tensor, 128B swizzle) in a ring of stages; one thread issues tcgen05.mma (kind::f8f6f4 / kind::f16 /
// kind::i8) into a TMEM accumulator; four epilogue warps read it back with tcgen05.ld, apply the per-token and
// per-channel scales and write bf16/fp32.  When N/128 tiles cannot fill the SMs the K loop is split and fp32
// partials are reduced by a second (deterministic, fixed-order) kernel.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = epilogue.
#include "common.cuh"

namespace b200 {

constexpr int GEMM_BN = 128;      // weight rows per tile  (UMMA M)
constexpr int GEMM_BK_BYTES = 128;  // one 128B swizzle row of K per stage
constexpr int GEMM_THREADS = 192;

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

struct GemmEpilogue {
  const float* a_scale;
  const float* w_scale;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  void* out;
  int out_f32;
};

__device__ __forceinline__ void store_out(const GemmEpilogue& ep, float acc, int m, int n, int N) {
  float y = acc;
  if (ep.a_scale) y *= ep.a_scale[m];
  if (ep.w_scale) y *= ep.w_scale[n];
  if (ep.bias) y += __bfloat162float(ep.bias[n]);
  if (ep.residual) y += __bfloat162float(ep.residual[(int64_t)m * N + n]);
  if (ep.out_f32)
    reinterpret_cast<float*>(ep.out)[(int64_t)m * N + n] = y;
  else
    reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * N + n] = __float2bfloat16_rn(y);
}

template <int KIND, int BM, int NSTAGES>
__global__ void __launch_bounds__(GEMM_THREADS)
    gemm_tn_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap a_map, GemmEpilogue ep,
                   float* __restrict__ partial, int M, int N, int num_kb, int kb_per_split) {
  using T = KindTraits<KIND>;
  constexpr int BK_ELEMS = GEMM_BK_BYTES / T::ELEM;
  constexpr int W_TILE = GEMM_BN * GEMM_BK_BYTES;
  constexpr int A_TILE = BM * GEMM_BK_BYTES;
  constexpr int STAGE = W_TILE + A_TILE;
  constexpr uint32_t TMEM_COLS = BM < 32 ? 32 : BM;
  constexpr uint32_t IDESC = umma_idesc(T::CFMT, T::ABFMT, GEMM_BN, BM);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NSTAGES * STAGE);
  uint64_t* empty = full + NSTAGES;
  uint64_t* tmem_full = empty + NSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * GEMM_BN, m0 = blockIdx.y * BM;
  const int kb0 = blockIdx.z * kb_per_split;
  int nkb = num_kb - kb0;
  if (nkb > kb_per_split) nkb = kb_per_split;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&a_map);
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
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
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      const uint64_t pol_w = policy_evict_first();  // weights are streamed once per step
      const uint64_t pol_a = policy_evict_last();   // activations are re-read by every n-tile
      for (int i = 0; i < nkb; ++i) {
        const int s = i % NSTAGES;
        if (i >= NSTAGES) mbar_wait(&empty[s], ((i / NSTAGES) - 1) & 1);
        uint8_t* ws = smem + (size_t)s * STAGE;
        mbar_expect_tx(&full[s], STAGE);
        tma_load_2d_hint(ws, &w_map, &full[s], (kb0 + i) * BK_ELEMS, n0, pol_w);
        tma_load_2d_hint(ws + W_TILE, &a_map, &full[s], (kb0 + i) * BK_ELEMS, m0, pol_a);
      }
    }
  } else if (warp == 1) {
    for (int i = 0; i < nkb; ++i) {
      const int s = i % NSTAGES;
      mbar_wait(&full[s], (i / NSTAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t wa = smem_u32(smem + (size_t)s * STAGE);
        const uint32_t aa = wa + W_TILE;
#pragma unroll
        for (int k = 0; k < GEMM_BK_BYTES / 32; ++k)  // UMMA_K = 32 bytes of K
          umma_ss<KIND>(tmem_base, umma_desc_sw128(wa + k * 32), umma_desc_sw128(aa + k * 32), IDESC, (uint32_t)((i | k) != 0));
        tc_commit(&empty[s]);  // smem slot is free once these MMAs retire
      }
      __syncwarp();
    }
    if (lane == 0) tc_commit(tmem_full);
    __syncwarp();
  } else {
    // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31; lane = weight row of the tile
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int lg = warp & 3;
    const int n = n0 + lg * 32 + lane;
    const bool n_ok = n < N;
#pragma unroll 1
    for (int c = 0; c < BM / 16; ++c) {
      uint32_t v[16];
      tmem_ld_32x32b_x16(tmem_base + ((uint32_t)(lg * 32) << 16) + c * 16, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int m = m0 + c * 16 + j;
        if (m < M && n_ok) {
          float acc = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
          if (partial)
            partial[((int64_t)blockIdx.z * M + m) * N + n] = acc;
          else
            store_out(ep, acc, m, n, N);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// fixed-order reduction of split-K partials + epilogue
__global__ void splitk_reduce_kernel(const float* __restrict__ partial, GemmEpilogue ep, int M, int N, int splits) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)M * N;
  if (idx >= total) return;
  float acc = 0.f;
  for (int z = 0; z < splits; ++z) acc += partial[(int64_t)z * total + idx];
  store_out(ep, acc, (int)(idx / N), (int)(idx % N), N);
}

}  // namespace b200

using namespace b200;

namespace {

struct GemmPlan {
  int bm, stages, m_tiles, n_tiles, num_kb, splits, kb_per_split;
};

int elem_bytes(int dt) { return dt == B200_BF16 ? 2 : 1; }

GemmPlan plan_gemm(int M, int N, int K, int ab_dtype, int split_k) {
  GemmPlan p;
  p.bm = M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  p.m_tiles = (M + p.bm - 1) / p.bm;
  p.n_tiles = (N + GEMM_BN - 1) / GEMM_BN;
  int bk = GEMM_BK_BYTES / elem_bytes(ab_dtype);
  p.num_kb = (K + bk - 1) / bk;
  int tiles = p.m_tiles * p.n_tiles;
  int sms = sm_count();
  int s = split_k;
  if (s <= 0) {
    s = 1;
    if (tiles * 4 <= sms * 3) {  // tiles cannot fill the SMs: split K so that ~one CTA lands on every SM
      s = sms / tiles;
      int max_s = p.num_kb / 4;  // keep >= 4 k-blocks per split
      if (s > max_s) s = max_s;
      if (s < 1) s = 1;
    }
  }
  if (s > p.num_kb) s = p.num_kb;
  p.kb_per_split = (p.num_kb + s - 1) / s;
  p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;  // every split owns >= 1 k-block
  // HBM-bound shapes want every CTA resident at once (they then share bandwidth evenly): if one CTA per SM cannot
  // hold all tiles, halve the ring so two CTAs fit per SM.
  int ctas = tiles * p.splits;
  p.stages = (p.bm <= 64 && ctas <= sms) ? 8 : 4;
  if (p.bm <= 64 && p.kb_per_split < 8) p.stages = 4;
  return p;
}

template <int KIND, int BM, int NSTAGES>
int launch_gemm(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, float* partial, int M, int N,
                const GemmPlan& p, cudaStream_t stream) {
  auto kern = gemm_tn_kernel<KIND, BM, NSTAGES>;
  constexpr size_t smem = (size_t)NSTAGES * (GEMM_BN * GEMM_BK_BYTES + BM * GEMM_BK_BYTES) + (2 * NSTAGES + 1) * 8 + 16 + 1024;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  dim3 grid(p.n_tiles, p.m_tiles, p.splits);
  kern<<<grid, GEMM_THREADS, smem, stream>>>(wm, am, ep, partial, M, N, p.num_kb, p.kb_per_split);
  B200_CUDA_OK(cudaGetLastError());
  B200_LAUNCHED(1);
  return B200_OK;
}

template <int KIND>
int dispatch_gemm(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, float* partial, int M, int N,
                  const GemmPlan& p, cudaStream_t stream) {
  if (p.bm == 32) {
    if (p.stages == 8) return launch_gemm<KIND, 32, 8>(wm, am, ep, partial, M, N, p, stream);
    return launch_gemm<KIND, 32, 4>(wm, am, ep, partial, M, N, p, stream);
  }
  if (p.bm == 64) {
    if (p.stages == 8) return launch_gemm<KIND, 64, 8>(wm, am, ep, partial, M, N, p, stream);
    return launch_gemm<KIND, 64, 4>(wm, am, ep, partial, M, N, p, stream);
  }
  if (p.bm == 128) return launch_gemm<KIND, 128, 4>(wm, am, ep, partial, M, N, p, stream);
  return launch_gemm<KIND, 256, 4>(wm, am, ep, partial, M, N, p, stream);
}

}  // namespace

extern "C" {

size_t b200_gemm_workspace_bytes(int M, int N, int K, int ab_dtype, int split_k) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  GemmPlan p = plan_gemm(M, N, K, ab_dtype, split_k);
  return p.splits > 1 ? (size_t)p.splits * M * N * sizeof(float) : 0;
}

int b200_gemm(const void* a, const void* w, void* out, const float* a_scale, const float* w_scale, const void* bias,
              const void* residual, int M, int N, int K, int ab_dtype, int out_dtype, int split_k, void* workspace,
              size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && w && out, "gemm: null pointer");
  B200_REQUIRE(ab_dtype == B200_BF16 || ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8, "gemm: bad ab_dtype %d", ab_dtype);
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F32, "gemm: bad out_dtype %d", out_dtype);
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  int eb = elem_bytes(ab_dtype);
  B200_REQUIRE(((int64_t)K * eb) % 16 == 0, "gemm: K*elem_bytes must be a multiple of 16 (TMA row pitch), K=%d", K);
  GemmPlan p = plan_gemm(M, N, K, ab_dtype, split_k);
  float* partial = nullptr;
  if (p.splits > 1) {
    size_t need = (size_t)p.splits * M * N * sizeof(float);
    if (!workspace || workspace_bytes < need) {
      set_last_error("gemm: workspace %zu < %zu bytes (splits=%d)", workspace_bytes, need, p.splits);
      return B200_ERR_WORKSPACE;
    }
    partial = (float*)workspace;
  }
  CUtensorMapDataType dt = ab_dtype == B200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  int bk = GEMM_BK_BYTES / eb;
  CUtensorMap wm, am;
  int rc = make_tmap_2d(&wm, w, dt, eb, (uint64_t)K, (uint64_t)N, (uint64_t)K * eb, bk, GEMM_BN);
  if (rc) return rc;
  rc = make_tmap_2d(&am, a, dt, eb, (uint64_t)K, (uint64_t)M, (uint64_t)K * eb, bk, p.bm);
  if (rc) return rc;
  GemmEpilogue ep{a_scale, w_scale, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)residual, out, out_dtype == B200_F32};
  cudaStream_t s = (cudaStream_t)stream;
  if (ab_dtype == B200_BF16)
    rc = dispatch_gemm<KIND_F16>(wm, am, ep, partial, M, N, p, s);
  else if (ab_dtype == B200_FP8_E4M3)
    rc = dispatch_gemm<KIND_F8F6F4>(wm, am, ep, partial, M, N, p, s);
  else
    rc = dispatch_gemm<KIND_I8>(wm, am, ep, partial, M, N, p, s);
  if (rc) return rc;
  if (partial) {
    int64_t total = (int64_t)M * N;
    splitk_reduce_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(partial, ep, M, N, p.splits);
    B200_CUDA_OK(cudaGetLastError());
  B200_LAUNCHED(1);
  }
  return B200_OK;
}

}  // extern "C"
