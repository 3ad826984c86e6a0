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
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

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

// INT8 weight-only (W8A16): bf16 activations, int8 weights streamed from HBM as int8 and widened to bf16 inside the
// kernel (exact: |w| <= 127), then kind::f16.  The per-channel scale is applied in the epilogue like the other kinds.
constexpr int KIND_WO8 = 3;
template <>
struct KindTraits<KIND_WO8> {
  static constexpr int ELEM = 2, CFMT = 1, ABFMT = 1;
};
constexpr int RAW_TILE = GEMM_BN * (GEMM_BK_BYTES / 2);  // 128 rows x 64 int8 of one k-block (un-swizzled)

// Widen one raw int8 weight tile [128 rows][64 B] into the K-major 128B-swizzled bf16 tile the UMMA descriptor expects
// (row pitch 128 B, 8-row groups 1024 B apart, 16-byte chunk c of row r stored at chunk c ^ (r & 7)).
// 128 threads; each handles 4 (row, 16-byte) pieces: consecutive threads read consecutive 16 B and the two output
// chunks of neighbouring rows fall into disjoint bank groups.
__device__ __forceinline__ void widen_w_tile(const uint8_t* __restrict__ raw, uint8_t* __restrict__ dst, int t) {
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int id = t + it * 128;
    const int row = id >> 2, c4 = id & 3;
    const uint4 v = *reinterpret_cast<const uint4*>(raw + row * 64 + c4 * 16);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[8];
    // int8 -> bf16 without the conversion unit: byte b (two's complement) XOR 0x80 = b + 128 as unsigned; dropped into the
    // low mantissa byte of 2^23 it reads as the float 8388608 + 128 + b; subtracting 8388736 leaves float(b) exactly, and a
    // value of at most 7 significant bits is its own bf16 truncation, so the high half-words ARE the bf16 codes.
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t x = w[j] ^ 0x80808080u;
      const float f0 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7440)) - 8388736.f;
      const float f1 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7441)) - 8388736.f;
      const float f2 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7442)) - 8388736.f;
      const float f3 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7443)) - 8388736.f;
      o[2 * j] = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
      o[2 * j + 1] = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
    }
    uint8_t* rowp = dst + (row >> 3) * 1024 + (row & 7) * 128;
    *reinterpret_cast<uint4*>(rowp + (((2 * c4) ^ (row & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<uint4*>(rowp + (((2 * c4 + 1) ^ (row & 7)) << 4)) = make_uint4(o[4], o[5], o[6], o[7]);
  }
}

// EPI_ROPE: EPI_LINEAR whose split-K owner epilogue also rotates q/k and scatters k/v (decode kernel, S > 1 only); its own
// instantiation so the plain linears keep their register budget (80 vs 137)
// EPI_SILU_QUANT: EPI_SILU_MUL whose CTAs also agree on the per-token amax through a grid-wide barrier and write the
// quantised activation themselves (decode kernel, S == 1, all CTAs co-resident; b200_gemm_silu_mul_quant)
enum { EPI_LINEAR = 0, EPI_SILU_MUL = 1, EPI_ROPE = 2, EPI_SILU_QUANT = 3 };

struct GemmEpilogue {
  const float* a_scale;
  const float* w_scale;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  void* out;
  int out_f32;
  float* group_amax;  // EPI_SILU_MUL: [M][N/32] max |act| over the 16 channels of each interleaved row group
  // EPI_SILU_QUANT: `out` is the quantised activation [M][N/2] (one byte per element), out_scale [M] its per-token scale;
  // fq_state = { uint64 tok_amax[128] (epoch << 32 | float bits), uint32 arrive counter, uint32 epoch }: zero once, then
  // self-maintaining (the epoch only grows, so stale maxima of earlier launches always lose the 64-bit atomicMax)
  float* out_scale;
  unsigned long long* fq_state;
  // EPI_ROPE (decode kernel with a split-K cluster only; b200_gemm_rope_kv_write): the output row is a packed qkv row
  // [Hq*D | Hkv*D | Hkv*D]; the token's owner rotates q and k (rotate-half, same roundings as rope_kv_write_kernel)
  // and scatters k, v into the block-KV pools by slot.
  struct RopeKv {
    const int64_t* positions;
    const int64_t* slot_mapping;
    const float* cos_t;
    const float* sin_t;
    __nv_bfloat16* k_pool;
    __nv_bfloat16* v_pool;
    long long num_slots;
    int max_pos, Hq, Hkv, D, bs;
  } rope;
};

// out = ((acc * a_scale) * w_scale + bias) + residual, every step rounded on its own (no FMA contraction), so that the
// decode, prefill and fused-epilogue paths produce identical bits for the same accumulator
__device__ __forceinline__ float scaled_value(float acc, float a, float w, bool has_bias, float b, bool has_res, float r) {
  float y = __fmul_rn(__fmul_rn(acc, a), w);
  if (has_bias) y = __fadd_rn(y, b);
  if (has_res) y = __fadd_rn(y, r);
  return y;
}
// span of CTA c: units [U*c/P, U*(c+1)/P); owner(x) = the CTA whose span holds unit x
__device__ __forceinline__ long long span_begin(long long U, int P, int c) { return U * c / P; }
__device__ __forceinline__ int span_owner(long long U, int P, long long x) { return (int)(((x + 1) * P - 1) / U); }

// SiLU(gate) * up for one channel; both inputs already rounded to bf16 like the unfused path (b200_gemm + b200_silu_mul)
__device__ __forceinline__ float silu_mul_bf16(float g, float u) {
  return __bfloat162float(__float2bfloat16_rn(silu_f32(g) * u));
}

// Decode-kernel epilogue of one 16-token column chunk held in v[] by the lane that owns weight row n (EPI_LINEAR), or by
// the lane pair (l, l^16) that owns the gate and the up row of one channel (EPI_SILU_MUL; W rows interleaved 16 gate / 16
// up per warp), with every epilogue operand already staged: a_s[] per-token scales in shared memory, this lane's channel scale /
// bias in registers, the residual column of this lane in shared memory (resid[m * resid_stride]).
template <int EPI>
__device__ __forceinline__ void epilogue_chunk_staged(const GemmEpilogue& ep, const float (&v)[16], int m_base, int M, int n,
                                                      bool n_ok, int N, int lane, const float* a_s, float w_sc, float bias_v,
                                                      const __nv_bfloat16* resid, int resid_stride) {
  const bool hb = ep.bias != nullptr, hr = ep.residual != nullptr;
  if constexpr (EPI == EPI_LINEAR) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int m = m_base + j;
      const float y = scaled_value(v[j], a_s[m], w_sc, hb, bias_v, hr, hr ? __bfloat162float(resid[m * resid_stride]) : 0.f);
      if (m < M && n_ok) {
        if (ep.out_f32)
          reinterpret_cast<float*>(ep.out)[(int64_t)m * N + n] = y;
        else
          reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * N + n] = __float2bfloat16_rn(y);
      }
    }
  } else {
    const int I = N >> 1, n_groups = N >> 5;
    const int ch = (n >> 5) * 16 + (lane & 15);
    float mine[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int m = m_base + j;
      float y = 0.f;
      if (m < M && n_ok) y = __bfloat162float(__float2bfloat16_rn(scaled_value(v[j], a_s[m], w_sc, hb, bias_v, false, 0.f)));
      mine[j] = y;
    }
    const bool hi = lane >= 16;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float send = hi ? mine[j] : mine[j + 8];
      const float got = __shfl_xor_sync(0xffffffffu, send, 16);
      const float g = hi ? got : mine[j];
      const float u = hi ? mine[j + 8] : got;
      const int m = m_base + (hi ? 8 : 0) + j;
      const bool ok = m < M && n_ok;
      const float a = silu_mul_bf16(g, u);
      if (ok) reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * I + ch] = __float2bfloat16_rn(a);
      float x = ok ? fabsf(a) : 0.f;
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
      if ((lane & 15) == 0 && ok) ep.group_amax[(int64_t)m * n_groups + (n >> 5)] = x;
    }
  }
}

// Stream-K (prefill) epilogue of one 16-token chunk, transposed through a per-warp shared-memory tile so that global
// memory sees 16-byte vectors instead of one 2-byte element per lane:
//   TMEM hands lane l the weight row n0w + l of 16 tokens; the warp writes the scaled values (a_scale, w_scale, bias
//   applied, fp32) as xp[token][row] (pitch XP_PITCH floats), then lane (token = l / 4 + 8 it, rows 8 (l % 4) ..) adds the
//   residual it loads as one 16-byte vector and stores 8 bf16 (or 8 fp32).  Per lane and chunk: 16 st.shared + 2 x
//   (2 ld.shared.v4 + ld.global.v4 + st.global.v4) instead of 16 x (4 scalar global loads + a 2-byte store) -- measured
//   on Llama-3.1-8B prefill (profiles/r02_ttft_breakdown_*.txt): o_proj 322 us, down 392 us at M = 2048 before.
// Same roundings as scaled_value().  EPI_SILU_MUL: the activation of 16 channels x 16 tokens goes through the tile as
// bf16 and leaves as one 16-byte store per lane.
constexpr int XP_PITCH = 36;                       // floats; 36 keeps the float4 reads of a quarter-warp on distinct banks
constexpr int XP_WARP_FLOATS = 16 * XP_PITCH;      // one warp's tile
template <int EPI>
__device__ __forceinline__ void epilogue_chunk_xpose(const GemmEpilogue& ep, const float (&v)[16], int m_base, int ml_base,
                                                     int M, int n0w, int N, int lane, const float* a_s, float w_sc,
                                                     float bias_v, float* xp) {
  const bool hb = ep.bias != nullptr, hr = ep.residual != nullptr;
  const int n = n0w + lane;
  const bool n_ok = n < N;
  if constexpr (EPI == EPI_LINEAR) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float y = __fmul_rn(__fmul_rn(v[j], a_s[ml_base + j]), w_sc);
      if (hb) y = __fadd_rn(y, bias_v);
      xp[j * XP_PITCH + lane] = y;
    }
    __syncwarp();
    const bool vec = (N & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 &&
                     (!hr || (reinterpret_cast<uintptr_t>(ep.residual) & 15) == 0);
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int mm = it * 8 + (lane >> 2), c8 = (lane & 3) * 8;
      const int m = m_base + mm, nn = n0w + c8;
      if (m >= M || nn >= N) continue;
      const float4 lo = *reinterpret_cast<const float4*>(xp + mm * XP_PITCH + c8);
      const float4 hi = *reinterpret_cast<const float4*>(xp + mm * XP_PITCH + c8 + 4);
      float y[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      const int64_t off = (int64_t)m * N + nn;
      if (vec) {
        if (hr) {
          const uint4 rv = *reinterpret_cast<const uint4*>(ep.residual + off);
          const __nv_bfloat162* rh = reinterpret_cast<const __nv_bfloat162*>(&rv);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            y[2 * e] = __fadd_rn(y[2 * e], __low2float(rh[e]));
            y[2 * e + 1] = __fadd_rn(y[2 * e + 1], __high2float(rh[e]));
          }
        }
        if (ep.out_f32) {
          float* o = reinterpret_cast<float*>(ep.out) + off;
          *reinterpret_cast<float4*>(o) = make_float4(y[0], y[1], y[2], y[3]);
          *reinterpret_cast<float4*>(o + 4) = make_float4(y[4], y[5], y[6], y[7]);
        } else {
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + off) =
              make_uint4(pack_bf16(y[0], y[1]), pack_bf16(y[2], y[3]), pack_bf16(y[4], y[5]), pack_bf16(y[6], y[7]));
        }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          if (nn + e < N) {
            float z = y[e];
            if (hr) z = __fadd_rn(z, __bfloat162float(ep.residual[off + e]));
            if (ep.out_f32)
              reinterpret_cast<float*>(ep.out)[off + e] = z;
            else
              reinterpret_cast<__nv_bfloat16*>(ep.out)[off + e] = __float2bfloat16_rn(z);
          }
        }
      }
    }
    __syncwarp();  // the tile is rewritten by the next chunk
  } else {
    // this lane's row: gate (lane < 16) or up (lane >= 16) of channel ch; both rounded to bf16 like the unfused path
    const int I = N >> 1, n_groups = N >> 5;
    float mine[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float y = __fmul_rn(__fmul_rn(v[j], a_s[ml_base + j]), w_sc);
      if (hb) y = __fadd_rn(y, bias_v);
      mine[j] = (m_base + j < M && n_ok) ? __bfloat162float(__float2bfloat16_rn(y)) : 0.f;
    }
    const bool hi_half = lane >= 16;  // lanes < 16 finish tokens 0..7 of the chunk, lanes >= 16 finish tokens 8..15
    __nv_bfloat16* xb = reinterpret_cast<__nv_bfloat16*>(xp);  // [16 tokens][16 channels] bf16
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float send = hi_half ? mine[j] : mine[j + 8];
      const float got = __shfl_xor_sync(0xffffffffu, send, 16);
      const float g = hi_half ? got : mine[j];
      const float u = hi_half ? mine[j + 8] : got;
      const int mm = (hi_half ? 8 : 0) + j;
      const float a = silu_mul_bf16(g, u);
      xb[mm * 16 + (lane & 15)] = __float2bfloat16_rn((m_base + mm < M && n_ok) ? a : 0.f);
    }
    __syncwarp();
    {
      const int mm = lane >> 1, c8 = (lane & 1) * 8;
      const int m = m_base + mm;
      const int ch = (n0w >> 5) * 16 + c8;
      // the lane pair (2 mm, 2 mm + 1) holds the 16 channels of token mm: their maximum is the group's amax
      const uint4 val = *reinterpret_cast<const uint4*>(xb + mm * 16 + c8);
      const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&val);
      float x = 0.f;
#pragma unroll
      for (int e = 0; e < 4; ++e) x = fmaxf(x, fmaxf(fabsf(__low2float(h2[e])), fabsf(__high2float(h2[e]))));
      x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
      if (m < M && n0w < N) {  // N2 % 32 == 0: a 32-row group is inside N or outside as a whole
        if ((lane & 1) == 0) ep.group_amax[(int64_t)m * n_groups + (n0w >> 5)] = x;
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)m * I + ch;
        if ((I & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0) {
          *reinterpret_cast<uint4*>(dst) = val;
        } else {
          const __nv_bfloat16* sv = reinterpret_cast<const __nv_bfloat16*>(&val);
#pragma unroll
          for (int e = 0; e < 8; ++e) dst[e] = sv[e];
        }
      }
    }
    __syncwarp();
  }
}

template <int KIND, int BM, int NSTAGES, int EPI>
__global__ void __launch_bounds__(KIND == KIND_WO8 ? GEMM_THREADS + 128 : GEMM_THREADS, 1)
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
  constexpr bool WO = KIND == KIND_WO8;
  constexpr int MK = KIND == KIND_WO8 ? (int)KIND_F16 : KIND;
  static_assert(2 * BM <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM columns");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* raw_ring = smem + (size_t)NSTAGES * STAGE;  // weight-only: NSTAGES raw int8 weight tiles
  uint64_t* full = reinterpret_cast<uint64_t*>(raw_ring + (WO ? (size_t)NSTAGES * RAW_TILE : 0));
  uint64_t* empty = full + NSTAGES;
  uint64_t* full_raw = empty + NSTAGES;   // weight-only: raw tile landed
  uint64_t* empty_raw = full_raw + NSTAGES;  // weight-only: raw tile widened
  uint64_t* full_cvt = empty_raw + NSTAGES;  // weight-only: bf16 operand tile written
  uint64_t* tfull = full_cvt + NSTAGES;   // [2] accumulator stage ready for the epilogue
  uint64_t* tempty = tfull + 2;        // [2] accumulator stage drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* fix_flag = reinterpret_cast<int*>(tmem_slot + 1);
  float* a_s_tile = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(fix_flag + 1) + 15) & ~uintptr_t(15));  // [BM]
  float* xpose = a_s_tile + BM;  // [4 epilogue warps][16][XP_PITCH]

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
      if (WO) {
        mbar_init(&full_raw[s], 1);
        mbar_init(&empty_raw[s], 1);
        mbar_init(&full_cvt[s], 1);
      }
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
        if constexpr (WO) {
          mbar_expect_tx(&full[i], A_TILE);
          mbar_expect_tx(&full_raw[i], RAW_TILE);
          tma_load_2d_hint(raw_ring + (size_t)i * RAW_TILE, &w_map, &full_raw[i], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN,
                           pol_w);
        } else {
          mbar_expect_tx(&full[i], STAGE);
          tma_load_2d_hint(smem + (size_t)i * STAGE, &w_map, &full[i], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN, pol_w);
        }
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
        if constexpr (WO) {
          mbar_wait(&empty_raw[s], ((i / NSTAGES) - 1) & 1);
          mbar_expect_tx(&full_raw[s], RAW_TILE);
          tma_load_2d_hint(raw_ring + (size_t)s * RAW_TILE, &w_map, &full_raw[s], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN,
                           pol_w);
          mbar_expect_tx(&full[s], A_TILE);
        } else {
          mbar_expect_tx(&full[s], STAGE);
          tma_load_2d_hint(ws, &w_map, &full[s], kb * BK_ELEMS, (tile / m_tiles) * GEMM_BN, pol_w);
        }
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
        if constexpr (WO) mbar_wait(&full_cvt[s], (i / NSTAGES) & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t wa = smem_u32(smem + (size_t)s * STAGE);
          const uint32_t aa = wa + W_TILE;
#pragma unroll
          for (int kk = 0; kk < GEMM_BK_BYTES / 32; ++kk)  // UMMA_K = 32 bytes of K
            umma_ss<MK>(acc, umma_desc_sw128(wa + kk * 32), umma_desc_sw128(aa + kk * 32), IDESC, (uint32_t)((k | kk) != 0));
          tc_commit(&empty[s]);  // smem slot is free once these MMAs retire
        }
        __syncwarp();
      }
      if (lane == 0) tc_commit(&tfull[as]);
      __syncwarp();
      u += nk;
    }
  } else if (WO && warp >= 6) {
    // weight-only: warps 6..9 widen raw int8 weight tiles into the bf16 operand half of each stage
    const int ct = threadIdx.x - 192;  // 0..127
    const int total = (int)(u1 - u0);
    for (int i = 0; i < total; ++i) {
      const int s = i % NSTAGES;
      mbar_wait(&full_raw[s], (i / NSTAGES) & 1);
      if (i >= NSTAGES) mbar_wait(&empty[s], ((i / NSTAGES) - 1) & 1);
      widen_w_tile(raw_ring + (size_t)s * RAW_TILE, smem + (size_t)s * STAGE, ct);
      fence_proxy_async();
      named_bar_sync(2, 128);
      if (ct == 0) {
        mbar_arrive(&full_cvt[s]);
        mbar_arrive(&empty_raw[s]);
      }
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
      const int n0w = n_tile * GEMM_BN + lg * 32;
      float* xp = xpose + lg * XP_WARP_FLOATS;
      // per-tile epilogue operands: the token scales of this m-tile in shared memory, this lane's channel scale / bias in
      // registers (the MMAs of the tile are still running)
      named_bar_sync(1, 128);  // the previous segment's readers of a_s_tile are done
      for (int m = et; m < BM; m += 128) a_s_tile[m] = (ep.a_scale && m0 + m < M) ? ep.a_scale[m0 + m] : 1.f;
      const float w_sc = (ep.w_scale && n_ok) ? ep.w_scale[n] : 1.f;
      const float bias_v = (ep.bias && n_ok) ? __bfloat162float(ep.bias[n]) : 0.f;
      named_bar_sync(1, 128);
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
          if (m0 + cc * 16 < M) epilogue_chunk_xpose<EPI>(ep, f, m0 + cc * 16, cc * 16, M, n0w, N, lane, a_s_tile, w_sc, bias_v, xp);
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
            if (m0 + mb < M) epilogue_chunk_xpose<EPI>(ep, f, m0 + mb, mb, M, n0w, N, lane, a_s_tile, w_sc, bias_v, xp);
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
// grid = n_tiles * S CTAs, cluster = the S CTAs (S = 1 .. 8) that share weight tile blockIdx.x / S; cluster rank r
// accumulates k-blocks [num_kb*r/S, num_kb*(r+1)/S).
//
// Measured on B200 (profiles/r01_gemm_sweep_*.json): one SM ingests ~55 GB/s through TMA however deep its ring is, and the
// activation tiles (L2 hits) count against that like the weight tiles do.  So the time of a decode GEMM is a fixed ~6 us
// of dependent latencies plus (bytes received by the busiest SM) / 55 GB/s: the schedule's job is to spread the
// (tile, k-block) units evenly over all 148 SMs -- hence the free choice of S -- and to keep the fixed part short.
//
// Shared memory: a ring of `nst` stages, weight tile (16 KB) and activation tile in separate arrays with separate
// barriers; all weight stages are issued before griddepcontrol.wait (weights are constants of the step), the activation
// tiles after it.  While the ring streams, the four epilogue warps stage everything the epilogue needs from global
// memory (per-token scales, per-channel scale / bias, the residual tile, the DSMEM address of every token's owner),
// so nothing between the last MMA and the output stores waits on a global load.
//
// Split-K reduction (S > 1): the M tokens are divided among the cluster, rank q owns tokens [M*q/S, M*(q+1)/S).
// Every CTA stages its fp32 partial tile [token][128 rows] in its own (dead) ring memory and sends each owner its slab
// with ONE bulk DSMEM copy (cp.async.bulk.shared::cluster) into
//   red[src rank][owned token][128 rows]   of the owner, completing on the owner's mbarrier (measured: 64 scalar
// st.shared::cluster per thread took 2.8 us per GEMM); then the owner adds the S partials in rank order (bit-reproducible), applies the epilogue on
// float4 row groups and stores whole 256 B output rows.
constexpr int DEC_MAX_ST = 12;

// Split-K owner epilogue with RoPE + KV scatter fused (b200_gemm_rope_kv_write; host guarantees N == (Hq + 2 Hkv) * D,
// N % 128 == 0, D in {64, 128}, bf16 out, no residual).  Thread (tq, t4) finishes tokens tq, tq+4, ... of this CTA's
// slab and holds dims d0 .. d0+3 of one head; its rotate-half partner (d0 +- D/2) sits in lane ^ (D/8).  The position /
// slot words and the cos / sin rows of up to RI tokens are loaded up front, back to back, so the loop pays two L2 round
// trips in total instead of two per token (the output stores would otherwise fence the next token's loads).
// Roundings are those of rope_kv_write_kernel: bf16(y), products and sums rounded separately.
__device__ __forceinline__ void rope_reduce_store(const GemmEpilogue& ep, const float* __restrict__ red, int S, int tp_max,
                                                  int m_lo, int cnt, int tq, int t4, int n0, int N,
                                                  const float* __restrict__ a_s, const float (&w4)[4],
                                                  const float (&b4)[4], bool hb) {
  constexpr int RI = 4;
  const auto& rp = ep.rope;
  const int half = rp.D >> 1;
  const int head = n0 / rp.D, d0 = n0 - head * rp.D;
  const bool hi_half = d0 >= half, roped = head < rp.Hq + rp.Hkv, is_q = head < rp.Hq, is_k = roped && !is_q;
  const int dcol = hi_half ? d0 - half : d0;
  __nv_bfloat16* pool = is_k ? rp.k_pool : rp.v_pool;
  const int kh = head - rp.Hq - (is_k ? 0 : rp.Hkv);
  for (int base = tq; base < cnt; base += 4 * RI) {  // warp-uniform trip counts: every lane reaches the shuffles
    long long pos[RI], slot[RI];
    float4 c4[RI], s4[RI];
#pragma unroll
    for (int i = 0; i < RI; ++i) {
      const int ml = base + 4 * i, m = m_lo + (ml < cnt ? ml : base);
      pos[i] = __ldg(rp.positions + m);
      slot[i] = __ldg(rp.slot_mapping + m);
    }
#pragma unroll
    for (int i = 0; i < RI; ++i) {
      if (pos[i] < 0) pos[i] = 0;
      if (pos[i] >= rp.max_pos) pos[i] = rp.max_pos - 1;
      c4[i] = __ldg(reinterpret_cast<const float4*>(rp.cos_t + pos[i] * half + dcol));
      s4[i] = __ldg(reinterpret_cast<const float4*>(rp.sin_t + pos[i] * half + dcol));
    }
#pragma unroll
    for (int i = 0; i < RI; ++i) {
      const int ml = base + 4 * i;
      if (ml >= cnt) break;
      const int m = m_lo + ml;
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (q < S) {
          const float4 p = *reinterpret_cast<const float4*>(red + ((size_t)q * tp_max + ml) * GEMM_BN + 4 * t4);
          sum.x += p.x;
          sum.y += p.y;
          sum.z += p.z;
          sum.w += p.w;
        }
      }
      const float as = a_s[m];
      const float acc[4] = {sum.x, sum.y, sum.z, sum.w};
      const float cc[4] = {c4[i].x, c4[i].y, c4[i].z, c4[i].w}, ss[4] = {s4[i].x, s4[i].y, s4[i].z, s4[i].w};
      float xb[4], pr[4], o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e)  // what the qkv row holds after the plain GEMM
        xb[e] = __bfloat162float(__float2bfloat16_rn(scaled_value(acc[e], as, w4[e], hb, b4[e], false, 0.f)));
#pragma unroll
      for (int e = 0; e < 4; ++e) pr[e] = __shfl_xor_sync(0xffffffffu, xb[e], half >> 2);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float a = __fmul_rn(xb[e], cc[e]), b = __fmul_rn(pr[e], ss[e]);
        o[e] = !roped ? xb[e] : hi_half ? __fadd_rn(a, b) : __fsub_rn(a, b);
      }
      const uint2 rot = make_uint2(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]));
      // the qkv row keeps what the unfused pair leaves there: rotated q, un-rotated k, v
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)m * N + n0) =
          is_q ? rot : make_uint2(pack_bf16(xb[0], xb[1]), pack_bf16(xb[2], xb[3]));
      if (!is_q && slot[i] >= 0 && slot[i] < rp.num_slots) {
        const long long blk = slot[i] / rp.bs;
        const int off = (int)(slot[i] - blk * rp.bs);
        *reinterpret_cast<uint2*>(pool + ((blk * rp.Hkv + kh) * rp.bs + off) * rp.D + d0) = rot;
      }
    }
  }
}

// Weight-only (KIND_WO8) decode path, second version: the widened bf16 weight rows never touch shared memory.  Eight
// converter warps (the four epilogue warps plus warps 6..9, two per TMEM lane quarter, alternating k-blocks) read their
// weight row's 64 raw int8 out of the TMA tile (64B swizzle: conflict-free 16-byte row reads), widen them in registers and
// park the row as 32 packed bf16x2 columns in a TMEM slot with tcgen05.st; the MMA takes its A operand straight from TMEM
// (tcgen05.mma with a tensor-memory A).  Shared memory holds only the raw int8 ring (8 KB a stage) and the activation
// ring, so the rings are twice as deep as with a widened copy in shared memory, there is no generic-proxy -> async-proxy
// fence on the way to the tensor core, and the MMA reads half as many shared-memory bytes.
constexpr int WO_CONV_WARPS = 8;
constexpr int WO_THREADS = 64 + 32 * WO_CONV_WARPS + 32;  // + the activation producer warp (last warp)
constexpr int DEC_THREADS = GEMM_THREADS + 32;            // decode kernel: + the activation producer warp (warp 6)
constexpr uint32_t WO_TMEM_COLS = 256;  // accumulator (BM columns, >= 32) + widened-weight slots of 32 columns each
template <int BM>
__host__ __device__ constexpr int wo_slots() { return (int)(WO_TMEM_COLS - (BM < 32 ? 32 : BM)) / 32; }

// 64 raw int8 of one weight row (already in registers) -> 32 packed bf16x2 words, K order
__device__ __forceinline__ void widen_row64(const uint4 (&raw)[4], uint32_t (&o)[32]) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const uint32_t w[4] = {raw[c].x, raw[c].y, raw[c].z, raw[c].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {  // same exact bit trick as widen_w_tile
      const uint32_t x = w[j] ^ 0x80808080u;
      const float f0 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7440)) - 8388736.f;
      const float f1 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7441)) - 8388736.f;
      const float f2 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7442)) - 8388736.f;
      const float f3 = __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7443)) - 8388736.f;
      o[c * 8 + 2 * j] = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
      o[c * 8 + 2 * j + 1] = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
    }
  }
}

// One converter warp: k-blocks first, first + 2, ... of this CTA's stream; `quarter` = warp id % 4 = the TMEM lane quarter
// (and so the 32 weight rows) this warp may touch.  Raw tile s = i % nst in shared memory (TMA, 64B swizzle: the 16-byte
// chunk c of row r sits at chunk c ^ ((r >> 1) & 3)), TMEM slot t = i % NW.
template <int BM>
__device__ __forceinline__ void wo_convert(const uint8_t* __restrict__ raw_ring, uint64_t* full_raw, uint64_t* empty_raw,
                                           uint64_t* full_w, uint64_t* empty_w, uint32_t tmem_w0, int nk, int nst,
                                           int first, int quarter, int lane) {
  constexpr int NW = wo_slots<BM>();
  const int row = quarter * 32 + lane;
  const int sw = (row >> 1) & 3;
  for (int i = first; i < nk; i += 2) {
    const int s = i % nst, t = i % NW;
    mbar_wait(&full_raw[s], (uint32_t)(i / nst) & 1u);
    if (i >= NW) mbar_wait(&empty_w[t], (uint32_t)(i / NW - 1) & 1u);  // the MMAs that read slot t have retired
    tc_fence_after();
    const uint8_t* rp = raw_ring + (size_t)s * RAW_TILE + row * 64;
    uint4 raw[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) raw[c] = *reinterpret_cast<const uint4*>(rp + ((c ^ sw) << 4));
    uint32_t o[32];
    widen_row64(raw, o);
    tmem_st_32x32b_x32(tmem_w0 + ((uint32_t)(quarter * 32) << 16) + (uint32_t)t * 32u, o);
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {
      mbar_arrive(&full_w[t]);
      mbar_arrive(&empty_raw[s]);
    }
  }
}

template <int KIND, int BM, int EPI>
__global__ void __launch_bounds__(KIND == KIND_WO8 ? WO_THREADS : DEC_THREADS, KIND == KIND_WO8 ? 1 : 2)
    gemm_decode_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap a_map,
                       GemmEpilogue ep, int M, int N, int num_kb, int nst, int ring_bytes, int l2_ahead,
                       unsigned long long* __restrict__ trace) {
  using T = KindTraits<KIND>;
  constexpr int BK_ELEMS = GEMM_BK_BYTES / T::ELEM;
  constexpr int W_TILE = GEMM_BN * GEMM_BK_BYTES;
  constexpr int A_TILE = BM * GEMM_BK_BYTES;
  constexpr bool WO = KIND == KIND_WO8;
  constexpr uint32_t ACC_COLS = BM < 32 ? 32 : BM;
  constexpr uint32_t TMEM_COLS = WO ? WO_TMEM_COLS : ACC_COLS;
  constexpr int NW = wo_slots<BM>();  // weight-only: widened-weight slots in TMEM
  constexpr int AP_WARP = WO ? 2 + WO_CONV_WARPS : 6;  // the activation producer: the last warp
  // optional per-CTA time stamps (b200_gemm_set_trace; tools/gemm_trace.py): 0 start, 1 producer past the dependency,
  // 2 first stage landed, 3 last MMA retired, 4 partials pushed, 5 cluster barrier passed, 6 outputs stored
#define B200_STAMP(k) do { if (trace) trace[(size_t)blockIdx.x * 8 + (k)] = globaltimer_ns(); } while (0)
  if (threadIdx.x == 0) B200_STAMP(0);
  constexpr uint32_t IDESC = umma_idesc(T::CFMT, T::ABFMT, GEMM_BN, BM);
  constexpr int MK = KIND == KIND_WO8 ? (int)KIND_F16 : KIND;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = (int)cluster_nctarank(), r = (int)cluster_ctarank();
  const int tp_max = (M + S - 1) / S;                                   // most tokens one rank owns
  const int m_lo = (int)((long long)M * r / S), m_hi = (int)((long long)M * (r + 1) / S);  // tokens this rank finishes
  // ring_bytes = max(the rings the staging tile may alias once they are dead, the staging tile):
  //   W8A8 / bf16:  [w_ring: nst x 16 KB | ring_bytes][a_ring: nst x A_TILE]
  //   weight-only:  [a_ring: nst x A_TILE][raw_ring: nst x 8 KB | together ring_bytes]   (no widened copy in smem)
  uint8_t* w_ring = smem;
  uint8_t* a_ring = WO ? smem : smem + ring_bytes;
  uint8_t* raw_ring = a_ring + (size_t)nst * A_TILE;  // weight-only: nst raw int8 tiles
  float* red = reinterpret_cast<float*>(WO ? smem + ring_bytes : raw_ring);  // [S][tp_max][128] (S > 1)
  float* a_s = red + (S > 1 ? (size_t)S * tp_max * GEMM_BN : 0);         // [BM] per-token scales
  __nv_bfloat16* resid_s = reinterpret_cast<__nv_bfloat16*>(a_s + BM);  // [owned tokens][128] (iff residual)
  uint64_t* full_w = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(resid_s) +
                                                 (ep.residual ? (size_t)(S > 1 ? tp_max : BM) * GEMM_BN * 2 : 0));
  uint64_t* empty_w = full_w + DEC_MAX_ST;
  uint64_t* full_a = empty_w + DEC_MAX_ST;
  uint64_t* empty_a = full_a + DEC_MAX_ST;
  uint64_t* full_raw = empty_a + DEC_MAX_ST;   // weight-only: raw int8 tile landed / consumed by the widening warps
  uint64_t* empty_raw = full_raw + DEC_MAX_ST;
  uint64_t* tfull = empty_raw + DEC_MAX_ST;
  uint64_t* red_full = tfull + 1;  // all S partials of this CTA's tokens have landed in `red`
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(red_full + 1);
  float* stage = reinterpret_cast<float*>(smem);  // [BM][128] fp32 partial, aliases the (dead) ring after the last MMA

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x / S;
  const int kb0 = (int)((long long)num_kb * r / S), kb1 = (int)((long long)num_kb * (r + 1) / S);
  const int nk = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&a_map);
    for (int s = 0; s < (WO && NW > nst ? NW : nst); ++s) {
      // weight-only: full_w / empty_w guard the TMEM weight slots (one arrival per quarter warp of the converting group)
      mbar_init(&full_w[s], WO ? 4 : 1);
      mbar_init(&empty_w[s], 1);
      mbar_init(&full_a[s], 1);
      mbar_init(&empty_a[s], 1);
      if (WO) {
        mbar_init(&full_raw[s], 1);
        mbar_init(&empty_raw[s], 4);
      }
    }
    mbar_init(tfull, 1);
    mbar_init(red_full, 1);
    fence_mbar_init();
    fence_proxy_async();
    // the S incoming bulk copies (one per rank, this CTA's own included) complete on red_full
    if (S > 1 && m_hi > m_lo) mbar_expect_tx(red_full, (uint32_t)(S * (m_hi - m_lo) * GEMM_BN * 4));
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_launch();
  // phase 1 of the cluster barrier: "this CTA runs, its red_full barrier is armed" -- peers wait for it before their
  // first copy into this CTA's shared memory
  if (S > 1) cluster_arrive_release();
  const uint32_t tmem_base = *tmem_slot;

  // per-thread epilogue constants (epilogue warps only; filled while the ring streams)
  const int lg = warp & 3;
  const int nl = lg * 32 + lane;  // weight row of the tile this lane reads from TMEM
  const int et = threadIdx.x - 64;
  const int t4 = et & 31, tq = et >> 5;  // reduction phase: rows 4*t4 .. 4*t4+3, owned tokens tq, tq+4, ...
  float w_sc = 1.f, bias_v = 0.f;        // S == 1: row nl
  float w4[4] = {1.f, 1.f, 1.f, 1.f}, b4[4] = {0.f, 0.f, 0.f, 0.f};  // S > 1: rows 4*t4 + e

  if (warp == 0) {
    if (lane == 0 && nk > 0) {
      const uint64_t pol_w = policy_evict_first();  // every weight byte is read once per step
      const int npre = nk < nst ? nk : nst;
      // weights do not depend on any kernel of the step: fill the ring with W tiles before the dependency resolves
      for (int i = 0; i < npre; ++i) {
        if constexpr (WO) {
          mbar_expect_tx(&full_raw[i], RAW_TILE);
          tma_load_2d_hint(raw_ring + (size_t)i * RAW_TILE, &w_map, &full_raw[i], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
        } else {
          mbar_expect_tx(&full_w[i], W_TILE);
          tma_load_2d_hint(w_ring + (size_t)i * W_TILE, &w_map, &full_w[i], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
        }
      }
      // ... and pull the next `l2_ahead` weight tiles of this CTA's stream into L2: HBM is idle while the predecessor
      // (a small glue kernel, or the tail of a GEMM) drains, so by the time the ring recycles these loads are L2 hits
      {
        const int pf_end = nk < npre + l2_ahead ? nk : npre + l2_ahead;
        for (int i = npre; i < pf_end; ++i) tma_prefetch_l2_2d(&w_map, (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN);
      }
      // refills: weights only -- the activation tiles come from their own producer warp (AP_WARP).  One issuing thread
      // sustains about one TMA box row per 5.5 cycles (tools/tma_ingest.cu: 128x64B boxes 21 GB/s per SM from one
      // producer warp, 40 GB/s from two), so W and A boxes issued by the same thread queue behind each other.
      int s = 0, ph = 0;  // stage / parity of the empty barriers for refills
      for (int i = npre; i < nk; ++i) {
        if constexpr (WO) {
          mbar_wait(&empty_raw[s], ph);
          mbar_expect_tx(&full_raw[s], RAW_TILE);
          tma_load_2d_hint(raw_ring + (size_t)s * RAW_TILE, &w_map, &full_raw[s], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
        } else {
          mbar_wait(&empty_w[s], ph);
          mbar_expect_tx(&full_w[s], W_TILE);
          tma_load_2d_hint(w_ring + (size_t)s * W_TILE, &w_map, &full_w[s], (kb0 + i) * BK_ELEMS, n_tile * GEMM_BN, pol_w);
        }
        if (++s == nst) {
          s = 0;
          ph ^= 1;
        }
      }
    }
    __syncwarp();  // lanes 1..31 park here
  } else if (warp == 1) {
    int s = 0, ph = 0;
    int t = 0, pht = 0;  // weight-only: TMEM weight slot / its parity
    for (int i = 0; i < nk; ++i) {
      mbar_wait(&full_w[WO ? t : s], WO ? pht : ph);
      if (trace && lane == 0 && i < 8) trace[(size_t)gridDim.x * 8 + (size_t)blockIdx.x * 16 + i] = globaltimer_ns();
      mbar_wait(&full_a[s], ph);
      if (trace && lane == 0 && i < 8) trace[(size_t)gridDim.x * 8 + (size_t)blockIdx.x * 16 + 8 + i] = globaltimer_ns();
      tc_fence_after();
      if (lane == 0) {
        if (i == 0) B200_STAMP(2);
        const uint32_t aa = smem_u32(a_ring + (size_t)s * A_TILE);
        if constexpr (WO) {
          // A operand = the widened weight rows in TMEM slot t: lane = weight row, 8 columns (16 bf16 of K) per UMMA_K
          const uint32_t wt = tmem_base + ACC_COLS + (uint32_t)t * 32u;
#pragma unroll
          for (int kk = 0; kk < GEMM_BK_BYTES / 32; ++kk)
            umma_ts_f16(tmem_base, wt + kk * 8, umma_desc_sw128(aa + kk * 32), IDESC, (uint32_t)((i | kk) != 0));
          tc_commit(&empty_a[s]);
          tc_commit(&empty_w[t]);
        } else {
          const uint32_t wa = smem_u32(w_ring + (size_t)s * W_TILE);
#pragma unroll
          for (int kk = 0; kk < GEMM_BK_BYTES / 32; ++kk)  // UMMA_K = 32 bytes of K
            umma_ss<MK>(tmem_base, umma_desc_sw128(wa + kk * 32), umma_desc_sw128(aa + kk * 32), IDESC,
                        (uint32_t)((i | kk) != 0));
          tc_commit(&empty_a[s]);
          tc_commit(&empty_w[s]);
        }
      }
      __syncwarp();
      if (++s == nst) {
        s = 0;
        ph ^= 1;
      }
      if (++t == NW) {
        t = 0;
        pht ^= 1;
      }
    }
    if (lane == 0) tc_commit(tfull);
    __syncwarp();
  } else if (warp == AP_WARP) {
    // ---- activation producer: everything it loads was written by the previous kernel of the stream
    if (lane == 0 && nk > 0) {
      const uint64_t pol_a = policy_evict_last();  // activations are re-read by every tile
      griddep_wait();
      B200_STAMP(1);
      int s = 0, ph = 1;  // first pass over the ring: the stages are free (parity-1 wait on a fresh barrier returns at once)
      for (int i = 0; i < nk; ++i) {
        mbar_wait(&empty_a[s], ph);
        mbar_expect_tx(&full_a[s], A_TILE);
        tma_load_2d_hint(a_ring + (size_t)s * A_TILE, &a_map, &full_a[s], (kb0 + i) * BK_ELEMS, 0, pol_a);
        if (++s == nst) {
          s = 0;
          ph ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (WO && warp >= 6) {
    // weight-only: the second converter group (odd k-blocks); see wo_convert below
    wo_convert<BM>(raw_ring, full_raw, empty_raw, full_w, empty_w, tmem_base + ACC_COLS, nk, nst, 1, warp & 3, lane);
  } else {
    // ---- epilogue warps, part 1: stage the epilogue operands while the ring streams
    griddep_wait();
    for (int m = et; m < BM; m += 128) a_s[m] = (ep.a_scale && m < M) ? ep.a_scale[m] : 1.f;
    const int n0 = n_tile * GEMM_BN;
    if (S == 1) {
      if (n0 + nl < N) {
        if (ep.w_scale) w_sc = ep.w_scale[n0 + nl];
        if (ep.bias) bias_v = __bfloat162float(ep.bias[n0 + nl]);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int n = n0 + 4 * t4 + e;
        if (n < N) {
          if (ep.w_scale) w4[e] = ep.w_scale[n];
          if (ep.bias) b4[e] = __bfloat162float(ep.bias[n]);
        }
      }
    }
    if (ep.residual) {
      // rows [n0, n0+128) of the tokens this CTA finishes: S == 1 all of them, else [m_lo, m_hi)
      const int t_lo = S == 1 ? 0 : m_lo, t_cnt = S == 1 ? (M < BM ? M : BM) : m_hi - m_lo;
      const bool vec = (N & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.residual) & 15) == 0;
#pragma unroll 4
      for (int idx = et; idx < t_cnt * 16; idx += 128) {  // 16 vectors of 8 bf16 per token row
        const int ml = idx >> 4, c8 = (idx & 15) * 8;
        const __nv_bfloat16* src = ep.residual + (int64_t)(t_lo + ml) * N + n0 + c8;
        uint4 val = make_uint4(0u, 0u, 0u, 0u);
        if (vec) {
          if (n0 + c8 < N) val = *reinterpret_cast<const uint4*>(src);
        } else {
          __nv_bfloat16 tmp[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) tmp[e] = n0 + c8 + e < N ? src[e] : __float2bfloat16_rn(0.f);
          val = *reinterpret_cast<uint4*>(tmp);
        }
        *reinterpret_cast<uint4*>(resid_s + ml * GEMM_BN + c8) = val;
      }
    }
    named_bar_sync(1, 128);

    if constexpr (WO) {
      // ---- weight-only: the first converter group (even k-blocks).  The raw tiles were requested before the dependency
      // resolved, so the first ones are usually waiting here.
      wo_convert<BM>(raw_ring, full_raw, empty_raw, full_w, empty_w, tmem_base + ACC_COLS, nk, nst, 0, lg, lane);
    }

    // ---- part 2: the accumulator is final
    mbar_wait(tfull, 0);
    tc_fence_after();
    if (et == 0) B200_STAMP(3);
    const uint32_t acc = tmem_base + ((uint32_t)(lg * 32) << 16);
    if constexpr (EPI == EPI_SILU_QUANT) {
      // ---- gate/up GEMM + SiLU*mul + dynamic per-token quantisation in ONE kernel (host guarantees S == 1 and that the
      // whole grid is resident).  Every lane keeps its 8 activations per 16-token chunk in registers; the CTAs agree on
      // the per-token amax with one 64-bit atomicMax per token and a grid-wide barrier; then each lane quantises what it
      // holds.  Same bits as b200_gemm_silu_mul + b200_quantize_per_token (max of the same bf16 values, same division).
      constexpr int QD = KIND == KIND_I8 ? (int)B200_INT8 : (int)B200_FP8_E4M3;
      unsigned long long* tok_amax = ep.fq_state;
      unsigned int* bar_cnt = reinterpret_cast<unsigned int*>(ep.fq_state + 128);
      unsigned int* bar_epoch = bar_cnt + 1;
      unsigned int* s_amax = reinterpret_cast<unsigned int*>(smem);  // [BM] (+ the epoch at [BM]); the rings are dead
      float* s_scale = reinterpret_cast<float*>(smem) + BM + 4;       // [BM]
      const int n = n0 + nl;
      const bool n_ok = n < N;
      const int I = N >> 1;
      const int ch = (n >> 5) * 16 + (lane & 15);
      const bool hi = lane >= 16;
      const bool hb = ep.bias != nullptr;
      if (et < BM) s_amax[et] = 0u;
      if (et == 0) s_amax[BM] = *reinterpret_cast<volatile unsigned int*>(bar_epoch);  // stable until THIS CTA has arrived
      named_bar_sync(1, 128);
      float areg[BM / 16][8];
#pragma unroll
      for (int cc = 0; cc < BM / 16; ++cc) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
        float mine[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int m = cc * 16 + j;
          const float a0 = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
          mine[j] = (m < M && n_ok)
                        ? __bfloat162float(__float2bfloat16_rn(scaled_value(a0, a_s[m], w_sc, hb, bias_v, false, 0.f)))
                        : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float send = hi ? mine[j] : mine[j + 8];
          const float got = __shfl_xor_sync(0xffffffffu, send, 16);
          const float a = silu_mul_bf16(hi ? got : mine[j], hi ? mine[j + 8] : got);
          const int m = cc * 16 + (hi ? 8 : 0) + j;
          const bool ok = m < M && n_ok;
          areg[cc][j] = ok ? a : 0.f;
          float x = ok ? fabsf(a) : 0.f;
#pragma unroll
          for (int o = 1; o < 16; o <<= 1) x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
          if ((lane & 15) == 0 && m < BM) atomicMax(&s_amax[m], __float_as_uint(x));  // x >= 0: uint order = float order
        }
      }
      named_bar_sync(1, 128);
      const unsigned int e0 = s_amax[BM];
      if (et < M)
        atomicMax(&tok_amax[et], ((unsigned long long)(e0 + 1u) << 32) | (unsigned long long)s_amax[et]);
      __threadfence();
      named_bar_sync(1, 128);
      if (et == 0) {
        const unsigned int old = atomicAdd(bar_cnt, 1u);
        if (old == gridDim.x - 1) {
          atomicExch(bar_cnt, 0u);
          __threadfence();
          atomicAdd(bar_epoch, 1u);
        } else {
          const long long t0 = clock64();
          unsigned int e;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(e) : "l"(bar_epoch) : "memory");
            if (clock64() - t0 > 4000000000LL) {  // ~2 s: a grid that is not fully resident must not hang the box
              printf("b200: silu-quant grid barrier timeout block %d epoch %u\n", blockIdx.x, e0);
              __trap();
            }
          } while (e == e0);
        }
        __threadfence();
      }
      named_bar_sync(1, 128);
      if (et < BM) {
        float sc = 1.f;
        if (et < M) {
          const unsigned long long tv = __ldcg(&tok_amax[et]);
          const float amax = __uint_as_float((unsigned int)tv);
          const float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
          sc = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
          if (blockIdx.x == 0) ep.out_scale[et] = sc;
        }
        s_scale[et] = sc;
      }
      named_bar_sync(1, 128);
      uint8_t* oq = reinterpret_cast<uint8_t*>(ep.out);
#pragma unroll
      for (int cc = 0; cc < BM / 16; ++cc) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int m = cc * 16 + (hi ? 8 : 0) + j;
          if (m < M && n_ok) oq[(int64_t)m * I + ch] = quant_byte<QD>(areg[cc][j], s_scale[m]);
        }
      }
    } else if (S == 1) {
      const int n = n0 + nl;
#pragma unroll 1
      for (int cc = 0; cc < BM / 16; ++cc) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
        epilogue_chunk_staged<(EPI == EPI_ROPE || EPI == EPI_SILU_QUANT) ? (int)EPI_LINEAR : EPI>(ep, f, cc * 16, M, n, n < N, N, lane, a_s, w_sc, bias_v,
                                                                        resid_s + nl, GEMM_BN);
      }
    } else {
      // TMEM -> local staging tile [token][128 rows] (the ring is dead: every MMA of this CTA has retired) ...
#pragma unroll 1
      for (int cc = 0; cc < BM / 16; ++cc) {
        if (cc * 16 >= M) break;
        uint32_t v[16];
        tmem_ld_32x32b_x16(acc + cc * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j)
          stage[(cc * 16 + j) * GEMM_BN + nl] = KIND == KIND_I8 ? (float)(int32_t)v[j] : __uint_as_float(v[j]);
      }
      fence_proxy_async();  // generic-proxy writes -> visible to the bulk-copy (async proxy) reads below
      named_bar_sync(1, 128);
      // ... then ONE bulk DSMEM copy per owner: tokens [M*q/S, M*(q+1)/S) are a contiguous slab of the staging tile and
      // land in slot `r` of owner q's `red`, completing on the owner's red_full barrier
      cluster_wait();  // phase 1: every peer runs and has armed its red_full barrier
      if (et < S) {
        const int q = et;
        const int lo = (int)((long long)M * q / S), hi = (int)((long long)M * (q + 1) / S);
        if (hi > lo)
          bulk_copy_to_cluster(mapa_shared(smem_u32(red + (size_t)r * tp_max * GEMM_BN), (uint32_t)q),
                               smem_u32(stage + (size_t)lo * GEMM_BN), (uint32_t)((hi - lo) * GEMM_BN * 4),
                               mapa_shared(smem_u32(red_full), (uint32_t)q));
      }
    }
  }
  tc_fence_before();
  __syncwarp();
  if (threadIdx.x == 64) B200_STAMP(4);
  if (S > 1) {
    if (warp < 2 || warp >= 6) cluster_wait();  // phase 1 for the warps that do not push
    if (warp >= 2 && warp < 6 && m_hi > m_lo) mbar_wait(red_full, 0);  // the S partials of this CTA's tokens have landed
    if (threadIdx.x == 64) B200_STAMP(5);
    // phase 2: nobody may exit while a peer's copy still reads its staging tile; arriving here means "all copies INTO
    // me are complete", so once every CTA has arrived every copy OUT of everyone is complete too
    cluster_arrive_relaxed();
    if (warp >= 2 && warp < 6) {
      const int n0 = n_tile * GEMM_BN + 4 * t4;
      const bool hb = ep.bias != nullptr, hr = EPI != EPI_SILU_MUL && ep.residual != nullptr;
      bool rope_done = false;
      if constexpr (EPI == EPI_ROPE) {
        rope_reduce_store(ep, red, S, tp_max, m_lo, m_hi - m_lo, tq, t4, n0, N, a_s, w4, b4, hb);
        rope_done = true;
      }
#pragma unroll 2
      for (int ml = tq; ml < m_hi - m_lo && !rope_done; ml += 4) {
        const int m = m_lo + ml;
        float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < S) {
            const float4 p = *reinterpret_cast<const float4*>(red + ((size_t)q * tp_max + ml) * GEMM_BN + 4 * t4);
            sum.x += p.x;
            sum.y += p.y;
            sum.z += p.z;
            sum.w += p.w;
          }
        }
        const float as = a_s[m];
        float rs[4] = {0.f, 0.f, 0.f, 0.f};
        if (hr) {
          const uint2 rv = *reinterpret_cast<const uint2*>(resid_s + ml * GEMM_BN + 4 * t4);
          const __nv_bfloat162 r01 = *reinterpret_cast<const __nv_bfloat162*>(&rv.x);
          const __nv_bfloat162 r23 = *reinterpret_cast<const __nv_bfloat162*>(&rv.y);
          rs[0] = __low2float(r01);
          rs[1] = __high2float(r01);
          rs[2] = __low2float(r23);
          rs[3] = __high2float(r23);
        }
        float y[4];
        y[0] = scaled_value(sum.x, as, w4[0], hb, b4[0], hr, rs[0]);
        y[1] = scaled_value(sum.y, as, w4[1], hb, b4[1], hr, rs[1]);
        y[2] = scaled_value(sum.z, as, w4[2], hb, b4[2], hr, rs[2]);
        y[3] = scaled_value(sum.w, as, w4[3], hb, b4[3], hr, rs[3]);
        if constexpr (EPI != EPI_SILU_MUL) {
          if ((N & 3) == 0 && n0 + 3 < N) {
            if (ep.out_f32)
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.out) + (int64_t)m * N + n0) =
                  make_float4(y[0], y[1], y[2], y[3]);
            else
              *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)m * N + n0) =
                  make_uint2(pack_bf16(y[0], y[1]), pack_bf16(y[2], y[3]));
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (n0 + e < N) {
                if (ep.out_f32)
                  reinterpret_cast<float*>(ep.out)[(int64_t)m * N + n0 + e] = y[e];
                else
                  reinterpret_cast<__nv_bfloat16*>(ep.out)[(int64_t)m * N + n0 + e] = __float2bfloat16_rn(y[e]);
              }
            }
          }
        } else {
          // rows of one 32-row group: 16 gate rows (lanes 8g .. 8g+3) then the 16 up rows of the same channels
          // (lanes 8g+4 .. 8g+7); N2 % 32 == 0, so a group is inside N or outside as a whole
          const int I = N >> 1, n_groups = N >> 5;
          const bool ok = n0 < N;
          float mine[4], other[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            mine[e] = ok ? __bfloat162float(__float2bfloat16_rn(y[e])) : 0.f;
            other[e] = __shfl_xor_sync(0xffffffffu, mine[e], 4);
          }
          const bool is_gate = (lane & 4) == 0;
          float a[4], x = 0.f;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            a[e] = silu_mul_bf16(is_gate ? mine[e] : other[e], is_gate ? other[e] : mine[e]);
            x = fmaxf(x, ok ? fabsf(a[e]) : 0.f);
          }
          x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
          x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 2));
          if (ok && is_gate) {
            const int ch = (n0 >> 5) * 16 + (lane & 3) * 4;
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)m * I + ch) =
                make_uint2(pack_bf16(a[0], a[1]), pack_bf16(a[2], a[3]));
            if ((lane & 7) == 0) ep.group_amax[(int64_t)m * n_groups + (n0 >> 5)] = x;
          }
        }
      }
    }
  }
  if (S > 1) cluster_wait();  // phase 2
  __syncthreads();
  if (threadIdx.x == 0) B200_STAMP(6);
#undef B200_STAMP
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace b200

using namespace b200;

namespace {

unsigned long long* g_trace = nullptr;  // b200_gemm_set_trace

struct GemmPlan {
  bool decode;  // M <= 128: cluster split-K kernel; else persistent stream-K
  int bm, m_tiles, n_tiles, tiles, num_kb;
  int ctas;  // stream-K
  size_t smem, ws_bytes;
};

// bytes of one ACTIVATION element (weight-only: bf16 activations, int8 weights)
int elem_bytes(int dt) { return (dt == B200_BF16 || dt == B200_INT8_WO) ? 2 : 1; }
int w_elem_bytes(int dt) { return dt == B200_BF16 ? 2 : 1; }
constexpr int stages_for(int bm, bool wo = false) { return bm == 32 ? 10 : bm == 64 ? 8 : bm == 128 ? 6 : wo ? 3 : 4; }
constexpr size_t smem_for(int bm, bool wo = false) {
  return (size_t)stages_for(bm, wo) * (GEMM_BN * GEMM_BK_BYTES + bm * GEMM_BK_BYTES + (wo ? RAW_TILE : 0)) +
         (5 * stages_for(bm, wo) + 4) * 8 + 8 + 16 + (size_t)bm * 4 + 4 * XP_WARP_FLOATS * 4 + 1024;
}

// split_k: decode = cluster size along K (0 = auto; 1, 2, 4, 8); stream-K = cap on the persistent CTAs (0 = one per SM)
int plan_gemm(GemmPlan& p, int M, int N, int K, int ab_dtype, int split_k) {
  p.decode = M <= 128;
  p.bm = M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  p.m_tiles = (M + p.bm - 1) / p.bm;
  p.n_tiles = (N + GEMM_BN - 1) / GEMM_BN;
  p.tiles = p.m_tiles * p.n_tiles;
  int bk = GEMM_BK_BYTES / elem_bytes(ab_dtype);
  p.num_kb = (K + bk - 1) / bk;
  p.ctas = 0;
  p.smem = 0;
  p.ws_bytes = 0;
  if (p.decode) return B200_OK;  // the decode schedule is planned per kernel instantiation (occupancy queries)
  long long units = (long long)p.tiles * p.num_kb;
  int ctas = sm_count();
  if (split_k > 0 && split_k < ctas) ctas = split_k;
  if (ctas > GEMM_CNT_INTS) ctas = GEMM_CNT_INTS;
  if ((long long)ctas > units) ctas = (int)units;
  p.ctas = ctas;
  p.smem = smem_for(p.bm, ab_dtype == B200_INT8_WO);
  p.ws_bytes = (size_t)GEMM_CNT_INTS * sizeof(int) + (size_t)ctas * 2 * p.bm * GEMM_BN * sizeof(float);
  return B200_OK;
}

template <int KIND, int EPI>
int launch_streamk(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, void* workspace, int M, int N,
                   const GemmPlan& p, cudaStream_t stream) {
  constexpr int BM = 256;
  constexpr bool WO = KIND == KIND_WO8;
  auto kern = gemm_streamk_kernel<KIND, BM, stages_for(BM, WO), EPI>;
  constexpr size_t smem = smem_for(BM, WO);
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  int* counters = (int*)workspace;
  float* partials = (float*)((uint8_t*)workspace + GEMM_CNT_INTS * sizeof(int));
  B200_CUDA_OK(launch_kernel(kern, dim3(p.ctas), dim3(WO ? GEMM_THREADS + 128 : GEMM_THREADS), smem, stream, wm, am, ep,
                             partials, counters, M, N,
                             p.m_tiles, p.num_kb, p.tiles, p.m_tiles == 1 ? 1 : 0));
  return B200_OK;
}

// ---- decode schedule: (cluster size S, CTAs per SM, ring depths) per problem shape
constexpr size_t SMEM_ONE_PER_SM = 232448;  // 227 KB: the per-CTA maximum
constexpr size_t SMEM_TWO_PER_SM = 115712;  // (228 KB - 2 x 1 KB reserved) / 2
constexpr int W_TILE_BYTES = GEMM_BN * GEMM_BK_BYTES;

struct DecodeCfg {
  int S, nst, ring_bytes;
  size_t smem;
  int cap_ctas;  // CTAs the chip holds at once in this configuration
  double cost_us;
};

// ring depth that fits `budget` bytes
bool decode_rings(int bm, int M, int S, int nk, bool resid, size_t budget, DecodeCfg& c, bool wo = false) {
  // per stage: the activation tile plus the weight tile as it sits in shared memory -- 16 KB of operand-ready rows, or
  // (weight-only) the 8 KB raw int8 tile: the widened rows live in TMEM
  const int a_tile = bm * GEMM_BK_BYTES;
  const int w_tile = wo ? RAW_TILE : W_TILE_BYTES;
  const int tp_max = (M + S - 1) / S;
  const size_t red = S > 1 ? (size_t)S * tp_max * GEMM_BN * 4 : 0;
  const size_t fixed = 1024 + red + (size_t)bm * 8 + (resid ? (size_t)(S > 1 ? tp_max : bm) * GEMM_BN * 2 : 0) + 768;
  const size_t stage = S > 1 ? (size_t)bm * GEMM_BN * 4 : 0;  // the partial tile is staged over the dead ring(s)
  if (budget < fixed + (size_t)(a_tile + w_tile) || budget < fixed + stage) return false;
  int n = (int)((budget - fixed) / (size_t)(a_tile + w_tile));
  if (n > DEC_MAX_ST) n = DEC_MAX_ST;
  if (n > nk) n = nk;
  if (n < 1) return false;
  c.nst = n;
  if (wo) {  // [a_ring][raw_ring] together are what the staging tile aliases
    size_t ring = (size_t)n * (a_tile + w_tile);
    if (ring < stage) ring = stage;
    c.ring_bytes = (int)ring;
    c.smem = fixed + ring;
    return c.smem <= budget;
  }
  size_t ring = (size_t)n * W_TILE_BYTES;
  if (ring < stage) ring = stage;
  c.ring_bytes = (int)ring;
  c.smem = fixed + ring + (size_t)n * a_tile;
  return c.smem <= budget;
}

template <int KIND, int BM, int EPI>
int plan_decode(DecodeCfg& best, int M, int n_tiles, int num_kb, int split_k, bool resid) {
  auto kern = gemm_decode_kernel<KIND, BM, EPI>;
  static std::mutex mu;
  static std::map<std::tuple<int, int, int, int, bool>, DecodeCfg> cache;
  const auto key = std::make_tuple(M, n_tiles, num_kb, split_k, resid);
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      best = it->second;
      return B200_OK;
    }
  }
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_ONE_PER_SM));
    attr_done = true;
  }
  if (split_k < 0 || split_k > 8 || split_k > num_kb) {
    set_last_error("gemm: split_k %d invalid here (0 = auto, or 1 .. 8 and at most %d k-blocks)", split_k, num_kb);
    return B200_ERR_INVALID_ARG;
  }
  static const int force_occ = getenv("B200_GEMM_FORCE_OCC") ? atoi(getenv("B200_GEMM_FORCE_OCC")) : 0;  // tuning aid
  const int sms = sm_count();
  bool found = false;
  // two CTAs per SM (half the shared memory each) or one with all of it: M <= 64 always fits two; M = 65..128 (32 KB
  // activation + weight per stage, 64 KB of split-K partials) is tried both ways and the cost model decides
  for (int per_sm = 2; per_sm >= 1; --per_sm) {
  if (per_sm == 2 && (force_occ == 1 || KIND == KIND_WO8)) continue;  // weight-only: 320 threads x ~126 registers, one CTA
  if (per_sm == 1 && BM <= 64 && force_occ != 1 && KIND != KIND_WO8) continue;
  const size_t budget = per_sm == 2 ? SMEM_TWO_PER_SM : SMEM_ONE_PER_SM;
  for (int S = 1; S <= 8; ++S) {
    if (split_k != 0 && S != split_k) continue;
    if (S > num_kb) break;
    const int nk = (num_kb + S - 1) / S;
    DecodeCfg c{};
    c.S = S;
    if (!decode_rings(BM, M, S, nk, resid, budget, c, KIND == KIND_WO8)) continue;
    // clusters of 4 and 8 strand SMs (GPC packing): ask the runtime how many clusters one wave holds
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_tiles * S);
    cfg.blockDim = dim3(KIND == KIND_WO8 ? WO_THREADS : DEC_THREADS);
    cfg.dynamicSmemBytes = c.smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)S;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg) != cudaSuccess || clusters <= 0) {
      cudaGetLastError();
      continue;
    }
    c.cap_ctas = clusters * S;
    // cost model (us), fitted to the sweeps in profiles/r01_gemm_sweep_v3.json:
    //   fixed latency chain  +  max(busiest SM's stages at its latency-bound rate, chip ingest, HBM)  +  reduction.
    // An SM's ingest rate grows with the stages it has in flight (one ring round trip ~1.55 us); CTAs spread
    // round-robin, so with more CTAs than SMs the busiest SM runs two rings side by side.
    const int ctas = n_tiles * S;
    const double stage_bytes = (double)((KIND == KIND_WO8 ? RAW_TILE : W_TILE_BYTES) + BM * GEMM_BK_BYTES);
    const int sm_used = ctas < sms ? ctas : sms;
    int on_busiest = (ctas + sm_used - 1) / sm_used;
    const int resident = on_busiest < per_sm ? on_busiest : per_sm;
    // ... and one CTA never ingests more than ~52 KB/us however deep its ring is (measured at M = 128, one CTA per SM,
    // 7 stages: 0.55 us per 32 KB stage; two co-resident CTAs reach ~94 KB/us together)
    double rate_sm = resident * c.nst * stage_bytes / 1.55;  // bytes per us
    if (rate_sm > resident * 52e3) rate_sm = resident * 52e3;
    const double t_sm = on_busiest * nk * stage_bytes / rate_sm;
    const double t_chip = (double)n_tiles * num_kb * stage_bytes / 9.5e6;
    const double t_hbm = (double)n_tiles * num_kb * (KIND == KIND_WO8 ? RAW_TILE : W_TILE_BYTES) / 6.2e6;
    double t = t_sm > t_chip ? t_sm : t_chip;
    if (t_hbm > t) t = t_hbm;
    t += 6.0;
    if (S > 1) t += 0.8 + 0.05 * ((M + S - 1) / S);
    const int capacity = c.cap_ctas * per_sm;  // the occupancy query counts one CTA per SM
    if (ctas > capacity) t += ((ctas - 1) / capacity) * (3.0 + nk * stage_bytes / (c.nst * stage_bytes / 1.55));
    c.cost_us = t;
    if (!found || c.cost_us < best.cost_us - 1e-9) {
      best = c;
      found = true;
    }
  }
  }
  if (!found) {
    set_last_error("gemm: no decode configuration fits (n_tiles %d, k-blocks %d, split_k %d)", n_tiles, num_kb, split_k);
    return B200_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> g(mu);
  cache[key] = best;
  return B200_OK;
}

template <int KIND, int BM, int EPI>
int launch_decode(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, int M, int N, const GemmPlan& p,
                  int split_k, cudaStream_t stream) {
  DecodeCfg c;
  int rc = plan_decode<KIND, BM, EPI>(c, M, p.n_tiles, p.num_kb, split_k, ep.residual != nullptr);
  if (rc) return rc;
  auto kern = gemm_decode_kernel<KIND, BM, EPI>;
  static const int l2_ahead = getenv("B200_GEMM_L2_AHEAD") ? atoi(getenv("B200_GEMM_L2_AHEAD")) : 0;  // tuning aid
  B200_CUDA_OK(launch_kernel_cluster(kern, dim3(p.n_tiles * c.S), dim3(KIND == KIND_WO8 ? WO_THREADS : DEC_THREADS), c.smem,
                                     c.S, stream, wm, am, ep, M, N, p.num_kb, c.nst, c.ring_bytes, l2_ahead, g_trace));
  return B200_OK;
}

template <int KIND, int EPI>
int dispatch_gemm(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, void* ws, int M, int N,
                  const GemmPlan& p, int split_k, cudaStream_t stream) {
  if (!p.decode) return launch_streamk<KIND, EPI>(wm, am, ep, ws, M, N, p, stream);
  if (p.bm == 32) return launch_decode<KIND, 32, EPI>(wm, am, ep, M, N, p, split_k, stream);
  if (p.bm == 64) return launch_decode<KIND, 64, EPI>(wm, am, ep, M, N, p, split_k, stream);
  return launch_decode<KIND, 128, EPI>(wm, am, ep, M, N, p, split_k, stream);
}

template <int KIND>
int launch_rope(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, int M, int N, const GemmPlan& p,
                int split_k, cudaStream_t stream) {
  if (p.bm == 32) return launch_decode<KIND, 32, EPI_ROPE>(wm, am, ep, M, N, p, split_k, stream);
  if (p.bm == 64) return launch_decode<KIND, 64, EPI_ROPE>(wm, am, ep, M, N, p, split_k, stream);
  return launch_decode<KIND, 128, EPI_ROPE>(wm, am, ep, M, N, p, split_k, stream);
}

int gemm_common(const void* a, const void* w, const GemmEpilogue& ep, int epi, int M, int N, int K, int ab_dtype,
                int split_k, void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && w && ep.out, "gemm: null pointer");
  B200_REQUIRE(ab_dtype == B200_BF16 || ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8 || ab_dtype == B200_INT8_WO,
               "gemm: bad ab_dtype %d", ab_dtype);
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  B200_REQUIRE(split_k >= 0, "gemm: bad split_k %d", split_k);
  const bool wo = ab_dtype == B200_INT8_WO;
  int eb = elem_bytes(ab_dtype);
  B200_REQUIRE(((int64_t)K * w_elem_bytes(ab_dtype)) % 16 == 0,
               "gemm: K*elem_bytes must be a multiple of 16 (TMA row pitch), K=%d", K);
  B200_REQUIRE(!wo || ep.a_scale == nullptr, "gemm: weight-only takes bf16 activations (a_scale must be NULL)");
  GemmPlan p;
  int rc = plan_gemm(p, M, N, K, ab_dtype, split_k);
  if (rc) return rc;
  if (p.ws_bytes && (!workspace || workspace_bytes < p.ws_bytes)) {
    set_last_error("gemm: workspace %zu < %zu bytes", workspace_bytes, p.ws_bytes);
    return B200_ERR_WORKSPACE;
  }
  CUtensorMapDataType dt = eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  int bk = GEMM_BK_BYTES / eb;
  CUtensorMap wm, am;
  if (wo && p.decode)  // raw int8 rows, 64 B of K per row and stage, 64B swizzle: a converter thread reads its whole row
    rc = make_tmap_2d_sw64(&wm, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)K, (uint64_t)N, (uint64_t)K, bk, GEMM_BN);
  else if (wo)  // stream-K: no swizzle, the widening warps read 16-byte pieces with plain loads
    rc = make_tmap_2d_plain(&wm, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)K, (uint64_t)N, (uint64_t)K, bk, GEMM_BN);
  else
    rc = make_tmap_2d(&wm, w, dt, eb, (uint64_t)K, (uint64_t)N, (uint64_t)K * eb, bk, GEMM_BN);
  if (rc) return rc;
  rc = make_tmap_2d(&am, a, dt, eb, (uint64_t)K, (uint64_t)M, (uint64_t)K * eb, bk, p.bm);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (epi == EPI_ROPE) {  // decode schedule only (b200_gemm_rope_kv_write checked the plan)
    B200_REQUIRE(p.decode, "gemm: the RoPE epilogue exists in the decode schedule only");
    if (ab_dtype == B200_BF16) return launch_rope<KIND_F16>(wm, am, ep, M, N, p, split_k, s);
    if (ab_dtype == B200_FP8_E4M3) return launch_rope<KIND_F8F6F4>(wm, am, ep, M, N, p, split_k, s);
    if (wo) return launch_rope<KIND_WO8>(wm, am, ep, M, N, p, split_k, s);
    return launch_rope<KIND_I8>(wm, am, ep, M, N, p, split_k, s);
  }
  if (epi == EPI_LINEAR) {
    if (ab_dtype == B200_BF16) return dispatch_gemm<KIND_F16, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, split_k, s);
    if (ab_dtype == B200_FP8_E4M3) return dispatch_gemm<KIND_F8F6F4, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, split_k, s);
    if (wo) return dispatch_gemm<KIND_WO8, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, split_k, s);
    return dispatch_gemm<KIND_I8, EPI_LINEAR>(wm, am, ep, workspace, M, N, p, split_k, s);
  }
  if (ab_dtype == B200_BF16) return dispatch_gemm<KIND_F16, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, split_k, s);
  if (ab_dtype == B200_FP8_E4M3) return dispatch_gemm<KIND_F8F6F4, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, split_k, s);
  if (wo) return dispatch_gemm<KIND_WO8, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, split_k, s);
  return dispatch_gemm<KIND_I8, EPI_SILU_MUL>(wm, am, ep, workspace, M, N, p, split_k, s);
}

}  // namespace

namespace {
template <int KIND, int EPI>
int query_plan(DecodeCfg& c, int bm, int M, int n_tiles, int num_kb, int split_k, bool resid) {
  if (bm == 32) return plan_decode<KIND, 32, EPI>(c, M, n_tiles, num_kb, split_k, resid);
  if (bm == 64) return plan_decode<KIND, 64, EPI>(c, M, n_tiles, num_kb, split_k, resid);
  return plan_decode<KIND, 128, EPI>(c, M, n_tiles, num_kb, split_k, resid);
}
}  // namespace

namespace {
// gate/up GEMM + SiLU*mul + per-token quantisation in one launch (decode schedule, W8A8 kinds).  Returns
// B200_ERR_UNSUPPORTED -- and the caller runs b200_gemm_silu_mul + b200_quantize_per_token -- unless the planner's own
// choice for this shape is one CTA per weight tile (S == 1) with the whole grid resident at once: the CTAs meet at a
// grid-wide barrier, which must never wait for a CTA that has not been scheduled.
template <int KIND>
static int launch_silu_quant(const CUtensorMap& wm, const CUtensorMap& am, const GemmEpilogue& ep, int M, int N,
                             const GemmPlan& p, cudaStream_t s) {
  DecodeCfg c;
  int rc = p.bm == 32   ? plan_decode<KIND, 32, EPI_SILU_QUANT>(c, M, p.n_tiles, p.num_kb, 0, false)
           : p.bm == 64 ? plan_decode<KIND, 64, EPI_SILU_QUANT>(c, M, p.n_tiles, p.num_kb, 0, false)
                        : plan_decode<KIND, 128, EPI_SILU_QUANT>(c, M, p.n_tiles, p.num_kb, 0, false);
  if (rc) return rc;
  const int per_sm = c.smem > SMEM_TWO_PER_SM ? 1 : 2;
  if (c.S != 1 || p.n_tiles > c.cap_ctas * per_sm) {
    set_last_error("gemm_silu_mul_quant: plan S=%d, %d tiles, capacity %d: not one resident wave", c.S, p.n_tiles,
                   c.cap_ctas * per_sm);
    return B200_ERR_UNSUPPORTED;
  }
  if (p.bm == 32) return launch_decode<KIND, 32, EPI_SILU_QUANT>(wm, am, ep, M, N, p, 0, s);
  if (p.bm == 64) return launch_decode<KIND, 64, EPI_SILU_QUANT>(wm, am, ep, M, N, p, 0, s);
  return launch_decode<KIND, 128, EPI_SILU_QUANT>(wm, am, ep, M, N, p, 0, s);
}

}  // namespace

extern "C" {

void b200_gemm_set_trace(void* buf) { g_trace = (unsigned long long*)buf; }

int b200_gemm_decode_plan(int M, int N, int K, int ab_dtype, int split_k, int silu_mul, int has_residual, int* out6) {
  B200_REQUIRE(out6 && M > 0 && M <= 128 && N > 0 && K > 0, "gemm_decode_plan: bad arguments");
  B200_REQUIRE(ab_dtype == B200_BF16 || ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8 || ab_dtype == B200_INT8_WO,
               "gemm_decode_plan: bad dtype");
  GemmPlan p;
  int rc = plan_gemm(p, M, N, K, ab_dtype, split_k);
  if (rc) return rc;
  DecodeCfg c{};
  const bool r = has_residual != 0;
  if (silu_mul == 2) {  // the qkv GEMM with the RoPE / KV-write epilogue
    rc = ab_dtype == B200_BF16 ? query_plan<KIND_F16, EPI_ROPE>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_FP8_E4M3 ? query_plan<KIND_F8F6F4, EPI_ROPE>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_INT8_WO  ? query_plan<KIND_WO8, EPI_ROPE>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
                                     : query_plan<KIND_I8, EPI_ROPE>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r);
  } else if (silu_mul) {
    rc = ab_dtype == B200_BF16 ? query_plan<KIND_F16, EPI_SILU_MUL>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_FP8_E4M3 ? query_plan<KIND_F8F6F4, EPI_SILU_MUL>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_INT8_WO  ? query_plan<KIND_WO8, EPI_SILU_MUL>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
                                     : query_plan<KIND_I8, EPI_SILU_MUL>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r);
  } else {
    rc = ab_dtype == B200_BF16 ? query_plan<KIND_F16, EPI_LINEAR>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_FP8_E4M3 ? query_plan<KIND_F8F6F4, EPI_LINEAR>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
         : ab_dtype == B200_INT8_WO  ? query_plan<KIND_WO8, EPI_LINEAR>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r)
                                     : query_plan<KIND_I8, EPI_LINEAR>(c, p.bm, M, p.n_tiles, p.num_kb, split_k, r);
  }
  if (rc) return rc;
  out6[0] = c.S;
  out6[1] = c.smem > SMEM_TWO_PER_SM ? 1 : 2;
  out6[2] = c.nst;
  out6[3] = c.nst;
  out6[4] = c.cap_ctas;
  out6[5] = (int)c.smem;
  return B200_OK;
}

size_t b200_gemm_workspace_bytes(int M, int N, int K, int ab_dtype, int split_k) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  GemmPlan p;
  if (plan_gemm(p, M, N, K, ab_dtype, split_k)) return 0;
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

int b200_gemm_rope_kv_write(const void* a, const void* w, void* qkv_out, const float* a_scale, const float* w_scale,
                            const void* bias, int M, int N, int K, int ab_dtype, int split_k, void* workspace,
                            size_t workspace_bytes, const int64_t* positions, const float* cos_table,
                            const float* sin_table, int max_pos, void* k_pool, void* v_pool, const int64_t* slot_mapping,
                            int Hq, int Hkv, int D, int block_size, int num_blocks, void* stream) {
  B200_REQUIRE(positions && cos_table && sin_table && k_pool && v_pool && slot_mapping, "gemm_rope_kv_write: null pointer");
  B200_REQUIRE(N == (Hq + 2 * Hkv) * D, "gemm_rope_kv_write: N=%d is not (Hq + 2 Hkv) * D = %d", N, (Hq + 2 * Hkv) * D);
  GemmEpilogue ep{a_scale, w_scale, (const __nv_bfloat16*)bias, nullptr, qkv_out, 0, nullptr};
  // The rotation lives in the split-K owner's epilogue of the decode kernel (a token's whole head in one warp).  Every
  // other schedule (one CTA per tile, the stream-K prefill kernel, odd head sizes) runs the GEMM, then the RoPE/scatter
  // kernel: same bits either way.
  bool fused = false;
  if (M <= 128 && (D == 64 || D == 128) && N % GEMM_BN == 0 && max_pos > 0 &&
      (ab_dtype == B200_BF16 || ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8 || ab_dtype == B200_INT8_WO)) {
    int plan6[6];
    if (b200_gemm_decode_plan(M, N, K, ab_dtype, split_k, 2, 0, plan6) == B200_OK && plan6[0] > 1) fused = true;
  }
  static const bool no_fuse = getenv("B200_NO_ROPE_FUSION") != nullptr;  // A/B aid
  if (fused && !no_fuse) {
    ep.rope = {positions, slot_mapping, cos_table, sin_table, (__nv_bfloat16*)k_pool, (__nv_bfloat16*)v_pool,
               (long long)num_blocks * block_size, max_pos, Hq, Hkv, D, block_size};
    return gemm_common(a, w, ep, EPI_ROPE, M, N, K, ab_dtype, split_k, workspace, workspace_bytes, stream);
  }
  int rc = gemm_common(a, w, ep, EPI_LINEAR, M, N, K, ab_dtype, split_k, workspace, workspace_bytes, stream);
  if (rc) return rc;
  return b200_rope_kv_write(qkv_out, positions, cos_table, sin_table, max_pos, k_pool, v_pool, slot_mapping, M, Hq, Hkv, D,
                            block_size, num_blocks, stream);
}

int b200_gemm_int8_wo(const void* a_bf16, const void* w_int8, void* out, const float* w_scale, const void* bias,
                      const void* residual, int M, int N, int K, int out_dtype, int split_k, void* workspace,
                      size_t workspace_bytes, void* stream) {
  return b200_gemm(a_bf16, w_int8, out, nullptr, w_scale, bias, residual, M, N, K, B200_INT8_WO, out_dtype, split_k, workspace,
                   workspace_bytes, stream);
}

int b200_gemm_silu_mul_quant(const void* a, const void* w_gu, void* out_q, float* out_scale, const float* a_scale,
                             const float* w_scale, const void* bias, void* state, int M, int N2, int K, int ab_dtype,
                             void* stream) {
  B200_REQUIRE(a && w_gu && out_q && out_scale && a_scale && w_scale && state, "gemm_silu_mul_quant: null pointer");
  B200_REQUIRE(ab_dtype == B200_FP8_E4M3 || ab_dtype == B200_INT8, "gemm_silu_mul_quant: W8A8 dtypes only (got %d)", ab_dtype);
  B200_REQUIRE(N2 % 32 == 0 && M > 0 && K > 0 && K % 16 == 0, "gemm_silu_mul_quant: bad shape M=%d N2=%d K=%d", M, N2, K);
  if (M > 128) {
    set_last_error("gemm_silu_mul_quant: decode schedule only (M <= 128), got %d", M);
    return B200_ERR_UNSUPPORTED;
  }
  GemmPlan p;
  int rc = plan_gemm(p, M, N2, K, ab_dtype, 0);
  if (rc) return rc;
  GemmEpilogue ep{a_scale, w_scale, (const __nv_bfloat16*)bias, nullptr, out_q, 0, nullptr, out_scale,
                  (unsigned long long*)state};
  CUtensorMap wm, am;
  rc = make_tmap_2d(&wm, w_gu, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)K, (uint64_t)N2, (uint64_t)K, GEMM_BK_BYTES, GEMM_BN);
  if (rc) return rc;
  rc = make_tmap_2d(&am, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)K, (uint64_t)M, (uint64_t)K, GEMM_BK_BYTES, p.bm);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (ab_dtype == B200_FP8_E4M3) return launch_silu_quant<KIND_F8F6F4>(wm, am, ep, M, N2, p, s);
  return launch_silu_quant<KIND_I8>(wm, am, ep, M, N2, p, s);
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
