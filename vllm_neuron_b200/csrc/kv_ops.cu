// Index construction on device, KV-cache write (plain and fused with RoPE), and the small
// HBM-bound glue kernels of the decoder step.  All byte/index work: 128-bit loads/stores, one CTA per token row.
#include "common.cuh"

namespace b200 {

// ============================================================ index construction (int work, bit-exact)
// model_runner.py:765-832: position = len(prompt)+len(output)-1, slot = bt[pos/bs]*bs + pos%bs
__global__ void build_decode_meta_kernel(const int32_t* __restrict__ block_table, int64_t bt_stride,
                                         const int32_t* __restrict__ row_ids, const int32_t* __restrict__ num_tokens,
                                         int64_t* __restrict__ slot_mapping, int64_t* __restrict__ positions,
                                         int32_t* __restrict__ ctx_lens, int B, int max_blocks, int bs, int num_spec) {
  griddep_launch();
  griddep_wait();
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  int r = row_ids ? row_ids[b] : b;
  int n = num_tokens[r];
  int pos = n - 1;
  int64_t slot = -1;
  if (pos >= 0 && pos / bs < max_blocks) slot = (int64_t)block_table[(int64_t)r * bt_stride + pos / bs] * bs + pos % bs;
  // speculation (model_runner.py:825-830): the row carries k consecutive slots, slots[0] + i -- NOT re-derived through the
  // block table, exactly as the reference builds them; a padded row (-1) stays -1 throughout
  for (int i = 0; i < num_spec; ++i) slot_mapping[(int64_t)b * num_spec + i] = slot < 0 ? -1 : slot + i;
  positions[b] = pos < 0 ? 0 : pos;
  ctx_lens[b] = n < 0 ? 0 : n;
}

// model_loader.py:308-334 (_remask_fused_spec_output): accepted tokens [B][T] arrive 0-padded; everything past the number
// of tokens generated this step (next_pos - position, clamped to [0, T]) becomes -1 so the runner can strip it
__global__ void remask_spec_output_kernel(const int64_t* __restrict__ accepted, const int64_t* __restrict__ next_pos,
                                          const int64_t* __restrict__ positions, int64_t* __restrict__ out, int B, int T) {
  griddep_launch();
  griddep_wait();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T) return;
  int b = i / T, t = i - b * T;
  long long cnt = next_pos[b] - positions[b];
  cnt = cnt < 0 ? 0 : cnt > T ? T : cnt;
  out[i] = t < cnt ? accepted[i] : -1;
}

// model_runner.py:739-763 + 869-874: slot list over range(max_model_len) sliced at num_computed, -1 padded
__global__ void build_prefill_meta_kernel(const int32_t* __restrict__ bt_row, int prompt_len, int num_computed,
                                          int64_t* __restrict__ slot_mapping, int64_t* __restrict__ positions,
                                          int max_model_len, int bs) {
  griddep_launch();
  griddep_wait();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= max_model_len) return;
  int tok = num_computed + i;
  bool live = tok < prompt_len && tok < max_model_len;
  slot_mapping[i] = live ? (int64_t)bt_row[tok / bs] * bs + tok % bs : -1;
  positions[i] = live ? tok : 0;
}

__global__ void block_table_scatter_kernel(int32_t* __restrict__ block_table, int64_t bt_stride,
                                           const int32_t* __restrict__ updates, int n, int rows, int max_blocks) {
  griddep_launch();
  griddep_wait();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int r = updates[3 * i], c = updates[3 * i + 1], v = updates[3 * i + 2];
  if (r >= 0 && r < rows && c >= 0 && c < max_blocks) block_table[(int64_t)r * bt_stride + c] = v;
}

__global__ void advance_decode_kernel(const int64_t* __restrict__ sampled, int64_t* __restrict__ tokens_io,
                                      int32_t* __restrict__ num_tokens, const int32_t* __restrict__ row_ids, int B) {
  griddep_launch();
  griddep_wait();
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  tokens_io[b] = sampled[b];
  int r = row_ids ? row_ids[b] : b;
  num_tokens[r] += 1;
}

// ============================================================ KV write
// pools [num_blocks][Hkv][bs][D]; one CTA per token, 16-byte vectors.
__global__ void kv_write_kernel(const uint4* __restrict__ k, const uint4* __restrict__ v, int64_t k_stride_v,
                                int64_t v_stride_v, uint4* __restrict__ k_pool, uint4* __restrict__ v_pool,
                                const int64_t* __restrict__ slot_mapping, int Hkv, int dvec, int bs, int64_t num_slots) {
  griddep_launch();
  griddep_wait();
  int t = blockIdx.x;
  int64_t slot = slot_mapping[t];
  if (slot < 0 || slot >= num_slots) return;  // -1 = padding (model_runner.py:76-81)
  int64_t blk = slot / bs;
  int off = (int)(slot % bs);
  int per_tok = Hkv * dvec;
  for (int i = threadIdx.x; i < per_tok; i += blockDim.x) {
    int h = i / dvec, d = i % dvec;
    int64_t dst = ((blk * Hkv + h) * bs + off) * dvec + d;
    k_pool[dst] = k[(int64_t)t * k_stride_v + i];
    v_pool[dst] = v[(int64_t)t * v_stride_v + i];
  }
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(p[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  uint32_t* p = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) p[i] = pack_bf16(f[2 * i], f[2 * i + 1]);
  return u;
}

// Fused RoPE(q,k) + KV scatter.  qkv row = [Hq*D | Hkv*D | Hkv*D].  Rotate-half; products and sums are
// rounded separately (no FMA contraction) so the cached K is bit-identical to the fp32 restatement.
__global__ void rope_kv_write_kernel(uint4* __restrict__ qkv, const int64_t* __restrict__ positions,
                                     const float* __restrict__ cos_t, const float* __restrict__ sin_t, int max_pos,
                                     uint4* __restrict__ k_pool, uint4* __restrict__ v_pool,
                                     const int64_t* __restrict__ slot_mapping, int Hq, int Hkv, int D, int bs,
                                     int64_t num_slots) {
  griddep_launch();
  griddep_wait();
  int t = blockIdx.x;
  int dvec = D / 8, hvec = D / 16;
  int64_t row_v = (int64_t)(Hq + 2 * Hkv) * dvec;
  uint4* row = qkv + (int64_t)t * row_v;
  int64_t pos = positions[t];
  if (pos < 0) pos = 0;
  if (pos >= max_pos) pos = max_pos - 1;
  int64_t slot = slot_mapping[t];
  bool live = slot >= 0 && slot < num_slots;
  int64_t blk = live ? slot / bs : 0;
  int off = live ? (int)(slot % bs) : 0;
  const float* c = cos_t + pos * (D / 2);
  const float* s = sin_t + pos * (D / 2);
  int n_rope = (Hq + Hkv) * hvec;
  for (int i = threadIdx.x; i < n_rope; i += blockDim.x) {
    int h = i / hvec, j = i % hvec;
    uint4* lo_p = row + (int64_t)h * dvec + j;
    uint4* hi_p = lo_p + hvec;
    uint4 lo = *lo_p, hi = *hi_p;
    float x1[8], x2[8], o1[8], o2[8];
    unpack8(lo, x1);
    unpack8(hi, x2);
    float4 c0 = *reinterpret_cast<const float4*>(c + 8 * j), c1 = *reinterpret_cast<const float4*>(c + 8 * j + 4);
    float4 s0 = *reinterpret_cast<const float4*>(s + 8 * j), s1 = *reinterpret_cast<const float4*>(s + 8 * j + 4);
    float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      o1[e] = __fsub_rn(__fmul_rn(x1[e], cc[e]), __fmul_rn(x2[e], ss[e]));
      o2[e] = __fadd_rn(__fmul_rn(x2[e], cc[e]), __fmul_rn(x1[e], ss[e]));
    }
    uint4 r1 = pack8(o1), r2 = pack8(o2);
    if (h < Hq) {
      *lo_p = r1;
      *hi_p = r2;
    } else if (live) {
      int kh = h - Hq;
      int64_t dst = ((blk * Hkv + kh) * bs + off) * dvec + j;
      k_pool[dst] = r1;
      k_pool[dst + hvec] = r2;
    }
  }
  if (live) {
    const uint4* vrow = row + (int64_t)(Hq + Hkv) * dvec;
    for (int i = threadIdx.x; i < Hkv * dvec; i += blockDim.x) {
      int h = i / dvec, d = i % dvec;
      v_pool[((blk * Hkv + h) * bs + off) * dvec + d] = vrow[i];
    }
  }
}

// ============================================================ block reductions
__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = (l < nw) ? sh[l] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* sh) {
  v = warp_max(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = (l < nw) ? sh[l] : 0.f;
  return warp_max(r);
}

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

template <int QD>
__device__ __forceinline__ uint2 quant8(const float (&y)[8], float scale) {
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) lo |= (uint32_t)quant_byte<QD>(y[e], scale) << (8 * e);
#pragma unroll
  for (int e = 0; e < 4; ++e) hi |= (uint32_t)quant_byte<QD>(y[4 + e], scale) << (8 * e);
  return make_uint2(lo, hi);
}

// ============================================================ RMSNorm (+ dynamic per-token quantisation)
constexpr int NORM_THREADS = 256;
constexpr int NORM_MAXV = 4;  // 4 x 8 bf16 per thread -> H <= 8192

template <int QD>
__global__ void __launch_bounds__(NORM_THREADS)
    rmsnorm_kernel(const uint4* __restrict__ x, const uint4* __restrict__ w, uint4* __restrict__ out_bf16,
                   uint2* __restrict__ out_q, float* __restrict__ out_scale, int H, float eps) {
  griddep_launch();
  griddep_wait();
  __shared__ float sh[32];
  int t = blockIdx.x, nvec = H / 8;
  const uint4* xr = x + (int64_t)t * nvec;
  float xv[NORM_MAXV][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NORM_MAXV; ++i) {
    int vi = threadIdx.x + i * NORM_THREADS;
    if (vi < nvec) {
      unpack8(xr[vi], xv[i]);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += xv[i][e] * xv[i][e];
    }
  }
  ss = block_sum(ss, sh);
  float rstd = rsqrtf(ss / (float)H + eps);
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < NORM_MAXV; ++i) {
    int vi = threadIdx.x + i * NORM_THREADS;
    if (vi < nvec) {
      float wv[8];
      unpack8(w[vi], wv);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        xv[i][e] = bf16_round(xv[i][e] * rstd * wv[e]);
        amax = fmaxf(amax, fabsf(xv[i][e]));
      }
      if (out_bf16) out_bf16[(int64_t)t * nvec + vi] = pack8(xv[i]);
    }
  }
  if constexpr (QD != B200_BF16) {
    amax = block_max(amax, sh);
    float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
    float scale = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
    if (threadIdx.x == 0) out_scale[t] = scale;
#pragma unroll
    for (int i = 0; i < NORM_MAXV; ++i) {
      int vi = threadIdx.x + i * NORM_THREADS;
      if (vi < nvec) out_q[(int64_t)t * nvec + vi] = quant8<QD>(xv[i], scale);
    }
  }
}

// ============================================================ SiLU(gate) * up (+ quantisation)
__device__ __forceinline__ void silu_mul8(const uint4& g, const uint4& u, float (&a)[8]) {
  float gf[8], uf[8];
  unpack8(g, gf);
  unpack8(u, uf);
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    a[e] = bf16_round(silu_f32(gf[e]) * uf[e]);
  }
}

template <int QD>
__global__ void __launch_bounds__(NORM_THREADS)
    silu_mul_kernel(const uint4* __restrict__ gate_up, uint4* __restrict__ out_bf16, uint2* __restrict__ out_q,
                    float* __restrict__ out_scale, int I) {
  griddep_launch();
  griddep_wait();
  __shared__ float sh[32];
  int t = blockIdx.x, nvec = I / 8;
  const uint4* g = gate_up + (int64_t)t * 2 * nvec;
  const uint4* u = g + nvec;
  float amax = 0.f;
  for (int vi = threadIdx.x; vi < nvec; vi += NORM_THREADS) {
    float a[8];
    silu_mul8(g[vi], u[vi], a);
    if (out_bf16) out_bf16[(int64_t)t * nvec + vi] = pack8(a);
#pragma unroll
    for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(a[e]));
  }
  if constexpr (QD != B200_BF16) {
    amax = block_max(amax, sh);
    float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
    float scale = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
    if (threadIdx.x == 0) out_scale[t] = scale;
    for (int vi = threadIdx.x; vi < nvec; vi += NORM_THREADS) {  // second read hits L1/L2
      float a[8];
      silu_mul8(g[vi], u[vi], a);
      out_q[(int64_t)t * nvec + vi] = quant8<QD>(a, scale);
    }
  }
}

// Dynamic per-token quantisation.  With group_amax == nullptr the CTA of token t scans its row for max|x| first.
// With group_amax [T][n_groups] (written by the fused gate/up GEMM epilogue) the row maximum is the max over the groups,
// so the row can be cut into gridDim.y column slices that quantise independently.
template <int QD>
__global__ void __launch_bounds__(NORM_THREADS)
    quantize_kernel(const uint4* __restrict__ x, uint2* __restrict__ out_q, float* __restrict__ out_scale, int K,
                    const float* __restrict__ group_amax, int n_groups) {
  griddep_launch();
  griddep_wait();
  __shared__ float sh[32];
  int t = blockIdx.x, nvec = K / 8;
  const uint4* xr = x + (int64_t)t * nvec;
  float amax = 0.f;
  int v0 = 0, v1 = nvec;
  if (group_amax) {
    for (int i = threadIdx.x; i < n_groups; i += NORM_THREADS) amax = fmaxf(amax, group_amax[(int64_t)t * n_groups + i]);
    int per = (nvec + gridDim.y - 1) / gridDim.y;
    v0 = blockIdx.y * per;
    v1 = v0 + per < nvec ? v0 + per : nvec;
  } else {
    for (int vi = threadIdx.x; vi < nvec; vi += NORM_THREADS) {
      float a[8];
      unpack8(xr[vi], a);
#pragma unroll
      for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(a[e]));
    }
  }
  amax = block_max(amax, sh);
  float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
  float scale = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
  if (threadIdx.x == 0 && blockIdx.y == 0) out_scale[t] = scale;
  for (int vi = v0 + threadIdx.x; vi < v1; vi += NORM_THREADS) {
    float a[8];
    unpack8(xr[vi], a);
    out_q[(int64_t)t * nvec + vi] = quant8<QD>(a, scale);
  }
}

__global__ void embedding_kernel(const int64_t* __restrict__ tokens, const uint4* __restrict__ table,
                                 uint4* __restrict__ out, int nvec, int V) {
  griddep_launch();
  griddep_wait();
  int t = blockIdx.x;
  int64_t tok = tokens[t];
  if (tok < 0) tok = 0;
  if (tok >= V) tok = V - 1;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) out[(int64_t)t * nvec + i] = table[tok * nvec + i];
}

// first index of the row maximum (ties -> lowest index).  A row (128 K fp32 logits, just written by lm_head, so mostly
// L2 hits) is cut into ARGMAX_CL slices scanned by the CTAs of one thread-block cluster with 128-bit loads; every CTA
// reduces its slice and stores (max, index) into the shared memory of cluster rank 0, which picks the row's winner.
// (One 1024-thread CTA per row with scalar loads took 37 us in the decode graph: 64 CTAs cannot pull 33 MB quickly.)
constexpr int ARGMAX_CL = 8;
constexpr int ARGMAX_THREADS = 256;

__device__ __forceinline__ void argmax_take(float& best, int& bi, float x, int i) {
  if (x > best || (x == best && i < bi)) {
    best = x;
    bi = i;
  }
}

__global__ void __launch_bounds__(ARGMAX_THREADS) argmax_kernel(const float* __restrict__ logits, int64_t row_stride,
                                                                float* __restrict__ out_val, int64_t* __restrict__ out_idx,
                                                                int V, int64_t idx_offset) {
  griddep_launch();
  griddep_wait();
  __shared__ float sv[32];
  __shared__ int si[32];
  __shared__ float cv[ARGMAX_CL];  // rank 0: one candidate per cluster rank
  __shared__ int ci[ARGMAX_CL];
  const int nr = (int)cluster_nctarank(), rk = (int)cluster_ctarank();
  const int t = blockIdx.x / nr;
  const float* r = logits + (int64_t)t * row_stride;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  // slice [lo, hi) of this rank, in units of 4 floats when the row is 16-byte aligned; the ragged tail goes to the last rank
  const bool vec = (reinterpret_cast<uintptr_t>(r) & 15) == 0;
  const int nv = vec ? V >> 2 : 0;
  const int lo = (int)((long long)nv * rk / nr), hi = (int)((long long)nv * (rk + 1) / nr);
  const float4* r4 = reinterpret_cast<const float4*>(r);
#pragma unroll 4
  for (int i = lo + threadIdx.x; i < hi; i += ARGMAX_THREADS) {
    const float4 x = r4[i];
    argmax_take(best, bi, x.x, 4 * i);
    argmax_take(best, bi, x.y, 4 * i + 1);
    argmax_take(best, bi, x.z, 4 * i + 2);
    argmax_take(best, bi, x.w, 4 * i + 3);
  }
  if (rk == nr - 1)
    for (int i = 4 * nv + threadIdx.x; i < V; i += ARGMAX_THREADS) argmax_take(best, bi, r[i], i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    argmax_take(best, bi, ov, oi);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) {
    sv[w] = best;
    si[w] = bi;
  }
  __syncthreads();
  if (w == 0) {
    constexpr int NW = ARGMAX_THREADS / 32;
    best = l < NW ? sv[l] : -INFINITY;
    bi = l < NW ? si[l] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      argmax_take(best, bi, ov, oi);
    }
    if (l == 0) {  // candidate of this rank -> rank 0's shared memory (DSMEM store; rank 0 stores locally the same way)
      st_cluster_f32(mapa_shared(smem_u32(&cv[rk]), 0u), best);
      st_cluster_f32(mapa_shared(smem_u32(&ci[rk]), 0u), __int_as_float(bi));
    }
  }
  cluster_sync_all();  // release / acquire: the candidates are visible to rank 0
  if (rk == 0 && threadIdx.x == 0) {
    best = -INFINITY;
    bi = 0x7fffffff;
    for (int q = 0; q < nr; ++q) argmax_take(best, bi, cv[q], ci[q]);
    if (out_val) out_val[t] = best;
    out_idx[t] = idx_offset + (bi == 0x7fffffff ? 0 : bi);
  }
}

}  // namespace b200

// ============================================================ C ABI
using namespace b200;

extern "C" {

int b200_build_decode_meta(const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                           const int32_t* num_tokens, int64_t* slot_mapping, int64_t* positions, int32_t* ctx_lens, int B,
                           int max_blocks, int block_size, int num_spec, void* stream) {
  B200_REQUIRE(block_table && num_tokens && slot_mapping && positions && ctx_lens, "build_decode_meta: null pointer");
  B200_REQUIRE(B >= 0 && max_blocks > 0 && block_size > 0, "build_decode_meta: bad sizes B=%d max_blocks=%d bs=%d", B,
               max_blocks, block_size);
  B200_REQUIRE(num_spec >= 1 && num_spec <= 64, "build_decode_meta: num_spec %d outside 1..64", num_spec);
  if (B == 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(build_decode_meta_kernel, dim3((B + 127) / 128), dim3(128), 0, (cudaStream_t)stream, block_table,
                             bt_row_stride, row_ids, num_tokens, slot_mapping, positions, ctx_lens, B, max_blocks, block_size,
                             num_spec));
  return B200_OK;
}

int b200_remask_spec_output(const int64_t* accepted, const int64_t* next_pos, const int64_t* positions, int64_t* out, int B,
                            int T, void* stream) {
  B200_REQUIRE(accepted && next_pos && positions && out, "remask_spec_output: null pointer");
  B200_REQUIRE(B >= 0 && T > 0, "remask_spec_output: bad sizes B=%d T=%d", B, T);
  if (B == 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(remask_spec_output_kernel, dim3((B * T + 127) / 128), dim3(128), 0, (cudaStream_t)stream, accepted,
                             next_pos, positions, out, B, T));
  return B200_OK;
}

int b200_build_prefill_meta(const int32_t* block_table_row, int prompt_len, int num_computed, int64_t* slot_mapping_out,
                            int64_t* positions_out, int max_model_len, int block_size, void* stream) {
  B200_REQUIRE(block_table_row && slot_mapping_out && positions_out, "build_prefill_meta: null pointer");
  B200_REQUIRE(prompt_len >= 0 && num_computed >= 0 && max_model_len > 0 && block_size > 0, "build_prefill_meta: bad sizes");
  B200_CUDA_OK(launch_kernel(build_prefill_meta_kernel, dim3((max_model_len + 255) / 256), dim3(256), 0, (cudaStream_t)stream, block_table_row, prompt_len, num_computed, slot_mapping_out, positions_out, max_model_len, block_size));
  return B200_OK;
}

int b200_block_table_scatter(int32_t* block_table, int64_t bt_row_stride, const int32_t* updates, int n, int rows,
                             int max_blocks, void* stream) {
  B200_REQUIRE(block_table && (updates || n == 0), "block_table_scatter: null pointer");
  if (n <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(block_table_scatter_kernel, dim3((n + 127) / 128), dim3(128), 0, (cudaStream_t)stream, block_table, bt_row_stride, updates, n,
                                                                                 rows, max_blocks));
  return B200_OK;
}

int b200_advance_decode(const int64_t* sampled, int64_t* tokens_io, int32_t* num_tokens, const int32_t* row_ids, int B,
                        void* stream) {
  B200_REQUIRE(sampled && tokens_io && num_tokens, "advance_decode: null pointer");
  if (B <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(advance_decode_kernel, dim3((B + 127) / 128), dim3(128), 0, (cudaStream_t)stream, sampled, tokens_io, num_tokens, row_ids, B));
  return B200_OK;
}

int b200_kv_write(const void* k, const void* v, int64_t k_token_stride, int64_t v_token_stride, void* k_pool, void* v_pool,
                  const int64_t* slot_mapping, int T, int Hkv, int D, int block_size, int num_blocks, void* stream) {
  B200_REQUIRE(k && v && k_pool && v_pool && slot_mapping, "kv_write: null pointer");
  B200_REQUIRE(D % 8 == 0 && k_token_stride % 8 == 0 && v_token_stride % 8 == 0, "kv_write: D and strides must be multiples of 8");
  B200_REQUIRE((((uintptr_t)k | (uintptr_t)v | (uintptr_t)k_pool | (uintptr_t)v_pool) & 15) == 0, "kv_write: 16-byte alignment");
  if (T <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(kv_write_kernel, dim3(T), dim3(128), 0, (cudaStream_t)stream, (const uint4*)k, (const uint4*)v, k_token_stride / 8,
                                                       v_token_stride / 8, (uint4*)k_pool, (uint4*)v_pool, slot_mapping,
                                                       Hkv, D / 8, block_size, (int64_t)num_blocks * block_size));
  return B200_OK;
}

int b200_rope_kv_write(void* qkv, const int64_t* positions, const float* cos_table, const float* sin_table, int max_pos,
                       void* k_pool, void* v_pool, const int64_t* slot_mapping, int T, int Hq, int Hkv, int D,
                       int block_size, int num_blocks, void* stream) {
  B200_REQUIRE(qkv && positions && cos_table && sin_table && k_pool && v_pool && slot_mapping, "rope_kv_write: null pointer");
  B200_REQUIRE(D % 16 == 0, "rope_kv_write: D must be a multiple of 16 (got %d)", D);
  if (T <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(rope_kv_write_kernel, dim3(T), dim3(256), 0, (cudaStream_t)stream, (uint4*)qkv, positions, cos_table, sin_table, max_pos,
                                                            (uint4*)k_pool, (uint4*)v_pool, slot_mapping, Hq, Hkv, D,
                                                            block_size, (int64_t)num_blocks * block_size));
  return B200_OK;
}

int b200_rmsnorm(const void* x, const void* w, void* out_bf16, void* out_q, float* out_scale, int T, int H, float eps,
                 int q_dtype, void* stream) {
  B200_REQUIRE(x && w, "rmsnorm: null pointer");
  B200_REQUIRE(H % 8 == 0 && H <= 8 * NORM_THREADS * NORM_MAXV, "rmsnorm: H=%d unsupported (multiple of 8, <= %d)", H,
               8 * NORM_THREADS * NORM_MAXV);
  if (T <= 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (q_dtype == B200_BF16) {
    B200_REQUIRE(out_bf16, "rmsnorm: out_bf16 required");
    B200_CUDA_OK(launch_kernel(rmsnorm_kernel<B200_BF16>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)x, (const uint4*)w, (uint4*)out_bf16, nullptr, nullptr, H, eps));
  } else if (q_dtype == B200_FP8_E4M3) {
    B200_REQUIRE(out_q && out_scale, "rmsnorm: out_q/out_scale required");
    B200_CUDA_OK(launch_kernel(rmsnorm_kernel<B200_FP8_E4M3>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)x, (const uint4*)w, (uint4*)out_bf16, (uint2*)out_q, out_scale, H, eps));
  } else if (q_dtype == B200_INT8) {
    B200_REQUIRE(out_q && out_scale, "rmsnorm: out_q/out_scale required");
    B200_CUDA_OK(launch_kernel(rmsnorm_kernel<B200_INT8>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)x, (const uint4*)w, (uint4*)out_bf16, (uint2*)out_q, out_scale, H, eps));
  } else {
    B200_REQUIRE(false, "rmsnorm: bad q_dtype %d", q_dtype);
  }
  return B200_OK;
}

int b200_silu_mul(const void* gate_up, void* out_bf16, void* out_q, float* out_scale, int T, int I, int q_dtype,
                  void* stream) {
  B200_REQUIRE(gate_up, "silu_mul: null pointer");
  B200_REQUIRE(I % 8 == 0, "silu_mul: I must be a multiple of 8");
  if (T <= 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (q_dtype == B200_BF16) {
    B200_REQUIRE(out_bf16, "silu_mul: out_bf16 required");
    B200_CUDA_OK(launch_kernel(silu_mul_kernel<B200_BF16>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)gate_up, (uint4*)out_bf16, nullptr, nullptr, I));
  } else if (q_dtype == B200_FP8_E4M3) {
    B200_REQUIRE(out_q && out_scale, "silu_mul: out_q/out_scale required");
    B200_CUDA_OK(launch_kernel(silu_mul_kernel<B200_FP8_E4M3>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)gate_up, (uint4*)out_bf16, (uint2*)out_q, out_scale, I));
  } else if (q_dtype == B200_INT8) {
    B200_REQUIRE(out_q && out_scale, "silu_mul: out_q/out_scale required");
    B200_CUDA_OK(launch_kernel(silu_mul_kernel<B200_INT8>, dim3(T), dim3(NORM_THREADS), 0, s, (const uint4*)gate_up, (uint4*)out_bf16, (uint2*)out_q, out_scale, I));
  } else {
    B200_REQUIRE(false, "silu_mul: bad q_dtype %d", q_dtype);
  }
  return B200_OK;
}

int b200_quantize_per_token(const void* x, void* out_q, float* out_scale, const float* group_amax, int n_groups, int T,
                            int K, int q_dtype, void* stream) {
  B200_REQUIRE(x && out_q && out_scale, "quantize_per_token: null pointer");
  B200_REQUIRE(K % 8 == 0, "quantize_per_token: K must be a multiple of 8");
  B200_REQUIRE(!group_amax || n_groups > 0, "quantize_per_token: n_groups must be positive with group_amax");
  if (T <= 0) return B200_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int slices = 1;
  if (group_amax) {  // no row scan needed: cut rows so that ~2 CTAs land on every SM
    slices = (2 * sm_count() + T - 1) / T;
    int max_slices = (K / 8 + NORM_THREADS - 1) / NORM_THREADS;
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
  }
  if (q_dtype == B200_FP8_E4M3)
    B200_CUDA_OK(launch_kernel(quantize_kernel<B200_FP8_E4M3>, dim3(T, slices), dim3(NORM_THREADS), 0, s, (const uint4*)x,
                               (uint2*)out_q, out_scale, K, group_amax, n_groups));
  else if (q_dtype == B200_INT8)
    B200_CUDA_OK(launch_kernel(quantize_kernel<B200_INT8>, dim3(T, slices), dim3(NORM_THREADS), 0, s, (const uint4*)x,
                               (uint2*)out_q, out_scale, K, group_amax, n_groups));
  else
    B200_REQUIRE(false, "quantize_per_token: bad q_dtype %d", q_dtype);
  return B200_OK;
}

int b200_embedding(const int64_t* tokens, const void* table, void* out, int T, int H, int V, void* stream) {
  B200_REQUIRE(tokens && table && out, "embedding: null pointer");
  B200_REQUIRE(H % 8 == 0, "embedding: H must be a multiple of 8");
  if (T <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(embedding_kernel, dim3(T), dim3(256), 0, (cudaStream_t)stream, tokens, (const uint4*)table, (uint4*)out, H / 8, V));
  return B200_OK;
}

int b200_argmax(const float* logits, int64_t row_stride, float* out_val, int64_t* out_idx, int T, int V,
                int64_t idx_offset, void* stream) {
  B200_REQUIRE(logits && out_idx, "argmax: null pointer");
  B200_REQUIRE(V > 0, "argmax: V must be positive");
  if (T <= 0) return B200_OK;
  // short rows (unit-test vocabularies) do not need the cluster; the headline vocabulary gets ARGMAX_CL CTAs per row
  const int cl = V >= 16384 ? ARGMAX_CL : 1;
  B200_CUDA_OK(launch_kernel_cluster(argmax_kernel, dim3(T * cl), dim3(ARGMAX_THREADS), 0, cl, (cudaStream_t)stream, logits,
                                     row_stride, out_val, out_idx, V, idx_offset));
  return B200_OK;
}

}  // extern "C"
