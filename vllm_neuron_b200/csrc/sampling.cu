// On-device sampling for the decode step: per-row (top_k, top_p, temperature) from a [B][3] tensor, as the reference hands
// to NxDI's on-device sampler (vllm_neuron/worker/neuronx_distributed_model_runner.py:1106-1140; top_k capped at 256,
// temperature 0 => top_k 1).  Two kernels:
//   topk_candidates_kernel : one CTA per row; exact radix select of the 256 largest fp32 logits of the (local) vocab shard,
//                            sorted by (value desc, index asc); ties broken towards the LOWER index like argmax_kernel.
//   sample_candidates_kernel: merges the candidate lists of all TP shards of a row, applies top-k / temperature / top-p and
//                            draws with the caller's uniform u[row] by inverse CDF (fp32, sequential, fixed order).
// Both are HBM/L2-bound integer+compare work (the logits row was just written by the lm_head GEMM); no tensor cores.
#include "common.cuh"

namespace b200 {

constexpr int SAMPLE_K = 256;        // candidates per (row, shard): the reference's _MAX_NEURON_SAMPLING_TOP_K (RUNNER:74)
constexpr int TOPK_THREADS = 1024;
constexpr int EQ_CAP = 1024;

// monotone map fp32 -> uint32 (larger float <=> larger key); -0 < +0 is irrelevant here
__device__ __forceinline__ uint32_t f2key(float x) {
  const uint32_t u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

// in-place bitonic sort of n (power of two) 64-bit keys in shared memory, DESCENDING; all threads of the CTA call it
__device__ __forceinline__ void bitonic_sort_desc(unsigned long long* a, int n) {
  for (int k = 2; k <= n; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = a[i], y = a[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (x < y) : (x > y)) {
            a[i] = y;
            a[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

// composite sort key: value key in the high word, inverted index in the low word (so that among equal values the lower
// index sorts first in a descending sort)
__device__ __forceinline__ unsigned long long pack_cand(uint32_t key, uint32_t idx) {
  return ((unsigned long long)key << 32) | (unsigned long long)(0xffffffffu - idx);
}

__global__ void __launch_bounds__(TOPK_THREADS)
    topk_candidates_kernel(const float* __restrict__ logits, int64_t row_stride, float* __restrict__ cand_val,
                           int64_t* __restrict__ cand_idx, int V, int64_t idx_offset) {
  griddep_launch();
  griddep_wait();
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long cand[SAMPLE_K];
  __shared__ unsigned int eq_idx[EQ_CAP];
  __shared__ unsigned int s_prefix, s_need, s_cnt, s_eq_cnt, s_eq_total;
  __shared__ unsigned int warp_tot[TOPK_THREADS / 32];
  const float* r = logits + (int64_t)blockIdx.x * row_stride;
  const int tid = threadIdx.x;
  const int K = V < SAMPLE_K ? V : SAMPLE_K;

  // ---- radix select of the K-th largest key: 4 passes of 8 bits, most significant first
  if (tid == 0) {
    s_prefix = 0;
    s_need = (unsigned)K;  // how many keys are still to be found among those matching the prefix
  }
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t pmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int i = tid; i < V; i += TOPK_THREADS) {
      const uint32_t k = f2key(r[i]);
      if ((k & pmask) == prefix) atomicAdd(&hist[(k >> shift) & 0xff], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      unsigned need = s_need, acc = 0;
      int b = 255;
      for (; b > 0; --b) {
        if (acc + hist[b] >= need) break;
        acc += hist[b];
      }
      s_prefix = prefix | ((uint32_t)b << shift);
      s_need = need - acc;  // keys wanted from bin b (>= 1)
      if (pass == 3) s_eq_total = hist[b];
    }
    __syncthreads();
  }
  const uint32_t T = s_prefix;          // key of the K-th largest element
  const unsigned need_eq = s_need;      // how many elements equal to T belong to the top K
  const unsigned eq_total = s_eq_total; // how many elements equal T in the row
  if (tid == 0) {
    s_cnt = 0;
    s_eq_cnt = 0;
  }
  __syncthreads();

  // ---- gather: every key > T (K - need_eq of them), then need_eq of the keys == T, lowest indices first
  const unsigned n_gt = (unsigned)K - need_eq;
  for (int i = tid; i < V; i += TOPK_THREADS) {
    const uint32_t k = f2key(r[i]);
    if (k > T) {
      const unsigned s = atomicAdd(&s_cnt, 1u);
      cand[s] = pack_cand(k, (uint32_t)i);
    } else if (k == T && eq_total <= EQ_CAP) {
      const unsigned s = atomicAdd(&s_eq_cnt, 1u);
      eq_idx[s] = (unsigned)i;
    }
  }
  __syncthreads();
  if (eq_total <= EQ_CAP) {
    if (eq_total == need_eq) {
      for (unsigned j = tid; j < eq_total; j += TOPK_THREADS) cand[n_gt + j] = pack_cand(T, eq_idx[j]);
    } else {
      // more ties than places: rank each tied index among the tied indices, keep the need_eq lowest
      for (unsigned j = tid; j < eq_total; j += TOPK_THREADS) {
        const unsigned me = eq_idx[j];
        unsigned rank = 0;
        for (unsigned q = 0; q < eq_total; ++q) rank += eq_idx[q] < me;
        if (rank < need_eq) cand[n_gt + rank] = pack_cand(T, me);
      }
    }
  } else {
    // pathological row (> EQ_CAP elements tie at the threshold, e.g. constant logits): ordered scan, chunk by chunk
    unsigned base = 0;  // ties seen in earlier chunks (uniform across the CTA)
    for (int c0 = 0; c0 < V && base < need_eq; c0 += TOPK_THREADS) {
      const int i = c0 + tid;
      const bool is_eq = i < V && f2key(r[i]) == T;
      const unsigned bal = __ballot_sync(0xffffffffu, is_eq);
      const int w = tid >> 5, l = tid & 31;
      if (l == 0) warp_tot[w] = __popc(bal);
      __syncthreads();
      unsigned before = 0, total = 0;
      for (int q = 0; q < TOPK_THREADS / 32; ++q) {
        const unsigned t = warp_tot[q];
        if (q < w) before += t;
        total += t;
      }
      const unsigned rank = base + before + __popc(bal & ((1u << l) - 1u));
      if (is_eq && rank < need_eq) cand[n_gt + rank] = pack_cand(T, (uint32_t)i);
      base += total;
      __syncthreads();
    }
  }
  for (int j = K + tid; j < SAMPLE_K; j += TOPK_THREADS) cand[j] = 0ull;  // V < 256: pad (sorts last)
  __syncthreads();
  bitonic_sort_desc(cand, SAMPLE_K);
  for (int j = tid; j < SAMPLE_K; j += TOPK_THREADS) {
    const unsigned long long c = cand[j];
    const bool ok = j < K;
    cand_val[(int64_t)blockIdx.x * SAMPLE_K + j] = ok ? key2f((uint32_t)(c >> 32)) : -INFINITY;
    cand_idx[(int64_t)blockIdx.x * SAMPLE_K + j] = ok ? idx_offset + (int64_t)(0xffffffffu - (uint32_t)c) : -1;
  }
}

// cand_val / cand_idx: [groups][B][256] (one group per TP shard, each sorted descending; -inf / -1 padding).
// params [B][3] fp32 = (top_k, top_p, temperature) (RUNNER:1106-1134).  u [B] uniform in [0, 1).
// Row algorithm (fp32, sequential, fixed order -> bit-reproducible; oracle/sampling_ref.py restates it):
//   merge the groups, sort by (value desc, index asc); k = clamp(top_k, 1, min(256, #finite));
//   w_j = expf((v_j - v_0) / temperature), j < k;  Z = sum_j w_j (ascending j);
//   top-p: keep the shortest prefix with cumulative weight >= top_p * Z (always at least one token);
//   draw: first kept j whose cumulative weight exceeds u * Z_kept (last kept token if rounding leaves none).
__global__ void __launch_bounds__(256)
    sample_candidates_kernel(const float* __restrict__ cand_val, const int64_t* __restrict__ cand_idx, int groups, int B,
                             const float* __restrict__ params, const float* __restrict__ u, int64_t* __restrict__ out_ids) {
  griddep_launch();
  griddep_wait();
  extern __shared__ unsigned long long merged[];  // [n_pow2] composite keys
  __shared__ float wgt[SAMPLE_K];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = groups * SAMPLE_K;
  int n2 = SAMPLE_K;
  while (n2 < n) n2 <<= 1;
  // composite key: value key (high word), inverted global token index (low word)
  for (int i = tid; i < n2; i += blockDim.x) {
    unsigned long long c = 0ull;
    if (i < n) {
      const int g = i / SAMPLE_K, j = i % SAMPLE_K;
      const int64_t off = ((int64_t)g * B + b) * SAMPLE_K + j;
      const int64_t id = cand_idx[off];
      if (id >= 0) c = pack_cand(f2key(cand_val[off]), (uint32_t)id);
    }
    merged[i] = c;
  }
  __syncthreads();
  if (groups > 1) bitonic_sort_desc(merged, n2);
  if (tid < SAMPLE_K) wgt[tid] = 0.f;
  __syncthreads();
  const float top_k_f = params[b * 3 + 0], top_p = params[b * 3 + 1], temp = params[b * 3 + 2];
  // number of real candidates among the first 256
  __shared__ int s_valid;
  if (tid == 0) s_valid = 0;
  __syncthreads();
  if (tid < SAMPLE_K && merged[tid] != 0ull && (tid == SAMPLE_K - 1 || merged[tid + 1] == 0ull)) s_valid = tid + 1;
  __syncthreads();
  int k = (int)top_k_f;
  if (k < 1) k = 1;
  if (k > s_valid) k = s_valid;
  const float v0 = key2f((uint32_t)(merged[0] >> 32));
  if (tid < k) wgt[tid] = expf(__fdiv_rn(key2f((uint32_t)(merged[tid] >> 32)) - v0, temp));
  __syncthreads();
  if (tid == 0) {
    int pick = 0;
    if (k > 1) {
      float z = 0.f;
      for (int j = 0; j < k; ++j) z += wgt[j];
      const float cut = top_p * z;
      float acc = 0.f;
      int kept = 0;
      for (int j = 0; j < k; ++j) {  // token j is kept iff the weight before it is still below the cut
        if (j > 0 && acc >= cut) break;
        acc += wgt[j];
        kept = j + 1;
      }
      const float target = u[b] * acc;
      float run = 0.f;
      pick = kept - 1;
      for (int j = 0; j < kept; ++j) {
        run += wgt[j];
        if (run > target) {
          pick = j;
          break;
        }
      }
    }
    const unsigned long long c = merged[pick];
    // global token index: idx_offset + V <= 2^32 is enforced by b200_topk_candidates, so the low word is the index
    out_ids[b] = (int64_t)(0xffffffffu - (uint32_t)c);
  }
}

// ---------------------------------------------------------------- log-probabilities (RUNNER:347-356, 1142-1239)
// The reference returns logprobs from vLLM's Sampler on full logits (raw log-softmax, the sampled token's rank, the top-n
// tokens).  On device: one pass per row for (max, sum exp(x - max)) of the local vocab shard, then everything else comes
// from the sorted 256-candidate lists b200_topk_candidates already produces.
constexpr int LSE_THREADS = 512;
__global__ void __launch_bounds__(LSE_THREADS) row_lse_kernel(const float* __restrict__ logits, int64_t row_stride,
                                                               float* __restrict__ out_max, float* __restrict__ out_sum,
                                                               int V) {
  griddep_launch();
  griddep_wait();
  __shared__ float red[LSE_THREADS / 32];
  __shared__ float bc;
  const float* r = logits + (int64_t)blockIdx.x * row_stride;
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < V; i += LSE_THREADS) m = fmaxf(m, r[i]);
  m = warp_max(m);
  if (l == 0) red[w] = m;
  __syncthreads();
  if (w == 0) {
    m = l < LSE_THREADS / 32 ? red[l] : -INFINITY;
    m = warp_max(m);
    if (l == 0) bc = m;
  }
  __syncthreads();
  m = bc;
  // fixed thread -> element map and a fixed reduction tree: the sum is bit-reproducible
  float s = 0.f;
  for (int i = threadIdx.x; i < V; i += LSE_THREADS) s += expf(r[i] - m);
  s = warp_sum(s);
  __syncthreads();
  if (l == 0) red[w] = s;
  __syncthreads();
  if (w == 0) {
    s = l < LSE_THREADS / 32 ? red[l] : 0.f;
    s = warp_sum(s);
    if (l == 0) {
      out_max[blockIdx.x] = m;
      out_sum[blockIdx.x] = s;
    }
  }
}

// One warp per row.  cand_* [groups][B][256] sorted (value desc, index asc) per group, row_max / row_sum [groups][B].
// out_ids / out_logprobs [B][1 + n]: column 0 = the sampled token, columns 1..n = the n most likely tokens;
// out_rank [B] = number of tokens whose logit is >= the sampled token's, itself included (vLLM's rank rule:
// batched_count_greater_than is `(x >= value).sum(-1)`).  The sampled token always is a
// candidate (the sampler draws from the candidates; the greedy token is candidate 0 of its shard).
__global__ void __launch_bounds__(128) gather_logprobs_kernel(const float* __restrict__ cand_val,
                                                              const int64_t* __restrict__ cand_idx, int groups, int B,
                                                              const int64_t* __restrict__ sampled,
                                                              const float* __restrict__ row_max,
                                                              const float* __restrict__ row_sum, int n,
                                                              int32_t* __restrict__ out_ids, float* __restrict__ out_lp,
                                                              int32_t* __restrict__ out_rank) {
  griddep_launch();
  griddep_wait();
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5), l = threadIdx.x & 31;
  if (b >= B) return;
  float m = -INFINITY;
  for (int g = 0; g < groups; ++g) m = fmaxf(m, row_max[(int64_t)g * B + b]);
  float s = 0.f;
  for (int g = 0; g < groups; ++g) s += row_sum[(int64_t)g * B + b] * expf(row_max[(int64_t)g * B + b] - m);
  const float lse = m + logf(s);
  const int64_t tok = sampled[b];
  float xs = -INFINITY;
  for (int g = 0; g < groups; ++g)
    for (int j = l; j < SAMPLE_K; j += 32) {
      const int64_t o = ((int64_t)g * B + b) * SAMPLE_K + j;
      if (cand_idx[o] == tok) xs = cand_val[o];
    }
  xs = warp_max(xs);
  int cnt = 0;
  for (int g = 0; g < groups; ++g)
    for (int j = l; j < SAMPLE_K; j += 32) {
      const int64_t o = ((int64_t)g * B + b) * SAMPLE_K + j;
      cnt += (cand_idx[o] >= 0 && cand_val[o] >= xs) ? 1 : 0;
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (l != 0) return;
  const int64_t row = (int64_t)b * (1 + n);
  out_ids[row] = (int32_t)tok;
  out_lp[row] = xs - lse;
  out_rank[b] = cnt;
  int head[16];
  for (int g = 0; g < groups; ++g) head[g] = 0;
  for (int j = 0; j < n; ++j) {  // k-way merge of the sorted shard lists: value desc, then index asc
    int bg = -1;
    float bv = -INFINITY;
    int64_t bi = 0;
    for (int g = 0; g < groups; ++g) {
      if (head[g] >= SAMPLE_K) continue;
      const int64_t o = ((int64_t)g * B + b) * SAMPLE_K + head[g];
      const int64_t ci = cand_idx[o];
      if (ci < 0) continue;
      const float cv = cand_val[o];
      if (bg < 0 || cv > bv || (cv == bv && ci < bi)) {
        bg = g;
        bv = cv;
        bi = ci;
      }
    }
    if (bg < 0) {
      out_ids[row + 1 + j] = -1;
      out_lp[row + 1 + j] = -INFINITY;
    } else {
      ++head[bg];
      out_ids[row + 1 + j] = (int32_t)bi;
      out_lp[row + 1 + j] = bv - lse;
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_topk_candidates(const float* logits, int64_t row_stride, float* cand_val, int64_t* cand_idx, int B, int V,
                         int64_t idx_offset, void* stream) {
  B200_REQUIRE(logits && cand_val && cand_idx, "topk_candidates: null pointer");
  B200_REQUIRE(V > 0 && idx_offset >= 0 && idx_offset + V <= 0xffffffffll, "topk_candidates: bad V / idx_offset");
  if (B <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(topk_candidates_kernel, dim3(B), dim3(TOPK_THREADS), 0, (cudaStream_t)stream, logits, row_stride,
                             cand_val, cand_idx, V, idx_offset));
  return B200_OK;
}

int b200_sample_candidates(const float* cand_val, const int64_t* cand_idx, int groups, int B, const float* params,
                           const float* u, int64_t* out_ids, void* stream) {
  B200_REQUIRE(cand_val && cand_idx && params && u && out_ids, "sample_candidates: null pointer");
  B200_REQUIRE(groups >= 1 && groups <= 16, "sample_candidates: groups %d outside 1..16", groups);
  if (B <= 0) return B200_OK;
  int n2 = SAMPLE_K;
  while (n2 < groups * SAMPLE_K) n2 <<= 1;
  B200_CUDA_OK(launch_kernel(sample_candidates_kernel, dim3(B), dim3(256), (size_t)n2 * 8, (cudaStream_t)stream, cand_val,
                             cand_idx, groups, B, params, u, out_ids));
  return B200_OK;
}

int b200_row_lse(const float* logits, int64_t row_stride, float* out_max, float* out_sumexp, int B, int V, void* stream) {
  B200_REQUIRE(logits && out_max && out_sumexp, "row_lse: null pointer");
  B200_REQUIRE(V > 0, "row_lse: V must be positive");
  if (B <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(row_lse_kernel, dim3(B), dim3(LSE_THREADS), 0, (cudaStream_t)stream, logits, row_stride, out_max,
                             out_sumexp, V));
  return B200_OK;
}

int b200_gather_logprobs(const float* cand_val, const int64_t* cand_idx, int groups, int B, const int64_t* sampled,
                         const float* row_max, const float* row_sumexp, int num_logprobs, int32_t* out_ids,
                         float* out_logprobs, int32_t* out_rank, void* stream) {
  B200_REQUIRE(cand_val && cand_idx && sampled && row_max && row_sumexp && out_ids && out_logprobs && out_rank,
               "gather_logprobs: null pointer");
  B200_REQUIRE(groups >= 1 && groups <= 16, "gather_logprobs: groups %d outside 1..16", groups);
  B200_REQUIRE(num_logprobs >= 0 && num_logprobs <= SAMPLE_K, "gather_logprobs: num_logprobs %d outside 0..256", num_logprobs);
  if (B <= 0) return B200_OK;
  B200_CUDA_OK(launch_kernel(gather_logprobs_kernel, dim3((B + 3) / 4), dim3(128), 0, (cudaStream_t)stream, cand_val, cand_idx,
                             groups, B, sampled, row_max, row_sumexp, num_logprobs, out_ids, out_logprobs, out_rank));
  return B200_OK;
}

}  // extern "C"
