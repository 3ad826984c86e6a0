// Block-KV paged attention for sm_100a.
//
// Decode (q_len = 1) is HBM-bound (~4 flop/byte): the design goal is to keep >= 100 KB of KV tiles in flight
// per SM.  KV pools are [num_blocks][Hkv][32][D] bf16, so one (block, kv head) tile is a contiguous 32 x D
// slab; it is fetched by TMA (cp.async.bulk.tensor, 128B swizzle, 64-column boxes) straight through the block
// table into a per-warp ring of stages.  Every warp is an independent worker on (sequence, kv head, context chunk)
// items (flash-decoding split along the context), so there is no CTA-wide synchronisation in the main loop.
// The math runs "swapped" on mma.sync m16n8k16: S^T = K Q^T with the <= 8 GQA query heads of the kv head in the
// N=8 slot, O^T = V^T P^T; P^T is obtained from the S^T accumulators with movmatrix.  A second kernel merges
// the per-chunk partials (log-sum-exp).
//
// Prefix-cached prefill (batch 1, q_len > 1) is a flash-attention-2 style kernel over the same TMA tiles with
// the causal mask shifted by the cached prefix length.
#include <mutex>
#include <unordered_map>

#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int BS = 32;  // pa_block_size (reference README.md:37; all BASELINE configs)

// byte address of 16-byte chunk `chunk` (0 .. D/8-1) of row `row` in a TMA-swizzled [D/64 boxes][32 rows][128 B] tile
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
  return base + ((chunk >> 3) << 12) + (row << 7) + (((chunk & 7) ^ (row & 7)) << 4);
}

template <int D>
__device__ __forceinline__ void issue_tile_pair(const CUtensorMap* kmap, const CUtensorMap* vmap, uint32_t kdst_generic_off,
                                                uint8_t* kdst, uint8_t* vdst, uint64_t* bar, int tile_row) {
  (void)kdst_generic_off;
  constexpr int NBOX = D / 64;
  mbar_expect_tx(bar, 2 * BS * D * 2);
#pragma unroll
  for (int bx = 0; bx < NBOX; ++bx) tma_load_2d(kdst + bx * 4096, kmap, bar, bx * 64, tile_row);
#pragma unroll
  for (int bx = 0; bx < NBOX; ++bx) tma_load_2d(vdst + bx * 4096, vmap, bar, bx * 64, tile_row);
}

// ---- balanced flash-decoding partition
// All (sequence, kv head) pairs are laid end to end as one list of 32-token blocks:
//   pair (b, kvh) owns global blocks [Hkv*pre[b] + kvh*nblk(b), +nblk(b)),  pre[b] = sum_{b' < b} nblk(b').
// Worker w (one warp) takes the contiguous range [w*R, (w+1)*R) of that list, R = max(ceil(total / workers), min_quota),
// so every warp streams the same number of KV bytes whatever the context lengths are (ragged batches, ctx % chunk != 0).
// A pair cut by a range boundary gets one partial (m, l, O) per worker that touched it: worker w writes part index
// w - first_worker(pair); the combine kernel recomputes the same arithmetic to know how many parts a pair has.
constexpr int DEC_MAX_B = 2048;  // sequences per launch (prefix table lives in shared memory)

struct PairPos {
  int b, kvh, blk, nblk;
};
// largest b with Hkv*pre[b] <= g (skips empty sequences), then the position inside the pair
__device__ __forceinline__ PairPos locate_block(const int* pre, int B, int Hkv, int g) {
  int lo = 0, hi = B - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (Hkv * pre[mid] <= g)
      lo = mid;
    else
      hi = mid - 1;
  }
  PairPos p;
  p.b = lo;
  p.nblk = pre[lo + 1] - pre[lo];
  int rem = g - Hkv * pre[lo];
  p.kvh = rem / p.nblk;
  p.blk = rem - p.kvh * p.nblk;
  return p;
}

// exclusive prefix of blocks-per-sequence into shared memory (pre[0..B]); all threads of the CTA call this
__device__ __forceinline__ int blocks_of(int ctx, int max_blocks) {
  int n = ctx > 0 ? (ctx + BS - 1) / BS : 0;
  return n < max_blocks ? n : max_blocks;
}
__device__ __forceinline__ void build_block_prefix(int* pre, const int32_t* __restrict__ ctx_lens, int B, int max_blocks) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    int carry = 0;
    for (int base = 0; base < B; base += 32) {
      int b = base + lane;
      int x = b < B ? blocks_of(ctx_lens[b], max_blocks) : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      if (b < B) pre[b + 1] = carry + x;
      carry += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) pre[0] = 0;
  }
  __syncthreads();
}

template <int D, int NSTAGES, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32, 1)
    paged_attn_decode_kernel(const __grid_constant__ CUtensorMap k_map, const __grid_constant__ CUtensorMap v_map,
                             const __nv_bfloat16* __restrict__ q, int64_t q_stride,
                             const int32_t* __restrict__ block_table, int64_t bt_stride,
                             const int32_t* __restrict__ row_ids, const int32_t* __restrict__ ctx_lens,
                             float* __restrict__ part_o, float* __restrict__ part_ml, int32_t* __restrict__ part_cnt,
                             int B, int Hq, int Hkv, int max_parts, int min_quota, int max_blocks, float scale_log2) {
  constexpr int TILE = BS * D * 2;  // bytes of one K (or V) tile
  constexpr int KSTEPS = D / 16;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* my_tiles = smem + (size_t)warp * NSTAGES * 2 * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)NWARPS * NSTAGES * 2 * TILE) + warp * NSTAGES;
  int* pre = reinterpret_cast<int*>(smem + (size_t)NWARPS * NSTAGES * 2 * TILE + NWARPS * NSTAGES * 8);

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  griddep_launch();
  griddep_wait();  // ctx_lens, q and the newest KV slot come from kernels earlier in the step
  build_block_prefix(pre, ctx_lens, B, max_blocks);

  const int G = Hq / Hkv;
  const int total = Hkv * pre[B];
  const int nworkers = gridDim.x * NWARPS;
  int R = (total + nworkers - 1) / nworkers;
  if (R < min_quota) R = min_quota;
  const int w = blockIdx.x * NWARPS + warp;
  const int g0 = w * R;
  if (g0 >= total) return;
  const int g1 = g0 + R < total ? g0 + R : total;

  uint32_t parity = 0;  // bit s = parity to wait for on stage s
  const int h0 = (lane & 3) * 2;
  const int qn = lane >> 2;  // head slot this lane feeds into the B operand

  // TMA row coordinate of global block g: ((block id) * Hkv + kvh) * 32, resolved through the block table
  auto tile_row_of = [&](int g) -> int {
    if (g >= g1) return 0;
    PairPos p = locate_block(pre, B, Hkv, g);
    const int row = row_ids ? row_ids[p.b] : p.b;
    const int blkid = block_table[(int64_t)row * bt_stride + p.blk];
    return (blkid * Hkv + p.kvh) * BS;
  };
  // lane l of tr_cur holds the coordinate of block (batch base + l); tr_next the following 32
  int tr_cur = tile_row_of(g0 + lane);
  int tr_next = tile_row_of(g0 + 32 + lane);
  auto issue = [&](int s, int rel) {  // rel = g - g0
    if ((rel & 31) == 0 && rel > 0) {
      tr_cur = tr_next;
      tr_next = tile_row_of(g0 + rel + 32 + lane);
    }
    const int tr = __shfl_sync(0xffffffffu, tr_cur, rel & 31);
    if (lane == 0)
      issue_tile_pair<D>(&k_map, &v_map, 0, my_tiles + (size_t)s * 2 * TILE, my_tiles + (size_t)s * 2 * TILE + TILE,
                         &bars[s], tr);
  };
#pragma unroll
  for (int s = 0; s < NSTAGES; ++s)
    if (g0 + s < g1) issue(s, s);

  PairPos cur = locate_block(pre, B, Hkv, g0);
  int ctx = ctx_lens[cur.b];
  bool fresh = true;
  uint32_t qb[KSTEPS][2];
  float oacc[KSTEPS][4];
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  for (int g = g0; g < g1; ++g) {
    if (fresh) {
      // Q^T as the B operand: lane holds head (lane/4), dims (lane%4)*2+{0,1} (+8) of every 16-dim step
      const __nv_bfloat16* qp = q + (int64_t)cur.b * q_stride + (int64_t)(cur.kvh * G + (qn < G ? qn : 0)) * D + h0;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t v0 = *reinterpret_cast<const uint32_t*>(qp + ks * 16);
        uint32_t v1 = *reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8);
        qb[ks][0] = qn < G ? v0 : 0u;
        qb[ks][1] = qn < G ? v1 : 0u;
      }
#pragma unroll
      for (int i = 0; i < KSTEPS; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
      m0 = m1 = -INFINITY;
      l0 = l1 = 0.f;
      fresh = false;
    }
    const int rel = g - g0;
    const int s = rel % NSTAGES;
    mbar_wait(&bars[s], (parity >> s) & 1u);
    parity ^= 1u << s;
    const uint32_t kbase = smem_u32(my_tiles + (size_t)s * 2 * TILE);
    const uint32_t vbase = kbase + TILE;
    const int nvalid = ctx - cur.blk * BS;  // >= 1; >= 32 for all but the last block

    // ---- S^T[32 tokens][8 heads] = K Q^T
    float sacc[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      sacc[mt][0] = sacc[mt][1] = sacc[mt][2] = sacc[mt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t a[4];
        ldmatrix_x4(a, tile_addr(kbase, mt * 16 + (lane & 15), ks * 2 + (lane >> 4)));
        mma_bf16_16816(sacc[mt], a, qb[ks][0], qb[ks][1]);
      }
    }
    // ---- online softmax, per head (2 heads per lane: h0, h0+1), log2 domain
    float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const bool ok = mt * 16 + hf * 8 + (lane >> 2) < nvalid;
        float x0 = ok ? sacc[mt][hf * 2] * scale_log2 : -INFINITY;
        float x1 = ok ? sacc[mt][hf * 2 + 1] * scale_log2 : -INFINITY;
        sacc[mt][hf * 2] = x0;
        sacc[mt][hf * 2 + 1] = x1;
        bm0 = fmaxf(bm0, x0);
        bm1 = fmaxf(bm1, x1);
      }
#pragma unroll
    for (int o = 4; o <= 16; o <<= 1) {
      bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, o));
      bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, o));
    }
    const float mn0 = fmaxf(m0, bm0), mn1 = fmaxf(m1, bm1);
    const float r0 = mn0 == -INFINITY ? 0.f : mn0, r1 = mn1 == -INFINITY ? 0.f : mn1;
    const float al0 = exp2f(m0 - r0), al1 = exp2f(m1 - r1);
    m0 = mn0;
    m1 = mn1;
    float ps0 = 0.f, ps1 = 0.f;
    uint32_t pb[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        float p0 = exp2f(sacc[mt][hf * 2] - r0);
        float p1 = exp2f(sacc[mt][hf * 2 + 1] - r1);
        ps0 += p0;
        ps1 += p1;
        pb[mt][hf] = movmatrix_trans(pack_bf16(p0, p1));  // -> (head = lane/4, tokens (lane%4)*2+{0,1})
      }
    l0 = l0 * al0 + ps0;
    l1 = l1 * al1 + ps1;
    // ---- O^T[D][8 heads] = alpha * O^T + V^T P^T
#pragma unroll
    for (int dm = 0; dm < KSTEPS; ++dm) {
      oacc[dm][0] *= al0;
      oacc[dm][1] *= al1;
      oacc[dm][2] *= al0;
      oacc[dm][3] *= al1;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        uint32_t a[4];
        ldmatrix_x4_trans(a, tile_addr(vbase, ks * 16 + (lane & 7) + ((lane >> 4) & 1) * 8, dm * 2 + ((lane >> 3) & 1)));
        mma_bf16_16816(oacc[dm], a, pb[ks][0], pb[ks][1]);
      }
    }
    __syncwarp();
    if (g + NSTAGES < g1) issue(s, rel + NSTAGES);

    // ---- end of this worker's part of the pair: park the partial (unnormalised O, running max in log2 units, sum)
    const bool pair_done = cur.blk + 1 == cur.nblk;
    if (pair_done || g + 1 == g1) {
      float t0 = l0, t1 = l1;
#pragma unroll
      for (int o = 4; o <= 16; o <<= 1) {
        t0 += __shfl_xor_sync(0xffffffffu, t0, o);
        t1 += __shfl_xor_sync(0xffffffffu, t1, o);
      }
      const int pair_start = Hkv * pre[cur.b] + cur.kvh * cur.nblk;
      const int part = w - pair_start / R;
      const int d0 = lane >> 2;
      // the worker that finishes the pair holds its last part: tell the merge kernel how many there are
      if (pair_done && lane == 0) part_cnt[cur.b * Hkv + cur.kvh] = part + 1;
      if (h0 < G) {
        int64_t base = ((int64_t)(cur.b * Hq + cur.kvh * G + h0) * max_parts + part);
#pragma unroll
        for (int dm = 0; dm < KSTEPS; ++dm) {
          part_o[base * D + dm * 16 + d0] = oacc[dm][0];
          part_o[base * D + dm * 16 + d0 + 8] = oacc[dm][2];
        }
        if (d0 == 0) {
          part_ml[base * 2] = m0;
          part_ml[base * 2 + 1] = t0;
        }
      }
      if (h0 + 1 < G) {
        int64_t base = ((int64_t)(cur.b * Hq + cur.kvh * G + h0 + 1) * max_parts + part);
#pragma unroll
        for (int dm = 0; dm < KSTEPS; ++dm) {
          part_o[base * D + dm * 16 + d0] = oacc[dm][1];
          part_o[base * D + dm * 16 + d0 + 8] = oacc[dm][3];
        }
        if (d0 == 0) {
          part_ml[base * 2] = m1;
          part_ml[base * 2 + 1] = t1;
        }
      }
      fresh = true;
    }
    // ---- advance the cursor
    if (pair_done) {
      cur.blk = 0;
      if (++cur.kvh == Hkv) {
        cur.kvh = 0;
        do {
          ++cur.b;
        } while (cur.b < B && pre[cur.b + 1] == pre[cur.b]);
        if (cur.b < B) {
          cur.nblk = pre[cur.b + 1] - pre[cur.b];
          ctx = ctx_lens[cur.b];
        }
      }
    } else {
      ++cur.blk;
    }
  }
}

// Merge the per-worker partials of one token and (optionally) quantise the token's attention output for the o_proj
// GEMM in the same pass:  out = sum_i 2^(m_i - M) O_i / sum_i 2^(m_i - M) l_i, rounded to bf16; then per-token
// scale = max|out| / qmax, q = rn(out / scale) exactly as quantize_per_token does on the bf16 tensor.
// Latency is all that matters here (a few MB out of L2): one thread per 8 output elements, and every load of the first
// COMB_CHUNK parts -- the part count, the (m, l) pairs and the partial rows -- is issued before any of them is used, so
// the common case (<= 4 parts per pair) costs one memory round trip.  Parts are always added in index order.
constexpr int COMB_THREADS = 512;
constexpr int COMB_MAXV = 4;   // 8-element vectors per thread -> Hq*D <= 16384
constexpr int COMB_CHUNK = 4;  // parts merged per round trip

template <int QD>
__global__ void __launch_bounds__(COMB_THREADS)
    paged_attn_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                              const int32_t* __restrict__ part_cnt, const int32_t* __restrict__ ctx_lens,
                              __nv_bfloat16* __restrict__ out, uint2* __restrict__ out_q, float* __restrict__ out_scale,
                              int B, int Hq, int Hkv, int D, int max_parts) {
  griddep_launch();
  griddep_wait();
  __shared__ float sh_f[COMB_THREADS / 32];
  const int b = blockIdx.x;
  const int wid = threadIdx.x >> 5, ln = threadIdx.x & 31;
  const int G = Hq / Hkv;
  const int dvec = D / 8, nvec = Hq * dvec;
  const bool empty = ctx_lens[b] <= 0;

  float y[COMB_MAXV][8];
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < COMB_MAXV; ++i) {
    const int vi = threadIdx.x + i * COMB_THREADS;
    if (vi < nvec) {
      const int h = vi / dvec, dv = vi % dvec;
      const float2* pml = reinterpret_cast<const float2*>(part_ml) + (int64_t)(b * Hq + h) * max_parts;
      const float* po = part_o + (int64_t)(b * Hq + h) * max_parts * D + dv * 8;
      // ---- first chunk: everything in flight at once (slots >= nparts hold stale data and are masked below)
      const int cnt_raw = part_cnt[b * Hkv + h / G];
      float2 ml0[COMB_CHUNK];
      float4 a0[COMB_CHUNK], a1[COMB_CHUNK];
#pragma unroll
      for (int c = 0; c < COMB_CHUNK; ++c) {
        const int cc = c < max_parts ? c : 0;
        ml0[c] = pml[cc];
        a0[c] = *reinterpret_cast<const float4*>(po + (int64_t)cc * D);
        a1[c] = *reinterpret_cast<const float4*>(po + (int64_t)cc * D + 4);
      }
      const int nparts = empty ? 0 : (cnt_raw < max_parts ? cnt_raw : max_parts);
      float Mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < COMB_CHUNK; ++c)
        if (c < nparts) Mx = fmaxf(Mx, ml0[c].x);
      for (int c0 = COMB_CHUNK; c0 < nparts; c0 += COMB_CHUNK) {
        float mm[COMB_CHUNK];
#pragma unroll
        for (int c = 0; c < COMB_CHUNK; ++c) mm[c] = c0 + c < nparts ? pml[c0 + c].x : -INFINITY;
#pragma unroll
        for (int c = 0; c < COMB_CHUNK; ++c) Mx = fmaxf(Mx, mm[c]);
      }
      float acc[8], L = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
      for (int c = 0; c < COMB_CHUNK; ++c) {
        if (c < nparts) {
          const float wgt = exp2f(ml0[c].x - Mx);
          acc[0] += wgt * a0[c].x; acc[1] += wgt * a0[c].y; acc[2] += wgt * a0[c].z; acc[3] += wgt * a0[c].w;
          acc[4] += wgt * a1[c].x; acc[5] += wgt * a1[c].y; acc[6] += wgt * a1[c].z; acc[7] += wgt * a1[c].w;
          L += wgt * ml0[c].y;
        }
      }
      for (int c0 = COMB_CHUNK; c0 < nparts; c0 += COMB_CHUNK) {
        float2 ml[COMB_CHUNK];
        float4 b0[COMB_CHUNK], b1[COMB_CHUNK];
#pragma unroll
        for (int c = 0; c < COMB_CHUNK; ++c) {
          const int cc = c0 + c < nparts ? c0 + c : c0;
          ml[c] = pml[cc];
          b0[c] = *reinterpret_cast<const float4*>(po + (int64_t)cc * D);
          b1[c] = *reinterpret_cast<const float4*>(po + (int64_t)cc * D + 4);
        }
#pragma unroll
        for (int c = 0; c < COMB_CHUNK; ++c) {
          if (c0 + c < nparts) {
            const float wgt = exp2f(ml[c].x - Mx);
            acc[0] += wgt * b0[c].x; acc[1] += wgt * b0[c].y; acc[2] += wgt * b0[c].z; acc[3] += wgt * b0[c].w;
            acc[4] += wgt * b1[c].x; acc[5] += wgt * b1[c].y; acc[6] += wgt * b1[c].z; acc[7] += wgt * b1[c].w;
            L += wgt * ml[c].y;
          }
        }
      }
      const bool ok = nparts > 0 && L > 0.f;
      uint32_t packed[4];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float v = ok ? acc[e] / L : 0.f;
        y[i][e] = __bfloat162float(__float2bfloat16_rn(v));
        amax = fmaxf(amax, fabsf(y[i][e]));
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) packed[e] = pack_bf16(y[i][2 * e], y[i][2 * e + 1]);
      if (out) *reinterpret_cast<uint4*>(out + ((int64_t)b * Hq + h) * D + dv * 8) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    }
  }
  if constexpr (QD != B200_BF16) {
    amax = warp_max(amax);
    if (ln == 0) sh_f[wid] = amax;
    __syncthreads();
    amax = 0.f;
#pragma unroll
    for (int i = 0; i < COMB_THREADS / 32; ++i) amax = fmaxf(amax, sh_f[i]);
    const float qmax = QD == B200_FP8_E4M3 ? 448.f : 127.f;
    const float scale = amax > 0.f ? __fdiv_rn(amax, qmax) : 1.f;
    if (threadIdx.x == 0) out_scale[b] = scale;
#pragma unroll
    for (int i = 0; i < COMB_MAXV; ++i) {
      const int vi = threadIdx.x + i * COMB_THREADS;
      if (vi < nvec) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) lo |= (uint32_t)quant_byte<QD>(y[i][e], scale) << (8 * e);
#pragma unroll
        for (int e = 0; e < 4; ++e) hi |= (uint32_t)quant_byte<QD>(y[i][4 + e], scale) << (8 * e);
        out_q[(int64_t)b * nvec + vi] = make_uint2(lo, hi);
      }
    }
  }
}

// ============================================================ prefix-cached prefill (one sequence)
// CTA = 4 warps x 16 query positions of one q head; KV tiles double-buffered by TMA and shared by the CTA.
template <int D>
__global__ void __launch_bounds__(128)
    paged_attn_prefill_kernel(const __grid_constant__ CUtensorMap k_map, const __grid_constant__ CUtensorMap v_map,
                              const __nv_bfloat16* __restrict__ q, int64_t q_stride,
                              const int32_t* __restrict__ bt_row, __nv_bfloat16* __restrict__ out, int S_new,
                              int computed_len, int Hq, int Hkv, float scale_log2) {
  constexpr int TILE = BS * D * 2;
  constexpr int KSTEPS = D / 16;
  constexpr int NST = 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + NST * 2 * TILE);
  uint64_t* empty = full + NST;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.y, kvh = head / (Hq / Hkv);
  const int q0 = blockIdx.x * 64;
  griddep_launch();
  griddep_wait();
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 4);
    }
    fence_mbar_init();
    fence_proxy_async();
  }
  __syncthreads();

  int q_last = q0 + 63 < S_new - 1 ? q0 + 63 : S_new - 1;
  const int nblk = (computed_len + q_last) / BS + 1;
  auto issue = [&](int s, int j) {
    issue_tile_pair<D>(&k_map, &v_map, 0, smem + (size_t)s * 2 * TILE, smem + (size_t)s * 2 * TILE + TILE, &full[s],
                       (bt_row[j] * Hkv + kvh) * BS);
  };
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST && s < nblk; ++s) issue(s, s);
  }

  // Q as the A operand (rows = query positions)
  const int r_lo = q0 + warp * 16 + (lane >> 2), r_hi = r_lo + 8;
  const bool v_lo = r_lo < S_new, v_hi = r_hi < S_new;
  uint32_t qa[KSTEPS][4];
  {
    const __nv_bfloat16* p_lo = q + (int64_t)(v_lo ? r_lo : 0) * q_stride + (int64_t)head * D + (lane & 3) * 2;
    const __nv_bfloat16* p_hi = q + (int64_t)(v_hi ? r_hi : 0) * q_stride + (int64_t)head * D + (lane & 3) * 2;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      qa[ks][0] = v_lo ? *reinterpret_cast<const uint32_t*>(p_lo + ks * 16) : 0u;
      qa[ks][1] = v_hi ? *reinterpret_cast<const uint32_t*>(p_hi + ks * 16) : 0u;
      qa[ks][2] = v_lo ? *reinterpret_cast<const uint32_t*>(p_lo + ks * 16 + 8) : 0u;
      qa[ks][3] = v_hi ? *reinterpret_cast<const uint32_t*>(p_hi + ks * 16 + 8) : 0u;
    }
  }
  float oacc[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
  float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.f, l_hi = 0.f;
  const int qpos_lo = computed_len + r_lo, qpos_hi = computed_len + r_hi;

  for (int j = 0; j < nblk; ++j) {
    const int s = j % NST;
    mbar_wait(&full[s], (j / NST) & 1);
    const uint32_t kbase = smem_u32(smem + (size_t)s * 2 * TILE), vbase = kbase + TILE;
    float sacc[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
#pragma unroll
      for (int np = 0; np < 2; ++np) {  // pairs of 8-token n-tiles
        uint32_t bfr[4];
        ldmatrix_x4(bfr, tile_addr(kbase, (np * 2 + (lane >> 4)) * 8 + (lane & 7), ks * 2 + ((lane >> 3) & 1)));
        mma_bf16_16816(sacc[np * 2], qa[ks], bfr[0], bfr[1]);
        mma_bf16_16816(sacc[np * 2 + 1], qa[ks], bfr[2], bfr[3]);
      }
    }
    float bm_lo = -INFINITY, bm_hi = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        int kpos = j * BS + nt * 8 + (lane & 3) * 2 + e;
        float x_lo = kpos <= qpos_lo ? sacc[nt][e] * scale_log2 : -INFINITY;
        float x_hi = kpos <= qpos_hi ? sacc[nt][2 + e] * scale_log2 : -INFINITY;
        sacc[nt][e] = x_lo;
        sacc[nt][2 + e] = x_hi;
        bm_lo = fmaxf(bm_lo, x_lo);
        bm_hi = fmaxf(bm_hi, x_hi);
      }
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
      bm_lo = fmaxf(bm_lo, __shfl_xor_sync(0xffffffffu, bm_lo, o));
      bm_hi = fmaxf(bm_hi, __shfl_xor_sync(0xffffffffu, bm_hi, o));
    }
    const float mn_lo = fmaxf(m_lo, bm_lo), mn_hi = fmaxf(m_hi, bm_hi);
    const float rf_lo = mn_lo == -INFINITY ? 0.f : mn_lo, rf_hi = mn_hi == -INFINITY ? 0.f : mn_hi;
    const float al_lo = exp2f(m_lo - rf_lo), al_hi = exp2f(m_hi - rf_hi);
    m_lo = mn_lo;
    m_hi = mn_hi;
    float ps_lo = 0.f, ps_hi = 0.f;
    uint32_t pa[2][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      float p0 = exp2f(sacc[nt][0] - rf_lo), p1 = exp2f(sacc[nt][1] - rf_lo);
      float p2 = exp2f(sacc[nt][2] - rf_hi), p3 = exp2f(sacc[nt][3] - rf_hi);
      ps_lo += p0 + p1;
      ps_hi += p2 + p3;
      pa[nt >> 1][(nt & 1) * 2] = pack_bf16(p0, p1);
      pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
    l_lo = l_lo * al_lo + ps_lo;
    l_hi = l_hi * al_hi + ps_hi;
#pragma unroll
    for (int dn = 0; dn < D / 8; dn += 2) {
      oacc[dn][0] *= al_lo;
      oacc[dn][1] *= al_lo;
      oacc[dn][2] *= al_hi;
      oacc[dn][3] *= al_hi;
      oacc[dn + 1][0] *= al_lo;
      oacc[dn + 1][1] *= al_lo;
      oacc[dn + 1][2] *= al_hi;
      oacc[dn + 1][3] *= al_hi;
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        uint32_t bfr[4];
        ldmatrix_x4_trans(bfr, tile_addr(vbase, kt * 16 + ((lane >> 3) & 1) * 8 + (lane & 7), dn + (lane >> 4)));
        mma_bf16_16816(oacc[dn], pa[kt], bfr[0], bfr[1]);
        mma_bf16_16816(oacc[dn + 1], pa[kt], bfr[2], bfr[3]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
    if (threadIdx.x == 0 && j + NST < nblk) {
      mbar_wait(&empty[s], (j / NST) & 1);
      issue(s, j + NST);
    }
  }
#pragma unroll
  for (int o = 1; o <= 2; o <<= 1) {
    l_lo += __shfl_xor_sync(0xffffffffu, l_lo, o);
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, o);
  }
  const float inv_lo = l_lo > 0.f ? 1.f / l_lo : 0.f, inv_hi = l_hi > 0.f ? 1.f / l_hi : 0.f;
#pragma unroll
  for (int dn = 0; dn < D / 8; ++dn) {
    int col = dn * 8 + (lane & 3) * 2;
    if (v_lo)
      *reinterpret_cast<uint32_t*>(out + ((int64_t)r_lo * Hq + head) * D + col) =
          pack_bf16(oacc[dn][0] * inv_lo, oacc[dn][1] * inv_lo);
    if (v_hi)
      *reinterpret_cast<uint32_t*>(out + ((int64_t)r_hi * Hq + head) * D + col) =
          pack_bf16(oacc[dn][2] * inv_hi, oacc[dn][3] * inv_hi);
  }
}

}  // namespace b200

// ============================================================ host side / C ABI
using namespace b200;

namespace {

struct MapKey {
  const void* p;
  int nb, hkv, d;
  bool operator==(const MapKey& o) const { return p == o.p && nb == o.nb && hkv == o.hkv && d == o.d; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.p) ^ (size_t)k.nb * 1315423911u ^ (size_t)k.hkv * 2654435761u ^ (size_t)k.d;
  }
};
std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

// pool viewed as a 2D bf16 tensor [num_blocks*Hkv*32 rows][D cols]; box = 64 cols x 32 rows, 128B swizzle
int pool_tmap(const void* pool, int num_blocks, int Hkv, int D, CUtensorMap* out) {
  MapKey key{pool, num_blocks, Hkv, D};
  {
    std::lock_guard<std::mutex> g(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
      *out = it->second;
      return 0;
    }
  }
  int rc = make_tmap_2d(out, pool, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)D, (uint64_t)num_blocks * Hkv * BS,
                        (uint64_t)D * 2, 64, BS);
  if (rc) return rc;
  std::lock_guard<std::mutex> g(g_map_mu);
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[key] = *out;
  return 0;
}

constexpr int DEC_WARPS = 4;
constexpr int DEC_STAGES = 3;
constexpr int DEC_MIN_QUOTA = 4;  // blocks per worker below which spreading further only adds partials

struct DecodePlan {
  int min_quota, max_parts, grid;
  size_t ws_bytes;
};
DecodePlan plan_decode(int B, int Hq, int Hkv, int D, int max_blocks, int min_quota) {
  DecodePlan p;
  p.min_quota = min_quota > 0 ? min_quota : DEC_MIN_QUOTA;
  p.max_parts = (max_blocks + p.min_quota - 1) / p.min_quota + 1;
  long max_total = (long)B * Hkv * max_blocks;
  long want = (max_total + (long)p.min_quota * DEC_WARPS - 1) / ((long)p.min_quota * DEC_WARPS);
  p.grid = (int)(want < sm_count() ? want : sm_count());
  if (p.grid < 1) p.grid = 1;
  p.ws_bytes = (size_t)B * Hq * p.max_parts * (D + 2) * sizeof(float) + (size_t)B * Hkv * sizeof(int32_t);
  return p;
}

template <int D>
int launch_decode(const CUtensorMap& km, const CUtensorMap& vm, const void* q, int64_t q_stride, const int32_t* bt,
                  int64_t bt_stride, const int32_t* row_ids, const int32_t* ctx_lens, void* out, void* out_q,
                  float* out_scale, int q_dtype, float* part_o, float* part_ml, int32_t* part_cnt, int B, int Hq, int Hkv,
                  int max_blocks, const DecodePlan& p, float scale, cudaStream_t stream) {
  auto kern = paged_attn_decode_kernel<D, DEC_STAGES, DEC_WARPS>;
  const size_t smem = (size_t)DEC_WARPS * DEC_STAGES * 2 * BS * D * 2 + DEC_WARPS * DEC_STAGES * 8 + (size_t)(B + 1) * 4 + 1024;
  constexpr size_t smem_max = (size_t)DEC_WARPS * DEC_STAGES * 2 * BS * D * 2 + DEC_WARPS * DEC_STAGES * 8 + (size_t)(DEC_MAX_B + 1) * 4 + 1024;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
    attr_done = true;
  }
  B200_CUDA_OK(launch_kernel(kern, dim3(p.grid), dim3(DEC_WARPS * 32), smem, stream, km, vm, (const __nv_bfloat16*)q,
                             q_stride, bt, bt_stride, row_ids, ctx_lens, part_o, part_ml, part_cnt, B, Hq, Hkv, p.max_parts,
                             p.min_quota, max_blocks, scale * 1.4426950408889634f));
  if (q_dtype == B200_FP8_E4M3)
    B200_CUDA_OK(launch_kernel(paged_attn_combine_kernel<B200_FP8_E4M3>, dim3(B), dim3(COMB_THREADS), 0, stream, part_o,
                               part_ml, part_cnt, ctx_lens, (__nv_bfloat16*)out, (uint2*)out_q, out_scale, B, Hq, Hkv, D,
                               p.max_parts));
  else if (q_dtype == B200_INT8)
    B200_CUDA_OK(launch_kernel(paged_attn_combine_kernel<B200_INT8>, dim3(B), dim3(COMB_THREADS), 0, stream, part_o,
                               part_ml, part_cnt, ctx_lens, (__nv_bfloat16*)out, (uint2*)out_q, out_scale, B, Hq, Hkv, D,
                               p.max_parts));
  else
    B200_CUDA_OK(launch_kernel(paged_attn_combine_kernel<B200_BF16>, dim3(B), dim3(COMB_THREADS), 0, stream, part_o,
                               part_ml, part_cnt, ctx_lens, (__nv_bfloat16*)out, (uint2*)nullptr, (float*)nullptr, B, Hq, Hkv,
                               D, p.max_parts));
  return B200_OK;
}

template <int D>
int launch_prefill(const CUtensorMap& km, const CUtensorMap& vm, const void* q, int64_t q_stride, const int32_t* bt_row,
                   void* out, int S_new, int computed_len, int Hq, int Hkv, float scale, cudaStream_t stream) {
  auto kern = paged_attn_prefill_kernel<D>;
  constexpr size_t smem = (size_t)2 * 2 * BS * D * 2 + 64 + 1024;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  dim3 grid((S_new + 63) / 64, Hq);
  B200_CUDA_OK(launch_kernel(kern, grid, dim3(128), smem, stream, km, vm, (const __nv_bfloat16*)q, q_stride, bt_row,
                             (__nv_bfloat16*)out, S_new, computed_len, Hq, Hkv, scale * 1.4426950408889634f));
  return B200_OK;
}

}  // namespace

extern "C" {

size_t b200_paged_attn_decode_workspace_bytes(int B, int Hq, int Hkv, int D, int max_blocks, int min_quota) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || max_blocks <= 0) return 0;
  return plan_decode(B, Hq, Hkv, D, max_blocks, min_quota).ws_bytes;
}

int b200_paged_attn_decode(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool, int num_blocks,
                           const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                           const int32_t* ctx_lens, void* out, void* out_q, float* out_scale, int q_dtype, void* workspace,
                           size_t workspace_bytes, int B, int Hq, int Hkv, int D, int block_size, int max_blocks,
                           int min_quota, float scale, void* stream) {
  B200_REQUIRE(q && k_pool && v_pool && block_table && ctx_lens && workspace, "paged_attn_decode: null pointer");
  B200_REQUIRE(q_dtype == B200_BF16 || q_dtype == B200_FP8_E4M3 || q_dtype == B200_INT8, "paged_attn_decode: bad q_dtype %d", q_dtype);
  B200_REQUIRE(q_dtype == B200_BF16 ? out != nullptr : (out_q && out_scale), "paged_attn_decode: output pointer(s) missing");
  B200_REQUIRE(block_size == BS, "paged_attn_decode: block_size must be 32 (pa_block_size), got %d", block_size);
  B200_REQUIRE(Hkv > 0 && Hq % Hkv == 0 && Hq / Hkv <= 8, "paged_attn_decode: need Hq %% Hkv == 0 and group <= 8 (Hq=%d Hkv=%d)", Hq, Hkv);
  B200_REQUIRE(q_token_stride % 2 == 0 && ((uintptr_t)q & 3) == 0, "paged_attn_decode: q must be 4-byte aligned");
  B200_REQUIRE(B <= DEC_MAX_B, "paged_attn_decode: at most %d sequences per call (got %d)", DEC_MAX_B, B);
  if (D != 64 && D != 128) {
    set_last_error("paged_attn_decode: head_dim %d unsupported (64 or 128)", D);
    return B200_ERR_UNSUPPORTED;
  }
  if (Hq * D > 8 * COMB_THREADS * COMB_MAXV) {
    set_last_error("paged_attn_decode: Hq*D = %d exceeds %d", Hq * D, 8 * COMB_THREADS * COMB_MAXV);
    return B200_ERR_UNSUPPORTED;
  }
  if (B <= 0) return B200_OK;
  B200_REQUIRE(max_blocks > 0, "paged_attn_decode: max_blocks must be positive");
  DecodePlan p = plan_decode(B, Hq, Hkv, D, max_blocks, min_quota);
  if (workspace_bytes < p.ws_bytes) {
    set_last_error("paged_attn_decode: workspace %zu < %zu bytes", workspace_bytes, p.ws_bytes);
    return B200_ERR_WORKSPACE;
  }
  CUtensorMap km, vm;
  int rc = pool_tmap(k_pool, num_blocks, Hkv, D, &km);
  if (rc) return rc;
  rc = pool_tmap(v_pool, num_blocks, Hkv, D, &vm);
  if (rc) return rc;
  float* part_o = (float*)workspace;
  float* part_ml = part_o + (size_t)B * Hq * p.max_parts * D;
  int32_t* part_cnt = reinterpret_cast<int32_t*>(part_ml + (size_t)B * Hq * p.max_parts * 2);
  cudaStream_t s = (cudaStream_t)stream;
  if (D == 128)
    return launch_decode<128>(km, vm, q, q_token_stride, block_table, bt_row_stride, row_ids, ctx_lens, out, out_q,
                              out_scale, q_dtype, part_o, part_ml, part_cnt, B, Hq, Hkv, max_blocks, p, scale, s);
  return launch_decode<64>(km, vm, q, q_token_stride, block_table, bt_row_stride, row_ids, ctx_lens, out, out_q, out_scale,
                           q_dtype, part_o, part_ml, part_cnt, B, Hq, Hkv, max_blocks, p, scale, s);
}

int b200_paged_attn_prefill(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool, int num_blocks,
                            const int32_t* block_table_row, void* out, int S_new, int computed_len, int Hq, int Hkv,
                            int D, int block_size, float scale, void* stream) {
  B200_REQUIRE(q && k_pool && v_pool && block_table_row && out, "paged_attn_prefill: null pointer");
  B200_REQUIRE(block_size == BS, "paged_attn_prefill: block_size must be 32, got %d", block_size);
  B200_REQUIRE(Hkv > 0 && Hq % Hkv == 0, "paged_attn_prefill: Hq %% Hkv != 0");
  B200_REQUIRE(computed_len >= 0 && q_token_stride % 2 == 0, "paged_attn_prefill: bad computed_len/stride");
  if (D != 64 && D != 128) {
    set_last_error("paged_attn_prefill: head_dim %d unsupported (64 or 128)", D);
    return B200_ERR_UNSUPPORTED;
  }
  if (S_new <= 0) return B200_OK;
  CUtensorMap km, vm;
  int rc = pool_tmap(k_pool, num_blocks, Hkv, D, &km);
  if (rc) return rc;
  rc = pool_tmap(v_pool, num_blocks, Hkv, D, &vm);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // prompts / chunks of >= 64 new tokens: the tcgen05 kernel (128-row query tiles in TMEM); shorter ones would leave most of
  // a 128-row tile empty and keep the mma.sync kernel (64-row CTAs).  B200_PREFILL_LEGACY=1 forces the latter (A/B aid).
  static const bool legacy = getenv("B200_PREFILL_LEGACY") != nullptr;
  if (S_new >= 64 && !legacy && ((uintptr_t)q & 15) == 0 && (q_token_stride * 2) % 16 == 0)
    return launch_prefill_tc(km, vm, q, q_token_stride, block_table_row, out, S_new, computed_len, Hq, Hkv, D, scale, s);
  if (D == 128)
    return launch_prefill<128>(km, vm, q, q_token_stride, block_table_row, out, S_new, computed_len, Hq, Hkv, scale, s);
  return launch_prefill<64>(km, vm, q, q_token_stride, block_table_row, out, S_new, computed_len, Hq, Hkv, scale, s);
}

}  // extern "C"

// ---------------------------------------------------------------- decode attention for callers that cannot allocate
// The flash-decoding partials normally live in a caller-supplied workspace.  Some hosts have no way to get one: a
// foreign-language binding that can only hand over the tensors its framework already owns, a plugin running inside a
// serving process whose device-memory budget was fixed (and fully claimed by the KV pool) before the plugin was loaded,
// or code that runs where cudaMalloc is forbidden.  For them the partials go into the array below, which belongs to the
// loaded library image: every GPU that loads libb200_paged.so carries these bytes until the process exits.
constexpr size_t DECODE_STATIC_BYTES = (size_t)B200_DECODE_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_decode_static_scratch[DECODE_STATIC_BYTES];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mu;
void* g_static_ptr[STATIC_MAX_DEVICES] = {};  // device address of g_decode_static_scratch, per device ordinal

// In host code `g_decode_static_scratch` is only a handle for the runtime's symbol table, not memory a kernel can touch:
// ask the runtime for the array's address on the CURRENT device (each device has its own copy), once per device.
int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0) {
    (void)cudaGetLastError();
    set_last_error("paged_attn_decode_static: no current CUDA device (%s)", cudaGetErrorString(e));
    return B200_ERR_NO_DEVICE;
  }
  B200_REQUIRE(dev < STATIC_MAX_DEVICES, "paged_attn_decode_static: device ordinal %d >= %d", dev, STATIC_MAX_DEVICES);
  std::lock_guard<std::mutex> g(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_decode_static_scratch);
    if (e != cudaSuccess || !ptr) {
      (void)cudaGetLastError();
      set_last_error("paged_attn_decode_static: cudaGetSymbolAddress failed on device %d (%s)", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : B200_ERR_DRIVER;
    }
    g_static_ptr[dev] = ptr;
  }
  *out = g_static_ptr[dev];
  return B200_OK;
}
}  // namespace

extern "C" {

size_t b200_paged_attn_decode_static_max_bytes(void) { return DECODE_STATIC_BYTES; }

int b200_paged_attn_decode_static(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool,
                                  int num_blocks, const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                                  const int32_t* ctx_lens, void* out, void* out_q, float* out_scale, int q_dtype, int B,
                                  int Hq, int Hkv, int D, int block_size, int max_blocks, int min_quota, float scale,
                                  void* stream) {
  // the size check needs no device, so an oversized shape is refused the same way on every machine
  if (B > 0 && Hq > 0 && Hkv > 0 && D > 0 && max_blocks > 0) {
    const size_t need = plan_decode(B, Hq, Hkv, D, max_blocks, min_quota).ws_bytes;
    if (need > DECODE_STATIC_BYTES) {
      set_last_error(
          "paged_attn_decode_static: this shape needs %zu bytes of scratch but the built-in array holds %zu "
          "(B200_DECODE_STATIC_SCRATCH_BYTES); raise min_quota, split the batch, or use b200_paged_attn_decode with "
          "your own workspace",
          need, DECODE_STATIC_BYTES);
      return B200_ERR_WORKSPACE;
    }
  }
  void* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // Nothing in the scratch has to start at zero: the decode kernel writes every partial and every part count that the merge
  // kernel reads for this call's shapes, and entries left over from earlier calls lie beyond those counts.
  return b200_paged_attn_decode(q, q_token_stride, k_pool, v_pool, num_blocks, block_table, bt_row_stride, row_ids, ctx_lens,
                                out, out_q, out_scale, q_dtype, scratch, DECODE_STATIC_BYTES, B, Hq, Hkv, D, block_size,
                                max_blocks, min_quota, scale, stream);
}

}  // extern "C"
