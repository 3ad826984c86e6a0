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

template <int D, int NSTAGES, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32, 1)
    paged_attn_decode_kernel(const __grid_constant__ CUtensorMap k_map, const __grid_constant__ CUtensorMap v_map,
                             const __nv_bfloat16* __restrict__ q, int64_t q_stride,
                             const int32_t* __restrict__ block_table, int64_t bt_stride,
                             const int32_t* __restrict__ row_ids, const int32_t* __restrict__ ctx_lens,
                             float* __restrict__ part_o, float* __restrict__ part_ml, int B, int Hq, int Hkv,
                             int max_chunks, int chunk_blocks, float scale_log2) {
  constexpr int TILE = BS * D * 2;  // bytes of one K (or V) tile
  constexpr int KSTEPS = D / 16;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* my_tiles = smem + (size_t)warp * NSTAGES * 2 * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)NWARPS * NSTAGES * 2 * TILE) + warp * NSTAGES;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  __syncwarp();

  const int G = Hq / Hkv;
  const int worker = blockIdx.x * NWARPS + warp, nworkers = gridDim.x * NWARPS;
  const int total_items = B * Hkv * max_chunks;
  uint32_t parity = 0;  // bit s = parity to wait for on stage s
  const int h0 = (lane & 3) * 2;
  const int qn = lane >> 2;  // head slot this lane feeds into the B operand

  for (int item = worker; item < total_items; item += nworkers) {
    const int chunk = item % max_chunks;
    const int kvh = (item / max_chunks) % Hkv;
    const int b = item / (max_chunks * Hkv);
    const int ctx = ctx_lens[b];
    const int nblk_total = (ctx + BS - 1) / BS;
    const int blk0 = chunk * chunk_blocks;
    int nblk = nblk_total - blk0;
    if (nblk > chunk_blocks) nblk = chunk_blocks;
    if (nblk <= 0) continue;
    const int row = row_ids ? row_ids[b] : b;
    const int bt_reg = lane < nblk ? block_table[(int64_t)row * bt_stride + blk0 + lane] : 0;

    auto issue = [&](int s, int jj) {
      int blkid = __shfl_sync(0xffffffffu, bt_reg, jj);
      if (lane == 0)
        issue_tile_pair<D>(&k_map, &v_map, 0, my_tiles + (size_t)s * 2 * TILE, my_tiles + (size_t)s * 2 * TILE + TILE,
                           &bars[s], (blkid * Hkv + kvh) * BS);
    };
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s)
      if (s < nblk) issue(s, s);

    // Q^T as the B operand: lane holds head (lane/4), dims (lane%4)*2+{0,1} (+8) of every 16-dim step
    uint32_t qb[KSTEPS][2];
    {
      const __nv_bfloat16* qp = q + (int64_t)b * q_stride + (int64_t)(kvh * G + (qn < G ? qn : 0)) * D + h0;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t v0 = *reinterpret_cast<const uint32_t*>(qp + ks * 16);
        uint32_t v1 = *reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8);
        qb[ks][0] = qn < G ? v0 : 0u;
        qb[ks][1] = qn < G ? v1 : 0u;
      }
    }

    float oacc[KSTEPS][4];
#pragma unroll
    for (int i = 0; i < KSTEPS; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int j = 0; j < nblk; ++j) {
      const int s = j % NSTAGES;
      mbar_wait(&bars[s], (parity >> s) & 1u);
      parity ^= 1u << s;
      const uint32_t kbase = smem_u32(my_tiles + (size_t)s * 2 * TILE);
      const uint32_t vbase = kbase + TILE;
      const int nvalid = ctx - (blk0 + j) * BS;  // >= 1; >= 32 for all but the last block

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
      if (j + NSTAGES < nblk) issue(s, j + NSTAGES);
    }

    // ---- partial (unnormalised O, running max m in log2 units, sum l) for this chunk
#pragma unroll
    for (int o = 4; o <= 16; o <<= 1) {
      l0 += __shfl_xor_sync(0xffffffffu, l0, o);
      l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    }
    const int d0 = lane >> 2;
    if (h0 < G) {
      int64_t base = ((int64_t)(b * Hq + kvh * G + h0) * max_chunks + chunk);
#pragma unroll
      for (int dm = 0; dm < KSTEPS; ++dm) {
        part_o[base * D + dm * 16 + d0] = oacc[dm][0];
        part_o[base * D + dm * 16 + d0 + 8] = oacc[dm][2];
      }
      if (d0 == 0) {
        part_ml[base * 2] = m0;
        part_ml[base * 2 + 1] = l0;
      }
    }
    if (h0 + 1 < G) {
      int64_t base = ((int64_t)(b * Hq + kvh * G + h0 + 1) * max_chunks + chunk);
#pragma unroll
      for (int dm = 0; dm < KSTEPS; ++dm) {
        part_o[base * D + dm * 16 + d0] = oacc[dm][1];
        part_o[base * D + dm * 16 + d0 + 8] = oacc[dm][3];
      }
      if (d0 == 0) {
        part_ml[base * 2] = m1;
        part_ml[base * 2 + 1] = l1;
      }
    }
  }
}

// merge the per-chunk partials: out = sum_i 2^(m_i - M) O_i / sum_i 2^(m_i - M) l_i
template <int D>
__global__ void __launch_bounds__(D) paged_attn_combine_kernel(const float* __restrict__ part_o,
                                                               const float* __restrict__ part_ml,
                                                               const int32_t* __restrict__ ctx_lens,
                                                               __nv_bfloat16* __restrict__ out, int Hq, int max_chunks,
                                                               int chunk_blocks) {
  const int bh = blockIdx.x, b = bh / Hq, d = threadIdx.x;
  const int ctx = ctx_lens[b];
  const int nblk = (ctx + BS - 1) / BS;
  const int nch = (nblk + chunk_blocks - 1) / chunk_blocks;
  const float* po = part_o + (int64_t)bh * max_chunks * D;
  const float* pml = part_ml + (int64_t)bh * max_chunks * 2;
  float M = -INFINITY;
  for (int c = 0; c < nch; ++c) M = fmaxf(M, pml[2 * c]);
  float acc = 0.f, L = 0.f;
  for (int c = 0; c < nch; ++c) {
    float w = exp2f(pml[2 * c] - M);
    acc += w * po[(int64_t)c * D + d];
    L += w * pml[2 * c + 1];
  }
  out[(int64_t)bh * D + d] = __float2bfloat16_rn(nch > 0 && L > 0.f ? acc / L : 0.f);
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

int pick_chunk_blocks(int B, int Hkv, int max_blocks) {
  int workers = sm_count() * DEC_WARPS;
  int best = 32;
  double best_cost = 1e30;
  for (int cb = 32; cb >= 2; cb >>= 1) {
    int c = cb < max_blocks ? cb : max_blocks;
    long items = (long)B * Hkv * ((max_blocks + c - 1) / c);
    long rounds = (items + workers - 1) / workers;
    double cost = (double)rounds * (c + 1.5);
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = c;
    }
  }
  return best;
}

template <int D>
int launch_decode(const CUtensorMap& km, const CUtensorMap& vm, const void* q, int64_t q_stride, const int32_t* bt,
                  int64_t bt_stride, const int32_t* row_ids, const int32_t* ctx_lens, void* out, float* part_o,
                  float* part_ml, int B, int Hq, int Hkv, int max_chunks, int chunk_blocks, float scale,
                  cudaStream_t stream) {
  auto kern = paged_attn_decode_kernel<D, DEC_STAGES, DEC_WARPS>;
  constexpr size_t smem = (size_t)DEC_WARPS * DEC_STAGES * 2 * BS * D * 2 + DEC_WARPS * DEC_STAGES * 8 + 1024;
  static bool attr_done = false;
  if (!attr_done) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  long items = (long)B * Hkv * max_chunks;
  int grid = (int)((items + DEC_WARPS - 1) / DEC_WARPS);
  if (grid > sm_count()) grid = sm_count();
  if (grid < 1) grid = 1;
  kern<<<grid, DEC_WARPS * 32, smem, stream>>>(km, vm, (const __nv_bfloat16*)q, q_stride, bt, bt_stride, row_ids, ctx_lens,
                                               part_o, part_ml, B, Hq, Hkv, max_chunks, chunk_blocks,
                                               scale * 1.4426950408889634f);
  B200_CUDA_OK(cudaGetLastError());
  B200_LAUNCHED(1);
  paged_attn_combine_kernel<D><<<B * Hq, D, 0, stream>>>(part_o, part_ml, ctx_lens, (__nv_bfloat16*)out, Hq, max_chunks,
                                                        chunk_blocks);
  B200_CUDA_OK(cudaGetLastError());
  B200_LAUNCHED(1);
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
  kern<<<grid, 128, smem, stream>>>(km, vm, (const __nv_bfloat16*)q, q_stride, bt_row, (__nv_bfloat16*)out, S_new,
                                    computed_len, Hq, Hkv, scale * 1.4426950408889634f);
  B200_CUDA_OK(cudaGetLastError());
  B200_LAUNCHED(1);
  return B200_OK;
}

}  // namespace

extern "C" {

size_t b200_paged_attn_decode_workspace_bytes(int B, int Hq, int Hkv, int D, int max_blocks, int chunk_blocks) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || D <= 0 || max_blocks <= 0) return 0;
  // sized for the smallest chunk the heuristic may pick (2 blocks) unless the caller pins chunk_blocks
  int cb = chunk_blocks > 0 ? chunk_blocks : 2;
  size_t max_chunks = (size_t)(max_blocks + cb - 1) / cb;
  return (size_t)B * Hq * max_chunks * (D + 2) * sizeof(float);
}

int b200_paged_attn_decode(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool, int num_blocks,
                           const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                           const int32_t* ctx_lens, void* out, void* workspace, size_t workspace_bytes, int B, int Hq,
                           int Hkv, int D, int block_size, int max_blocks, int chunk_blocks, float scale, void* stream) {
  B200_REQUIRE(q && k_pool && v_pool && block_table && ctx_lens && out && workspace, "paged_attn_decode: null pointer");
  B200_REQUIRE(block_size == BS, "paged_attn_decode: block_size must be 32 (pa_block_size), got %d", block_size);
  B200_REQUIRE(Hkv > 0 && Hq % Hkv == 0 && Hq / Hkv <= 8, "paged_attn_decode: need Hq %% Hkv == 0 and group <= 8 (Hq=%d Hkv=%d)", Hq, Hkv);
  B200_REQUIRE(q_token_stride % 2 == 0 && ((uintptr_t)q & 3) == 0, "paged_attn_decode: q must be 4-byte aligned");
  if (D != 64 && D != 128) {
    set_last_error("paged_attn_decode: head_dim %d unsupported (64 or 128)", D);
    return B200_ERR_UNSUPPORTED;
  }
  if (B <= 0) return B200_OK;
  B200_REQUIRE(max_blocks > 0, "paged_attn_decode: max_blocks must be positive");
  int cb = chunk_blocks > 0 ? chunk_blocks : pick_chunk_blocks(B, Hkv, max_blocks);
  if (cb > 32) cb = 32;
  if (cb > max_blocks) cb = max_blocks;
  int max_chunks = (max_blocks + cb - 1) / cb;
  size_t need = (size_t)B * Hq * max_chunks * (D + 2) * sizeof(float);
  if (workspace_bytes < need) {
    set_last_error("paged_attn_decode: workspace %zu < %zu bytes", workspace_bytes, need);
    return B200_ERR_WORKSPACE;
  }
  CUtensorMap km, vm;
  int rc = pool_tmap(k_pool, num_blocks, Hkv, D, &km);
  if (rc) return rc;
  rc = pool_tmap(v_pool, num_blocks, Hkv, D, &vm);
  if (rc) return rc;
  float* part_o = (float*)workspace;
  float* part_ml = part_o + (size_t)B * Hq * max_chunks * D;
  cudaStream_t s = (cudaStream_t)stream;
  if (D == 128)
    return launch_decode<128>(km, vm, q, q_token_stride, block_table, bt_row_stride, row_ids, ctx_lens, out, part_o,
                              part_ml, B, Hq, Hkv, max_chunks, cb, scale, s);
  return launch_decode<64>(km, vm, q, q_token_stride, block_table, bt_row_stride, row_ids, ctx_lens, out, part_o, part_ml,
                           B, Hq, Hkv, max_chunks, cb, scale, s);
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
  if (D == 128)
    return launch_prefill<128>(km, vm, q, q_token_stride, block_table_row, out, S_new, computed_len, Hq, Hkv, scale, s);
  return launch_prefill<64>(km, vm, q, q_token_stride, block_table_row, out, S_new, computed_len, Hq, Hkv, scale, s);
}

}  // extern "C"
