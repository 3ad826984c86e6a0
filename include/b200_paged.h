/*
 * b200_paged.h -- C ABI of libb200_paged.so: the B200 (sm_100a) replacement for the
 * opaque device call on vllm-neuron's decode hot path.
 *
 * What it replaces.  In the reference every numeric op of a step happens inside ONE call,
 *   self.model(input_ids, position_ids, input_block_ids, slot_mapping, block_tables,
 *              full_context_lens, computed_context_lens, sampling_params, ...)
 *   vllm_neuron/worker/neuronx_distributed_model_runner.py:517-530
 *   -> vllm_neuron/worker/neuronx_distributed_model_loader.py:336-348  (NxDI forward)
 * The entry points below are the pieces of that call, plus the device-side version of the
 * host index construction that feeds it (model_runner.py:739-763, 798-832, 848-936).
 *
 * Conventions.  Plain pointers and sizes only (no torch types).  Every pointer is a DEVICE
 * pointer unless named host_*.  `stream` is a cudaStream_t passed as void*.  No entry point
 * allocates device memory or synchronises the device: scratch is passed in as `workspace`
 * (size from the matching *_workspace_bytes).  Return value 0 = ok; otherwise a B200_ERR_*
 * code (or a cudaError_t for launch failures) and b200_last_error() holds a message.
 * There is no CPU fallback: without an sm_100 device every compute call fails.
 *
 * Block-KV contract (reference: model_runner.py:76-81, platform.py:149-159):
 *   slot = block_id * block_size + offset;  slot == -1 is padding (never written);
 *   block 0 is vLLM's null block;  block tables are right-padded with 0.
 * Device pool layout (ours; NxDI's is not observable): K and V pools are separate arrays
 *   [num_blocks][Hkv][block_size][D] bf16, so one (block, kv head) tile is contiguous.
 */
#ifndef B200_PAGED_H_
#define B200_PAGED_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 8

enum {
  B200_OK = 0,
  B200_ERR_INVALID_ARG = -1,
  B200_ERR_NO_DEVICE = -2,      /* no CUDA device with compute capability 10.x */
  B200_ERR_WORKSPACE = -3,      /* workspace too small */
  B200_ERR_UNSUPPORTED = -4,    /* shape outside what the kernels cover */
  B200_ERR_DRIVER = -5          /* cuTensorMapEncodeTiled / driver entry point failure */
};

/* B200_INT8_WO is only valid as a GEMM `ab_dtype`: bf16 activations x int8 weights (weight-only quantisation) */
enum { B200_BF16 = 0, B200_FP8_E4M3 = 1, B200_INT8 = 2, B200_F32 = 3, B200_INT8_WO = 4 };

int b200_abi_version(void);
const char* b200_last_error(void);
/* 0 iff `device` is an sm_100 part (reference analogue: /dev/neuron* probe, vllm_neuron/__init__.py:8-23) */
int b200_device_check(int device);
/* number of kernel launches issued through this library by the calling process (bench.py's gpu_launches) */
uint64_t b200_launch_count(void);
/* Programmatic dependent launch for every kernel of the library (default on): a kernel may become resident while its
 * predecessor on the stream drains; the GEMM uses that window to prefetch weight tiles.  Returns the previous setting. */
int b200_set_pdl(int enabled);

/* ---------------------------------------------------------------- index construction on device
 * Replaces model_runner.py:765-832 (+ :887-917): per decode row b (persistent row r = row_ids[b]):
 *   position = num_tokens[r] - 1; full_ctx = position + 1;
 *   slot = block_table[r][position / bs] * bs + position % bs.
 * Outputs are int64 like the reference's tensors (ctx_lens int32 for the kernels). */
int b200_build_decode_meta(const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                           const int32_t* num_tokens, int64_t* slot_mapping, int64_t* positions, int32_t* ctx_lens,
                           int B, int max_blocks, int block_size, int num_spec, void* stream);
/* num_spec = 1: slot_mapping is [B].  num_spec = k > 1 (EAGLE / fused speculation, model_runner.py:825-830):
 * slot_mapping is [B][k] with slot_mapping[b][i] = slot_mapping[b][0] + i, built exactly like the reference's
 * `slots.append(slots[0] + i)` (not re-derived through the block table); a padded row stays -1.
 *
 * model_loader.py:308-334 (_remask_fused_spec_output): accepted [B][T] int64 is 0-padded; entries past
 * clamp(next_pos[b] - positions[b], 0, T) become -1 in out so that the runner strips them (model_runner.py:306-323). */
int b200_remask_spec_output(const int64_t* accepted, const int64_t* next_pos, const int64_t* positions, int64_t* out, int B,
                            int T, void* stream);
/* Replaces model_runner.py:739-763 (+ :853-885): batch-1 prefill with a prefix-cache hit.
 * slot_mapping_out[i] for i in [0, max_model_len): slot of token (num_computed + i) while that
 * token < prompt_len, else -1 (the reference's slice [num_computed:] right-padded with -1).
 * positions_out[i] = num_computed + i for the same range, else 0. */
int b200_build_prefill_meta(const int32_t* block_table_row, int prompt_len, int num_computed, int64_t* slot_mapping_out,
                            int64_t* positions_out, int max_model_len, int block_size, void* stream);
/* Mirror of InputBatch.block_table.append_row / add_row (model_runner.py:461-471, 484-486):
 * updates = n packed (row, col, block_id) int32 triples already on the device. */
int b200_block_table_scatter(int32_t* block_table, int64_t bt_row_stride, const int32_t* updates, int n, int rows,
                             int max_blocks, void* stream);
/* End of a decode step on device: tokens_io[b] = sampled[b]; num_tokens[row_ids[b]] += 1
 * (the device-side twin of model_runner.py:325-339, so a captured step can be replayed). */
int b200_advance_decode(const int64_t* sampled, int64_t* tokens_io, int32_t* num_tokens, const int32_t* row_ids, int B,
                        void* stream);

/* ---------------------------------------------------------------- KV-cache write
 * Scatter k, v [T][Hkv][D] (token stride in elements) into the pools by slot_mapping; slot < 0 skipped. */
int b200_kv_write(const void* k, const void* v, int64_t k_token_stride, int64_t v_token_stride, void* k_pool, void* v_pool,
                  const int64_t* slot_mapping, int T, int Hkv, int D, int block_size, int num_blocks, void* stream);
/* Fused: rotate q and k of a packed qkv row [T][(Hq+2Hkv)*D] (rotate-half RoPE, fp32 cos/sin tables
 * [max_pos][D/2]), write roped q back in place, and scatter roped k and v into the pools. */
int b200_rope_kv_write(void* qkv, const int64_t* positions, const float* cos_table, const float* sin_table, int max_pos,
                       void* k_pool, void* v_pool, const int64_t* slot_mapping, int T, int Hq, int Hkv, int D,
                       int block_size, int num_blocks, void* stream);

/* ---------------------------------------------------------------- paged attention
 * Decode (q_len = 1): out[b][h][:] = softmax(q K^T scale) V over ctx_lens[b] tokens gathered through
 * block_table[row_ids ? row_ids[b] : b].  Flash-decoding with a BALANCED split: all (sequence, kv head) pairs are laid
 * end to end as one list of 32-token blocks and every warp takes an equal contiguous share (>= min_quota blocks; 0 picks
 * the default), so ragged context lengths cost nothing; per-share partials live in `workspace` and a second kernel merges
 * them.  Outputs: q_dtype == B200_BF16 -> out [B][Hq][D] bf16;  FP8_E4M3 / INT8 -> out_q [B][Hq*D] with the dynamic
 * per-token scale out_scale [B] (same arithmetic as b200_quantize_per_token on the bf16 result), out optional.
 * ctx_lens[b] must be <= max_blocks * 32. */
size_t b200_paged_attn_decode_workspace_bytes(int B, int Hq, int Hkv, int D, int max_blocks, int min_quota);
int b200_paged_attn_decode(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool, int num_blocks,
                           const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                           const int32_t* ctx_lens, void* out, void* out_q, float* out_scale, int q_dtype, void* workspace,
                           size_t workspace_bytes, int B, int Hq, int Hkv, int D, int block_size, int max_blocks,
                           int min_quota, float scale, void* stream);
/* Decode attention for callers that cannot allocate device memory.  (No counterpart in the reference: NxDI owns all of
 * its device buffers, so vllm-neuron never faces the question.)
 * b200_paged_attn_decode wants a workspace for its flash-decoding partials.  Not every caller can produce one: a binding
 * from a language runtime that may only pass along tensors its framework already holds, a plugin loaded into a serving
 * process after the device-memory budget was handed out (vLLM gives what is left after the weights to the KV pool), or
 * any host that is not allowed to call cudaMalloc.  This variant takes the same inputs and outputs but no workspace: the
 * partials go into an array of B200_DECODE_STATIC_SCRATCH_BYTES that is part of the library image, so EVERY GPU that loads
 * the library gives up that much memory until the process exits, whether or not this entry point is ever called.
 *   - Limit: a shape whose b200_paged_attn_decode_workspace_bytes(...) exceeds the array is refused with
 *     B200_ERR_WORKSPACE (the message gives both numbers) and nothing is launched.  The headline shape (B=64, Hq=32,
 *     Hkv=8, D=128, 64 blocks per sequence) needs 18.1 MB; a larger min_quota shrinks the need.
 *   - It calls no cudaMalloc / cudaFree and no allocating library routine, and does not synchronise.
 *   - First call on each device: the library looks up the array's device address (cudaGetSymbolAddress, with that device
 *     current) and keeps it.  Make that first call on every device you use BEFORE you start capturing a stream.
 *   - There is ONE array per device and every call uses it.  Two calls that can overlap on the same GPU (different
 *     streams, or different host threads feeding different streams) corrupt each other's partials; keep them on one
 *     stream or order them with events.  The array keeps whatever the last call left there; nothing in it needs to be
 *     zero, the kernels overwrite every entry they later read.
 * Results are bit-identical to b200_paged_attn_decode with the same arguments. */
#define B200_DECODE_STATIC_SCRATCH_BYTES (32u << 20)
size_t b200_paged_attn_decode_static_max_bytes(void);
int b200_paged_attn_decode_static(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool,
                                  int num_blocks, const int32_t* block_table, int64_t bt_row_stride, const int32_t* row_ids,
                                  const int32_t* ctx_lens, void* out, void* out_q, float* out_scale, int q_dtype, int B,
                                  int Hq, int Hkv, int D, int block_size, int max_blocks, int min_quota, float scale,
                                  void* stream);
/* Prefix-cached prefill, one sequence: queries for absolute positions computed_len .. computed_len+S_new-1,
 * causal over cached prefix + new tokens, all K/V already in the pools. */
int b200_paged_attn_prefill(const void* q, int64_t q_token_stride, const void* k_pool, const void* v_pool, int num_blocks,
                            const int32_t* block_table_row, void* out, int S_new, int computed_len, int Hq, int Hkv,
                            int D, int block_size, float scale, void* stream);

/* ---------------------------------------------------------------- per-channel-scaled GEMM (tcgen05 / TMEM / TMA)
 * out[M][N] = (sum_k a[m][k] * w[n][k]) * a_scale[m] * w_scale[n] + bias[n] + residual[m][n]
 * a: [M][K], w: [N][K], both `ab_dtype` (BF16 | FP8_E4M3 | INT8); scales fp32 (NULL = 1), bias bf16 [N] or NULL,
 * residual bf16 [M][N] or NULL; out_dtype BF16 | F32.
 * M <= 128 (decode): one CTA per (128-row weight tile, K split); the `split_k` CTAs (0 = auto; 1, 2, 4 or 8) of a tile
 *   form a thread-block cluster and are summed through distributed shared memory in fixed rank order; needs no workspace.
 *   Weights stream through a deep TMA ring (up to 176 KB per SM in flight, issued before the previous kernel of the
 *   stream has finished), activations through a shallow one; the schedule (S, CTAs per SM, ring depths) is picked per
 *   shape from occupancy queries and a latency/bandwidth model (b200_gemm_decode_plan reports it).
 * M > 128 (prefill): persistent stream-K over min(split_k, #SM) CTAs (0 = one per SM), each an equal share of the
 *   (tile, k-block) space; tiles cut by a share boundary are summed in fixed order by the last contributor.
 *   `workspace` (b200_gemm_workspace_bytes) holds the fix-up counters in its first 4 KiB: it must be ZERO-FILLED before
 *   its first use and is left zeroed by every call; one workspace may serve all GEMMs issued on one stream. */
size_t b200_gemm_workspace_bytes(int M, int N, int K, int ab_dtype, int split_k);
/* The decode schedule (M <= 128) chosen on the current device for a shape: out6 = {cluster size S, CTAs per SM, weight
 * ring depth, activation ring depth, CTAs the chip holds at once, dynamic shared memory bytes}.  Needs a device.
 * silu_mul selects the epilogue variant planned for: 0 linear, 1 SiLU*mul (b200_gemm_silu_mul), 2 RoPE + KV write
 * (b200_gemm_rope_kv_write). */
int b200_gemm_decode_plan(int M, int N, int K, int ab_dtype, int split_k, int silu_mul, int has_residual, int* out6);
/* Profiling aid: with a non-NULL device buffer ([#CTAs][8] uint64) every decode-schedule GEMM CTA records %globaltimer
 * at its phase boundaries (tools/gemm_trace.py); NULL (default) switches it off. */
void b200_gemm_set_trace(void* device_buf);
int b200_gemm(const void* a, const void* w, void* out, const float* a_scale, const float* w_scale, const void* bias,
              const void* residual, int M, int N, int K, int ab_dtype, int out_dtype, int split_k, void* workspace,
              size_t workspace_bytes, void* stream);
/* INT8 weight-only linear (SURVEY 8(b) `b200_gemm_int8_wo`; reference config keys quantized / quantization_dtype=int8 with
 * bf16 activations, model_loader.py:885-898): out = (a_bf16 . w_int8^T) * w_scale[n] + bias + residual.  The int8 rows are
 * streamed from HBM as int8 (half the bytes of bf16) and widened to bf16 operand tiles in shared memory by the CTA's
 * epilogue warps (exact), then multiplied with kind::f16.  Same as b200_gemm with ab_dtype = B200_INT8_WO. */
int b200_gemm_int8_wo(const void* a_bf16, const void* w_int8, void* out, const float* w_scale, const void* bias,
                      const void* residual, int M, int N, int K, int out_dtype, int split_k, void* workspace,
                      size_t workspace_bytes, void* stream);
/* gate/up projection with SiLU(gate)*up fused into the epilogue.  w_gu [2I][K] holds, per group of 32 rows, 16 gate rows
 * followed by the 16 up rows of the same channels (w_scale / bias permuted alike); out_act [M][I] bf16 =
 * bf16(silu(bf16(gate)) * bf16(up)) exactly as b200_gemm + b200_silu_mul produce; group_amax [M][2I/32] fp32 receives
 * max|act| over the 16 channels of every row group, for the per-token quantisation that follows
 * (b200_quantize_per_token).  split_k as b200_gemm, at most 4 in the decode schedule. */
int b200_gemm_silu_mul(const void* a, const void* w_gu, void* out_act, const float* a_scale, const float* w_scale,
                       const void* bias, float* group_amax, int M, int N2, int K, int ab_dtype, int split_k,
                       void* workspace, size_t workspace_bytes, void* stream);
/* The same gate/up GEMM + SiLU*mul with the dynamic per-token quantisation of its result (the down_proj input) done in the
 * SAME launch: out_q [M][N2/2] e4m3 / int8, out_scale [M].  The CTAs agree on each token's amax through a grid-wide
 * barrier; `state` is 1040 bytes of device memory, zero-initialised ONCE and then owned by the library (any number of
 * launches on one stream).  Bit-identical to b200_gemm_silu_mul + b200_quantize_per_token.  Decode schedule only:
 * returns B200_ERR_UNSUPPORTED (nothing launched) when M > 128 or when the shape does not run as one resident wave of
 * one CTA per weight tile -- the caller then uses the two-call form.  "Resident" is worked out from the occupancy
 * query for a GPU that runs nothing else: if another stream or process holds SM slots while this kernel starts, CTAs wait
 * at the barrier for peers that cannot be scheduled and the kernel traps after about 2 s.  Use it only where this library's
 * stream owns the GPU; the product path keeps it off (B200_FUSE_SILU_QUANT unset) and uses the two-call form. */
int b200_gemm_silu_mul_quant(const void* a, const void* w_gu, void* out_q, float* out_scale, const float* a_scale,
                             const float* w_scale, const void* bias, void* state, int M, int N2, int K, int ab_dtype,
                             void* stream);

/* qkv projection with RoPE and the KV-cache write fused into its epilogue: b200_gemm (bf16 out, no residual) followed by
 * b200_rope_kv_write on its output, in ONE launch when the decode schedule splits K over a cluster (the token's owner
 * holds whole heads; rotate-half partner = lane ^ D/8), else as the two launches.  Same bits either way: qkv_out keeps
 * rotated q, un-rotated k and v; rotated k and v land in the pools at slot_mapping[t] (slot < 0 skipped).
 * Replaces the qkv linear + rotary + KV scatter inside NxDI's decoder layer (model_loader.py:336-348). */
int b200_gemm_rope_kv_write(const void* a, const void* w, void* qkv_out, const float* a_scale, const float* w_scale,
                            const void* bias, int M, int N, int K, int ab_dtype, int split_k, void* workspace,
                            size_t workspace_bytes, const int64_t* positions, const float* cos_table,
                            const float* sin_table, int max_pos, void* k_pool, void* v_pool, const int64_t* slot_mapping,
                            int Hq, int Hkv, int D, int block_size, int num_blocks, void* stream);

/* ---------------------------------------------------------------- glue kernels of the decoder step */
/* y = x * rsqrt(mean(x^2)+eps) * w (fp32 math).  Writes any of: out_bf16 [T][H]; out_q [T][H] (q_dtype FP8/INT8)
 * with dynamic per-token scale out_scale [T] (scale = amax/qmax, q = rn(y_bf16 / scale)). */
int b200_rmsnorm(const void* x, const void* w, void* out_bf16, void* out_q, float* out_scale, int T, int H, float eps,
                 int q_dtype, void* stream);
/* act = silu(gate) * up from gate_up [T][2I] (gate columns first); same output options as rmsnorm. */
int b200_silu_mul(const void* gate_up, void* out_bf16, void* out_q, float* out_scale, int T, int I, int q_dtype,
                  void* stream);
/* Dynamic per-token quantisation of x [T][K] bf16.  group_amax [T][n_groups] (optional, from b200_gemm_silu_mul)
 * supplies partial row maxima so rows can be cut across CTAs without a scan; NULL = scan the row. */
int b200_quantize_per_token(const void* x, void* out_q, float* out_scale, const float* group_amax, int n_groups, int T, int K,
                            int q_dtype, void* stream);
int b200_embedding(const int64_t* tokens, const void* table, void* out, int T, int H, int V, void* stream);
/* ---------------------------------------------------------------- on-device sampling (top-k / top-p / temperature)
 * Replaces NxDI's on-device sampler behind `self.model.sample(...)` (model_runner.py:1084-1098, model_loader.py:367-382)
 * fed by the [B][3] (top_k, top_p, temperature) tensor of get_nxd_sampling_params (model_runner.py:1106-1140).
 * b200_topk_candidates: per row the 256 (= _MAX_NEURON_SAMPLING_TOP_K, model_runner.py:74) largest fp32 logits of the
 *   local vocab shard, sorted by (value desc, index asc): cand_val [B][256] fp32, cand_idx [B][256] int64 (global token id
 *   = idx_offset + column; -inf / -1 padding when V < 256).  Exact (radix select), ties towards the lower index.
 * b200_sample_candidates: cand_* laid out [groups][B][256] (groups = TP shards, all-gathered by the caller); merges the
 *   groups, keeps top_k (1..256), w_j = exp((v_j - v_0)/temperature), keeps the shortest prefix with cumulative weight
 *   >= top_p * Z, and draws by inverse CDF with the caller's uniform u[b] in [0,1).  top_k == 1 is greedy (== b200_argmax).
 *   fp32, sequential, fixed order: the result is a pure function of (logits, params, u). */
int b200_topk_candidates(const float* logits, int64_t row_stride, float* cand_val, int64_t* cand_idx, int B, int V,
                         int64_t idx_offset, void* stream);
int b200_sample_candidates(const float* cand_val, const int64_t* cand_idx, int groups, int B, const float* params,
                           const float* u, int64_t* out_ids, void* stream);
/* Log-probabilities on device (the reference returns them from its CPU sampling path, model_runner.py:347-356, 1142-1239:
 * vLLM's Sampler = raw log-softmax of the logits, the sampled token's rank, the top-n tokens).
 * b200_row_lse: per row of the local vocab shard, out_max[b] = max_v logits[b][v], out_sumexp[b] = sum_v exp(logits - max)
 *   (fixed reduction tree: bit-reproducible).  TP: one (max, sumexp) pair per shard, laid out [groups][B] for the gather.
 * b200_gather_logprobs: with the sorted candidate lists of b200_topk_candidates ([groups][B][256]) and the sampled ids:
 *   out_ids / out_logprobs [B][1 + num_logprobs] (int32 / fp32; column 0 = the sampled token, then the num_logprobs most
 *   likely tokens, value desc / index asc), out_rank[b] = #tokens whose logit is >= the sampled token's (itself included). */
int b200_row_lse(const float* logits, int64_t row_stride, float* out_max, float* out_sumexp, int B, int V, void* stream);
int b200_gather_logprobs(const float* cand_val, const int64_t* cand_idx, int groups, int B, const int64_t* sampled,
                         const float* row_max, const float* row_sumexp, int num_logprobs, int32_t* out_ids,
                         float* out_logprobs, int32_t* out_rank, void* stream);
/* Greedy sampling (reference: top_k=1, temperature=1.0 for temperature 0, model_runner.py:1126-1128).
 * out_idx[t] = idx_offset + first index of the row maximum; out_val[t] = that maximum (for TP merges). */
int b200_argmax(const float* logits, int64_t row_stride, float* out_val, int64_t* out_idx, int T, int V,
                int64_t idx_offset, void* stream);

/* ---------------------------------------------------------------- tensor-parallel exchange (NVLink peer memory)
 * All-reduce (sum over `world` ranks, bf16) of the row-parallel GEMM outputs fused with the RMSNorm (+ dynamic per-token
 * quantisation) that consumes it; replaces the Neuron-runtime collectives NxDI issues after o_proj / down_proj
 * (tp_degree, model_loader.py:752-753).  host_peer_bufs[p] / host_peer_flags[p] are HOST arrays of `world` DEVICE
 * pointers: rank p's partial [T][H] bf16 and rank p's flag area ([>= T][8] uint32, zero-initialised once), both inside
 * peer-mapped (symmetric) memory.  epoch [>= T] uint32 is local, zero-initialised once.  Every rank must issue the same
 * sequence of calls; consecutive calls must alternate between two partial buffers (see csrc/tp_allreduce.cu).
 * Outputs (local): x_out [T][H] bf16 = the sum (ranks added in rank order, identical bits on every rank); with norm_w:
 * out_bf16 and/or out_q + out_scale as b200_rmsnorm.  norm_w == NULL: plain all-reduce. */
int b200_allreduce_rmsnorm(const void* const* host_peer_bufs, void* const* host_peer_flags, uint32_t* epoch, int rank,
                           int world, void* x_out, const void* norm_w, void* out_bf16, void* out_q, float* out_scale, int T,
                           int H, float eps, int q_dtype, void* stream);
/* Same exchange with the sum done INSIDE the NVSwitch (NVLS): mc_buf is the MULTICAST device address of the partial
 * [T][H] bf16 (the address that maps every rank's replica; torch symmetric memory: handle.multicast_ptr + offset); the
 * kernel issues one multimem.ld_reduce (bf16x2 add, fp32 accumulate) per 16 bytes, so a rank receives T*H*2 bytes
 * instead of (world-1) x that.  Flags, epoch, outputs and the alternation rule as above.  The order of the additions is
 * the switch's: identical bits on every rank, but not necessarily those of the rank-order sum (last bf16 bit). */
int b200_allreduce_rmsnorm_nvls(const void* mc_buf, void* const* host_peer_flags, uint32_t* epoch, int rank, int world,
                                void* x_out, const void* norm_w, void* out_bf16, void* out_q, float* out_scale, int T,
                                int H, float eps, int q_dtype, void* stream);

/* Greedy merge over the vocabulary shards of a TP group (the reference leaves sampling of the sharded logits to NxDI's
 * on-device sampler, model_loader.py:367-382).  host_peer_vals[p] / host_peer_idxs[p]: rank p's [B] fp32 local maxima and
 * [B] int64 GLOBAL indices inside peer-mapped memory, written by b200_argmax earlier on the stream; host_peer_flags[p]:
 * rank p's [8] uint32 flag words (zero-initialised once); epoch: one local uint32 (zero-initialised once).
 * out_ids[b] = index of the largest value over all ranks, ties to the lowest index; identical on every rank. */
int b200_tp_argmax_merge(const void* const* host_peer_vals, const void* const* host_peer_idxs,
                         void* const* host_peer_flags, uint32_t* epoch, int rank, int world, int64_t* out_ids, int B,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_H_ */
