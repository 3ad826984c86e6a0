"""B200 model runner: vllm-neuron's `execute_model(scheduler_output) -> ModelRunnerOutput` contract on B200.

Reference: vllm_neuron/worker/neuronx_distributed_model_runner.py
  * execute_model                       :252-291
  * _update_states                      :381-510   (request cache, block-id append / replace)
  * seq-id slot map                     :130-136, 263-268, 714-717
  * prefill / decode index construction :704-832   -> here: block ids are scattered into a persistent DEVICE
    block table and positions / slot_mapping / context lengths are derived ON DEVICE by
    b200_build_decode_meta / b200_build_prefill_meta (no per-step Python list building, no deepcopy).
  * _generate_model_runner_output       :293-359
  * get_kv_cache_spec                   :361-379
  * initialize_kv_cache                 :142-150   (no-op in the reference; here it allocates the pools)

The runner keeps its own light request cache instead of vLLM's InputBatch; the scheduler's block-id lists
stay owned by vLLM's KV manager and are only mirrored (host list + device table).
"""
from __future__ import annotations

import logging
from dataclasses import dataclass, field

import torch

from .. import ops
from ..admission import check_sampling_params, unsupported_sampling_params  # noqa: F401 (re-exported)
from ..config import B200Config, ModelShape, build_config
from ..model import B200CausalLM, bucket_batch, bucket_ctx_blocks

logger = logging.getLogger(__name__)

_MAX_SAMPLING_TOP_K = 256  # RUNNER:74 (_MAX_NEURON_SAMPLING_TOP_K)
_SLOT_MAPPING_PAD = -1  # RUNNER:80
_BLOCK_TABLE_PAD = 0  # RUNNER:81


@dataclass
class CachedRequest:
    req_id: str
    prompt_token_ids: list[int]
    output_token_ids: list[int] = field(default_factory=list)
    block_ids: list[int] = field(default_factory=list)
    num_computed_tokens: int = 0
    row: int = -1  # persistent block-table row (the reference's "neuron seq id")
    sampling_params: object = None
    rng: object = None  # per-request numpy Generator (SamplingParams.seed), created on first stochastic draw
    prefilled: int = 0  # prompt tokens whose K/V are in the cache (prefix-cache hits + chunks computed so far)
    samp_row: tuple | None = None  # (top_k, top_p, temperature) by the reference's rule; fixed per request, computed once

    @property
    def num_tokens(self) -> int:
        return len(self.prompt_token_ids) + len(self.output_token_ids)


def _first_group(block_ids):
    """vLLM passes block ids per KV-cache group; this path has exactly one group (RUNNER:743)."""
    if block_ids is None:
        return None
    if isinstance(block_ids, (tuple, list)) and block_ids and isinstance(block_ids[0], (list, tuple)):
        assert len(block_ids) == 1, "exactly one KV cache group is supported"
        return list(block_ids[0])
    return list(block_ids)


def _logprob_bucket(n: int) -> int:
    """Requested top-n log-probabilities run at 1 / 5 / 20 / 256 columns (few graph variants; vLLM reads what it needs)."""
    for c in (1, 5, 20):
        if n <= c:
            return c
    return _MAX_SAMPLING_TOP_K


class B200ModelRunner:
    def __init__(self, vllm_config, device: torch.device, *, model: B200CausalLM | None = None,
                 tp_rank: int = 0, tp_size: int = 1, tp_group=None):
        self.vllm_config = vllm_config
        self.model_config = vllm_config.model_config
        self.cache_config = vllm_config.cache_config
        self.parallel_config = vllm_config.parallel_config
        self.scheduler_config = vllm_config.scheduler_config
        self.device = device
        self.block_size = self.cache_config.block_size
        self.max_num_reqs = self.scheduler_config.max_num_seqs
        self.max_model_len = self.model_config.max_model_len
        self.tp_rank, self.tp_size, self.tp_group = tp_rank, tp_size, tp_group
        self.requests: dict[str, CachedRequest] = {}
        # request id -> persistent row; free rows (RUNNER:130-136)
        self.req_to_row: dict[str, int] = {}
        self.free_rows = set(range(self.max_num_reqs))
        self.model = model
        self.b200_config: B200Config | None = None
        self.num_blocks: int | None = None
        self.use_cuda_graph = not getattr(self.model_config, "enforce_eager", False)
        self._stage = None
        self.h2d_bytes_last = 0
        self.d2h_bytes_last = 0
        import numpy as np

        add = getattr(vllm_config, "additional_config", None) or {}
        ods = (add.get("override_neuron_config") or {}).get("on_device_sampling_config") or {}
        self.global_topk = int(ods.get("global_topk", 0)) if isinstance(ods, dict) else int(getattr(ods, "global_topk", 0))
        self._max_topk: int | None = None
        self._rng = np.random.Generator(np.random.PCG64(getattr(self.model_config, "seed", 0) or 0))

    # ------------------------------------------------------------------ config / load / KV allocation
    def build_b200_config(self) -> B200Config:
        add = getattr(self.vllm_config, "additional_config", None) or {}
        override = add.get("override_neuron_config") or add.get("override_b200_config") or {}
        self.b200_config = build_config(
            tensor_parallel_size=self.parallel_config.tensor_parallel_size,
            max_num_seqs=self.max_num_reqs,
            max_model_len=self.max_model_len,
            block_size=self.block_size,
            enable_prefix_caching=bool(self.cache_config.enable_prefix_caching),
            num_gpu_blocks_override=self.cache_config.num_gpu_blocks_override,
            override=override,
        )
        return self.b200_config

    def model_shape(self) -> ModelShape:
        return ModelShape.from_hf(self.model_config.hf_config)

    def load_model(self, *, load_dummy_weights: bool = False) -> None:
        """Weights are materialised in initialize_kv_cache (the model object owns weights + pools); this
        validates the config exactly where the reference does (get_neuron_model, LOADER:662-721)."""
        cfg = self.build_b200_config()
        self.shape = self.model_shape()
        self.is_block_kv_layout = cfg.is_block_kv_layout
        self.is_prefix_caching = cfg.is_prefix_caching
        self._load_dummy = True if load_dummy_weights else getattr(self.vllm_config.load_config, "load_format", "dummy") == "dummy"
        from ..loader import check_modules_to_not_convert

        self._quant_lm_head = check_modules_to_not_convert(cfg.modules_to_not_convert, self.shape)
        self._weights = None if self._load_dummy else self._open_checkpoint(cfg)

    def _open_checkpoint(self, cfg: B200Config):
        """HF checkpoint directory -> lazy weight source (loader.py).  With `quantized` + `quantized_checkpoints_path`
        the weights are quantised once, saved there and read back from there on later starts (LOADER:237-239)."""
        import os

        from ..loader import HFCheckpointWeights, save_quantized_checkpoint

        path = self.model_config.model
        tie = bool(getattr(self.model_config.hf_config, "tie_word_embeddings", False))
        qpath = cfg.quantized_checkpoints_path if cfg.quantized else None
        if qpath:
            # ONE decision for all ranks (a rank arriving after rank 0 has written the index must not skip the barrier):
            # rank 0 looks, the answer is broadcast, every rank then takes the same branch
            need = not os.path.exists(os.path.join(qpath, "model.safetensors.index.json"))
            if self.tp_group is not None:
                flag = torch.tensor([int(need)], dtype=torch.int32, device=self.device)
                torch.distributed.broadcast(flag, src=torch.distributed.get_global_rank(self.tp_group, 0), group=self.tp_group)
                need = bool(flag.item())
            if need:
                if self.tp_rank == 0:
                    kind = "fp8" if cfg.quant_mode == "fp8" else "int8"
                    logger.info("quantising %s (%s, %s) into %s", path, kind, cfg.quantization_type, qpath)
                    save_quantized_checkpoint(HFCheckpointWeights(path, self.shape, tie_word_embeddings=tie), qpath, kind,
                                              cfg.modules_to_not_convert, cfg.quantization_type)
                if self.tp_group is not None:
                    torch.distributed.barrier(group=self.tp_group)
            path = qpath
        return HFCheckpointWeights(path, self.shape, tie_word_embeddings=tie)

    def weight_bytes_estimate(self) -> int:
        """Per-rank bytes of the weights that initialize_kv_cache will materialise (used to size the KV budget)."""
        sh, tp = self.model_shape(), self.tp_size
        cfg = self.b200_config or self.build_b200_config()
        wb = 2 if cfg.quant_mode == "bf16" else 1
        qd, kd = sh.num_q_heads * sh.head_dim, sh.num_kv_heads * sh.head_dim
        per_layer = ((qd + 2 * kd) * sh.hidden + sh.hidden * qd + 3 * sh.intermediate * sh.hidden) * wb
        return (per_layer * sh.layers + sh.vocab * sh.hidden * wb) // tp + sh.vocab * sh.hidden * 2

    def warm_up(self) -> None:
        """Capture the decode graphs the first requests will hit: the full-batch bucket at the smallest and the largest
        context bound (reference: compile / warm-up is NxDI's).  Other (batch bucket, context bound) pairs are captured on
        first use into the same memory pool; the cache is an LRU of model.GRAPH_CACHE_CAP graphs."""
        if self.model is None or not self.use_cuda_graph:
            return
        from ..model import bucket_batch, bucket_ctx_blocks

        m = self.model
        saved = m.num_tokens.clone()
        m.num_tokens.fill_(0)  # pos = -1 -> slot -1: nothing is written
        bb = bucket_batch(self.max_num_reqs, self.max_num_reqs)
        for mcb in sorted({bucket_ctx_blocks(1, m.max_blocks), m.max_blocks}):
            m.decode_step(bb, max_ctx_blocks=mcb, use_graph=True)
        m.num_tokens.copy_(saved)
        torch.cuda.synchronize()

    def graph_pool_reserve_bytes(self) -> int:
        """HBM the shared CUDA-graph pool may grow to (activations + fp32 logits of the largest decode batch, with slack
        for fragmentation across bucket sizes); taken out of the KV budget in B200Worker.determine_available_memory."""
        sh, b = self.model_shape(), self.max_num_reqs
        tp = self.tp_size
        per_step = b * (sh.vocab // tp) * 4 + b * (4 * sh.hidden + 3 * sh.intermediate // tp + 2 * sh.hidden) * 2
        return 3 * per_step + (256 << 20)

    def get_kv_cache_spec(self):
        """RUNNER:361-379 hands vLLM ONE spec under 'layer'; we register one per decoder layer with the per-rank
        kv-head count so vLLM's block budget is right without --num-gpu-blocks-override (documented deviation)."""
        from vllm.v1.kv_cache_interface import FullAttentionSpec

        shape = self.model_shape()
        spec = FullAttentionSpec(block_size=self.block_size, num_kv_heads=shape.num_kv_heads // self.tp_size,
                                 head_size=shape.head_dim, dtype=torch.bfloat16, sliding_window=None)
        return {f"layers.{i}.attn": spec for i in range(shape.layers)}

    def initialize_kv_cache(self, kv_cache_config=None, num_blocks: int | None = None) -> None:
        """The reference's is a no-op (NxDI allocates pa_num_blocks x pa_block_size itself, LOADER:775-778).
        Here the pools are ours: [layers][2][num_blocks][Hkv/TP][block_size][D] bf16, block 0 = null block."""
        if num_blocks is None:
            num_blocks = getattr(kv_cache_config, "num_blocks", None) or self.b200_config.pa_num_blocks
        self.num_blocks = int(num_blocks)
        if self.model is None:
            self.model = B200CausalLM(
                self.shape, self._weights, quant=self.b200_config.quant_mode, quant_lm_head=self._quant_lm_head,
                quant_type=self.b200_config.quantization_type if self.b200_config.quantized else "per_channel_symmetric",
                num_blocks=self.num_blocks,
                max_num_seqs=self.max_num_reqs, max_model_len=self.max_model_len, block_size=self.block_size,
                device=self.device, tp_rank=self.tp_rank, tp_size=self.tp_size, tp_group=self.tp_group,
                seed=getattr(self.model_config, "seed", 0) or 0,
            )
        self._init_staging()

    def _init_staging(self):
        n = self.max_num_reqs
        self._pin_tokens = torch.zeros(n, dtype=torch.int64).pin_memory()
        self._pin_rows = torch.zeros(n, dtype=torch.int32).pin_memory()
        self._pin_ntok = torch.zeros(n + 1, dtype=torch.int32).pin_memory()  # + the padding row (always 0 tokens)
        self._dev_upd = torch.zeros((max(64, 4 * n), 3), dtype=torch.int32, device=self.device)
        self._pin_out = torch.zeros(n, dtype=torch.int64).pin_memory()
        self._pin_samp = torch.ones((n, 3), dtype=torch.float32).pin_memory()
        self._pin_u = torch.zeros(n, dtype=torch.float32).pin_memory()
        self._np_samp, self._np_u = self._pin_samp.numpy(), self._pin_u.numpy()
        # numpy views of the pinned staging buffers: one vectorised store per field instead of per-request tensor
        # __setitem__ calls (the reference builds Python lists and calls torch.tensor() on them, RUNNER:887-917)
        self._np_tokens, self._np_rows = self._pin_tokens.numpy(), self._pin_rows.numpy()
        self._np_ntok, self._np_out = self._pin_ntok.numpy(), self._pin_out.numpy()

    # ------------------------------------------------------------------ state tracking (RUNNER:381-510)
    def _update_states(self, so) -> list[tuple[int, int, int]]:
        """Returns the (row, col, block_id) triples to scatter into the device block table."""
        updates: list[tuple[int, int, int]] = []
        for req_id in so.finished_req_ids:
            self.requests.pop(req_id, None)
            row = self.req_to_row.pop(req_id, None)
            if row is not None:
                self.free_rows.add(row)  # RUNNER:263-268
        for nr in so.scheduled_new_reqs:
            assert nr.req_id not in self.req_to_row, "Encountered an existing request ID while prefilling a new request"
            assert self.free_rows, "No free sequence ID available!"
            # Unsupported sampling parameters are refused by the front end (B200Platform.validate_request).  A request that
            # reaches the worker anyway (an engine driven without the input processor) must not take the step down -- an
            # exception here would end every in-flight request -- so it is sampled from (top_k, top_p, temperature) alone,
            # exactly what the reference's device path does with it (RUNNER:1106-1140), and says so.
            bad = unsupported_sampling_params(getattr(nr, "sampling_params", None))
            if bad:
                logger.warning("request %s: ignoring sampling parameters without a device implementation: %s",
                               nr.req_id, ", ".join(bad))
            row = min(self.free_rows)
            self.free_rows.discard(row)
            self.req_to_row[nr.req_id] = row
            blocks = _first_group(nr.block_ids) or []
            st = CachedRequest(nr.req_id, list(nr.prompt_token_ids), [], blocks, nr.num_computed_tokens, row,
                               getattr(nr, "sampling_params", None))
            self.requests[nr.req_id] = st
            updates += [(row, c, b) for c, b in enumerate(blocks)]
        cached = so.scheduled_cached_reqs
        resumed = getattr(cached, "resumed_req_ids", None)
        legacy_resumed = getattr(cached, "resumed_from_preemption", None)
        requests, new_ids = self.requests, cached.new_block_ids
        if not resumed and not legacy_resumed and (not new_ids or new_ids.count(None) == len(new_ids)):
            return updates  # steady decode step: nobody got a block (31 of 32 steps at block size 32), nothing to mirror
        for i, req_id in enumerate(cached.req_ids):
            st = requests[req_id]
            nb = new_ids[i] if new_ids else None
            new_blocks = None if nb is None else _first_group(nb)
            is_resumed = (req_id in resumed) if resumed else bool(legacy_resumed and legacy_resumed[i])
            if not is_resumed:
                if new_blocks:
                    start = len(st.block_ids)
                    st.block_ids.extend(new_blocks)  # RUNNER:461-466
                    updates += [(st.row, start + c, b) for c, b in enumerate(new_blocks)]
            else:
                assert new_blocks is not None
                st.block_ids = list(new_blocks)  # RUNNER:467-471: replace after preemption
                updates += [(st.row, c, b) for c, b in enumerate(st.block_ids)]
                # vLLM preempts by recomputation: the new blocks are empty, so everything generated so far joins what the
                # next step prefills (the scheduler schedules all of those tokens) and the step samples the next token.
                # (Cannot happen under the validated block budget -- max_num_seqs x max_model_len always fits, LOADER:845-859
                # -- but a mirrored block list without its K/V would decode garbage.)
                st.prompt_token_ids = st.prompt_token_ids + st.output_token_ids
                st.output_token_ids = []
                nct = getattr(cached, "num_computed_tokens", None)
                st.prefilled = st.num_computed_tokens = int(nct[i]) if nct else 0
            if st.prefilled >= len(st.prompt_token_ids):
                st.num_computed_tokens = st.num_tokens - 1
        return updates

    def _push_block_updates(self, updates):
        if not updates:
            return
        n = len(updates)
        if n > self._dev_upd.shape[0]:
            self._dev_upd = torch.zeros((2 * n, 3), dtype=torch.int32, device=self.device)
        # a FRESH pinned tensor per push: the copy is asynchronous and a step made only of partial prefill chunks never
        # synchronises, so a reused staging buffer could be rewritten by the host before the previous step's DMA has read
        # it (the caching host allocator keeps this block alive until the copy has executed on the stream)
        host = torch.tensor(updates, dtype=torch.int32).pin_memory()
        self._dev_upd[:n].copy_(host, non_blocking=True)
        ops.block_table_scatter(self.model.block_table, self._dev_upd[:n])
        self.h2d_bytes_last += n * 12

    # ------------------------------------------------------------------ sampling parameters (RUNNER:1106-1140)
    def sampling_row(self, st: CachedRequest) -> tuple[float, float, float]:
        """(top_k, top_p, temperature) of one request by the reference's rule: top_k outside (0, max_topk) -> max_topk
        (= min(global_topk or vocab, 256)); temperature 0 -> top_k 1, temperature 1.0 (greedy)."""
        sp = st.sampling_params
        if sp is None:
            return (1.0, 1.0, 1.0)
        if self._max_topk is None:
            gv = getattr(self.model_config, "get_vocab_size", None)
            vocab = gv() if callable(gv) else self.model_shape().vocab
            self._max_topk = min(self.global_topk or vocab, _MAX_SAMPLING_TOP_K)
        max_topk = self._max_topk
        k = getattr(sp, "top_k", 0) or 0
        top_k = k if 0 < k < max_topk else max_topk
        top_p = float(getattr(sp, "top_p", 1.0))
        temp = float(getattr(sp, "temperature", 1.0))
        if temp == 0.0:
            top_k, temp = 1, 1.0
        return (float(top_k), top_p, temp)

    def _draw_uniform(self, st: CachedRequest) -> float:
        """One uniform in [0,1) for the request's next draw: its own generator when SamplingParams.seed is set, else the
        runner's (seeded from model_config.seed).  The device kernel is a pure function of (logits, params, u)."""
        import numpy as np

        seed = getattr(st.sampling_params, "seed", None)
        if seed is not None:
            if st.rng is None:
                st.rng = np.random.Generator(np.random.PCG64(seed))
            g = st.rng
        else:
            g = self._rng
        return float(np.float32(g.random(dtype=np.float32)))

    def _sampling_inputs(self, sts: list[CachedRequest]):
        """None when every request of the batch is greedy (argmax path); else host lists (params rows, uniforms)."""
        rows = []
        greedy = True
        for st in sts:  # a request's SamplingParams never change: its row is computed on first use (256 us -> 6 us at B=64)
            r = st.samp_row
            if r is None:
                r = st.samp_row = self.sampling_row(st)
            if r[0] != 1.0:
                greedy = False
            rows.append(r)
        if greedy:
            return None
        return rows, [self._draw_uniform(st) if r[0] != 1.0 else 0.0 for st, r in zip(sts, rows)]

    @staticmethod
    def _num_logprobs(sts) -> int | None:
        """None when no request of this part asked for log-probabilities, else the bucketed column count."""
        want = [int(lp) for lp in (getattr(st.sampling_params, "logprobs", None) for st in sts) if lp is not None]
        return _logprob_bucket(max(want)) if want else None

    def _logprobs_to_host(self, parts):
        """Start the D2H copies of the parts' LogprobsTensors triples (before the step's single synchronize)."""
        if not any(t is not None for _, t in parts):
            return None
        out = []
        for rows, t in parts:
            if t is None:
                out.append((rows, None))
                continue
            host = tuple(x[:rows].to("cpu", non_blocking=True) for x in t)
            self.d2h_bytes_last += sum(h.numel() * h.element_size() for h in host)
            out.append((rows, host))
        return out

    @staticmethod
    def _logprobs_lists(parts):
        """vLLM's LogprobsLists over the step's sampled rows (RUNNER:347-349 `logprobs_tensors.tolists()`).  A mixed step
        has parts of different widths (a prefill that asked for 1 log-probability next to a decode batch that asked for 20);
        all are widened to the widest.  vLLM's front end detokenises EVERY id of a row before it cuts the row down to what
        the request asked for, so the filler must be a real token id: the row's own sampled token (column 0) with
        log-probability -inf -- never -1, which fast tokenizers refuse.  The same goes for the -1 ids the gather kernel
        leaves when a vocabulary has fewer tokens than columns; those repeat column 0 outright (id and log-probability), so
        a duplicate key cannot change the sampled token's entry.  Parts that computed no log-probabilities (nobody in them
        asked, nobody reads them) are rows of token 0 at -inf."""
        if parts is None:
            return None
        import numpy as np
        from vllm.v1.outputs import LogprobsLists

        width = max(h[0].shape[1] for _, h in parts if h is not None)
        ids, lps, rks = [], [], []
        for rows, h in parts:
            i = np.zeros((rows, width), dtype=np.int32)
            l = np.full((rows, width), -np.inf, dtype=np.float32)
            r = np.zeros((rows,), dtype=np.int32)
            if h is not None:
                w = h[0].shape[1]
                i[:, :w], l[:, :w], r[:] = h[0].numpy(), h[1].numpy(), h[2].numpy()
                i[:, w:] = i[:, :1]
                short = i < 0  # only inside [:, :w]: the candidate list ran out (vocabulary smaller than the bucket)
                if short.any():
                    i = np.where(short, i[:, :1], i)
                    l = np.where(short, l[:, :1], l)
            ids.append(i); lps.append(l); rks.append(r)
        return LogprobsLists(np.concatenate(ids), np.concatenate(lps), np.concatenate(rks))

    # ------------------------------------------------------------------ the step
    @torch.inference_mode()
    def execute_model(self, scheduler_output, intermediate_tensors=None):
        from vllm.v1.outputs import EMPTY_MODEL_RUNNER_OUTPUT, ModelRunnerOutput

        self.h2d_bytes_last = self.d2h_bytes_last = 0
        updates = self._update_states(scheduler_output)
        if not scheduler_output.total_num_scheduled_tokens:
            return EMPTY_MODEL_RUNNER_OUTPUT  # RUNNER:272-274
        self._push_block_updates(updates)

        req_ids: list[str] = []
        pending: list[torch.Tensor] = []
        lp_parts: list[tuple] = []  # (rows, LogprobsTensors triple or None) per pending part
        no_sample: list[str] = []  # requests whose chunk did not finish the prompt (reference: -1 rows, RUNNER:1060-1064)
        # ---- prefill: one request at a time (SCHED:116 guarantees batch 1; several are processed in turn).  A chunk
        # (vLLM's native scheduler, RUNNER:654-679) is a prefill of tokens [prefilled, prefilled + n) against the cached
        # prefix; only the chunk that completes the prompt samples (prefill_completion_state, RUNNER:938-1051).
        cached_ids = list(scheduler_output.scheduled_cached_reqs.req_ids)
        chunk_reqs = [self.requests[nr.req_id] for nr in scheduler_output.scheduled_new_reqs]
        for st in chunk_reqs:
            st.prefilled = st.num_computed_tokens
        chunk_reqs += [self.requests[r] for r in cached_ids if self.requests[r].prefilled < len(self.requests[r].prompt_token_ids)]
        for st in chunk_reqs:
            n_sched = scheduler_output.num_scheduled_tokens[st.req_id]
            plen = len(st.prompt_token_ids)
            end = st.prefilled + n_sched
            assert 0 < n_sched and end <= plen <= self.max_model_len, (st.req_id, st.prefilled, n_sched, plen)
            new = torch.tensor(st.prompt_token_ids[st.prefilled:end], dtype=torch.int64).pin_memory()
            new_d = new.to(self.device, non_blocking=True)
            self.h2d_bytes_last += new.numel() * 8
            last = end == plen
            samp = self._sampling_inputs([st]) if last else None
            if samp is not None:
                sp_d = torch.tensor(samp[0], dtype=torch.float32).pin_memory().to(self.device, non_blocking=True)
                u_d = torch.tensor(samp[1], dtype=torch.float32).pin_memory().to(self.device, non_blocking=True)
                self.h2d_bytes_last += 16
                samp = (sp_d, u_d)
            nlp = self._num_logprobs([st]) if last else None
            tok = self.model.prefill(st.row, new_d, end, st.prefilled, samp, sample=last, num_logprobs=nlp)
            st.prefilled = end
            st.num_computed_tokens = end
            if last:
                pending.append(tok)
                lp_parts.append((1, self.model.last_logprobs))
                req_ids.append(st.req_id)
            else:
                no_sample.append(st.req_id)
        # ---- decode batch (RUNNER:765-832)
        chunk_set = {st.req_id for st in chunk_reqs}
        dec_ids = [r for r in cached_ids if r not in chunk_set]
        if dec_ids:
            b = len(dec_ids)
            reqs = self.requests
            sts = [reqs[rid] for rid in dec_ids]
            rows = [st.row for st in sts]
            ntok = [len(st.prompt_token_ids) + len(st.output_token_ids) for st in sts]
            # the batch runs at its bucket size bb >= b: rows b.. point at the model's padding row (0 tokens: slot -1,
            # context 0), so one captured graph serves every batch size of the bucket
            bb = bucket_batch(b, self.max_num_reqs) if self.use_cuda_graph else b
            self._np_tokens[:b] = [st.output_token_ids[-1] if st.output_token_ids else st.prompt_token_ids[-1] for st in sts]
            self._np_rows[:b] = rows
            if bb > b:
                self._np_tokens[b:bb] = 0
                self._np_rows[b:bb] = self.model.pad_row
            self._np_ntok[rows] = ntok
            max_ctx = max(ntok)
            bufs = self.model.decode_buffers(bb)
            bufs.tokens.copy_(self._pin_tokens[:bb], non_blocking=True)
            bufs.row_ids.copy_(self._pin_rows[:bb], non_blocking=True)
            self.model.num_tokens.copy_(self._pin_ntok, non_blocking=True)
            self.h2d_bytes_last += bb * 12 + self._pin_ntok.numel() * 4
            samp = self._sampling_inputs(sts)
            if samp is not None:
                self._np_samp[:b] = samp[0]
                self._np_u[:b] = samp[1]
                if bb > b:
                    self._np_samp[b:bb] = (1.0, 1.0, 1.0)
                    self._np_u[b:bb] = 0.0
                bufs.samp_params.copy_(self._pin_samp[:bb], non_blocking=True)
                bufs.samp_u.copy_(self._pin_u[:bb], non_blocking=True)
                self.h2d_bytes_last += bb * 16
            nlp = self._num_logprobs(sts)
            # bound the context split by the longest live context, bucketed so few graphs are captured
            mcb = bucket_ctx_blocks(-(-max_ctx // self.block_size), self.model.max_blocks)
            out = self.model.decode_step(bb, max_ctx_blocks=mcb, use_graph=self.use_cuda_graph, sampled=samp is not None,
                                         num_logprobs=nlp)
            pending.append(out[:b])
            lp_parts.append((b, self.model.last_logprobs))
            req_ids += dec_ids
        # ---- one D2H read of all sampled ids (the step's only sync point; RUNNER:304-306 `.tolist()`)
        sampled: list[list[int]] = []
        logprobs = None
        if pending:
            ids = torch.cat(pending) if len(pending) > 1 else pending[0]
            n = ids.numel()
            self._pin_out[:n].copy_(ids, non_blocking=True)
            lp_host = self._logprobs_to_host(lp_parts)
            torch.cuda.current_stream().synchronize()
            self.d2h_bytes_last += n * 8
            sampled = [[x] if x != -1 else [] for x in self._np_out[:n].tolist()]  # RUNNER:304-306: -1 is padding
            logprobs = self._logprobs_lists(lp_host)
        else:
            # a step made only of partial prefill chunks reads nothing back: wait for its copies anyway, so the host can
            # never run a whole step ahead of the device's view of the staging buffers
            torch.cuda.current_stream().synchronize()
        for rid, row in zip(req_ids, sampled):
            self.requests[rid].output_token_ids.extend(row)  # RUNNER:325-339
        req_ids += no_sample  # partial chunks report no token this step
        sampled += [[] for _ in no_sample]
        return ModelRunnerOutput(
            req_ids=req_ids,
            req_id_to_index={r: i for i, r in enumerate(req_ids)},
            sampled_token_ids=sampled,
            logprobs=logprobs,  # RUNNER:347-356 (the reference fills them on its CPU-sampling path)
            prompt_logprobs_dict={},
            pooler_output=[],
        )
