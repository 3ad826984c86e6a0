"""Llama / Qwen2 decoder over the block-KV cache, wired on the C-ABI kernels.

This is the B200 replacement for what the reference obtains from NxDI behind
`NeuronCausalLM.forward` (vllm_neuron/worker/neuronx_distributed_model_loader.py:336-365): given the
index tensors of a step it runs embedding -> L x (RMSNorm -> qkv GEMM -> RoPE + KV write by slot ->
paged attention through the block table -> o GEMM (+residual) [all-reduce] -> RMSNorm -> gate/up GEMM ->
SiLU*mul -> down GEMM (+residual) [all-reduce]) -> RMSNorm -> lm_head GEMM -> greedy argmax.
Host code only sequences kernel launches on the current CUDA stream; the decode step is captured into a
CUDA graph.  Tensor parallelism: q/kv heads and MLP columns are sharded, one NCCL all-reduce after o_proj and
after down_proj.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

from . import ops
from .config import ModelShape

# activation quantisation per mode: W8A8 modes quantise activations per token; weight-only keeps them bf16
_QKIND = {"bf16": None, "fp8": "fp8", "int8": "int8", "int8_wo": None}


def rope_tables(head_dim: int, max_pos: int, theta: float, scaling: dict | None, device):
    """fp32 cos/sin tables [max_pos, head_dim/2]; 'llama3' frequency scaling as published by Meta."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float64) / head_dim))
    if scaling and scaling.get("rope_type", scaling.get("type")) == "llama3":
        factor, lo, hi = scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"]
        old = scaling["original_max_position_embeddings"]
        wavelen = 2 * math.pi / inv_freq
        smooth = (old / wavelen - lo) / (hi - lo)
        scaled = torch.where(wavelen > old / lo, inv_freq / factor, inv_freq)
        mid = (1 - smooth) * inv_freq / factor + smooth * inv_freq
        inv_freq = torch.where((wavelen <= old / lo) & (wavelen >= old / hi), mid, scaled)
    ang = torch.outer(torch.arange(max_pos, dtype=torch.float64), inv_freq)
    return ang.cos().float().contiguous().to(device), ang.sin().float().contiguous().to(device)


QUANT_TYPES = ("per_channel_symmetric", "per_tensor_symmetric")  # LOADER:893-895 (the reference's default is per tensor)


def quantize_weight(w: torch.Tensor, kind: str, tensor_rows=None):
    """Offline symmetric weight quantisation (load time, not the hot path).
    scale[n] = amax|W[n,:]| / qmax;  fp8: cast RNE saturating to e4m3fn (qmax 448);  int8: round-half-even, clamp 127.
    tensor_rows = None: per output channel.  tensor_rows = [(r0, r1), ...]: 'per_tensor_symmetric' -- every listed row
    range is one original tensor (q / k / v, gate / up of a fused matrix) and shares ONE scale, still stored as scale[n]
    so the GEMM epilogue is the same code."""
    w32 = w.float()
    amax = w32.abs().amax(dim=-1)
    if tensor_rows is not None:
        amax = amax.clone()
        for a, b in tensor_rows:
            amax[a:b] = amax[a:b].max()
    qmax = 448.0 if kind == "fp8" else 127.0
    # tensor / tensor is a true IEEE division on both CPU and CUDA (tensor / python-scalar may become a reciprocal multiply)
    scale = torch.where(amax > 0, torch.div(amax, torch.full_like(amax, qmax)), torch.ones_like(amax)).float()
    y = w32 / scale.unsqueeze(-1)
    if kind == "fp8":
        q = y.clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    else:
        q = torch.round(y).clamp(-127, 127).to(torch.int8)
    return q.contiguous(), scale.contiguous()


def random_init_full_weights(shape: ModelShape, seed: int = 0, device="cpu", std: float = 0.02,
                             resid_std: float | None = None):
    """Random-init weights of the architecture (HF init: N(0, 0.02^2), norms = 1), full (un-sharded), bf16.
    Generated on CPU with a seeded generator so every rank / the oracle sees identical values.
    resid_std: std of the two projections that write into the residual stream (o_proj, down_proj); None = std.
    (GPT-2 / Megatron / lit-gpt initialise them with std / sqrt(2 L); the same draws are scaled, so a given seed
    names the same network up to that factor.)"""
    g = torch.Generator(device="cpu").manual_seed(seed)

    def rnd(*s, sd=None):
        return (torch.randn(*s, generator=g, dtype=torch.float32) * (std if sd is None else sd)).to(torch.bfloat16).to(device)

    qd, kd = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim
    w = {"embed": rnd(shape.vocab, shape.hidden), "lm_head": rnd(shape.vocab, shape.hidden),
         "final_norm": torch.ones(shape.hidden, dtype=torch.bfloat16, device=device), "layers": []}
    for _ in range(shape.layers):
        lw = {
            "input_norm": torch.ones(shape.hidden, dtype=torch.bfloat16, device=device),
            "post_norm": torch.ones(shape.hidden, dtype=torch.bfloat16, device=device),
            "wqkv": rnd(qd + 2 * kd, shape.hidden),
            "wo": rnd(shape.hidden, qd, sd=resid_std),
            "wgu": rnd(2 * shape.intermediate, shape.hidden),
            "wdown": rnd(shape.hidden, shape.intermediate, sd=resid_std),
        }
        if shape.qkv_bias:
            lw["bqkv"] = rnd(qd + 2 * kd)
        w["layers"].append(lw)
    return w


def tp_shard_plan(shape: ModelShape, rank: int, size: int) -> dict:
    """Row/column ranges of every weight owned by TP rank `rank`: q/kv heads and MLP columns are
    column-parallel (rows of W[N,K]), o_proj/down_proj are row-parallel (columns of W[N,K]), lm_head by vocab rows."""
    d = shape.head_dim
    qd, kd, I = shape.num_q_heads * d, shape.num_kv_heads * d, shape.intermediate
    r, t = rank, size
    return {
        "qkv_rows": [(r * qd // t, (r + 1) * qd // t), (qd + r * kd // t, qd + (r + 1) * kd // t),
                     (qd + kd + r * kd // t, qd + kd + (r + 1) * kd // t)],
        "o_cols": (r * qd // t, (r + 1) * qd // t),
        "gu_rows": [(r * I // t, (r + 1) * I // t), (I + r * I // t, I + (r + 1) * I // t)],
        "down_cols": (r * I // t, (r + 1) * I // t),
        "vocab_rows": (r * shape.vocab // t, (r + 1) * shape.vocab // t),
    }


def merge_sharded_argmax(vals: torch.Tensor, idxs: torch.Tensor, vocab: int) -> torch.Tensor:
    """vals/idxs [tp, B]: each rank's local (max, global index).  Global greedy token = lowest index among the
    ranks holding the global maximum (same tie rule as the single-GPU argmax kernel)."""
    best = vals.max(dim=0, keepdim=True).values
    cand = torch.where(vals == best, idxs, torch.full_like(idxs, vocab))
    return cand.min(dim=0).values


class _Linear:
    """One (possibly TP-sharded, possibly quantised) weight with its per-channel scale."""

    def __init__(self, w_full: torch.Tensor, quant: str, device, row_slices=None, col_slice=None, bias=None,
                 tensor_rows=None):
        # quantise the FULL weight first (per-channel scales are properties of the un-sharded matrix), then shard
        if isinstance(w_full, tuple):  # (q, scale) from a quantised checkpoint (loader.py): not quantised again
            wq, ws = w_full
            want = {"fp8": torch.float8_e4m3fn, "int8": torch.int8, "int8_wo": torch.int8}.get(quant)
            if wq.dtype != want:
                raise ValueError(f"quantised checkpoint holds {wq.dtype} weights but the configured mode is {quant!r}")
            wq, ws = wq.to(device), ws.float().to(device)
        elif quant in ("fp8", "int8", "int8_wo"):
            wq, ws = quantize_weight(w_full, "fp8" if quant == "fp8" else "int8", tensor_rows)  # on the master weight's device
            wq, ws = wq.to(device), ws.to(device)
        else:
            wq, ws = w_full.to(device), None
        if row_slices is not None:
            wq = torch.cat([wq[a:b] for a, b in row_slices], dim=0)
            if ws is not None:
                ws = torch.cat([ws[a:b] for a, b in row_slices], dim=0)
            if bias is not None:
                bias = torch.cat([bias[a:b] for a, b in row_slices], dim=0)
        if col_slice is not None:
            wq = wq[:, col_slice[0] : col_slice[1]]
        self.w = wq.contiguous()
        self.scale = None if ws is None else ws.contiguous()
        self.bias = None if bias is None else bias.to(device).contiguous()
        self.quant = quant
        self.n, self.k = self.w.shape
        self.interleaved = False

    def interleave_gate_up(self):
        """Store [gate; up] rows as 16 gate / 16 up groups (what the fused SiLU*mul GEMM epilogue reads)."""
        self.w, self.scale, self.bias = ops.interleave_gate_up(self.w, self.scale, self.bias)
        self.interleaved = True
        return self


@dataclass
class StepBuffers:
    tokens: torch.Tensor
    row_ids: torch.Tensor
    slot_mapping: torch.Tensor
    positions: torch.Tensor
    ctx_lens: torch.Tensor
    sampled: torch.Tensor
    samp_params: torch.Tensor  # [b, 3] fp32 (top_k, top_p, temperature) -- RUNNER:1106-1140
    samp_u: torch.Tensor       # [b] fp32 uniforms in [0, 1) for the inverse-CDF draw
    logprobs: tuple | None = None  # (ids int32 [b, 1+n], logprobs fp32 [b, 1+n], ranks int32 [b]) of the last step that asked


GRAPH_CACHE_CAP = 24  # captured decode graphs kept (LRU); all share one memory pool


def bucket_batch(b: int, max_num_seqs: int) -> int:
    """Decode batches run at the next power of two (capped at max_num_seqs): few graphs, rows b.. are padding."""
    bb = 1
    while bb < b:
        bb <<= 1
    return min(bb, max_num_seqs)


def bucket_ctx_blocks(n: int, max_blocks: int) -> int:
    """Context bound (in blocks) handed to the attention split, rounded up with <= 12.5 % slack: multiples of 8 up to
    128 blocks, then of 1/8 of the enclosing power of two."""
    n = max(1, n)
    step = 8
    while step * 16 < n:
        step <<= 1
    return min(max_blocks, -(-n // step) * step)


class B200CausalLM:
    def __init__(self, shape: ModelShape, full_weights: dict | None, *, quant: str = "bf16", num_blocks: int,
                 max_num_seqs: int, max_model_len: int, block_size: int = 32, device="cuda", tp_rank: int = 0,
                 tp_size: int = 1, tp_group=None, quant_lm_head: bool = True, seed: int = 0,
                 use_peer_exchange: bool = True, quant_type: str = "per_channel_symmetric"):
        assert quant in ("bf16", "fp8", "int8", "int8_wo"), f"unknown quant mode {quant!r}"
        if quant_type not in QUANT_TYPES:
            raise ValueError(f"quantization_type {quant_type!r} is not supported on the B200 path (one of {QUANT_TYPES})")
        self.quant_type = quant_type
        per_tensor = quant_type == "per_tensor_symmetric"
        assert shape.num_q_heads % tp_size == 0 and shape.num_kv_heads % tp_size == 0, "heads must divide TP"
        assert shape.intermediate % tp_size == 0 and shape.vocab % tp_size == 0
        ops.device_check(torch.device(device).index or 0)
        self.shape, self.quant, self.device = shape, quant, torch.device(device)
        self.tp_rank, self.tp_size, self.tp_group = tp_rank, tp_size, tp_group
        self.block_size, self.num_blocks = block_size, num_blocks
        self.max_num_seqs, self.max_model_len = max_num_seqs, max_model_len
        # ceil: vLLM allocates ceil(tokens / block_size) blocks, so a max_model_len that is not a multiple of the block size
        # still has a column for its last, partial block (the reference's floor, RUNNER:749,814, would drop it)
        self.max_blocks = -(-max_model_len // block_size)
        self.hq = shape.num_q_heads // tp_size
        self.hkv = shape.num_kv_heads // tp_size
        self.d = shape.head_dim
        self.inter = shape.intermediate // tp_size
        self.vocab_local = shape.vocab // tp_size
        self.scale = self.d**-0.5
        dev = self.device
        if full_weights is None:
            full_weights = _LazyRandomWeights(shape, seed)
        plan = tp_shard_plan(shape, tp_rank, tp_size)
        qkv_rows, gu_rows = plan["qkv_rows"], plan["gu_rows"]
        self.embed = full_weights["embed"].to(dev).contiguous()
        self.final_norm = full_weights["final_norm"].to(dev).contiguous()
        v0, v1 = plan["vocab_rows"]
        qd_f, kd_f, i_f = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim, shape.intermediate
        # per-tensor scales belong to the ORIGINAL tensors (q / k / v and gate / up are separate HF tensors), not to the
        # fused matrices or their TP shards
        tr_qkv = [(0, qd_f), (qd_f, qd_f + kd_f), (qd_f + kd_f, qd_f + 2 * kd_f)] if per_tensor else None
        tr_gu = [(0, i_f), (i_f, 2 * i_f)] if per_tensor else None
        tr_h = [(0, shape.hidden)] if per_tensor else None
        self.lm_head = _Linear(full_weights["lm_head"], quant if quant_lm_head else "bf16", dev, row_slices=[(v0, v1)],
                               tensor_rows=[(0, shape.vocab)] if per_tensor else None)
        self.vocab_offset = v0
        self.layers = []
        for li in range(shape.layers):
            lw = full_weights["layers"][li]
            self.layers.append(
                {
                    "input_norm": lw["input_norm"].to(dev).contiguous(),
                    "post_norm": lw["post_norm"].to(dev).contiguous(),
                    "qkv": _Linear(lw["wqkv"], quant, dev, row_slices=qkv_rows, bias=lw.get("bqkv"), tensor_rows=tr_qkv),
                    "o": _Linear(lw["wo"], quant, dev, col_slice=plan["o_cols"], tensor_rows=tr_h),
                    "gu": _Linear(lw["wgu"], quant, dev, row_slices=gu_rows, tensor_rows=tr_gu).interleave_gate_up(),
                    "down": _Linear(lw["wdown"], quant, dev, col_slice=plan["down_cols"], tensor_rows=tr_h),
                }
            )
            if hasattr(full_weights, "drop"):  # lazy sources (random init, checkpoint reader) free the layer's host copy
                full_weights.drop(li)
        self.cos, self.sin = rope_tables(self.d, max_model_len, shape.rope_theta, shape.rope_scaling, dev)
        # ---- KV pools: per layer [2(K,V)][num_blocks][Hkv_local][bs][D] bf16; block 0 is vLLM's null block
        self.kv = [torch.zeros((2, num_blocks, self.hkv, block_size, self.d), dtype=torch.bfloat16, device=dev)
                   for _ in range(shape.layers)]
        # ---- persistent per-sequence state (rows = the reference's seq-id slots, RUNNER:130-136)
        # one extra row (index max_num_seqs): the PADDING row of bucketed decode batches -- num_tokens 0, so position -1,
        # slot -1 (nothing written) and context length 0 (nothing read)
        self.block_table = torch.zeros((max_num_seqs + 1, self.max_blocks), dtype=torch.int32, device=dev)
        self.num_tokens = torch.zeros((max_num_seqs + 1,), dtype=torch.int32, device=dev)
        self.pad_row = max_num_seqs
        self._decode_bufs: dict[int, StepBuffers] = {}
        from collections import OrderedDict

        self._graphs: "OrderedDict[tuple, tuple]" = OrderedDict()
        self._graph_pool = None
        self._prefill_meta = None
        self.last_logprobs = None
        self._alloc_workspaces()
        # state of the fused gate/up + quantise kernel (token maxima, grid-barrier counter and epoch): zeroed once.
        # OFF by default: measured on the headline workload the one-launch form is SLOWER (5.83 vs 5.58 ms per step) --
        # CTAs that wait at the grid barrier keep their SM slots, so the down_proj GEMM can no longer become resident and
        # prefetch its weights while the gate/up GEMM drains (profiles/r02_summary_v1.md).  B200_SILU_QUANT_FUSION=1 enables.
        self._fq_state = torch.zeros(1040, dtype=torch.uint8, device=dev)
        import os as _os
        self._fuse_silu_quant = _os.environ.get("B200_SILU_QUANT_FUSION") == "1"
        # TP exchange over NVLink peer memory (our kernel); NCCL stays for prefill-sized batches and the greedy merge
        self.tp_comm = None
        if tp_size > 1 and use_peer_exchange:
            from .tp_comm import TpComm

            self.tp_comm = TpComm(tp_group, tp_rank, tp_size, max_num_seqs, shape.hidden, dev)

    # ------------------------------------------------------------------ workspaces
    def _linears(self):
        for L in self.layers[:1]:
            yield from (L["qkv"], L["o"], L["gu"], L["down"])
        yield self.lm_head

    def _alloc_workspaces(self):
        dev = self.device
        nbytes = 16
        for m in {self.max_num_seqs, min(self.max_model_len, 4096)}:
            for lin in self._linears():
                dt = {"bf16": 0, "fp8": 1, "int8": 2, "int8_wo": 4}[lin.quant]
                nbytes = max(nbytes, ops.gemm_workspace_bytes(m, lin.n, lin.k, dt, 0))
        # zero-filled once: the stream-K fix-up counters at its head are left at zero by every GEMM
        self.gemm_ws = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        self.attn_ws = torch.empty(
            max(16, ops.paged_attn_decode_workspace_bytes(self.max_num_seqs, self.hq, self.hkv, self.d, self.max_blocks, 0)),
            dtype=torch.uint8, device=dev)

    # ------------------------------------------------------------------ building blocks
    def _linear(self, lin: _Linear, x, xs=None, residual=None, out_dtype=torch.bfloat16, out=None):
        return ops.gemm(x, lin.w, xs, lin.scale, lin.bias, residual, out=out, out_dtype=out_dtype, workspace=self.gemm_ws)

    def _norm(self, x, w):
        q = _QKIND[self.quant]
        if q is None:
            return ops.rmsnorm(x, w, self.shape.rms_eps), None
        return ops.rmsnorm(x, w, self.shape.rms_eps, quant=q)

    def _row_parallel(self, lin: _Linear, a, a_s, x, next_norm_w):
        """Row-parallel linear (o_proj / down_proj) + residual, the TP exchange, and the RMSNorm that consumes the sum.
        Returns (x_new, h, h_scale); h is None when next_norm_w is None.
        TP=1: GEMM then rmsnorm.  TP>1, decode-sized batches: the GEMM writes its partial into peer-mapped memory and ONE
        kernel (b200_allreduce_rmsnorm) sums the ranks over NVLink and normalises/quantises.  Larger batches (prefill):
        NCCL all-reduce + rmsnorm."""
        t = a.shape[0]
        first = self.tp_rank == 0  # the residual is folded into rank 0's GEMM epilogue; the exchange sums it in once
        res = x if first else None
        if self.tp_size == 1:
            x = self._linear(lin, a, a_s, res)
        elif self.tp_comm is not None and t <= self.tp_comm.rows:
            i, part = self.tp_comm.next_partial(t)
            self._linear(lin, a, a_s, res, out=part)
            return self.tp_comm.exchange(i, t, next_norm_w, self.shape.rms_eps, _QKIND[self.quant])
        else:
            x = self._linear(lin, a, a_s, res)
            torch.distributed.all_reduce(x, group=self.tp_group)
        if next_norm_w is None:
            return x, None, None
        h, hs = self._norm(x, next_norm_w)
        return x, h, hs

    def _layers(self, x, positions, slot_mapping, attn_fn):
        t = x.shape[0]
        q = _QKIND[self.quant]
        qd = self.hq * self.d
        n_layers = len(self.layers)
        h, hs = self._norm(x, self.layers[0]["input_norm"])
        for li, L in enumerate(self.layers):
            kp, vp = self.kv[li][0], self.kv[li][1]
            # qkv projection with RoPE + KV-cache write in its epilogue (one launch in the decode schedule)
            lq = L["qkv"]
            qkv = ops.gemm_rope_kv_write(h, lq.w, positions, self.cos, self.sin, kp, vp, slot_mapping, self.hq, hs, lq.scale,
                                         lq.bias, workspace=self.gemm_ws)
            a = attn_fn(qkv[:, :qd].unflatten(1, (self.hq, self.d)), kp, vp, q)
            if q is None:
                a = (a.reshape(t, qd), None)
            elif not isinstance(a, tuple):  # prefill attention returns bf16; decode quantises in its merge kernel
                a = ops.quantize_per_token(a.reshape(t, qd), q)
            x, h2, h2s = self._row_parallel(L["o"], a[0], a[1], x, L["post_norm"])
            gu = L["gu"]
            # gate/up GEMM with SiLU(gate)*up in the epilogue; the per-group maxima feed the per-token quantisation
            fused = None
            if q is not None and t <= 128 and self._fuse_silu_quant:
                # decode-sized batches: SiLU*mul AND the per-token quantisation of its result in the GEMM's own launch
                fused = ops.gemm_silu_mul_quant(h2, gu.w, h2s, gu.scale, gu.bias, self._fq_state, q)
            if fused is not None:
                aq, a_s = fused
            else:
                act, group_amax = ops.gemm_silu_mul(h2, gu.w, h2s, gu.scale, gu.bias, workspace=self.gemm_ws)
                if q is None:
                    aq, a_s = act, None
                else:
                    aq, a_s = ops.quantize_per_token(act, q, group_amax=group_amax)
            nxt = self.layers[li + 1]["input_norm"] if li + 1 < n_layers else None
            x, h, hs = self._row_parallel(L["down"], aq, a_s, x, nxt)
        return x

    def _sample(self, x_rows, samp=None, num_logprobs: int | None = None):
        """final norm -> lm_head (fp32 logits of the local vocab shard) -> greedy argmax (TP: merge (val, idx)), or with
        samp = (params [T,3], u [T]) the on-device top-k / top-p / temperature sampler (TP: candidates all-gathered).
        num_logprobs = n: also returns vLLM's LogprobsTensors triple (ids [T,1+n], logprobs [T,1+n], ranks [T]) computed
        on the device from the candidate lists and one log-sum-exp pass (RUNNER:347-356; None = not requested)."""
        lm_q = _QKIND[self.lm_head.quant]
        if lm_q is None:
            h, hs = ops.rmsnorm(x_rows, self.final_norm, self.shape.rms_eps), None
        else:
            h, hs = ops.rmsnorm(x_rows, self.final_norm, self.shape.rms_eps, quant=lm_q)
        logits = self._linear(self.lm_head, h, hs, out_dtype=torch.float32)
        self.last_logits = logits  # kept for parity tests (a reference, no copy)
        self.last_logprobs = None
        cv = ci = None
        if samp is not None or num_logprobs is not None:
            cv, ci = ops.topk_candidates(logits, idx_offset=self.vocab_offset)
            if self.tp_size > 1:
                gv = torch.empty((self.tp_size, *cv.shape), dtype=cv.dtype, device=cv.device)
                gi = torch.empty((self.tp_size, *ci.shape), dtype=ci.dtype, device=ci.device)
                torch.distributed.all_gather_into_tensor(gv, cv, group=self.tp_group)
                torch.distributed.all_gather_into_tensor(gi, ci, group=self.tp_group)
                cv, ci = gv, gi
        if samp is not None:
            ids = ops.sample_candidates(cv, ci, samp[0], samp[1])
        elif self.tp_size == 1:
            ids = ops.argmax(logits)
        elif self.tp_comm is not None and logits.shape[0] <= self.tp_comm.rows:
            # per-shard (max, global index) straight into peer-mapped memory, merged by our kernel: no NCCL in the step
            ops.argmax(logits, idx_offset=self.vocab_offset, want_val=True, out_val=self.tp_comm.merge_vals,
                       out_idx=self.tp_comm.merge_idxs)
            ids = self.tp_comm.merge_argmax(logits.shape[0])
        else:
            val, idx = ops.argmax(logits, idx_offset=self.vocab_offset, want_val=True)
            vals = torch.empty((self.tp_size, val.numel()), dtype=val.dtype, device=val.device)
            idxs = torch.empty((self.tp_size, idx.numel()), dtype=idx.dtype, device=idx.device)
            torch.distributed.all_gather_into_tensor(vals, val, group=self.tp_group)
            torch.distributed.all_gather_into_tensor(idxs, idx, group=self.tp_group)
            ids = merge_sharded_argmax(vals, idxs, self.shape.vocab)
        if num_logprobs is not None:
            m, se = ops.row_lse(logits)
            if self.tp_size > 1:
                gm = torch.empty((self.tp_size, m.numel()), dtype=m.dtype, device=m.device)
                gs = torch.empty((self.tp_size, se.numel()), dtype=se.dtype, device=se.device)
                torch.distributed.all_gather_into_tensor(gm, m, group=self.tp_group)
                torch.distributed.all_gather_into_tensor(gs, se, group=self.tp_group)
                m, se = gm, gs
            self.last_logprobs = ops.gather_logprobs(cv, ci, ids, m, se, num_logprobs)
        return ids, logits

    # ------------------------------------------------------------------ decode
    def decode_buffers(self, b: int) -> StepBuffers:
        if b not in self._decode_bufs:
            dev = self.device
            self._decode_bufs[b] = StepBuffers(
                tokens=torch.zeros(b, dtype=torch.int64, device=dev),
                row_ids=torch.arange(b, dtype=torch.int32, device=dev),
                slot_mapping=torch.zeros(b, dtype=torch.int64, device=dev),
                positions=torch.zeros(b, dtype=torch.int64, device=dev),
                ctx_lens=torch.zeros(b, dtype=torch.int32, device=dev),
                sampled=torch.zeros(b, dtype=torch.int64, device=dev),
                samp_params=torch.tensor([[1.0, 1.0, 1.0]], dtype=torch.float32, device=dev).repeat(b, 1),
                samp_u=torch.zeros(b, dtype=torch.float32, device=dev),
            )
        return self._decode_bufs[b]

    def _decode_body(self, bufs: StepBuffers, max_ctx_blocks: int, advance: bool, sampled: bool = False,
                     num_logprobs: int | None = None):
        b = bufs.tokens.numel()
        # index construction on device (RUNNER:765-832): positions, slot_mapping, context lengths
        ops.build_decode_meta(self.block_table, bufs.row_ids, self.num_tokens, bufs.slot_mapping, bufs.positions,
                              bufs.ctx_lens, self.block_size)
        x = ops.embedding(bufs.tokens, self.embed)

        def attn(qv, kp, vp, quant):
            return ops.paged_attn_decode(qv, kp, vp, self.block_table, bufs.ctx_lens, self.scale, workspace=self.attn_ws,
                                         row_ids=bufs.row_ids, max_blocks=max_ctx_blocks, quant=quant)

        x = self._layers(x, bufs.positions, bufs.slot_mapping, attn)
        ids, _ = self._sample(x, (bufs.samp_params, bufs.samp_u) if sampled else None, num_logprobs)
        bufs.sampled.copy_(ids)
        if advance:  # device-side twin of RUNNER:325-339 so the captured step can be replayed back to back
            ops.advance_decode(bufs.sampled, bufs.tokens, self.num_tokens, bufs.row_ids)
        return bufs.sampled

    def decode_step(self, b: int, *, max_ctx_blocks: int | None = None, advance: bool = False, use_graph: bool = False,
                    sampled: bool = False, num_logprobs: int | None = None):
        """One decode step for the `b` rows in decode_buffers(b).row_ids; inputs already on the device:
        bufs.tokens (last sampled token per row), self.num_tokens[row] (= prompt + outputs incl. that token).
        sampled=True draws with bufs.samp_params / bufs.samp_u instead of the greedy argmax; num_logprobs=n leaves the
        LogprobsTensors triple in self.last_logprobs.
        Captured graphs are cached per (b, context bound, advance, sampled, num_logprobs) in an LRU of GRAPH_CACHE_CAP
        entries that share ONE memory pool; callers bucket b and the context bound (bucket_batch / bucket_ctx_blocks)."""
        bufs = self.decode_buffers(b)
        mcb = self.max_blocks if max_ctx_blocks is None else max_ctx_blocks
        if not use_graph:
            return self._decode_body(bufs, mcb, advance, sampled, num_logprobs)
        key = (b, mcb, advance, sampled, num_logprobs)
        hit = self._graphs.get(key)
        if hit is None:
            nt = self.num_tokens.clone()
            tk = bufs.tokens.clone()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):  # warm-up outside capture (function attributes, NCCL channels)
                self._decode_body(bufs, mcb, advance, sampled, num_logprobs)
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            self.num_tokens.copy_(nt)
            bufs.tokens.copy_(tk)
            while len(self._graphs) >= GRAPH_CACHE_CAP:  # least recently used graph goes; its memory returns to the pool
                self._graphs.popitem(last=False)
            if self._graph_pool is None:
                self._graph_pool = torch.cuda.graph_pool_handle()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=self._graph_pool):
                self._decode_body(bufs, mcb, advance, sampled, num_logprobs)
            self.num_tokens.copy_(nt)
            bufs.tokens.copy_(tk)
            hit = self._graphs[key] = (g, self.last_logits, self.last_logprobs)
        else:
            self._graphs.move_to_end(key)
        g, self.last_logits, self.last_logprobs = hit
        g.replay()
        return bufs.sampled

    # ------------------------------------------------------------------ prefill (batch 1, prefix-cached)
    def prefill(self, row: int, new_tokens: torch.Tensor, prompt_len: int, computed_len: int, samp=None,
                sample: bool = True, num_logprobs: int | None = None):
        """Compute the uncached suffix `new_tokens` (device int64 [prompt_len - computed_len]) of one request whose
        block ids are already in block_table[row]; returns the sampled next token (device int64 [1]).
        samp = (params [1,3], u [1]) selects the on-device sampler instead of greedy.
        Chunked prefill: `prompt_len` is the END of this chunk (tokens [computed_len, prompt_len) are computed against the
        cached prefix); sample=False for a chunk that does not finish the prompt (no lm_head, returns None)."""
        s_new = prompt_len - computed_len
        assert new_tokens.numel() == s_new and s_new > 0
        if self._prefill_meta is None:  # allocated once (max_model_len int64 each), reused by every prefill / chunk
            self._prefill_meta = (torch.empty(self.max_model_len, dtype=torch.int64, device=self.device),
                                  torch.empty(self.max_model_len, dtype=torch.int64, device=self.device))
        slot_full, pos_full = self._prefill_meta
        bt_row = self.block_table[row]
        ops.build_prefill_meta(bt_row, prompt_len, computed_len, slot_full, pos_full, self.block_size)
        slots, positions = slot_full[:s_new], pos_full[:s_new]
        x = ops.embedding(new_tokens, self.embed)

        def attn(qv, kp, vp, quant):
            return ops.paged_attn_prefill(qv, kp, vp, bt_row, computed_len, self.scale)

        x = self._layers(x, positions, slots, attn)
        self.num_tokens[row] = prompt_len
        if not sample:
            return None
        ids, _ = self._sample(x[-1:].contiguous(), samp, num_logprobs)
        return ids

    def kv_bytes_per_block(self) -> int:
        return 2 * self.hkv * self.block_size * self.d * 2 * self.shape.layers


class _LazyRandomWeights:
    """Generates random-init weights layer by layer (8B+ models do not fit a full CPU copy comfortably).
    Values depend only on (seed, tensor name, layer), never on the TP rank."""

    def __init__(self, shape: ModelShape, seed: int, std: float = 0.02):
        self.shape, self.seed, self.std = shape, seed, std
        self._cache: dict = {}
        self.layers = _LazyLayers(self)

    def _gen(self, tag: int, *s):
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        g = torch.Generator(device=dev).manual_seed(self.seed * 1000003 + tag)
        return (torch.randn(*s, generator=g, dtype=torch.float32, device=dev) * self.std).to(torch.bfloat16)

    def __getitem__(self, k):
        sh = self.shape
        if k == "layers":
            return self.layers
        if k == "embed":
            return self._gen(1, sh.vocab, sh.hidden)
        if k == "lm_head":
            return self._gen(2, sh.vocab, sh.hidden)
        if k == "final_norm":
            return torch.ones(sh.hidden, dtype=torch.bfloat16)
        raise KeyError(k)

    def layer(self, li: int):
        if li not in self._cache:
            sh = self.shape
            qd, kd = sh.num_q_heads * sh.head_dim, sh.num_kv_heads * sh.head_dim
            lw = {
                "input_norm": torch.ones(sh.hidden, dtype=torch.bfloat16),
                "post_norm": torch.ones(sh.hidden, dtype=torch.bfloat16),
                "wqkv": self._gen(100 + li * 8, qd + 2 * kd, sh.hidden),
                "wo": self._gen(101 + li * 8, sh.hidden, qd),
                "wgu": self._gen(102 + li * 8, 2 * sh.intermediate, sh.hidden),
                "wdown": self._gen(103 + li * 8, sh.hidden, sh.intermediate),
            }
            if sh.qkv_bias:
                lw["bqkv"] = self._gen(104 + li * 8, qd + 2 * kd)
            self._cache[li] = lw
        return self._cache[li]

    def drop(self, li: int):
        self._cache.pop(li, None)


class _LazyLayers:
    def __init__(self, parent):
        self.parent = parent

    def __getitem__(self, li):
        return self.parent.layer(li)
