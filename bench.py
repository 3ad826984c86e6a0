#!/usr/bin/env python
"""Headline benchmark: Llama-3.1-8B FP8 (per-channel W8A8) block-KV paged decode, batch 64, block_size 32, 2k context.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU, TP=N)
    python bench.py --impl reference --gpus N --steps K --warmup W   (the CPU restatement of the reference path)

One "step" = one decode step of the whole batch (embedding -> 32 x [norm, qkv GEMM, RoPE + KV write by slot, paged
attention through the block table, o GEMM, norm, gate/up GEMM, SiLU*mul, down GEMM] -> norm -> lm_head -> greedy).
  value : tokens/s, inputs resident in HBM, CUDA-graph replay, device-timed with CUDA events, max over ranks.
  e2e   : the same metric through the plugin's `execute_model(scheduler_output)` with HOST inputs: every step
          copies token ids / rows / lengths from pinned host memory and reads the sampled ids back.
  roofline : paged-attention decode kernel, algorithmic bytes / CUDA-event time per launch vs measured HBM peak.
  cpu_baseline : oracle/ (torch-CPU restatement) timed on the host cores on a bounded sample (N=1, rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"
BS = 32


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def attn_algorithmic_bytes(ctx_lens, hq, hkv, d, kv_bytes=2):
    """sum_seq ctx*2*Hkv*D*sizeof(kv) + 2*B*Hq*D*2 (q in, out) + B*ceil(ctx/bs)*4 (int32 block table)."""
    b = len(ctx_lens)
    return (sum(int(c) * 2 * hkv * d * kv_bytes for c in ctx_lens) + 2 * b * hq * d * 2
            + sum(-(-int(c) // BS) * 4 for c in ctx_lens))


def step_algorithmic_bytes(shape, quant, b, ctx, tp=1):
    """Weights + KV that one decode step must stream from HBM per GPU (lower bound used for the step roofline)."""
    wb = 1 if quant in ("fp8", "int8", "int8_wo") else 2
    qd, kd = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim
    per_layer = ((qd + 2 * kd) * shape.hidden + shape.hidden * qd + 2 * shape.intermediate * shape.hidden
                 + shape.hidden * shape.intermediate) * wb
    weights = (per_layer * shape.layers + shape.vocab * shape.hidden * wb) // tp
    kv = shape.layers * b * ctx * 2 * (shape.num_kv_heads // tp) * shape.head_dim * 2
    return weights, kv


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock / power / throttle reasons of this rank's GPU, sampled through NVML from a thread of THIS process every
    few milliseconds across warm-up and the timed region (nvidia-smi -lms cannot place a sample inside a 60 ms region).
    `mark()` brackets the timed region; the record reports the samples taken inside it (falling back to all samples)."""

    REASONS = (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown"), ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown"),
               ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown"), ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap"),
               ("hw_power_brake", "nvmlClocksEventReasonHwPowerBrakeSlowdown"))

    def __init__(self, gpu_index: int, period_s: float = 0.004):
        self.idx, self.period = gpu_index, period_s
        self.samples, self.marks, self.err, self._stop, self._th = [], [], None, False, None

    def _loop(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            # LOCAL_RANK indexes the visible devices; NVML indexes physical ones
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.idx
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if self.idx < len(ids) and ids[self.idx].isdigit():
                    phys = int(ids[self.idx])
            h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            masks = [(n, getattr(nv, a)) for n, a in self.REASONS if hasattr(nv, a)]
            while not self._stop:
                t = time.perf_counter()
                mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((t, mhz, pw, tuple(n for n, m in masks if r & m)))
                time.sleep(self.period)
        except Exception as e:  # no NVML: the record says so, the bench goes on
            self.err = f"{type(e).__name__}: {e}"

    def start(self):
        import threading

        self._th = threading.Thread(target=self._loop, daemon=True)
        self._th.start()
        return self

    def mark(self):
        self.marks.append(time.perf_counter())

    def stop(self):
        self._stop = True
        if self._th is not None:
            self._th.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err or "no samples"]}
        inside = self.samples
        if len(self.marks) >= 2:
            sel = [x for x in self.samples if self.marks[0] <= x[0] <= self.marks[-1]]
            inside = sel or self.samples
        sm = [x[1] for x in inside]
        pw = [x[2] for x in inside if x[2] is not None]
        reasons = sorted({n for x in inside for n in x[3]})
        return {"sm_mhz": statistics.median(sm), "sm_min_mhz": min(sm), "sm_max_mhz": getattr(self, "max_mhz", max(sm)),
                "power_w_max": max(pw) if pw else None, "samples": len(inside), "samples_total": len(self.samples),
                "source": "NVML, in-process thread, timed region" if inside is not self.samples else "NVML, whole run",
                "reasons": reasons}


# ---------------------------------------------------------------------------------------------- CPU leg (oracle)
def cpu_sample(shape, quant, b, ctx, threads, seed=0):
    """Builds ONE decoder layer + lm_head of the oracle with random-init weights and a KV pool holding `ctx` tokens per
    sequence.  Returns a closure that runs one sample (B tokens through 1 layer) and one that runs norm+lm_head+argmax."""
    import torch

    from oracle import llama_ref

    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(seed)

    def rnd(*s):
        return (torch.randn(*s, generator=g, dtype=torch.float32) * 0.02).to(torch.bfloat16)

    qd, kd = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim
    ones = torch.ones(shape.hidden, dtype=torch.bfloat16)
    w = {"embed": rnd(1024, shape.hidden), "lm_head": rnd(shape.vocab, shape.hidden), "final_norm": ones,
         "layers": [{"input_norm": ones, "post_norm": ones, "wqkv": rnd(qd + 2 * kd, shape.hidden), "wo": rnd(shape.hidden, qd),
                     "wgu": rnd(2 * shape.intermediate, shape.hidden), "wdown": rnd(shape.hidden, shape.intermediate)}]}
    nblk = -(-(ctx + 1) // BS)
    nb = b * nblk + 1
    cfg = llama_ref.RefConfig(shape.hidden, 1, shape.num_q_heads, shape.num_kv_heads, shape.head_dim, shape.intermediate,
                              shape.vocab, shape.rope_theta, shape.rope_scaling, shape.rms_eps, shape.qkv_bias, quant,
                              max_pos=nblk * BS, block_size=BS)
    ref = llama_ref.RefLlama(cfg, w, nb)
    for lin in (ref.lm_head, *[ref.layers[0][k] for k in ("qkv", "o", "gate_up", "down")]):
        lin.prepare_f64()
    ref.k_pools[0].copy_(torch.randn(ref.k_pools[0].shape, generator=g).to(torch.bfloat16))
    ref.v_pools[0].copy_(torch.randn(ref.v_pools[0].shape, generator=g).to(torch.bfloat16))
    perm = torch.randperm(nb - 1, generator=torch.Generator().manual_seed(7)) + 1
    bt = perm[: b * nblk].reshape(b, nblk)
    tokens = torch.randint(0, 1024, (b,), generator=g)
    pos = torch.full((b,), ctx - 1, dtype=torch.int64)
    slots = bt[:, (ctx - 1) // BS] * BS + (ctx - 1) % BS
    ctxl = torch.full((b,), ctx, dtype=torch.int64)
    from oracle import attn_ref, kv_ref

    def layer_once():
        def attn(q, kp, vp, scale):
            return attn_ref.paged_attention_decode_ref(q, kp, vp, bt, ctxl, scale)

        return ref._forward(tokens, pos, slots, attn)

    def head_once(x):
        return kv_ref.argmax_ref(ref._logits(x))

    return layer_once, head_once


def time_cpu_step(shape, quant, b, ctx, threads, reps):
    """Returns (list of extrapolated full-step seconds per rep, description).  One rep = 1 decoder layer on the full
    batch + lm_head/argmax; full step = L x layer + head ('1 layer x L extrapolation')."""
    layer_once, head_once = cpu_sample(shape, quant, b, ctx, threads)
    x = layer_once()  # warm-up (also first-touch of the widened weights)
    head_once(x)
    out = []
    for _ in range(reps):
        t0 = time.perf_counter()
        x = layer_once()
        t1 = time.perf_counter()
        head_once(x)
        t2 = time.perf_counter()
        out.append(shape.layers * (t1 - t0) + (t2 - t1))
    desc = (f"1 of {shape.layers} decoder layers (batch {b}, ctx {ctx}) + lm_head/argmax per rep, extrapolated "
            f"x{shape.layers}; {reps} reps; torch-CPU oracle, fp64-accumulate GEMMs, {threads} threads")
    return out, desc


# ---------------------------------------------------------------------------------------------- reference arm
def run_reference(args, shape):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # only rank 0 runs the CPU arm
    threads = os.cpu_count() or 1
    layer_once, head_once = cpu_sample(shape, args.quant, args.batch, args.ctx, threads)
    x = None
    for _ in range(max(1, min(args.warmup, 2))):  # CPU warm-up is capped: each rep is seconds, not microseconds
        x = layer_once()
        head_once(x)
    # bounded sample: each "step" runs 1 of L layers (+ the head); steps beyond the time budget reuse the running mean
    budget_s, t_layer, t_head, done = args.cpu_budget_s, [], [], 0
    t_start = time.perf_counter()
    for _ in range(args.steps):
        if done >= 4 and time.perf_counter() - t_start > budget_s:  # at least 4 layers' worth, whatever the budget
            break
        t0 = time.perf_counter()
        x = layer_once()
        t1 = time.perf_counter()
        head_once(x)
        t2 = time.perf_counter()
        t_layer.append(t1 - t0)
        t_head.append(t2 - t1)
        done += 1
    # MEDIAN of the timed layers (a shared host's stragglers moved the mean by +-25 % between runs), and the time actually
    # spent next to the extrapolated one, so that a reader can check the line against a wall clock
    step_s = shape.layers * statistics.median(t_layer) + statistics.median(t_head)
    value = args.batch / step_s
    timed_s = sum(t_layer) + sum(t_head)
    sample = (f"each step = 1 of {shape.layers} decoder layers on the full batch ({args.batch} seqs, ctx {args.ctx}) + "
              f"lm_head/argmax, extrapolated x{shape.layers}; {done} of {args.steps} steps timed inside a "
              f"{budget_s:.0f} s budget (median used; {timed_s:.1f} s actually timed, layer min/median/max "
              f"{min(t_layer):.3f}/{statistics.median(t_layer):.3f}/{max(t_layer):.3f} s); torch-CPU oracle "
              f"(oracle/llama_ref.py), fp64-accumulate GEMMs")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": args.quant, "data": "synthetic (random-init weights, random KV, random tokens)",
        "config": workload_config(args, shape, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        # ms_per_step is the EXTRAPOLATED full step; this is what the clock saw
        "timed": {"timed_s": round(timed_s, 3), "steps_timed": done, "layers_per_step_timed": 1,
                  "extrapolation_factor_layers": shape.layers},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, shape, tp):
    kind = {"bf16": "bf16", "fp8": "fp8 per-channel W8A8", "int8": "int8 per-channel W8A8",
            "int8_wo": "int8 per-channel weight-only"}[args.quant]
    return {
        "workload": f"{shape.name} {kind} block-KV paged decode",
        "global_batch": args.batch, "context": args.ctx, "block_size": BS, "parallelism": f"tp{tp}",
        "kv_dtype": "bf16", "l2": "inputs larger than L2 (weights + KV streamed per step >> 126 MB)",
    }


def fill_kv_pools(model, shape, rank, world, blocks=None):
    """Synthetic KV ~ N(0,1) (seed 11), drawn per (layer, GLOBAL kv head): a rank of a TP=N run holds exactly
    the values the TP=1 pool holds for its heads, so runs at different N work on the same data (and can be compared).
    blocks: only these block ids (the parity check's scratch blocks); None = the whole pool.  Block 0 stays zero."""
    import torch

    dev = model.device
    g = torch.Generator(device=dev)
    for li, kv in enumerate(model.kv):
        for hl in range(model.hkv):
            g.manual_seed(11 + 1009 * li + (rank * model.hkv + hl))
            n = kv.shape[1] if blocks is None else len(blocks)
            vals = torch.randn((2, n, kv.shape[3], kv.shape[4]), generator=g, device=dev, dtype=torch.float32).to(kv.dtype)
            if blocks is None:
                kv[:, :, hl] = vals
            else:
                kv[:, blocks, hl] = vals
        if blocks is None:
            kv[:, 0].zero_()  # null block


def tp_parity_check(args, shape, model, bufs, rank, world, dev, group, steps=4, ctx=96):
    """TP=N against TP=1 in this very process, before the timed region: rank 0 also builds the SAME weights un-sharded
    (weights depend on (seed, tensor, layer) only); both models run `steps` teacher-forced decode steps of the whole
    batch from the same state (context `ctx`, scratch blocks 1 .. B*nblk filled with the same per-global-head values).
    Two passes, each on the first n layers + the head of both models:
      shallow (n = 2), TIGHT bar -- a sharding / exchange / merge error is O(logit std) at any depth, while two layers of
        rounding noise stay small: logit rms difference < 3 % of the logit std (bf16, weight-only) or 30 % (W8A8: each rank
        quantises its row-parallel GEMM input over its own columns, so 8-bit codes differ from TP=1's);
      deep (n = all layers, or as many as fit beside rank 0's shard and KV pool), depth-aware bar -- two differently
        rounded computations decorrelate like sqrt(n): measured 0.42 std (8B fp8, 32 layers), 0.63 (70B fp8, 78 layers),
        0.08 (32B weight-only, 64 layers) at TP=8; bar = min(1.0, 1.6 x base x sqrt(n / 2)), base 0.10 (W8A8) / 0.012 (0.016 with the NVLS exchange).
        For W8A8 that bar is 0.64 std at 32 layers and 1.0 std at 78, where two UNRELATED logit vectors sit at 1.41 std: it
        cannot catch a wrong shard, so the W8A8 deep pass is reported (`"gate": "informational"`) and does not decide the
        exit status -- only a non-finite distance fails it.  The gates are the shallow pass (any quantisation) and the deep
        pass of bf16 / weight-only models.
    In both passes the TP=N sampled ids must equal TP=1's wherever TP=1's own top-2 margin is outside 8x the measured rms
    difference.  On failure every rank exits with status 3.  Returns the record that goes into the JSON line."""
    import math

    import torch
    import torch.distributed as dist
    from dataclasses import replace as _replace

    from vllm_neuron_b200.model import B200CausalLM

    if args.tp_check == "off":
        return {"tp_check": "off"}
    b = args.batch
    nblk = -(-(ctx + steps + 1) // BS)
    blocks = list(range(1, b * nblk + 1))
    bt = torch.zeros_like(model.block_table[:b])
    bt[:, :nblk] = torch.tensor(blocks, dtype=torch.int32, device=dev).reshape(b, nblk)
    wb = 1 if args.quant != "bf16" else 2
    qd, kd = shape.num_q_heads * shape.head_dim, shape.num_kv_heads * shape.head_dim
    # bytes of the un-sharded model on rank 0: per layer (weights + its scratch KV), the head (lm_head + bf16 embedding), and
    # room for the loader's transients (one full bf16 + fp32 copy of the largest matrix while it is quantised)
    per_layer = ((qd + 2 * kd + qd) * shape.hidden + 3 * shape.intermediate * shape.hidden) * wb \
        + 2 * (b * nblk + 1) * kd * BS * 2
    head = shape.vocab * shape.hidden * (wb + 2)
    transient = 2 * shape.intermediate * shape.hidden * 6 + (8 << 30)
    status = torch.zeros(2, dtype=torch.int32, device=dev)
    if rank == 0:
        free, _ = torch.cuda.mem_get_info(dev)
        fit = int((free - head - transient) // per_layer) if free > head + transient else 0
        status[1] = max(0, min(shape.layers, fit))
        if int(status[1].item()) < 1:
            status[0] = 2
    dist.broadcast(status, src=0)
    if int(status[0].item()) == 2:
        return {"tp_check": "skipped: not even one layer of the un-sharded model fits beside rank 0's shard"}
    n_fit = int(status[1].item())
    w8a8 = args.quant in ("fp8", "int8")
    tight = 0.3 if w8a8 else 0.03
    # the in-switch (NVLS) sum rounds bf16 ties differently from the rank-order fp32 sum (<= 1 ulp, ~20 % of the elements,
    # profiles/r02_tp_exchange_nvls_v1.md): measured deep-pass distance 0.084 std (pull) vs 0.109 std (NVLS) on the 64-layer
    # weight-only model, 0.054 vs 0.067 on the 32-layer bf16 one
    nvls = bool(getattr(getattr(model, "tp_comm", None), "nvls", False))
    base = 0.10 if w8a8 else (0.016 if nvls else 0.012)
    all_layers, all_kv = model.layers, model.kv
    fill_kv_pools(model, shape, rank, world, blocks=blocks)
    model.block_table[:b].copy_(bt)
    rec = {"tp_check": "ok", "steps": steps, "context": ctx, "layers": shape.layers,
           "exchange": "nvls" if nvls else "pull", "passes": []}
    n_shallow = min(2, n_fit)
    for n_chk, bar in ((n_shallow, tight), (n_fit, max(tight, min(1.0, 1.6 * base * math.sqrt(n_fit / 2))))):
        if rec["passes"] and n_chk == rec["passes"][0]["layers_checked"]:
            break
        gating = not (w8a8 and rec["passes"])  # the deep W8A8 pass is a record, not a gate (docstring)
        model.layers, model.kv = all_layers[:n_chk], all_kv[:n_chk]
        ref = None
        if rank == 0:
            ref = B200CausalLM(_replace(shape, layers=n_chk), None, quant=args.quant, num_blocks=b * nblk + 1, max_num_seqs=b,
                               max_model_len=nblk * BS, block_size=BS, device=dev, seed=0)
            fill_kv_pools(ref, shape, 0, 1, blocks=blocks)
            ref.block_table[:b].copy_(bt[:, :ref.max_blocks])
        tok = torch.randint(0, shape.vocab, (b,), generator=torch.Generator().manual_seed(4321)).to(dev)
        pr = {"layers_checked": n_chk, "gate": "pass/fail" if gating else "informational",
              "bar_frac_of_std": round(bar, 3), "tokens_compared": 0, "tokens_equal": 0, "logit_rms_err": [], "logit_std": []}
        rec["passes"].append(pr)
        for st in range(steps):
            bufs.tokens.copy_(tok)
            model.num_tokens[:b].fill_(ctx + st + 1)
            got = model.decode_step(b, advance=False, use_graph=False).clone()
            l_tp = model.last_logits.float()
            if ref is not None:
                rb = ref.decode_buffers(b)
                rb.tokens.copy_(tok)
                ref.num_tokens[:b].fill_(ctx + st + 1)
                want = ref.decode_step(b, advance=False, use_graph=False).clone()
                l_ref = ref.last_logits.float()
                shard = l_ref[:, model.vocab_offset: model.vocab_offset + l_tp.shape[1]]
                err = (l_tp - shard).pow(2).mean().sqrt().item()
                std = l_ref.std().item()
                top2 = l_ref.topk(2, dim=-1).values
                margin = top2[:, 0] - top2[:, 1]
                bad = (got != want) & (margin > 8 * err)
                pr["tokens_compared"] += b
                pr["tokens_equal"] += int((got == want).sum().item())
                pr["logit_rms_err"].append(round(err, 5))
                pr["logit_std"].append(round(std, 4))
                if (bad.any() or not err < bar * std) if gating else not math.isfinite(err):
                    status[0] = 3
                    rec["tp_check"] = (f"FAILED ({n_chk} layers) at step {st}: {int(bad.sum().item())} tokens differ outside the "
                                       f"margin guard, logit rms err {err:.3e} vs std {std:.3f} (bar {bar:.2f} std)")
                tok = want  # teacher-force both models with the TP=1 tokens
            dist.broadcast(tok, src=0)
            dist.broadcast(status, src=0)
            if int(status[0].item()) == 3:
                if rank == 0:
                    print(json.dumps({"error": "tp parity check failed", "parity": rec}), flush=True)
                sys.stdout.flush()
                os._exit(3)
        del ref
        torch.cuda.empty_cache()
    model.layers, model.kv = all_layers, all_kv
    rec["layers_checked"] = rec["passes"][-1]["layers_checked"]
    return rec


def traffic_capture(args, shape, b, world):
    """What the committed `ncu --set full` capture behind roofline.traffic measured: DRAM bytes of one launch, the
    algorithmic bytes of THAT launch (its context differs from the timed one's by a few tokens) and their ratio."""
    cap = getattr(args, "attn_traffic_capture", None)
    if not (cap and cap.get("algorithmic_bytes") and args.attn_traffic
            and shape.name == "Llama-3.1-8B" and b == 64 and args.ctx == 2048 and world == 1):
        return None
    return {"dram_bytes": args.attn_traffic, "algorithmic_bytes": cap["algorithmic_bytes"],
            "dram_over_algorithmic": round(args.attn_traffic / cap["algorithmic_bytes"], 4), "source": cap["source"]}


def parity_summary(rec) -> str:
    """The TP parity record in one short string for the `data` field (the full record is the line's `parity` object, which
    a reader that keeps only the head of the line loses): per pass `<layers>L <worst rms distance>/<bar> std <equal>/<compared>
    tok`, the bar replaced by `info` for a pass that does not gate."""
    if not rec:
        return "tp parity: n/a (1 GPU)"
    if "passes" not in rec:
        return f"tp parity: {rec.get('tp_check', '?')[:60]}"
    parts = []
    for p in rec["passes"]:
        fr = max((e / s for e, s in zip(p["logit_rms_err"], p["logit_std"]) if s > 0), default=float("nan"))
        bar = f"{p['bar_frac_of_std']:.2f}" if p.get("gate", "pass/fail") == "pass/fail" else "info"
        parts.append(f"{p['layers_checked']}L {fr:.3f}/{bar} std {p['tokens_equal']}/{p['tokens_compared']} tok")
    head = "ok" if rec.get("tp_check") == "ok" else "FAILED"
    return f"tp parity vs TP=1 {head} ({rec.get('exchange', '?')}): " + "; ".join(parts)


def measure_at(args, shape, model, bufs, bt_host, dev, rank, world, local, group, barrier, parity, max_model_len, num_blocks,
               sweep=False):
    """One bench line at context args.ctx on an already built model (KV pool, block table and weights resident)."""
    import torch
    import torch.distributed as dist

    from vllm_neuron_b200 import ops
    from vllm_neuron_b200.model import bucket_ctx_blocks

    b, k_steps, w_steps, e_steps, p_steps = args.batch, args.steps, args.warmup, args.e2e_steps, 2
    n0 = args.ctx - w_steps  # tokens per sequence before the first warm-up step; first timed step has ctx = args.ctx
    # a sweep shares one pool sized for the longest context: bound the attention split by THIS context's blocks, as the
    # model runner does for every step (bucket_ctx_blocks); the single-context run sizes max_model_len to the context
    mcb = bucket_ctx_blocks(-(-(args.ctx + k_steps + w_steps + e_steps + p_steps + 2) // BS), model.max_blocks) if sweep else None
    # ---- device-resident decode: CUDA-graph replay, the sampled token and the length advance on the device
    clocks = ClockSampler(local).start()
    l0 = ops.launch_count()
    model.num_tokens[:b].fill_(n0)
    model.decode_step(b, max_ctx_blocks=mcb, advance=True, use_graph=True)  # captures (one eager warm-up pass + one capture pass) and replays
    launches_per_step = (ops.launch_count() - l0) // 2
    model.num_tokens[:b].fill_(n0)
    for _ in range(w_steps):
        model.decode_step(b, max_ctx_blocks=mcb, advance=True, use_graph=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.mark()
    e0.record()
    for _ in range(k_steps):
        model.decode_step(b, max_ctx_blocks=mcb, advance=True, use_graph=True)
    e1.record()
    barrier()
    clocks.mark()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / k_steps
    value = b * k_steps / (ms * 1e-3)
    mean_ctx = args.ctx + (k_steps - 1) / 2

    # ---- end to end through execute_model (host inputs each step, sampled ids read back each step)
    e2e = run_e2e(args, shape, model, bt_host, dev, rank, world, group, barrier, n_start=args.ctx + k_steps)

    # ---- per-kernel device times (eager, CUDA events around every launch) for the roofline of the dominant kernel
    n_prof = int(model.num_tokens[0].item())
    with ops.kernel_timer() as kt:
        for _ in range(p_steps):
            model.decode_step(b, max_ctx_blocks=mcb, advance=True, use_graph=False)
    ksum = kt.summary()
    hbm_peak, peak_src = load_peaks()
    attn_cnt, attn_ms_eager = ksum["paged_attn_decode"]
    # The dominant kernel as it runs in the step: one launch per decoder layer over that layer's own KV pool (L x 545 MB
    # streamed per replay, far beyond L2), captured once and replayed; CUDA events on the launching stream.
    n_attn = int(model.num_tokens[0].item())
    attn_ms, attn_launches = time_attention_in_graph(model, bufs, b, reps=8, max_blocks=mcb)
    ctxs = [n_attn] * b
    attn_bytes = attn_algorithmic_bytes(ctxs, model.hq, model.hkv, model.d)
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9
    kernels = {}
    per_step_ms = 0.0
    for name, (cnt, mean_ms) in sorted(ksum.items(), key=lambda kv: -kv[1][0] * kv[1][1]):
        kernels[name] = {"launches_per_step": cnt / p_steps, "mean_us": round(mean_ms * 1e3, 2),
                         "us_per_step": round(cnt * mean_ms * 1e3 / p_steps, 1)}
        per_step_ms += cnt * mean_ms / p_steps
    wbytes, kvbytes = step_algorithmic_bytes(shape, args.quant, b, mean_ctx, world)
    roofline = {
        "bound": "hbm", "kernel": "paged_attn_decode_kernel (+ split combine)", "achieved": round(achieved, 1),
        "peak": hbm_peak, "unit": "GB/s", "frac": round(achieved / hbm_peak, 4),
        # the committed ncu --set full capture is of the headline shape (8B, B=64, 2k, TP=1) only
        "traffic": args.attn_traffic if (shape.name == "Llama-3.1-8B" and b == 64 and args.ctx == 2048 and world == 1) else None,
        # `traffic` is an ncu capture of an EARLIER launch of the same kernel at a slightly shorter context than the one timed
        # here: compare it with the capture's own algorithmic bytes (traffic_capture), not with algorithmic_bytes_per_launch
        "traffic_capture": traffic_capture(args, shape, b, world),
        "peak_source": peak_src, "algorithmic_bytes_per_launch": int(attn_bytes), "mean_us_per_launch": round(attn_ms * 1e3, 2),
        "timing": f"CUDA events around {attn_launches} graph-replayed launches (one per layer's KV pool per replay)",
        "mean_us_per_launch_eager": round(attn_ms_eager * 1e3, 2),
        # serialised basis (every kernel timed alone, like the ncu launch list): comparable with profiles/*launches*.csv
        "share_of_step": round(attn_cnt * attn_ms_eager / p_steps / per_step_ms, 4),
        # in-graph basis: L launches at the graph-replayed duration over the measured step (PDL overlaps the other kernels)
        "share_of_step_in_graph": round(shape.layers * attn_ms / ms_per_step, 4),
        "step": {"weights_bytes": int(wbytes), "kv_bytes": int(kvbytes),
                 "min_ms_at_peak": round((wbytes + kvbytes) / hbm_peak / 1e6, 4),
                 "frac_of_step_roofline": round((wbytes + kvbytes) / hbm_peak / 1e6 / ms_per_step, 4)},
    }

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        steps_s, desc = time_cpu_step(shape, args.quant, b, args.ctx, threads, reps=args.cpu_reps)
        cpu = {"value": b / statistics.median(steps_s), "unit": UNIT, "cores": threads, "kind": "port", "sample": desc}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": k_steps, "warmup": w_steps,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": args.quant, "data": "synthetic (random-init weights, N(0,1) KV pool, random tokens, permuted block ids); " + parity_summary(parity),
        "config": workload_config(args, shape, world),  # the same dict the reference arm prints
        "run": {"mean_context_in_timed_region": mean_ctx, "max_model_len": max_model_len, "num_blocks": num_blocks,
                "cuda_graph": True},
        "clocks": clk, "parity": parity, "e2e": e2e, "gpu_launches": launches_per_step * k_steps,
        "launches_per_step": launches_per_step, "roofline": roofline, "cpu_baseline": cpu, "kernels": kernels,
    }
    return line


# ---------------------------------------------------------------------------------------------- B200 arm
def run_b200(args, shape):
    import torch
    import torch.distributed as dist

    from vllm_neuron_b200 import ops
    from vllm_neuron_b200.model import B200CausalLM

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torchrun (one rank per GPU)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    ops.device_check(local)  # fails loudly without an sm_100 device / the CUDA library

    b, k_steps, w_steps, e_steps, p_steps = args.batch, args.steps, args.warmup, args.e2e_steps, 2
    # --sweep-ctx: ONE model / KV pool sized for the longest context, one JSON line per context (BASELINE configs[4])
    sweep = [int(c) for c in args.sweep_ctx.split(",")] if args.sweep_ctx else [args.ctx]
    assert min(sweep) - w_steps >= 1
    n_end = max(sweep) + k_steps + w_steps + e_steps + p_steps + 2
    max_model_len = -(-n_end // BS) * BS
    max_blocks = max_model_len // BS
    num_blocks = b * max_blocks + 1
    model = B200CausalLM(shape, None, quant=args.quant, num_blocks=num_blocks, max_num_seqs=b, max_model_len=max_model_len,
                         block_size=BS, device=dev, tp_rank=rank, tp_size=world, tp_group=group, seed=0)
    # synthetic state: block ids = random permutation of [1, num_blocks) (seed 7), KV ~ N(0,1) (seed 11)
    perm = torch.randperm(num_blocks - 1, generator=torch.Generator().manual_seed(7)) + 1
    bt_host = perm.reshape(b, max_blocks).to(torch.int32)
    model.block_table[:b].copy_(bt_host)
    fill_kv_pools(model, shape, rank, world)
    tok_host = torch.randint(0, shape.vocab, (b,), generator=torch.Generator().manual_seed(1234))
    bufs = model.decode_buffers(b)
    bufs.tokens.copy_(tok_host)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N > 1: correctness of this very model before anything is timed (TP=N vs TP=1 on rank 0, same weights)
    parity = tp_parity_check(args, shape, model, bufs, rank, world, dev, group) if world > 1 else None
    model.block_table[:b].copy_(bt_host)

    for ctx in sweep:
        args.ctx = ctx
        bufs.tokens.copy_(tok_host)
        line = measure_at(args, shape, model, bufs, bt_host, dev, rank, world, local, group, barrier, parity,
                          max_model_len, num_blocks, sweep=len(sweep) > 1)
        if rank == 0:
            print(json.dumps(line), flush=True)
    if world > 1:
        # Leave without tearing NCCL down: destroy_process_group() with captured graphs that hold NCCL kernels alive
        # hung the 2-GPU run (all ranks had finished and printed).  Everything measured is already out.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def time_attention_in_graph(model, bufs, b, reps, max_blocks=None):
    """Mean device time (ms) of one paged-attention decode launch (split kernel + merge) at the current context lengths:
    a CUDA graph holding one launch per layer (each on its own KV pool, as in the step) is replayed `reps` times between
    two events on the launching stream."""
    import torch

    from vllm_neuron_b200 import ops

    q = torch.randn((b, model.hq, model.d), generator=torch.Generator(device=model.device).manual_seed(13),
                    device=model.device, dtype=torch.float32).to(torch.bfloat16)
    quant = {"bf16": None, "int8_wo": None}.get(model.quant, model.quant)  # activation quantisation of the merge kernel
    ops.build_decode_meta(model.block_table, bufs.row_ids, model.num_tokens, bufs.slot_mapping, bufs.positions, bufs.ctx_lens,
                          model.block_size)

    def body():
        for kv in model.kv:
            ops.paged_attn_decode(q, kv[0], kv[1], model.block_table, bufs.ctx_lens, model.scale, workspace=model.attn_ws,
                                  row_ids=bufs.row_ids, max_blocks=max_blocks or model.max_blocks, quant=quant)

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        body()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        body()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    n = reps * len(model.kv)
    return e0.elapsed_time(e1) / n, n


def run_e2e(args, shape, model, bt_host, dev, rank, world, group, barrier, n_start):
    """Decode steps through B200ModelRunner.execute_model: the call vLLM's worker makes (reference WORKER:65-69 ->
    RUNNER:252-291).  Host state in, sampled ids out; H2D/D2H bytes counted by the runner from the tensors it copies."""
    import torch

    from tests_support import make_vllm_config, sched_output
    from vllm_neuron_b200.worker.b200_model_runner import B200ModelRunner, CachedRequest

    b = args.batch
    override = {} if args.quant == "bf16" else {
        "quantized": True, "quantization_dtype": "f8e4m3" if args.quant == "fp8" else "int8",
        "quantization_type": "per_channel_symmetric"}
    if args.quant == "int8_wo":
        override["weight_only"] = True
    cfg = make_vllm_config(shape, max_num_seqs=b, max_model_len=model.max_model_len,
                           num_gpu_blocks_override=model.num_blocks, tp=world, override=override)
    runner = B200ModelRunner(cfg, dev, model=model, tp_rank=rank, tp_size=world, tp_group=group)
    runner.load_model(load_dummy_weights=True)
    runner.initialize_kv_cache(num_blocks=model.num_blocks)
    # requests already mid-generation: n_start tokens each (synthetic prompt ids), every block pre-assigned
    gen = torch.Generator().manual_seed(1234)
    for i in range(b):
        rid = f"req-{i}"
        prompt = torch.randint(0, shape.vocab, (n_start,), generator=gen).tolist()
        runner.requests[rid] = CachedRequest(rid, prompt, [], bt_host[i].tolist(), n_start - 1, i, None, prefilled=n_start)
        runner.req_to_row[rid] = i
    runner.free_rows.clear()
    so = sched_output(cached=[(f"req-{i}", None) for i in range(b)])
    for _ in range(3):
        runner.execute_model(so)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    h2d = d2h = 0
    for _ in range(args.e2e_steps):
        out = runner.execute_model(so)
        h2d += runner.h2d_bytes_last
        d2h += runner.d2h_bytes_last
    e1.record()
    barrier()
    assert len(out.sampled_token_ids) == b
    ms = e0.elapsed_time(e1)
    if world > 1:
        import torch.distributed as dist

        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return {"value": b * args.e2e_steps / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d // args.e2e_steps,
            "d2h_bytes_per_step": d2h // args.e2e_steps, "steps": args.e2e_steps, "ms_per_step": ms / args.e2e_steps,
            "api": "B200ModelRunner.execute_model(scheduler_output)"}


def main():
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "ERROR")  # stdout carries ONE JSON line: no plugin-loader chatter
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--quant", default="fp8", choices=["bf16", "fp8", "int8", "int8_wo"])
    ap.add_argument("--tp-check", default="auto", choices=["auto", "off"],
                    help="N>1: before timing, 4 teacher-forced decode steps of the TP=N model against a TP=1 model of the "
                         "same weights on rank 0 (exit 3 on a mismatch outside the margin guard); off skips it")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--sweep-ctx", default="", help="comma-separated contexts measured on ONE model / KV pool sized for the "
                    "longest (BASELINE configs[4]: 2048,4096,8192,16384,32768); one JSON line per context")
    ap.add_argument("--e2e-steps", type=int, default=32)
    ap.add_argument("--cpu-reps", type=int, default=3)
    ap.add_argument("--cpu-budget-s", type=float, default=150.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layers", type=int, default=0, help="debug: override the layer count (result is then NOT the headline)")
    ap.add_argument("--attn-traffic", type=float, default=None,
                    help="dram bytes read+written per attention launch from the committed ncu --set full capture")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    from dataclasses import replace

    from vllm_neuron_b200.config import PRESETS

    shape = PRESETS[args.model]
    if args.layers:
        shape = replace(shape, layers=args.layers, name=f"{shape.name}[{args.layers} layers, DEBUG]")
    if args.attn_traffic is None:
        try:
            with open(os.path.join(ROOT, "profiles", "attn_traffic.json")) as f:
                cap = json.load(f)
            args.attn_traffic = cap.get("dram_bytes_per_launch")
            args.attn_traffic_capture = {"algorithmic_bytes": cap.get("algorithmic_bytes_same_launch"),
                                         "source": cap.get("source")}
        except Exception:
            pass
    if args.impl == "reference":
        run_reference(args, shape)
    else:
        run_b200(args, shape)


if __name__ == "__main__":
    main()
