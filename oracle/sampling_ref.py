"""Oracle (test infrastructure): the on-device sampler, restated on the CPU.

Reference side that IS visible and pinned (tests/golden/index_golden.json "sampling" cases, generated from
RUNNER:1106-1140): the [B,3] (top_k, top_p, temperature) rows -> oracle/index_ref.sampling_params_rows.
The sampler itself lives in NxDI (`self.model.sample`, LOADER:367-382; absent here); the kept-token set / probabilities are pinned to vLLM's own top-k / top-p filter
(tests/test_oracle_sampling.py), the inverse-CDF draw has no reference (device RNG); this file
restates the published top-k / temperature / top-p / multinomial scheme in the exact fp32, sequential order the CUDA
kernels use (csrc/sampling.cu), with the caller's uniform u[b] in place of the device RNG:
  candidates = the 256 largest logits, (value desc, index asc);  k = clamp(top_k, 1, #candidates)
  w_j = exp((v_j - v_0) / temperature) for j < k;  Z = sum w_j
  keep the shortest prefix whose cumulative weight reaches top_p * Z (at least one token)
  pick the first kept j whose cumulative weight exceeds u * Z_kept (else the last kept token)
"""
from __future__ import annotations

import numpy as np

SAMPLE_K = 256


def topk_candidates_ref(logits: np.ndarray, idx_offset: int = 0):
    """logits [B, V] fp32 -> (val [B,256] fp32, idx [B,256] int64), -inf / -1 padded."""
    b, v = logits.shape
    k = min(SAMPLE_K, v)
    val = np.full((b, SAMPLE_K), -np.inf, dtype=np.float32)
    idx = np.full((b, SAMPLE_K), -1, dtype=np.int64)
    for r in range(b):
        order = np.lexsort((np.arange(v), -logits[r].astype(np.float64)))[:k]  # value desc, then index asc
        val[r, :k] = logits[r, order]
        idx[r, :k] = order + idx_offset
    return val, idx


def sample_candidates_ref(cand_val: np.ndarray, cand_idx: np.ndarray, params: np.ndarray, u: np.ndarray):
    """cand_* [groups, B, 256] (or [B,256]); returns (ids [B] int64, margin [B]) where margin is the distance of the draw
    from the nearest decision boundary, relative to Z (tests skip rows whose margin is below fp32 noise)."""
    if cand_val.ndim == 2:
        cand_val, cand_idx = cand_val[None], cand_idx[None]
    g, b, _ = cand_val.shape
    ids = np.zeros(b, dtype=np.int64)
    margin = np.full(b, np.inf)
    f32 = np.float32
    for r in range(b):
        v = cand_val[:, r].reshape(-1)
        i = cand_idx[:, r].reshape(-1)
        ok = i >= 0
        v, i = v[ok], i[ok]
        order = np.lexsort((i, -v.astype(np.float64)))[:SAMPLE_K]
        v, i = v[order], i[order]
        k = int(params[r, 0])
        k = max(1, min(k, len(v)))
        top_p, temp = f32(params[r, 1]), f32(params[r, 2])
        if k == 1:
            ids[r] = i[0]
            continue
        w = np.exp(((v[:k] - v[0]).astype(f32) / temp).astype(f32)).astype(f32)
        z = f32(0)
        for j in range(k):
            z = f32(z + w[j])
        cut = f32(top_p * z)
        acc, kept = f32(0), 0
        for j in range(k):
            if j > 0:
                margin[r] = min(margin[r], abs(float(acc) - float(cut)) / float(z))
                if acc >= cut:
                    break
            acc = f32(acc + w[j])
            kept = j + 1
        target = f32(f32(u[r]) * acc)
        run, pick = f32(0), kept - 1
        for j in range(kept):
            run = f32(run + w[j])
            margin[r] = min(margin[r], abs(float(run) - float(target)) / float(z))
            if run > target:
                pick = j
                break
        ids[r] = i[pick]
    return ids, margin


def logprobs_ref(logits: np.ndarray, sampled: np.ndarray, num_logprobs: int):
    """vLLM's Sampler.compute_logprobs + gather_logprobs (vllm/v1/sample/sampler.py), which the reference's CPU sampling
    path returns through ModelRunnerOutput.logprobs (RUNNER:347-356, 1142-1239): raw log-softmax of the fp32 logits,
    column 0 = the sampled token, columns 1..n = the n most likely tokens (value desc, index asc on ties), rank = number
    of entries >= the sampled token's log-probability.  Pinned to vLLM's own function in tests/test_oracle_sampling.py."""
    x = logits.astype(np.float64)
    m = x.max(axis=1, keepdims=True)
    lse = m + np.log(np.exp(x - m).sum(axis=1, keepdims=True))
    lp = x - lse
    b = x.shape[0]
    order = np.lexsort((np.arange(x.shape[1])[None, :].repeat(b, 0), -x), axis=1)[:, :num_logprobs]
    tok = sampled.astype(np.int64)
    ids = np.concatenate([tok[:, None], order], axis=1).astype(np.int32)
    vals = np.take_along_axis(lp, ids.astype(np.int64), axis=1).astype(np.float32)
    rank = (x >= np.take_along_axis(x, tok[:, None], axis=1)).sum(axis=1).astype(np.int32)
    return ids, vals, rank
