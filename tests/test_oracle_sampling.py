"""CPU tests of the sampler oracle (oracle/sampling_ref.py): the restated top-k / temperature / top-p / inverse-CDF rule
against first-principles properties (no GPU)."""
import numpy as np

from oracle import sampling_ref as S


def _logits(b, v, seed):
    return np.random.default_rng(seed).standard_normal((b, v)).astype(np.float32) * 3


def test_candidates_are_the_sorted_top_256_with_low_index_ties():
    x = _logits(3, 5000, 0)
    x[1, 10] = x[1, 4000] = x[1].max() + 1  # tie at the top: index 10 first
    val, idx = S.topk_candidates_ref(x, idx_offset=7)
    assert idx[1, 0] == 17 and idx[1, 1] == 4007
    for r in range(3):
        assert np.array_equal(val[r], np.sort(x[r])[::-1][:256])
        assert np.array_equal(x[r, idx[r] - 7], val[r])
    v2, i2 = S.topk_candidates_ref(x[:, :100])
    assert np.isinf(v2[:, 100:]).all() and (i2[:, 100:] == -1).all()


def test_greedy_and_degenerate_rows():
    x = _logits(4, 1000, 1)
    val, idx = S.topk_candidates_ref(x)
    params = np.array([[1, 1.0, 1.0], [50, 1e-6, 1.0], [50, 1.0, 1e-3], [256, 1.0, 1.0]], dtype=np.float32)
    u = np.array([0.99, 0.99, 0.5, 0.0], dtype=np.float32)
    ids, _ = S.sample_candidates_ref(val, idx, params, u)
    am = x.argmax(axis=1)
    assert ids[0] == am[0]  # top_k 1
    assert ids[1] == am[1]  # top_p -> 0 keeps one token
    assert ids[2] == am[2]  # temperature -> 0 concentrates on the max
    assert ids[3] == am[3]  # u = 0 picks the most likely token


def test_empirical_distribution_matches_truncated_softmax():
    x = _logits(1, 300, 2)
    val, idx = S.topk_candidates_ref(x)
    k, top_p, temp = 8, 0.8, 0.7
    n = 4000
    u = ((np.arange(n) + 0.5) / n).astype(np.float32)  # stratified uniforms
    params = np.tile(np.array([[k, top_p, temp]], dtype=np.float32), (n, 1))
    ids, _ = S.sample_candidates_ref(np.repeat(val, n, 0), np.repeat(idx, n, 0), params, u)
    w = np.exp((val[0, :k] - val[0, 0]) / temp).astype(np.float64)
    p = w / w.sum()
    kept = int(np.searchsorted(np.cumsum(p), top_p - 1e-9) + 1)
    q = p[:kept] / p[:kept].sum()
    counts = np.array([(ids == idx[0, j]).sum() for j in range(kept)]) / n
    assert counts.sum() == 1.0 and np.abs(counts - q).max() < 2e-3


def test_merging_shards_equals_unsharded():
    x = _logits(5, 4096, 3)
    val, idx = S.topk_candidates_ref(x)
    parts = [S.topk_candidates_ref(x[:, s * 1024:(s + 1) * 1024], idx_offset=s * 1024) for s in range(4)]
    gv, gi = np.stack([p[0] for p in parts]), np.stack([p[1] for p in parts])
    params = np.tile(np.array([[40, 0.9, 1.1]], dtype=np.float32), (5, 1))
    u = np.random.default_rng(4).random(5).astype(np.float32)
    a, _ = S.sample_candidates_ref(val, idx, params, u)
    b, _ = S.sample_candidates_ref(gv, gi, params, u)
    assert np.array_equal(a, b)


def test_kept_distribution_matches_vllm_sampler_filter():
    """Pin the sampler oracle to vLLM's own top-k / top-p filter -- what the reference's CPU sampling path runs
    (RUNNER:1142-1239 -> vllm.v1.sample.sampler.Sampler -> apply_top_k_top_p): for rows of (top_k, top_p, temperature)
    the oracle, swept over a grid of uniforms, must draw exactly the tokens vLLM's filter keeps, each with the
    probability of vLLM's masked softmax (the inverse-CDF draw itself has no reference: the device RNG is NxDI's)."""
    import torch
    from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch

    b, v, grid = 6, 3000, 2048
    x = _logits(b, v, 5)
    params = np.array([[40, 0.9, 0.8], [256, 0.5, 1.0], [5, 1.0, 1.3], [100, 0.95, 0.6], [256, 1.0, 1.0], [17, 0.3, 2.0]],
                      dtype=np.float32)
    val, idx = S.topk_candidates_ref(x)
    want = apply_top_k_top_p_pytorch(torch.tensor(x) / torch.tensor(params[:, 2:3]), torch.tensor(params[:, 0]).long(),
                                     torch.tensor(params[:, 1]), allow_cpu_sync=True).softmax(dim=-1).numpy()
    us = ((np.arange(grid) + 0.5) / grid).astype(np.float32)
    for r in range(b):
        ids, _ = S.sample_candidates_ref(np.repeat(val[r:r + 1], grid, 0), np.repeat(idx[r:r + 1], grid, 0),
                                         np.repeat(params[r:r + 1], grid, 0), us)
        freq = np.bincount(ids, minlength=v) / grid
        kept = np.nonzero(want[r] > 0)[0]
        # support: every drawn token is kept by vLLM; every kept token heavier than one grid cell is drawn
        assert set(np.nonzero(freq)[0]) <= set(kept), (r, set(np.nonzero(freq)[0]) - set(kept))
        assert all(freq[t] > 0 for t in kept if want[r, t] > 2.0 / grid)
        assert np.abs(freq - want[r]).max() <= 2.0 / grid, (r, np.abs(freq - want[r]).max())


def test_logprobs_oracle_matches_vllm_sampler():
    """Pin logprobs_ref to vLLM's Sampler.compute_logprobs + gather_logprobs -- what the reference returns on its CPU
    sampling path (RUNNER:347-356)."""
    import torch
    from vllm.v1.sample.sampler import Sampler

    b, v, n = 9, 5000, 7
    x = _logits(b, v, 21)
    x[0, 3] = x[0, 77] = x[0].max() + 0.5  # tie at the top: both count towards the rank
    sampled = np.array([3, 10, 4999, 0, 17, 100, 2500, 42, 77], dtype=np.int64)
    sampled[1] = int(x[1].argmax())
    lp = Sampler.compute_logprobs(torch.tensor(x))
    with torch.compiler.set_stance("force_eager"):  # vLLM torch.compiles its rank count; no OpenMP toolchain needed here
        want = Sampler.gather_logprobs(lp, n, torch.tensor(sampled))
    ids, vals, rank = S.logprobs_ref(x, sampled, n)
    assert np.array_equal(rank, want.selected_token_ranks.numpy())
    assert np.array_equal(ids[:, 0], want.logprob_token_ids[:, 0].numpy())
    np.testing.assert_allclose(vals, want.logprobs.numpy(), rtol=0, atol=2e-6)
    # top-n ids: identical wherever the values are distinct (torch.topk's tie order is unspecified)
    wi = want.logprob_token_ids.numpy()
    for r in range(b):
        distinct = np.concatenate([[True], np.diff(vals[r, 1:]) != 0]) & np.concatenate([np.diff(vals[r, 1:]) != 0, [True]])
        assert np.array_equal(ids[r, 1:][distinct], wi[r, 1:][distinct])
