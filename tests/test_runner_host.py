"""CPU tests of the model runner's host-side state tracking (no GPU, no kernels): request cache, persistent rows,
block-id append / replace -- the behaviours the reference pins in test/unit/worker/test_model_runner.py:984-1018,
1124-1161 -- and the (row, col, block) triples handed to the device block table."""
import pytest
import torch

from parity_util import make_vllm_config, sched_output
from vllm_neuron_b200.config import PRESETS
from vllm_neuron_b200.worker.b200_model_runner import B200ModelRunner


def _runner(mns=4):
    cfg = make_vllm_config(PRESETS["tiny-test"], max_num_seqs=mns, max_model_len=256, num_gpu_blocks_override=41)
    return B200ModelRunner(cfg, torch.device("cpu"))


def test_new_request_takes_lowest_free_row_and_mirrors_blocks():
    r = _runner()
    upd = r._update_states(sched_output(new_reqs=[("a", list(range(40)), [7, 3], 0)]))
    assert r.req_to_row == {"a": 0} and r.requests["a"].block_ids == [7, 3]
    assert upd == [(0, 0, 7), (0, 1, 3)]
    upd = r._update_states(sched_output(new_reqs=[("b", list(range(5)), [9], 0)]))
    assert r.req_to_row["b"] == 1 and upd == [(1, 0, 9)]


def test_cached_request_block_append_and_resume_replace():
    r = _runner()
    r._update_states(sched_output(new_reqs=[("a", list(range(40)), [7, 3], 0)]))
    r.requests["a"].prefilled = 40                                                  # what execute_model's prefill leaves
    r.requests["a"].output_token_ids.append(11)
    upd = r._update_states(sched_output(cached=[("a", [5])]))
    assert r.requests["a"].block_ids == [7, 3, 5] and upd == [(0, 2, 5)]          # RUNNER:461-466
    assert r.requests["a"].num_computed_tokens == 40                                # num_tokens - 1
    upd = r._update_states(sched_output(cached=[("a", None)]))
    assert upd == [] and r.requests["a"].block_ids == [7, 3, 5]
    so = sched_output(cached=[("a", [20, 21])])
    so.scheduled_cached_reqs.resumed_req_ids = {"a"}
    upd = r._update_states(so)
    assert r.requests["a"].block_ids == [20, 21] and upd == [(0, 0, 20), (0, 1, 21)]  # RUNNER:467-471 (replace)
    # preemption = recompute: the generated token joins the tokens the next step prefills into the new (empty) blocks
    st = r.requests["a"]
    assert st.prompt_token_ids == list(range(40)) + [11] and st.output_token_ids == [] and st.prefilled == 0
    assert st.num_tokens == 41


def test_finished_request_frees_its_row():
    r = _runner(mns=2)
    r._update_states(sched_output(new_reqs=[("a", [1, 2, 3], [4], 0)]))
    r._update_states(sched_output(new_reqs=[("b", [1, 2, 3], [5], 0)]))
    assert not r.free_rows
    with pytest.raises(AssertionError, match="No free sequence ID"):
        r._update_states(sched_output(new_reqs=[("c", [1], [6], 0)]))
    r._update_states(sched_output(finished=["a"]))
    assert r.free_rows == {0} and "a" not in r.requests
    r._update_states(sched_output(new_reqs=[("c", [1], [6], 0)]))
    assert r.req_to_row["c"] == 0
    with pytest.raises(AssertionError, match="existing request ID"):
        r._update_states(sched_output(new_reqs=[("c", [1], [6], 0)]))


def test_empty_step_returns_empty_output():
    from vllm.v1.outputs import EMPTY_MODEL_RUNNER_OUTPUT

    r = _runner()
    assert r.execute_model(sched_output()) is EMPTY_MODEL_RUNNER_OUTPUT  # RUNNER:272-274


def test_config_and_kv_spec_follow_the_reference_keys():
    r = _runner()
    cfg = r.build_b200_config()
    assert cfg.pa_block_size == 32 and cfg.pa_num_blocks == 41 and cfg.is_block_kv_layout and cfg.is_prefix_caching
    spec = r.get_kv_cache_spec()
    assert len(spec) == PRESETS["tiny-test"].layers
    s0 = next(iter(spec.values()))
    assert s0.block_size == 32 and s0.num_kv_heads == 2 and s0.head_size == 64 and s0.dtype == torch.bfloat16


def test_sampling_rows_follow_the_reference_rule(golden_index):
    """RUNNER:1106-1140, pinned by the golden vectors generated from the reference's own get_nxd_sampling_params:
    top_k outside (0, max_topk) -> max_topk = min(global_topk | vocab, 256); temperature 0 -> (1, top_p, 1.0)."""
    from types import SimpleNamespace

    from vllm_neuron_b200.worker.b200_model_runner import CachedRequest

    for c in golden_index["sampling"]:
        r = _runner()
        r.global_topk = c["global_topk"] or 0
        r.model_config.get_vocab_size = lambda v=c["vocab"]: v
        for i, (k, p, t) in enumerate(c["params"]):
            st = CachedRequest(f"s{i}", [1], sampling_params=SimpleNamespace(top_k=k, top_p=p, temperature=t, seed=None))
            got = r.sampling_row(st)
            assert got == (float(c["top_k"][i]), pytest.approx(c["top_p"][i]), pytest.approx(c["temperature"][i])), (c, i, got)


def test_all_greedy_batch_takes_the_argmax_path_and_seeded_draws_repeat():
    from types import SimpleNamespace

    from vllm_neuron_b200.worker.b200_model_runner import CachedRequest

    r = _runner()
    greedy = CachedRequest("g", [1], sampling_params=SimpleNamespace(top_k=0, top_p=1.0, temperature=0.0, seed=None))
    assert r._sampling_inputs([greedy, greedy]) is None
    a = CachedRequest("a", [1], sampling_params=SimpleNamespace(top_k=50, top_p=0.9, temperature=0.8, seed=123))
    b = CachedRequest("b", [1], sampling_params=SimpleNamespace(top_k=50, top_p=0.9, temperature=0.8, seed=123))
    rows, u = r._sampling_inputs([greedy, a])
    assert rows[0] == (1.0, 1.0, 1.0) and rows[1] == (50.0, 0.9, 0.8) and u[0] == 0.0 and 0.0 <= u[1] < 1.0
    _, u2 = r._sampling_inputs([b])
    assert u2[0] == u[1]  # same SamplingParams.seed -> same stream of uniforms


def test_unsupported_sampling_params_do_not_take_the_step_down(caplog):
    """A worker exception ends every in-flight request (vLLM v1 engine core), so the runner must not raise on a request
    whose parameters the device sampler lacks: it warns, keeps the row bookkeeping consistent and samples from
    (top_k, top_p, temperature) alone, like the reference's device path (RUNNER:1106-1140).  Refusal happens at the front
    end (tests/test_platform_plugin.py::test_validate_request_refuses_what_the_device_sampler_lacks)."""
    import logging

    from vllm import SamplingParams

    r = _runner(mns=2)
    so = sched_output(new_reqs=[("a", [1, 2, 3], [4], 0)])
    so.scheduled_new_reqs[0].sampling_params = SamplingParams(repetition_penalty=1.1, logit_bias={5: 2.0}, temperature=0.7)
    with caplog.at_level(logging.WARNING):
        upd = r._update_states(so)
    assert upd == [(0, 0, 4)] and r.req_to_row == {"a": 0} and r.free_rows == {1}
    assert any("repetition_penalty=1.1" in m and "logit_bias" in m for m in caplog.messages)
    assert r.requests["a"].sampling_params.temperature == 0.7


def test_logprob_rows_of_a_mixed_step_hold_only_real_token_ids():
    """A prefill part that asked for 1 log-probability (2 columns) next to a decode part that asked for 5 (6 columns), plus
    a part nobody asked for: every id handed to vLLM must be detokenisable (>= 0), the filler is the row's sampled token
    at -inf, and the columns a request asked for are untouched."""
    import numpy as np

    narrow = (torch.tensor([[7, 9]], dtype=torch.int32), torch.tensor([[-0.5, -0.25]]), torch.tensor([2], dtype=torch.int32))
    wide_ids = torch.tensor([[3, 3, 8, 1, 4, 6], [5, 2, 5, -1, -1, -1]], dtype=torch.int32)  # 2nd row: 3-token vocabulary
    wide_lp = torch.tensor([[-0.125, -0.125, -2.5, -3.0, -4.0, -5.0], [-1.0, -0.5, -1.0, -np.inf, -np.inf, -np.inf]])
    wide = (wide_ids, wide_lp, torch.tensor([1, 2], dtype=torch.int32))
    out = B200ModelRunner._logprobs_lists([(1, narrow), (2, wide), (3, None)])
    ids, lps, ranks = np.asarray(out[0]), np.asarray(out[1]), np.asarray(out[2])
    assert ids.shape == lps.shape == (6, 6) and (ids >= 0).all()
    assert ids[0].tolist() == [7, 9, 7, 7, 7, 7] and lps[0, :2].tolist() == [-0.5, -0.25] and np.isinf(lps[0, 2:]).all()
    assert ids[1].tolist() == [3, 3, 8, 1, 4, 6]
    assert ids[2].tolist() == [5, 2, 5, 5, 5, 5] and lps[2].tolist() == [-1.0, -0.5, -1.0, -1.0, -1.0, -1.0]
    assert (ids[3:] == 0).all() and np.isinf(lps[3:]).all()
    assert ranks.tolist() == [2, 1, 2, 0, 0, 0]
    # what vLLM's front end does with a row: detokenise all ids, then keep num_logprobs + 1 entries
    from vllm.logprobs import append_logprobs_for_next_position

    acc = []
    append_logprobs_for_next_position(acc, ids[0].tolist(), lps[0].tolist(), [None] * 6, int(ranks[0]), 1)
    assert {k: v.logprob for k, v in acc[0].items()} == {7: -0.5, 9: -0.25}
