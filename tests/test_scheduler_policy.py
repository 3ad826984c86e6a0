"""CPU tests of the continuous-batching policy (reference behaviours: test/unit/core/test_scheduler.py:147-164,
404-422 for admission; :430-663 for the min_tokens stop rules)."""
import pytest

from vllm_neuron_b200.core.scheduler import admit_count, stop_reason


@pytest.mark.parametrize("running,pending,cap,want", [
    (0, 0, 4, 0),
    (0, 1, 4, 1),   # empty engine: start a new batch
    (0, 3, 4, 1),   # prefill batch is 1 (SCHED:116)
    (2, 3, 4, 1),   # a free decode line: one more prefill
    (4, 3, 4, 0),   # decode batch full: nothing admitted
    (3, 1, 4, 1),
    (0, 2, 1, 1),
])
def test_admit_count(running, pending, cap, want):
    assert admit_count(running, pending, cap) == want


def _stop(**kw):
    base = dict(num_tokens=10, num_output_tokens=3, max_tokens=16, max_model_len=2048, min_tokens=0, last_token_id=5,
                eos_token_id=2, ignore_eos=False, stop_token_ids=())
    base.update(kw)
    return stop_reason(**base)


def test_stop_rules():
    assert _stop() is None
    assert _stop(last_token_id=2) == "stop"
    assert _stop(last_token_id=2, ignore_eos=True) is None
    assert _stop(last_token_id=2, min_tokens=5) is None            # EOS suppressed until min_tokens reached
    assert _stop(last_token_id=2, min_tokens=3) == "stop"
    assert _stop(last_token_id=7, stop_token_ids=(7,)) == ("stop", 7)
    assert _stop(last_token_id=7, stop_token_ids=(7,), min_tokens=9) is None
    assert _stop(num_output_tokens=16) == "length"                 # max_tokens wins over min_tokens
    assert _stop(num_output_tokens=16, min_tokens=32) == "length"
    assert _stop(num_tokens=2048) == "length"
    assert _stop(pooling=True) is None and _stop(pooling=True, has_pooler_output=True) == "stop"


def test_stop_check_on_a_real_vllm_request():
    """The installed vLLM (0.2x) keeps the EOS id on SamplingParams, 0.11 kept it on the Request (SCHED:149): the
    wrapper must work on the real object.  min_tokens holds a stop token back; the cap still wins."""
    from vllm import SamplingParams
    from vllm.v1.request import Request, RequestStatus

    from vllm_neuron_b200.core.scheduler import check_stop_with_min_tokens

    sp = SamplingParams(temperature=0, max_tokens=3, min_tokens=2, stop_token_ids=[7])
    r = Request(request_id="x", prompt_token_ids=[1, 2, 3], sampling_params=sp, pooling_params=None)
    r.append_output_token_ids(7)
    assert check_stop_with_min_tokens(r, 100) is False
    r.append_output_token_ids(7)
    assert check_stop_with_min_tokens(r, 100) is True and r.status == RequestStatus.FINISHED_STOPPED and r.stop_reason == 7
    r2 = Request(request_id="y", prompt_token_ids=[1, 2, 3], sampling_params=SamplingParams(max_tokens=50), pooling_params=None)
    r2.append_output_token_ids(5)
    assert check_stop_with_min_tokens(r2, 4) is True and r2.status == RequestStatus.FINISHED_LENGTH_CAPPED


def test_request_wrapper_equals_the_pure_rule():
    """check_stop_with_min_tokens (written out on the Request's fields for speed) == stop_reason over a grid of states."""
    import itertools
    from types import SimpleNamespace

    from vllm.v1.request import RequestStatus

    from vllm_neuron_b200.core.scheduler import check_stop_with_min_tokens, stop_reason

    for n_out, max_tokens, min_tokens, last, ignore_eos, stop_ids, n_prompt in itertools.product(
            (0, 1, 3, 5), (3, 100), (0, 4), (2, 7, 9), (False, True), ((), (7,)), (4, 98)):
        outs = [1] * (n_out - 1) + [last] if n_out else []
        sp = SimpleNamespace(min_tokens=min_tokens, ignore_eos=ignore_eos, stop_token_ids=list(stop_ids), eos_token_id=2)
        req = SimpleNamespace(sampling_params=sp, pooling_params=None, num_output_tokens=n_out, num_tokens=n_prompt + n_out,
                              max_tokens=max_tokens, output_token_ids=outs, status=None, stop_reason=None)
        want = stop_reason(num_tokens=n_prompt + n_out, num_output_tokens=n_out, max_tokens=max_tokens, max_model_len=100,
                           min_tokens=min_tokens, last_token_id=outs[-1] if outs else -1, eos_token_id=2,
                           ignore_eos=ignore_eos, stop_token_ids=stop_ids)
        got = check_stop_with_min_tokens(req, 100)
        assert got == (want is not None), (n_out, max_tokens, min_tokens, last, ignore_eos, stop_ids, n_prompt)
        if want == "length":
            assert req.status == RequestStatus.FINISHED_LENGTH_CAPPED
        elif want is not None:
            assert req.status == RequestStatus.FINISHED_STOPPED
            assert req.stop_reason == (want[1] if isinstance(want, tuple) else None)


# ---------------------------------------------------------------- against the reference's own scheduler (goldens)
def _golden_sched():
    import json
    import os

    with open(os.path.join(os.path.dirname(__file__), "golden", "index_golden.json")) as f:
        return json.load(f)["scheduler"]


def test_stop_rule_matches_reference_goldens():
    """check_stop_with_min_tokens against the outputs of the reference's own function (SCHED:130-166) on 160 request
    states (tests/golden/gen_from_reference.py ran it)."""
    from types import SimpleNamespace

    from vllm.v1.request import RequestStatus

    from vllm_neuron_b200.core.scheduler import check_stop_with_min_tokens

    cases = _golden_sched()["stop"]
    assert len(cases) >= 100 and {c["status"] for c in cases} >= {None, "FINISHED_STOPPED", "FINISHED_LENGTH_CAPPED"}
    for c in cases:
        n_out = c["num_output_tokens"]
        req = SimpleNamespace(
            num_tokens=c["num_tokens"], num_output_tokens=n_out, max_tokens=c["max_tokens"], pooling_params=None,
            eos_token_id=c["eos_token_id"], output_token_ids=[5] * (n_out - 1) + [c["last_token"]], status=None, stop_reason=None,
            sampling_params=SimpleNamespace(min_tokens=c["min_tokens"], ignore_eos=c["ignore_eos"],
                                            stop_token_ids=c["stop_token_ids"], eos_token_id=c["eos_token_id"]))
        assert check_stop_with_min_tokens(req, c["max_model_len"]) == c["stopped"], c
        assert (None if req.status is None else RequestStatus(req.status).name) == c["status"], c
        assert req.stop_reason == c["stop_reason"], c


def test_can_schedule_and_queue_movement_match_reference_goldens():
    """can_schedule and schedule()'s hold-back / hide-the-decodes queue movement against the reference's own class run on
    plain deques with the base scheduler stubbed (SCHED:69-127)."""
    from collections import deque
    from types import SimpleNamespace

    import vllm_neuron_b200.core.scheduler as S

    g = _golden_sched()
    cls = S.ContinuousBatchingB200Scheduler
    for c in g["can_schedule"]:
        fake = SimpleNamespace(running=[0] * c["running"], waiting=[0] * c["waiting"], max_num_running_reqs=c["max_running"])
        assert cls.can_schedule(fake, None) == c["can"], c
    seen = {}
    base = cls.__mro__[1]
    orig = base.schedule

    def base_schedule(self):
        seen["waiting"], seen["running"] = list(self.waiting), list(self.running)
        return "OUT"

    base.schedule = base_schedule
    try:
        for c in g["schedule"]:
            obj = object.__new__(cls)
            obj.waiting = deque(f"w{i}" for i in range(c["waiting"]))
            obj.running = [f"r{i}" for i in range(c["running"])]
            obj.holdback_queue = deque()
            obj.max_num_running_reqs = c["max_running"]
            assert obj.schedule() == "OUT"
            assert seen["waiting"] == c["base_saw_waiting"] and seen["running"] == c["base_saw_running"], c
            assert list(obj.waiting) == c["waiting_after"] and list(obj.running) == c["running_after"], c
    finally:
        base.schedule = orig
