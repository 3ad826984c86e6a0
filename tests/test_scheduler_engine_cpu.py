"""The continuous-batching scheduler inside the REAL vLLM 0.2x Scheduler / KV-cache manager, on the CPU (the reference's
test/unit/core/test_scheduler.py:147-164, 404-422 does the same against vLLM 0.11 with mocks): tools/scheduler_cpu_sim.py
feeds six requests through four batch lines with a fake one-token-per-request model runner and prints every
SchedulerOutput.  Run in a subprocess because vLLM resolves its platform once per process."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def sim():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "scheduler_cpu_sim.py")], cwd=ROOT, capture_output=True,
                       text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_platform_rewrite_reaches_the_real_vllm_config(sim):
    assert sim["platform"] == "B200Platform"
    assert sim["scheduler_cls"].endswith("ContinuousBatchingB200Scheduler")      # PLAT:176-187
    assert sim["worker_cls"] == "vllm_neuron_b200.worker.b200_worker.B200Worker"  # PLAT:162-164
    assert sim["num_gpu_blocks_override"] == 49                                  # 48 + the null block, PLAT:149-159
    assert sim["async_scheduling"] is False and sim["chunked_prefill"] is False


def test_a_step_is_one_prefill_or_a_decode_batch(sim):
    steps = [s for s in sim["steps"] if s["num_scheduled_tokens"]]
    assert steps
    for s in steps:
        assert (len(s["new"]) == 1 and not s["cached"]) or (not s["new"] and s["cached"]), s  # SCHED:116: never mixed
        assert len(s["cached"]) <= sim["max_num_seqs"]
        if s["cached"]:
            assert all(s["num_scheduled_tokens"][r] == 1 for r in s["cached"])
    # the first four requests are prefilled back to back, one per step, before anything decodes
    assert [s["new"][0][0] for s in steps[:4]] == ["r0", "r1", "r2", "r3"]
    assert steps[4]["cached"] and sorted(steps[4]["cached"]) == ["r0", "r1", "r2", "r3"]


def test_waiting_requests_enter_when_a_line_frees_and_everything_finishes(sim):
    assert sim["produced"] == sim["plan"]  # every request produced exactly max_tokens tokens, none lost or repeated
    steps = sim["steps"]
    admit = {s["new"][0][0]: i for i, s in enumerate(steps) if s["new"]}
    # r4 had to wait for r1 (3 tokens) to finish: it is prefilled right after r1's last decode step
    last_r1 = max(i for i, s in enumerate(steps) if "r1" in s["cached"])
    assert admit["r4"] == last_r1 + 1
    running_before_r4 = set(steps[last_r1]["cached"])
    assert len(running_before_r4) == sim["max_num_seqs"]  # all four lines were busy until then


def test_prefix_cache_hit_is_reported_to_the_runner(sim):
    r5 = next(s["new"][0] for s in sim["steps"] if s["new"] and s["new"][0][0] == "r5")
    r0 = next(s["new"][0] for s in sim["steps"] if s["new"] and s["new"][0][0] == "r0")
    assert r0[1] == 0 and r5[1] == 64            # num_computed_tokens: two full 32-token blocks of r0's prompt
    assert r5[2][0][:2] == r0[2][0][:2]           # ... served by the very same block ids (RUNNER:739-763)
    assert 0 not in r5[2][0] and 0 not in r0[2][0]  # block 0 is the null block, never handed out
