"""The byte accounting behind bench.py's roofline object (SURVEY 8(d) / DESIGN 3): algorithmic bytes of the attention launch
and of the whole step must be the documented figures for the headline configuration."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["bench_mod"] = mod
    spec.loader.exec_module(mod)
    return mod


def test_attention_bytes_headline():
    b = _bench()
    # 64 sequences x 2048 tokens x (K + V) x 8 kv heads x 128 dims x 2 B  +  q in / out  +  int32 block table
    got = b.attn_algorithmic_bytes([2048] * 64, 32, 8, 128)
    assert got == 64 * 2048 * 2 * 8 * 128 * 2 + 2 * 64 * 32 * 128 * 2 + 64 * 64 * 4 == 537_935_872
    # ragged contexts: ceil(ctx / 32) table entries per sequence
    assert b.attn_algorithmic_bytes([1, 33], 8, 2, 64) == (1 + 33) * 2 * 2 * 64 * 2 + 2 * 2 * 8 * 64 * 2 + (1 + 2) * 4


def test_step_bytes_headline_and_tp():
    b = _bench()
    from vllm_neuron_b200.config import PRESETS

    shape = PRESETS["llama-3.1-8b"]
    w, kv = b.step_algorithmic_bytes(shape, "fp8", 64, 2048, 1)
    assert w == 7_504_658_432 and kv == 32 * 64 * 2048 * 2 * 8 * 128 * 2 == 17_179_869_184
    w8, kv8 = b.step_algorithmic_bytes(shape, "fp8", 64, 2048, 8)
    assert w8 == w // 8 and kv8 == kv // 8
    wb, _ = b.step_algorithmic_bytes(shape, "bf16", 64, 2048, 1)
    assert wb == 2 * w


def test_context_bucket_bounds():
    from vllm_neuron_b200.model import bucket_batch, bucket_ctx_blocks

    for n in (1, 7, 8, 9, 64, 65, 129, 1000, 1030):
        got = bucket_ctx_blocks(n, 1031)
        assert n <= got <= min(1031, max(8, int(n * 1.125) + 8)), (n, got)
    assert [bucket_batch(b, 64) for b in (1, 2, 3, 5, 33, 64)] == [1, 2, 4, 8, 64, 64]


def test_parity_summary_fits_the_head_of_the_line():
    """The one-line TP parity summary that goes into `data`: worst distance per pass over its bar, `info` for the W8A8 deep
    pass (reported, not gating), token agreement; short enough to survive a reader that truncates the JSON line."""
    b = _bench()
    assert b.parity_summary(None) == "tp parity: n/a (1 GPU)"
    assert b.parity_summary({"tp_check": "off"}) == "tp parity: off"
    rec = {"tp_check": "ok", "exchange": "nvls", "passes": [
        {"layers_checked": 2, "gate": "pass/fail", "bar_frac_of_std": 0.3, "tokens_compared": 256, "tokens_equal": 240,
         "logit_rms_err": [0.01, 0.011], "logit_std": [0.1, 0.1]},
        {"layers_checked": 32, "gate": "informational", "bar_frac_of_std": 0.64, "tokens_compared": 256, "tokens_equal": 61,
         "logit_rms_err": [0.043, 0.04], "logit_std": [0.1, 0.1]}]}
    s = b.parity_summary(rec)
    assert s == "tp parity vs TP=1 ok (nvls): 2L 0.110/0.30 std 240/256 tok; 32L 0.430/info std 61/256 tok"
    assert len(s) < 120


def test_traffic_is_paired_with_the_capture_it_came_from():
    """roofline.traffic comes from one committed ncu capture; the line must carry that capture's own algorithmic bytes so
    that nobody divides it by the (different) launch timed in the bench."""
    from types import SimpleNamespace

    b = _bench()
    from vllm_neuron_b200.config import PRESETS

    args = SimpleNamespace(attn_traffic=542090000, ctx=2048,
                           attn_traffic_capture={"algorithmic_bytes": 538700000, "source": "profiles/x.json"})
    got = b.traffic_capture(args, PRESETS["llama-3.1-8b"], 64, 1)
    assert got == {"dram_bytes": 542090000, "algorithmic_bytes": 538700000, "dram_over_algorithmic": 1.0063,
                   "source": "profiles/x.json"}
    assert b.traffic_capture(args, PRESETS["llama-3.1-8b"], 64, 2) is None      # capture is of the TP=1 headline shape only
    assert b.traffic_capture(SimpleNamespace(attn_traffic=None, ctx=2048), PRESETS["llama-3.1-8b"], 64, 1) is None
