"""How far do the CUDA path's bf16 logits drift from the CPU oracle's as the model gets deeper?  Runs the model parity
driver of tests/test_gpu_model.py at the TinyLlama-1.1B shape (22 layers) for two init scales and prints the logit error
against the logit spread.  Measured on B200: std 0.02 -> rms 2.4 %, max 12.7 % of the spread (tiny-test, 2 layers:
1.0 % / 4.1 %); std 0.05 -> 6.4 % / 34 % (every linear then has gain 2.3 and the random network amplifies rounding
differences).  This is what sets the near-tie band of tests/test_gpu_vllm_engine.py::...cfg1...  (test tooling: imports oracle/ via tests/)."""
import sys, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import test_gpu_model as t
for std in (0.02, 0.05):
    runner, model, ref, outs, ref_outs, margins, live, blocks = t._run("bf16", steps=4, preset="tinyllama-1.1b", std=std)
    st = model.stats
    sd = t._rms(st["ref"] - st["ref"].mean())
    e = st["gpu"] - st["ref"]
    flat_got = [x for k in sorted(outs) for x in outs[k]]; flat_ref = [x for k in sorted(outs) for x in ref_outs[k]]
    print(f"std {std}: logits rms err {t._rms(e):.4e} max {e.abs().max().item():.4e} spread {sd:.4f} -> rms/spread {t._rms(e)/sd:.4f} max/spread {e.abs().max().item()/sd:.4f}; token mismatches {sum(a!=b for a,b in zip(flat_got,flat_ref))}/{len(flat_got)}", flush=True)
