"""Measured dense FP8 (e4m3) tensor peak of this B200: the denominator BASELINE.md 2 leaves to the builder.

Same protocol as the driver's bf16 figure in MEASURED_PEAKS.json (torch.matmul 8192^3: best of 10 = burst, back to back for
4 s = sustained), with cuBLASLt's FP8 GEMM (torch._scaled_mm, e4m3 x e4m3 -> bf16, per-tensor scales, fp32 accumulate) in
place of the bf16 matmul; bf16 is re-measured beside it so the two come from one box and one clock state.  LIBRARY code, used
only as a yardstick -- nothing on the product path calls it.  Writes one JSON object (default gpurun_out/fp8_peak.json).

    python tools/fp8_peak.py [--out gpurun_out/fp8_peak.json]
"""
import argparse
import json
import time

import torch


def _time(fn, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters  # ms


def measure(fn, flops):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    burst = max(flops / (_time(fn, 1) * 1e-3) / 1e12 for _ in range(10))
    t_end, n, ms = time.time() + 4.0, 0, 0.0
    while time.time() < t_end:  # sustained: back to back for 4 s, in chunks of 20 launches
        ms += _time(fn, 20) * 20
        n += 20
    return round(burst, 1), round(flops * n / (ms * 1e-3) / 1e12, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/fp8_peak.json")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    res = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__, "shapes": []}
    one = torch.ones((), device=dev, dtype=torch.float32)
    for n in (4096, 8192, 16384):
        x = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        w = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        x8, w8 = x.to(torch.float8_e4m3fn), w.to(torch.float8_e4m3fn)
        flops = 2.0 * n * n * n
        bf = measure(lambda: torch.matmul(x, w.t()), flops)
        f8 = measure(lambda: torch._scaled_mm(x8, w8.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16), flops)
        res["shapes"].append({"M=N=K": n, "bf16_tflops_burst": bf[0], "bf16_tflops_sustained": bf[1],
                              "fp8_tflops_burst": f8[0], "fp8_tflops_sustained": f8[1]})
        del x, w, x8, w8
    best = max(res["shapes"], key=lambda r: r["fp8_tflops_sustained"])
    res["fp8_tflops"] = max(r["fp8_tflops_burst"] for r in res["shapes"])
    res["fp8_tflops_sustained"] = best["fp8_tflops_sustained"]
    res["bf16_tflops_sustained_same_run"] = max(r["bf16_tflops_sustained"] for r in res["shapes"])
    res["how"] = ("torch._scaled_mm e4m3 x e4m3 -> bf16 (cuBLASLt), per-tensor scale 1.0, N^3 in {4096, 8192, 16384}: best of 10 "
                  "single launches (burst) and back to back for 4 s (sustained), CUDA events; bf16 torch.matmul beside it")
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
