#!/usr/bin/env python
"""In-graph kernel timeline of one decode step (torch.profiler / CUPTI activity records over a CUDA-graph replay).

    python tools/timeline.py [--quant fp8] [--layers 8] [--json out.json]

Unlike the ncu launch list (serialised, cold caches) this shows the kernels as they run inside the replayed graph, with
programmatic-dependent-launch overlap: per kernel name the mean duration, the mean gap to the next kernel start, and the
share of the step.  A profiler is attached, so the numbers explain the step; they are never a bench value.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from dataclasses import replace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_neuron_b200 import ops  # noqa: E402
from vllm_neuron_b200.config import PRESETS  # noqa: E402
from vllm_neuron_b200.model import B200CausalLM  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--quant", default="fp8")
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        group = dist.group.WORLD
    ops.device_check(local)
    shape = replace(PRESETS[a.model], layers=a.layers)
    bs = 32
    mml = -(-(a.ctx + 64) // bs) * bs
    mb = mml // bs
    nb = a.batch * mb + 1
    model = B200CausalLM(shape, None, quant=a.quant, num_blocks=nb, max_num_seqs=a.batch, max_model_len=mml, device=f"cuda:{local}",
                         tp_rank=rank, tp_size=world, tp_group=group)
    perm = torch.randperm(nb - 1, generator=torch.Generator().manual_seed(7)) + 1
    model.block_table[:a.batch].copy_(perm.reshape(a.batch, mb).to(torch.int32))
    for kv in model.kv:
        kv.normal_()
    bufs = model.decode_buffers(a.batch)
    bufs.tokens.copy_(torch.randint(0, shape.vocab, (a.batch,)))
    model.num_tokens[:a.batch].fill_(a.ctx)
    for _ in range(4):
        model.decode_step(a.batch, advance=True, use_graph=True)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            model.decode_step(a.batch, advance=True, use_graph=True)
        torch.cuda.synchronize()
    if rank != 0:
        torch.distributed.barrier()
        os._exit(0)
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
           and "memset" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    n = len(evs) // 3
    evs = evs[n: 2 * n]  # middle replay
    t0 = evs[0].time_range.start
    t_end = max(e.time_range.end for e in evs)
    rows = []
    agg: dict[str, list] = {}
    for i, e in enumerate(evs):
        nxt = evs[i + 1].time_range.start if i + 1 < len(evs) else e.time_range.end
        name = e.name.split("(")[0].replace("b200::", "")
        dur = e.time_range.end - e.time_range.start
        pitch = nxt - e.time_range.start  # start-to-start: what the kernel costs the step
        prev_end = max((x.time_range.end for x in evs[:i]), default=e.time_range.start)
        tail = max(0.0, e.time_range.end - prev_end)  # end-to-end: what the kernel adds after everything before it ended
        rows.append((e.time_range.start - t0, dur, pitch, name, tail))
        s = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
        s[0] += 1
        s[1] += dur
        s[2] += pitch
        s[3] += tail
    step = t_end - t0
    print(f"step {step:.1f} us, {len(evs)} kernels, layers={a.layers}")
    print(f"{'kernel':70s} {'n':>4s} {'dur us':>8s} {'pitch us':>9s} {'tail us':>8s} {'share':>6s}")
    out = []
    for name, (cnt, dur, pitch, tail) in sorted(agg.items(), key=lambda kv: -kv[1][3]):
        print(f"{name[:70]:70s} {cnt:4d} {dur / cnt:8.2f} {pitch / cnt:9.2f} {tail / cnt:8.2f} {tail / step:6.1%}")
        out.append({"kernel": name, "n": cnt, "mean_dur_us": dur / cnt, "mean_pitch_us": pitch / cnt,
                    "mean_tail_us": tail / cnt, "share": tail / step})
    # one layer in order (the second layer of the step)
    per_layer = (len(evs) - 6) // a.layers if a.layers else 0
    if per_layer > 0:
        print("\none layer, in launch order (start us, dur us, start-to-start us, end-to-end us):")
        base = 3 + per_layer
        for st, dur, pitch, name, tail in rows[base: base + per_layer]:
            print(f"  {st:9.1f} {dur:7.2f} {pitch:7.2f} {tail:7.2f}  {name[:80]}")
    if world > 1:
        sys.stdout.flush()
        torch.distributed.barrier()
        if a.json:
            json.dump({"step_us": step, "layers": a.layers, "kernels": out}, open(a.json, "w"), indent=1)
        os._exit(0)
    if a.json:
        json.dump({"step_us": step, "layers": a.layers, "kernels": out,
                   "layer_trace": [{"start_us": r[0], "dur_us": r[1], "pitch_us": r[2], "tail_us": r[4], "kernel": r[3]}
                                   for r in rows[3 + per_layer: 3 + 2 * per_layer]] if per_layer > 0 else []},
                  open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
