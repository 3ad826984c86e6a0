"""Turn the raw ncu artefacts a gpurun call brings back into the small summaries kept under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv            -> per-kernel share table (markdown, stdout)
  python tools/ncu_summary.py full gpurun_out/prof_attn.ncu-rep OUT.json  -> selected metrics per launch (needs `ncu -i`)

The launch list comes from `ncu --metrics gpu__time_duration.sum --clock-control none` (cold caches, serialised: shares
only); the full capture from `ncu --set full --clock-control none --import-source on` (tools/gpu_round.sh).
"""
from __future__ import annotations

import csv
import io
import json
import re
import subprocess
import sys
from collections import defaultdict

FULL_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "smsp__cycles_active.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
]


def _rows(text: str):
    lines = [ln for ln in text.splitlines() if ln.startswith('"')]
    return list(csv.DictReader(io.StringIO("\n".join(lines))))


def short(name: str) -> str:
    name = re.sub(r"\(.*", "", name).replace("void ", "").replace("b200::", "")
    return name.strip()


def launches(path: str) -> None:
    rows = _rows(open(path).read())
    per = defaultdict(list)
    for r in rows:
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
            per[short(r["Kernel Name"])].append(us)
    total = sum(sum(v) for v in per.values())
    print(f"{sum(len(v) for v in per.values())} launches, {total / 1e3:.2f} ms serialised\n")
    print("| share | n | mean us | kernel |\n|---|---|---|---|")
    for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
        print(f"| {100 * sum(v) / total:5.1f}% | {len(v)} | {sum(v) / len(v):.1f} | {k} |")


def full(rep: str, out: str) -> None:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = _rows(txt)
    units = rows[0] if rows and rows[0].get("ID", "") == "" else None
    res = []
    for r in rows:
        if not r.get("ID", "").strip().isdigit():
            continue
        d = {"Kernel Name": r["Kernel Name"]}
        for m in FULL_METRICS:
            if m in r:
                d[m] = f"{r[m]} {units[m]}".strip() if units and units.get(m) else r[m]
        res.append(d)
    json.dump({"source": " ".join(sys.argv), "launches": res}, open(out, "w"), indent=1)
    for d in res:
        print({k: d.get(k) for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                     "dram__bytes_read.sum.per_second", "launch__registers_per_thread")})


if __name__ == "__main__":
    if len(sys.argv) >= 3 and sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif len(sys.argv) >= 4 and sys.argv[1] == "full":
        full(sys.argv[2], sys.argv[3])
    else:
        sys.exit(__doc__)
