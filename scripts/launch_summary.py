"""Markdown summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list (per kernel: launches, time, share).

    python scripts/launch_summary.py gpurun_out/launches.csv STEPS > profiles/rN_bench_launches_summary.md
"""
import csv
import re
import sys
from collections import OrderedDict

path, steps = sys.argv[1], int(sys.argv[2])
rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
hdr = rows[0]
name_i, val_i, unit_i = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rows[1:]:
    t = float(r[val_i].replace(",", ""))
    t = {"ns": t / 1e3, "us": t, "ms": t * 1e3}.get(r[unit_i], t)
    name = re.sub(r"\(.*", "", r[name_i])[:100]
    n, s = agg.get(name, (0, 0.0))
    agg[name] = (n + 1, s + t)
total = sum(s for _, s in agg.values())
count = sum(n for n, _ in agg.values())
print(f"{steps} steps, {count} launches ({count // steps} per tracked frame), {total / 1e3:.2f} ms summed kernel time.\n")
print("| kernel | launches | total us | share |\n|---|---|---|---|")
for name, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if s / total < 0.002:
        continue
    print(f"| `{name}` | {n} | {s:.1f} | {100 * s / total:.1f}% |")
own = [(n, s) for k, (n, s) in agg.items() if "woft::" in k]
print(f"\nHand-written kernels (woft::*): {sum(n for n, _ in own)} launches ({sum(n for n, _ in own) // steps} per frame), "
      f"{sum(s for _, s in own):.1f} us = {100 * sum(s for _, s in own) / total:.1f}% of the step; the rest is whatever "
      "else the step launched (ATen element-wise / copy kernels; cuDNN / cuBLAS kernels would be listed above by name).")
