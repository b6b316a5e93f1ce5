#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel.

    python scripts/summarize_launches.py gpurun_out/launches.csv > profiles/r01_launches.md
"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = val * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        name = name.replace("ts::<unnamed>::", "")
        rows.append((name, r["Grid Size"], r["Block Size"], ns))
    agg = OrderedDict()
    for name, grid, block, ns in rows:
        k = (name, grid, block)
        a = agg.setdefault(k, [0, 0.0, 1e30, 0.0])
        a[0] += 1
        a[1] += ns
        a[2] = min(a[2], ns)
        a[3] = max(a[3], ns)
    total = sum(a[1] for a in agg.values())
    print(f"source: {path}; {len(rows)} launches, total {total / 1e3:.1f} us "
          "(ncu-serialised, cold-cache: shares matter, not absolutes)\n")
    print("| kernel | grid | block | launches | avg us | min us | max us | share |")
    print("|---|---|---|---:|---:|---:|---:|---:|")
    for (name, grid, block), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {grid} | {block} | {a[0]} | {a[1] / a[0] / 1e3:.2f} | "
              f"{a[2] / 1e3:.2f} | {a[3] / 1e3:.2f} | {100 * a[1] / total:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1])
