#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel -> markdown table.

    python tools/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md
"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        name = r["Kernel Name"]
        name = re.sub(r"\(.*$", "", name)
        rows.append((name, r.get("Grid Size", ""), r.get("Block Size", ""), ns))
    agg = OrderedDict()
    for name, grid, block, ns in rows:
        k = (name, grid, block)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print(f"total launches {len(rows)}, total device time {total / 1e6:.3f} ms (cold-cache, serialised under ncu)\n")
    print("| kernel | grid | block | launches | total us | avg us | share |")
    print("|---|---|---|---:|---:|---:|---:|")
    for (name, grid, block), (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {grid} | {block} | {n} | {ns / 1e3:.1f} | {ns / 1e3 / n:.2f} | {100 * ns / total:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1])
