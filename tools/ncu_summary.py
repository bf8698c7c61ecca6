#!/usr/bin/env python
"""Turns an `ncu --metrics gpu__time_duration.sum --csv` launch list into the per-kernel table kept under profiles/.
python tools/ncu_summary.py gpurun_out/launches.csv > profiles/launches_rN_summary.md"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for rec in csv.DictReader(lines):
        if rec.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(rec["Metric Value"].replace(",", ""))
        unit = rec.get("Metric Unit", "ns")
        us = val / 1000.0 if unit in ("ns", "nsecond") else val * (1000.0 if unit in ("ms", "msecond") else 1.0)
        name = re.sub(r"\(.*$", "", rec["Kernel Name"])
        name = re.sub(r"^void\s+", "", name).replace("b200::", "")
        rows.append((name, us))
    tot = sum(us for _, us in rows)
    agg = defaultdict(lambda: [0, 0.0])
    for name, us in rows:
        agg[name][0] += 1
        agg[name][1] += us
    print(f"launches: {len(rows)}, total {tot:.1f} us\n")
    print("| kernel | launches | total us | share | avg us |")
    print("|---|---:|---:|---:|---:|")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {n} | {us:.1f} | {100 * us / tot:.1f}% | {us / n:.2f} |")


if __name__ == "__main__":
    main()
