#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.

usage: python tools/ncu_launch_summary.py launches.csv [skip_first_n]
"""
import csv
import re
import sys
from collections import defaultdict


def main(path, skip=0):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        rows.append((int(r["ID"]), r["Kernel Name"], val * scale))
    rows = rows[skip:]
    agg = defaultdict(lambda: [0, 0.0])
    for _, name, us in rows:
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        name = re.sub(r"^void ", "", name)
        short = name.split("(")[0][:90]
        agg[short][0] += 1
        agg[short][1] += us
    total = sum(v[1] for v in agg.values())
    print(f"{len(rows)} launches, {total/1e3:.2f} ms total (cold-cache, serialised: compare shares)")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
        print(f"{us/total*100:6.2f}%  {us/1e3:9.3f} ms  {n:6d} x {us/n:9.1f} us  {name}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
