#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md"""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        ns = val * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        rows.append((name, ns))
    agg = defaultdict(lambda: [0, 0.0])
    for name, ns in rows:
        agg[name][0] += 1
        agg[name][1] += ns
    total = sum(v[1] for v in agg.values()) or 1.0
    print("| kernel | launches | total ms | avg us | share |")
    print("|---|---|---|---|---|")
    for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| %s | %d | %.3f | %.1f | %.1f%% |" % (name, n, ns / 1e6, ns / n / 1e3, 100 * ns / total))
    print("\n%d launches, %.3f ms total (ncu per-launch times: cold-cache, serialised)" % (len(rows), total / 1e6))


if __name__ == "__main__":
    main(sys.argv[1])
