#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel family.
Usage: python tools/summarize_launches.py gpurun_out/launches_X.csv > profiles/launches_X.md"""
import csv
import re
import sys
from collections import defaultdict


def family(name):
    m = re.search(r"ldm::(\w+)", name)
    base = m.group(1) if m else name.split("(")[0][:60]
    t = re.search(r"<(.*?)>\(", name)
    return base + (f"<{t.group(1)}>" if t else "")


def main(path):
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        if r.get("Metric Unit") == "us":
            v *= 1e3
        rows.append((family(r["Kernel Name"]), v, r["Grid Size"], r["Block Size"]))
    tot = sum(v for _, v, _, _ in rows)
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for k, v, _, _ in rows:
        a = agg[k]
        a[0] += 1
        a[1] += v
        a[2] = max(a[2], v)
    print(f"# ncu launch list summary: {path}\n")
    print(f"{len(rows)} launches, {tot/1e6:.3f} ms total device time (cold-cache, serialised: compare SHARES)\n")
    print("| kernel | launches | total ms | share | mean us | max us |")
    print("|---|---:|---:|---:|---:|---:|")
    for k, (n, s, mx) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {s/1e6:.3f} | {100*s/tot:.1f}% | {s/n/1e3:.1f} | {mx/1e3:.1f} |")


if __name__ == "__main__":
    main(sys.argv[1])
