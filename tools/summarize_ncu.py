#!/usr/bin/env python
"""Markdown summary of `ncu -i X.ncu-rep --page raw --csv` exports (tools/gpu_ncu_full.sh): one row per profiled
launch with duration, DRAM bytes and GB/s against the measured copy peak, DRAM / tensor-pipe / XU-pipe / issue
utilisation, occupancy, registers and shared memory.
Usage: python tools/summarize_ncu.py gpurun_out/ncu_*_TAG.csv > profiles/ncu_full_TAG.md"""
import csv
import json
import os
import re
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))

# (column title, substrings that must all occur in the ncu metric name, first match wins)
WANTED = [
    ("us", ["gpu__time_duration.sum"]),
    ("dram rd MB", ["dram__bytes_read.sum"]),
    ("dram wr MB", ["dram__bytes_write.sum"]),
    ("dram %", ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]),
    ("tensor %", ["sm__pipe_tensor", "cycles_active", "pct_of_peak_sustained_active"]),
    ("xu %", ["sm__inst_executed_pipe_xu", "pct_of_peak_sustained_active"]),
    ("fma %", ["sm__pipe_fma_cycles_active", "pct_of_peak_sustained_active"]),
    ("issue %", ["sm__inst_issued", "pct_of_peak_sustained_active"]),
    ("L2 %", ["lts__t_sectors.avg.pct_of_peak_sustained_elapsed"]),
    ("warps %", ["sm__warps_active.avg.pct_of_peak_sustained_active"]),
    ("regs", ["launch__registers_per_thread"]),
    ("smem KB", ["launch__shared_mem_per_block_dynamic"]),
]


def short(name):
    name = re.sub(r"^void\s+", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("ldm::", "").replace("<unnamed>::", "")
    return name[:48]


def num(s):
    try:
        return float(s.replace(",", ""))
    except (ValueError, AttributeError):
        return None


def to_base(v, unit, kind):
    """ncu prints per-column units; normalise times to us and byte counts to MB."""
    if v is None:
        return None
    u = (unit or "").strip().lower().split("/")[0]
    if kind == "time":
        return v * {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6,
                    "nsecond": 1e-3}.get(u, 1e-3)
    if kind == "bytes":
        return v * {"byte": 1e-6, "bytes": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3, "b": 1e-6,
                    "kb": 1e-3, "mb": 1.0, "gb": 1e3}.get(u, 1e-6)
    if kind == "kb":
        return v * {"byte": 1e-3, "kbyte": 1.0, "mbyte": 1e3}.get(u, 1e-3)
    return v


def main(paths):
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = float(peaks.get("hbm_gbs", 6454.6))
    print(f"# ncu --set full --clock-control none: per-launch summary (HBM peak used for the GB/s share: {hbm:.0f} GB/s measured copy)\n")
    print("Per-launch times under ncu are cold-cache and serialised; utilisation percentages are ncu's own "
          "(`pct_of_peak_sustained_*`). `GB/s` = (dram read + write) / duration.\n")
    for path in paths:
        rows = list(csv.reader(open(path, newline="")))
        hdr_i = next((i for i, r in enumerate(rows) if r and r[0] == "ID"), None)
        if hdr_i is None:
            print(f"## {os.path.basename(path)}: no kernel rows\n")
            continue
        hdr, units, body = rows[hdr_i], rows[hdr_i + 1], rows[hdr_i + 2:]
        col = {}
        for title, subs in WANTED:
            for j, h in enumerate(hdr):
                if all(s in h for s in subs):
                    col[title] = j
                    break
        kname = hdr.index("Kernel Name")
        grid = hdr.index("Grid Size") if "Grid Size" in hdr else None
        block = hdr.index("Block Size") if "Block Size" in hdr else None
        titles = [t for t, _ in WANTED if t in col]
        print(f"## {os.path.basename(path)} ({len(body)} launches)\n")
        print("| # | kernel | grid | block | " + " | ".join(titles) + " | GB/s | of HBM peak |")
        print("|---:|---|---|---|" + "---:|" * (len(titles) + 2))
        for r in body:
            if len(r) <= kname:
                continue
            vals = {}
            for t in titles:
                v = num(r[col[t]])
                kind = "time" if t == "us" else "bytes" if t.startswith("dram ") and t.endswith("MB") else "kb" if t == "smem KB" else ""
                vals[t] = to_base(v, units[col[t]], kind)
            cells = []
            for t in titles:
                v = vals[t]
                cells.append("-" if v is None else f"{v:.0f}" if t == "regs" else f"{v:.1f}")
            gbs = None
            if vals.get("us") and vals.get("dram rd MB") is not None and vals.get("dram wr MB") is not None:
                gbs = (vals["dram rd MB"] + vals["dram wr MB"]) * 1e6 / (vals["us"] * 1e-6) / 1e9
            tail = f" | {gbs:.0f} | {100 * gbs / hbm:.0f}% |" if gbs is not None else " | - | - |"
            print(f"| {r[0]} | `{short(r[kname])}` | {r[grid] if grid is not None else ''} | "
                  f"{r[block] if block is not None else ''} | " + " | ".join(cells) + tail)
        print()
        # warp-stall breakdown (stall cycles per issue-active cycle): what the warps wait for
        stall = [(h.split("issue_stalled_")[1].split("_per_issue")[0], j) for j, h in enumerate(hdr)
                 if "average_warps_issue_stalled_" in h and h.endswith("_per_issue_active.ratio")]
        if stall:
            keep = [n for n, _ in stall if n not in ("selected",)]
            print("| # | kernel | " + " | ".join(keep) + " |")
            print("|---:|---|" + "---:|" * len(keep))
            for r in body:
                if len(r) <= kname:
                    continue
                cells = []
                for n, j in stall:
                    if n in keep:
                        v = num(r[j])
                        cells.append("-" if v is None else f"{v:.2f}")
                print(f"| {r[0]} | `{short(r[kname])}` | " + " | ".join(cells) + " |")
            print()


if __name__ == "__main__":
    main(sys.argv[1:])
