#!/usr/bin/env python
"""ncu metric CSV (dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum per launch) of ONE UNet eval ->
per-family DRAM traffic JSON read by bench.py (`roofline.traffic`, `traffic_by_family`).

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
        -s <launches of the first eval> -c <launches of one eval> --csv --log-file gpurun_out/traffic_all_TAG.csv \
        python tools/op_table.py --reps 1
    python tools/summarize_traffic.py gpurun_out/traffic_all_TAG.csv profiles/traffic_TAG.json
"""
import csv
import json
import re
import sys

FAMILY = [  # kernel-name regex -> C-ABI family as bench.py names it
    (r"igemm_tc_kernel", "ldmb200_igemm_tc"), (r"attention_tc_kernel", "ldmb200_attention_tc"),
    (r"gn_fused_apply|gn_finalize|gn_apply|gn_stats", "ldmb200_groupnorm_fused"), (r"layernorm", "ldmb200_layernorm"), (r"cast_rows|upsample2x|space_to_depth2|nchw_to_nhwc|nhwc_to_nchw", "layout"),
    (r"small_linear|timestep_embedding", "small dense"),
]


def main():
    src, dst = sys.argv[1], sys.argv[2]
    evals = int(sys.argv[3]) if len(sys.argv) > 3 else 1      # the capture holds this many identical evals: keep the last
    rows = {}
    with open(src) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        d = rows.setdefault(int(r["ID"]), {"kernel": r["Kernel Name"]})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
        d[r["Metric Name"] + "_unit"] = r["Metric Unit"]
    rows = {i: d for i, d in rows.items() if "pack_weights" not in d["kernel"]}
    if evals > 1:
        ids = sorted(rows)
        per = len(ids) // evals
        assert per * evals == len(ids), (len(ids), evals)
        rows = {i: rows[i] for i in ids[-per:]}
    fam = {}
    for i, d in sorted(rows.items()):
        name = "other"
        for rx, fn in FAMILY:
            if re.search(rx, d["kernel"]):
                name = fn
                break
        t = d.get("gpu__time_duration.sum", 0.0)
        unit = d.get("gpu__time_duration.sum_unit", "ns")
        t_ns = t * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd = d.get("dram__bytes_read.sum", 0.0) * scale.get(d.get("dram__bytes_read.sum_unit", "byte"), 1.0)
        wr = d.get("dram__bytes_write.sum", 0.0) * scale.get(d.get("dram__bytes_write.sum_unit", "byte"), 1.0)
        f = fam.setdefault(name, {"launches": 0, "dram_bytes_read": 0.0, "dram_bytes_write": 0.0, "ncu_time_ns": 0.0})
        f["launches"] += 1
        f["dram_bytes_read"] += rd
        f["dram_bytes_write"] += wr
        f["ncu_time_ns"] += t_ns
    for f in fam.values():
        f["dram_bytes_total"] = f["dram_bytes_read"] + f["dram_bytes_write"]
    out = {"families": fam, "launches": len(rows), "source": src.replace("gpurun_out/", "profiles/"),
           "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none over "
                  "every launch of one UNet eval (B=64, 64x64, f16 mode; tools/op_table.py --reps 1); cold-cache serialised launches"}
    json.dump(out, open(dst, "w"), indent=1)
    for k, f in sorted(fam.items(), key=lambda kv: -kv[1]["ncu_time_ns"]):
        print(f"{k:28s} {f['launches']:4d} launches  {f['dram_bytes_total'] / 1e9:8.2f} GB  {f['ncu_time_ns'] / 1e6:8.3f} ms")


if __name__ == "__main__":
    main()
