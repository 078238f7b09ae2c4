#!/usr/bin/env python
"""Summarises an ncu pass (gpu__time_duration.sum, sm__pipe_tensor_subpipe_hmma_cycles_active..., dram bytes; CSV) over
tools/ncu_target.py into profiles/r02_tensor_activity.json: per kernel of the LAST recurrent step -- launches, time, share,
tensor-pipe activity, DRAM bytes -- and the time-weighted tensor-pipe activity of the conv stack and of the whole step.

    python tools/summarize_activity.py gpurun_out/r02m_activity_c5chunk.csv gpurun_out/r02m_activity_1080p.csv
"""
import csv
import json
import re
import sys

A = "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active"
out = {}
for tag, fn in zip(("configs4_chunk_4x512x512", "frame_1080p"), sys.argv[1:3]):
    lines = [l for l in open(fn) if not l.startswith("==")]
    rows, order = {}, []
    for row in csv.DictReader(lines):
        i = int(row["ID"])
        if i not in rows:
            rows[i] = {"name": re.sub(r"\(.*", "", row["Kernel Name"]).replace("void isr::", "").replace("isr::", "")}
            order.append(i)
        v = float(row["Metric Value"].replace(",", "")); u = row["Metric Unit"]; m = row["Metric Name"]
        if m == "gpu__time_duration.sum":
            v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
        if m.startswith("dram__bytes"):
            v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        rows[i][m] = v
    ks = [rows[i] for i in order]
    last = ks[[i for i, k in enumerate(ks) if "pack_recurrent" in k["name"]][-1]:]
    T = sum(k["gpu__time_duration.sum"] for k in last)
    conv = [k for k in last if "conv3x3" in k["name"]]
    agg = {}
    for k in last:
        a = agg.setdefault(k["name"], {"launches": 0, "us": 0.0, "tensor_us": 0.0, "dram_bytes": 0.0})
        a["launches"] += 1; a["us"] += k["gpu__time_duration.sum"]; a["tensor_us"] += k["gpu__time_duration.sum"] * k[A] / 100
        a["dram_bytes"] += k.get("dram__bytes_read.sum", 0) + k.get("dram__bytes_write.sum", 0)
    for a in agg.values():
        a["tensor_pipe_active_pct"] = round(100 * a["tensor_us"] / a["us"], 1); a["share_pct"] = round(100 * a["us"] / T, 1)
        a["us"] = round(a["us"], 1); a["dram_GB"] = round(a["dram_bytes"] / 1e9, 3); del a["tensor_us"]; del a["dram_bytes"]
    out[tag] = {"launches": len(last), "kernel_sum_us": round(T, 1),
                "tensor_pipe_active_pct_time_weighted_whole_step": round(sum(k["gpu__time_duration.sum"] * k[A] for k in last) / T, 1),
                "tensor_pipe_active_pct_time_weighted_conv_stack": round(
                    sum(k["gpu__time_duration.sum"] * k[A] for k in conv) / sum(k["gpu__time_duration.sum"] for k in conv), 1),
                "by_kernel": agg}
    print(tag, len(last), "launches", round(T, 1), "us; tensor pipe active: conv stack",
          out[tag]["tensor_pipe_active_pct_time_weighted_conv_stack"], "% whole step",
          out[tag]["tensor_pipe_active_pct_time_weighted_whole_step"], "%", file=sys.stderr)
    for n, a in agg.items():
        print("   ", n, a, file=sys.stderr)
out["what"] = ("ncu --metrics gpu__time_duration.sum, sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active, "
               "dram bytes; --clock-control none; one recurrent step (tools/ncu_target.py), precision hybrid (K = 10, 8-bit low "
               "plane in the trunk); time-weighted = sum(time x activity) / sum(time)")
json.dump(out, open("profiles/r02_tensor_activity.json", "w"), indent=1)
