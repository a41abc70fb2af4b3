#!/usr/bin/env python
"""Summarise an `ncu --set full` report into a small CSV for profiles/ (one row per kernel launch, or --group to
aggregate launches by kernel name).  Usage: python tools/ncu_summary.py in.ncu-rep|raw.csv out.csv [--group]"""
import csv
import subprocess
import sys
from collections import OrderedDict

COLS = OrderedDict([
    ("Kernel Name", "kernel"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"),
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor_pipe_pct"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor_pipe_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occupancy_pct"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"),
    ("smsp__cycles_active.avg", "smsp_cycles_active"),
])

UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12,
        "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6, "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


def main():
    src, dst = sys.argv[1], sys.argv[2]
    group = "--group" in sys.argv
    if src.endswith(".csv"):  # raw page already exported on the GPU box (ncu -i rep --page raw --csv)
        out = open(src).read()
    else:
        out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True,
                             check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, body = rows[0], rows[1], rows[2:]
    sel = []
    seen = set()
    for name, short in COLS.items():
        if name in hdr and short not in seen:
            sel.append((hdr.index(name), short))
            seen.add(short)
    recs = []
    for r in body:
        d = OrderedDict()
        for i, short in sel:
            v = r[i]
            if short == "kernel":
                d[short] = v.split("(")[0]
                continue
            try:
                f = float(v.replace(",", ""))
            except ValueError:
                d[short] = v
                continue
            if short in ("time", "dram_read", "dram_write"):
                f *= UNIT.get(units[i], 1.0)
            d[short] = f
        recs.append(d)
    for d in recs:
        if "time" in d and d["time"]:
            d["dram_GBps"] = round((d.get("dram_read", 0) + d.get("dram_write", 0)) / d["time"] / 1e3, 1)
    if group:
        agg = OrderedDict()
        for d in recs:
            a = agg.setdefault(d["kernel"], {"kernel": d["kernel"], "launches": 0, "time": 0.0, "dram_read": 0.0,
                                             "dram_write": 0.0, "_w": {}})
            a["launches"] += 1
            t = d.get("time", 0.0)
            a["time"] += t
            a["dram_read"] += d.get("dram_read", 0.0)
            a["dram_write"] += d.get("dram_write", 0.0)
            for k in ("dram_pct", "sm_pct", "tensor_pipe_pct", "achieved_occupancy_pct", "l2_hit_pct", "regs"):
                if isinstance(d.get(k), float):
                    a["_w"][k] = a["_w"].get(k, 0.0) + d[k] * t
        recs = []
        for a in agg.values():
            w = a.pop("_w")
            for k, v in w.items():
                a[k + "_timeweighted"] = round(v / a["time"], 2) if a["time"] else 0
            a["dram_GBps"] = round((a["dram_read"] + a["dram_write"]) / a["time"] / 1e3, 1) if a["time"] else 0
            recs.append(a)
        recs.sort(key=lambda a: -a["time"])
    keys = []
    for d in recs:
        for k in d:
            if k not in keys:
                keys.append(k)
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["# units: time us, dram_read/dram_write bytes, dram_GBps = (read+write)/time"])
        w.writerow(keys)
        for d in recs:
            w.writerow([d.get(k, "") for k in keys])
    print(f"{dst}: {len(recs)} rows")


if __name__ == "__main__":
    main()
