"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals and shares."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
tot = defaultdict(lambda: [0, 0.0])
with open(path, newline="") as f:
    rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki])
    name = re.sub(r"<.*", "", name).replace("void ", "")
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(r[ui], 1.0)
    tot[name][0] += 1
    tot[name][1] += v
total = sum(v[1] for v in tot.values())
print(f"# {path}: {sum(v[0] for v in tot.values())} launches, {total / 1e3:.2f} ms total device time (cold-cache, serialised)")
print("kernel,launches,total_us,share")
for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k},{n},{us:.1f},{us / total:.4f}")
