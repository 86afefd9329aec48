"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list:
    python profiles/ncu_summary.py launches.csv "<command that was profiled>" > summary.json"""
import csv
import json
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"]
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("b200::", "")
    val = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = val / 1e3 if unit in ("ns", "nsecond") else val if unit in ("us", "usecond") else val * 1e3
    rows.append((name, us))
# one factorization = from one k_scatter_A launch up to the next (the capture may hold warm-up kernels before
# it and the beginning of another factorization after it)
starts = [i for i, (n, _) in enumerate(rows) if n.startswith("k_scatter_A")]
if starts:
    rows = rows[starts[0]:(starts[1] if len(starts) > 1 else len(rows))]
tot = sum(u for _, u in rows)
agg = defaultdict(lambda: [0, 0.0])
for n, u in rows:
    agg[n][0] += 1
    agg[n][1] += u
kernels = [dict(kernel=n, launches=c, total_us=round(u, 1), share=round(u / tot, 4), avg_us=round(u / c, 2))
           for n, (c, u) in sorted(agg.items(), key=lambda kv: -kv[1][1])]
print(json.dumps(dict(command=sys.argv[2] if len(sys.argv) > 2 else "",
                      note="ncu --metrics gpu__time_duration.sum --clock-control none; per-launch times are serialised "
                           "and cold-cache: compare shares",
                      total_us=round(tot, 1), launches=len(rows), kernels=kernels), indent=1))
