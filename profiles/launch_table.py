"""Per-launch device times of one profiled factorization, joined with the plan tables:
    python profiles/launch_table.py [workload] > table.json
For every update launch: tiles, executed (padded-tile) flops, mean K, duration, TFLOP/s; plus the panel-chain
launches.  Diagnostic only (event overhead inside the run): never a bench number."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                              # noqa: E402
from suitesparse_b200 import b200                        # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
kind, size, ordering, desc = bench.WORKLOADS[name]
ref, t_an = bench.analyze(kind, size, ordering)
sym = ref.symbolic()
Sp, Si, Sx = ref.seam()
F = b200.Factor(sym, device=0)
F.upload_matrix(Sp, Si, Sx)
for _ in range(2):
    F.factor_device()
ms_plain = F.getd("ms_factor")
F.seti("profile", 1)
F.factor_device()
launches = F.plan_table("launches", b200.LAUNCH_DT)
tiles = F.plan_table("tiles", b200.TILE_DT)
items = F.plan_table("items", b200.ITEM_DT)
rows = []
for i, L in enumerate(launches):
    k = int(L["kind"])
    if k >= len(b200.KIND_NAMES) or L["count"] <= 0:
        continue
    ms = F.getd("ms_launch%d" % i)
    row = dict(i=i, kind=b200.KIND_NAMES[k], level=int(L["level"]), stream=int(L["stream"]), count=int(L["count"]),
               start=round(F.getd("ms_launch_start%d" % i), 4), ms=round(ms, 4))
    if k in b200.TILE_SHAPE:
        TM, TN = b200.TILE_SHAPE[k]
        t = tiles[int(L["first"]):int(L["first"]) + int(L["count"])]
        kk = (t["k1"] - t["k0"]).astype(np.float64)
        row["flops_exec"] = float(np.sum(kk) * TM * TN * 2)
        row["kmean"] = float(kk.mean()); row["kmax"] = float(kk.max())
        row["tflops_exec"] = round(row["flops_exec"] / (ms * 1e9), 2) if ms > 0 else None
        row["nitems"] = int(np.unique(t["item"]).size)
    rows.append(row)
d128 = [r for r in rows if r["kind"] == "upd_dmma128"]
tot = sum(r["ms"] for r in d128)
buckets = {}
for r in d128:
    key = ("K<=64" if r["kmax"] <= 64 else "K<=256" if r["kmax"] <= 256 else "K>256",
           "tiles<296" if r["count"] < 296 else "tiles<1184" if r["count"] < 1184 else "tiles>=1184")
    b = buckets.setdefault("%s %s" % key, dict(ms=0.0, flops=0.0, n=0))
    b["ms"] += r["ms"]; b["flops"] += r["flops_exec"]; b["n"] += 1
for b in buckets.values():
    b["tflops_exec"] = round(b["flops"] / (b["ms"] * 1e9), 2); b["ms"] = round(b["ms"], 3); del b["flops"]
print(json.dumps(dict(workload=name, ms_unprofiled=ms_plain, ms_profiled=F.getd("ms_factor"), d128_ms=round(tot, 3),
                      d128_buckets=buckets, peak=b200.dmma_peak_tflops(0))))
print(json.dumps(dict(launches=rows)))
