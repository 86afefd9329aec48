"""Where the device solve spends its time: python profiles/solve_table.py [workload] [nrhs]
(profile mode: events around every solve launch; diagnostic only)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                              # noqa: E402
from suitesparse_b200 import b200, matrices as M         # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
nrhs = int(sys.argv[2]) if len(sys.argv) > 2 else 1
kind, size, ordering, desc = bench.WORKLOADS[name]
ref, _ = bench.analyze(kind, size, ordering)
sym = ref.symbolic()
Sp, Si, Sx = ref.seam()
F = b200.Factor(sym, device=0)
F.upload_matrix(Sp, Si, Sx)
F.factor_device()
b = np.asfortranarray(np.tile(M.rhs(sym["n"])[:, None], (1, nrhs)))
ts = []
for _ in range(4):
    F.solve(b)
    ts.append(F.getd("ms_solve"))
F.seti("profile", 1)
F.solve(b)
parts = {k: (round(F.getd("ms_solve_part%d" % i), 3), int(F.getd("solve_launches%d" % i)))
         for i, k in enumerate(["small_fwd", "wide_fwd", "small_bwd", "wide_bwd"])}
print(json.dumps(dict(workload=name, nrhs=nrhs, solve_cols=int(F.getd("solve_cols")), solve_ms=ts, profiled_ms=F.getd("ms_solve"), parts_ms_launches=parts,
                      xsize_gb=sym["xsize"] * 8 / 1e9)))
