"""Phase breakdown of the wide-supernode solve kernels (diagnostic build only: make NVFLAGS+=-DB200_SOLVE_TRACE):
average SM-clock cycles between the stamps of the CTA that ends up solving the diagonal block, over all launches.
python profiles/solve_trace.py [workload]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                              # noqa: E402
from suitesparse_b200 import b200, matrices as M         # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
kind, size, ordering, desc = bench.WORKLOADS[name]
ref, _ = bench.analyze(kind, size, ordering)
sym = ref.symbolic()
Sp, Si, Sx = ref.seam()
F = b200.Factor(sym, device=0)
F.upload_matrix(Sp, Si, Sx)
F.factor_device()
b = M.rhs(sym["n"])
for _ in range(4):
    F.solve(b)
tr = [F.getd("solve_trace%d" % k) for k in range(32)]
fw = ["descriptor+y load", "L loop", "x update", "fence+sync", "counter", "x_blk load", "diag solve", "y store"]
bw = ["descriptor+x load", "L passes+atomics", "fence+sync", "counter", "y_blk load", "diag solve"]
out = dict(workload=name, solve_ms=F.getd("ms_solve"),
           forward={k: round(tr[i] / max(tr[15], 1), 1) for i, k in enumerate(fw)}, forward_launches=tr[15],
           backward={k: round(tr[16 + i] / max(tr[31], 1), 1) for i, k in enumerate(bw)}, backward_launches=tr[31])
out["forward_total_cycles"] = round(sum(out["forward"].values()), 1)
out["backward_total_cycles"] = round(sum(out["backward"].values()), 1)
print(json.dumps(out))
