"""Device time of one factorization with and without the subtree-group fork of the single-GPU plan:
    python profiles/measure_groups.py [workload ...]
CHOLMOD_B200_GROUPS is read when a context is created (1 = every level in lock-step, the old plan)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                              # noqa: E402
from suitesparse_b200 import b200                        # noqa: E402

for name in (sys.argv[1:] or ["cfg2"]):
    kind, size, ordering, desc = bench.WORKLOADS[name]
    ref, _ = bench.analyze(kind, size, ordering)
    sym = ref.symbolic()
    Sp, Si, Sx = ref.seam()
    out = dict(workload=name, fl=sym["fl"])
    Lfirst = None
    for groups in (1, 8, 4, 16):
        os.environ["CHOLMOD_B200_GROUPS"] = str(groups)
        F = b200.Factor(sym, device=0)
        F.upload_matrix(Sp, Si, Sx)
        ts = []
        for _ in range(6):
            assert F.factor_device() == b200.OK
            ts.append(F.getd("ms_factor"))
        Lx = F.download_factor()
        if Lfirst is None:
            Lfirst = Lx
        err = float(np.max(np.abs(Lx - Lfirst)) / np.max(np.abs(Lfirst)))
        out["groups%d" % groups] = dict(ms=round(float(np.median(ts[1:])), 3), ngroups=F.geti("ngroups"),
                                        fork=[F.geti("fork_lo"), F.geti("fork_hi")], nlaunches=F.geti("nlaunches"),
                                        tflops=round(sym["fl"] / np.median(ts[1:]) / 1e9, 2), err_vs_lockstep=err)
        F.close()
    os.environ.pop("CHOLMOD_B200_GROUPS", None)
    print(json.dumps(out))
