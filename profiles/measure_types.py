"""Device time of one numeric factorization per entry type (real/complex x double/single) on the same
pattern: profiles/measure_types.py [size]  (3-D 7-point Laplacian size^3, METIS; complex = M.hermitian).
Prints one JSON line per type; complex flops are counted as 4x Common->fl (one complex multiply-add = 4 real)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refd                                  # noqa: E402  (host analysis = the reference's)
from suitesparse_b200 import b200, matrices as M         # noqa: E402

size = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n, Ap, Ai, Ax = M.lap3d7(size)
ref = refd.Ref()
ref.set_matrix(n, Ap, Ai, Ax, -1)
ref.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
sym = ref.symbolic()
Sp, Si, Sx = ref.seam()
_, _, _, Zx = M.hermitian(n, Sp, Si, Sx)                 # phases on the permuted matrix: still Hermitian PD
peak = b200.dmma_peak_tflops(0)
for dtype, vals, mult in [(np.float64, Sx, 1), (np.complex128, Zx, 4), (np.float32, Sx, 1), (np.complex64, Zx, 4)]:
    F = b200.Factor(sym, dtype=dtype, device=0)
    F.upload_matrix(Sp, Si, vals.astype(dtype))
    ts = []
    for it in range(5):
        assert F.factor_device() == b200.OK
        ts.append(F.getd("ms_factor"))
    ms = float(np.median(ts[1:]))
    b = M.rhs(n).astype(np.complex128 if mult == 4 else np.float64)
    x = F.solve(b)
    F.seti("profile", 1)
    F.factor_device()
    kinds = {b200.KIND_NAMES[k]: round(F.getd("ms_kind%d" % k), 3) for k in range(len(b200.KIND_NAMES))
             if F.geti("launches_kind%d" % k)}
    print(json.dumps({"dtype": np.dtype(dtype).name, "n": n, "fl": sym["fl"], "ms": ms,
                      "gflops_real_equiv": mult * sym["fl"] / ms / 1e6, "dmma_peak_tflops": peak,
                      "solve_ms": F.getd("ms_solve"), "x_finite": bool(np.all(np.isfinite(x))), "ms_kind": kinds}))
    F.close()
