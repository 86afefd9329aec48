"""GPU tests of the cholmod_common-level glue beyond the plain factorize call (tests/test_gpu_dropin.py):
the reference's own demo program under LD_PRELOAD, device sweeps behind cholmod_l_solve with L->x left on the
device, the reference's GPU memory settings and cholmod_gpu_stats counters, a NaN on the diagonal
(CHOLMOD/Tcov/t_null2.c:2263-2346), two threads factorizing the same structure, and several GPUs behind
cholmod_l_factorize (CHOLMOD_B200_NGPUS).  Child processes: LD_PRELOAD must be set at start-up."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from suitesparse_b200 import b200, matrices as M

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "suitesparse_b200", "lib", "libcholmod_b200_shim.so")
DEMO = os.path.join(ROOT, "oracle", "_ref", "cholmod_dl_demo")


def run_child(code, extra_env, timeout=900):
    if not os.path.exists(SHIM):
        pytest.skip("libcholmod_b200_shim.so not built (needs the reference headers at build time)")
    env = dict(os.environ, LD_PRELOAD=SHIM, CHOLMOD_USE_GPU="1", **extra_env)
    r = subprocess.run([sys.executable, "-c", code % dict(root=ROOT, shim=SHIM)], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=timeout)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def write_matrix_market(path, n, Ap, Ai, Ax):
    """Lower triangle of a symmetric matrix as a Matrix-Market coordinate file (what cholmod_dl_demo reads)."""
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real symmetric\n%d %d %d\n" % (n, n, len(Ai)))
        for k in range(n):
            for p in range(Ap[k], Ap[k + 1]):
                f.write("%d %d %.17g\n" % (Ai[p] + 1, k + 1, Ax[p]))


@pytest.mark.parametrize("which", ["bcsstk02", "config1", "lap3d7_20"])
def test_reference_demo_program_runs_on_the_gpu(which, tmp_path):
    """The reference's cholmod_dl_demo, unmodified, with the shim preloaded: analysis, factorization, solves,
    refinement, its own residual checks -- "All tests passed"."""
    if not os.path.exists(DEMO):
        pytest.skip("oracle/_ref/cholmod_dl_demo not built")
    path = str(tmp_path / (which + ".mtx"))
    if which == "bcsstk02":
        g = np.load(os.path.join(ROOT, "tests", "golden", "bcsstk02.npz"))
        assert int(g["stype"]) != 0
        # the fixture stores the reference's own ctest matrix; one triangle, either one
        n, Ap, Ai, Ax = int(g["n"]), g["Ap"], g["Ai"], g["Ax"]
        rows = np.asarray(Ai)
        cols = np.repeat(np.arange(n), np.diff(Ap))
        lo, hi = np.maximum(rows, cols), np.minimum(rows, cols)
        with open(path, "w") as f:
            f.write("%%MatrixMarket matrix coordinate real symmetric\n%d %d %d\n" % (n, n, len(rows)))
            for i, j, v in zip(lo, hi, Ax):
                f.write("%d %d %.17g\n" % (i + 1, j + 1, v))
    elif which == "config1":
        write_matrix_market(path, *M.lap2d5(100))
    else:
        write_matrix_market(path, *M.lap3d7(20))
    outs = {}
    for use_gpu in ("0", "1"):
        env = dict(os.environ, LD_PRELOAD=SHIM, CHOLMOD_USE_GPU=use_gpu, CHOLMOD_B200_TRACE="1")
        r = subprocess.run([DEMO, path], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
        assert "All tests passed" in r.stdout, r.stdout[-2000:]
        outs[use_gpu] = r
    # the GPU run really went through the module (the shim says so on stderr when CHOLMOD_B200_TRACE is set)
    assert "cholmod_b200: numeric factorization on device" in outs["1"].stderr
    assert "cholmod_b200: numeric factorization on device" not in outs["0"].stderr


LAZY = r"""
import json, os, sys, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refd
from suitesparse_b200 import matrices as M
n, Ap, Ai, Ax = M.lap3d7(20)
b = M.rhs(n)
os.environ["CHOLMOD_USE_GPU"] = "0"
r0 = refd.Ref(); r0.set_matrix(n, Ap, Ai, Ax, -1); r0.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
assert r0.factorize(0.0) == 0
Lcpu = r0.Lx().copy()
os.environ["CHOLMOD_USE_GPU"] = "1"
r = refd.Ref(); r.seti("useGPU", 1); r.set_matrix(n, Ap, Ai, Ax, -1); r.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
st = r.factorize(0.0)
host_after_factor = r.Lx().copy()           # lazy: the host array has not been written
x = r.solve(b)                              # cholmod_l_solve -> interposed sweeps on the device
resid = r.resid(x, b)
host_after_solve = r.Lx().copy()
rc = r.rcond()                              # a host reader of L->x: now the factor is fetched
host_after_rcond = r.Lx().copy()
scale = np.max(np.abs(Lcpu))
res = dict(status=st, resid=resid, rcond=rc, rcond_cpu=r0.rcond(),
           stale_after_factor=float(np.max(np.abs(host_after_factor - Lcpu)) / scale),
           stale_after_solve=float(np.max(np.abs(host_after_solve - Lcpu)) / scale),
           err_after_rcond=float(np.max(np.abs(host_after_rcond - Lcpu)) / scale),
           stats=dict(syrk=r.geti("gpu_syrk_calls"), gemm=r.geti("gpu_gemm_calls"), potrf=r.geti("gpu_potrf_calls"),
                      trsm=r.geti("gpu_trsm_calls"), t=r.getd("gpu_syrk_time") + r.getd("gpu_gemm_time")
                      + r.getd("gpu_potrf_time") + r.getd("gpu_trsm_time")), nsuper=r.geti("nsuper"))
print("RESULT " + json.dumps(res))
"""


@pytest.mark.parametrize("lazy", [True, False])
def test_solve_on_the_device_behind_cholmod_solve(lazy):
    """cholmod_l_factorize + cholmod_l_solve of an unmodified caller: the two supernodal sweeps run on the device;
    with CHOLMOD_B200_LAZY_LX=1 host L->x is not even written until a host reader (cholmod_l_rcond) needs it."""
    env = {"CHOLMOD_B200_LAZY_LX": "1"} if lazy else {"CHOLMOD_B200_DEVICE_SOLVE": "1"}
    res = run_child(LAZY, env)
    assert res["status"] == 0 and res["resid"] <= 1e-12
    if lazy:
        assert res["stale_after_factor"] > 1e-3 and res["stale_after_solve"] > 1e-3     # never copied ...
    else:
        assert res["stale_after_factor"] <= 1e-10                                        # ... or copied at once
    assert res["err_after_rcond"] <= 1e-10
    assert abs(res["rcond"] - res["rcond_cpu"]) <= 1e-10 * res["rcond_cpu"]
    # cholmod_gpu_stats: the reference's BLAS-call counts, on the GPU side
    s = res["stats"]
    assert s["potrf"] == res["nsuper"] and s["syrk"] >= s["gemm"] > 0 and 0 < s["trsm"] <= res["nsuper"] and s["t"] > 0


MISC = r"""
import json, os, sys, threading, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refd
from suitesparse_b200 import matrices as M
res = {}
# ---- Common->maxGpuMemBytes (cholmod_l_gpu_allocate's knob): a context that cannot fit is refused
n, Ap, Ai, Ax = M.lap3d7(16)
r = refd.Ref(); r.seti("useGPU", 1); r.seti("maxGpuMemBytes", 100000)
r.set_matrix(n, Ap, Ai, Ax, -1); r.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
res["limited_status"] = r.factorize(0.0)
r.seti("maxGpuMemBytes", 0)
res["unlimited_status"] = r.factorize(0.0)
b = M.rhs(n)
res["unlimited_resid"] = r.resid(r.solve(b), b)
# ---- NaN on the diagonal (CHOLMOD/Tcov/t_null2.c:2263-2346): factorize returns TRUE, status OK or NOT_POSDEF,
# rcond == 0
for k in (0, 2):
    Axn = Ax.copy(); Axn[Ap[k]] = np.nan
    rn = refd.Ref(); rn.seti("useGPU", 1); rn.set_matrix(n, Ap, Ai, Axn, -1)
    rn.analyze(refd.CHOLMOD_AMD, refd.CHOLMOD_SUPERNODAL)
    res["nan%%d" %% k] = dict(status=rn.factorize(0.0), rcond=rn.rcond())
# ---- two threads, two cholmod_commons, the same symbolic structure, different values
os.environ["CHOLMOD_USE_GPU"] = "0"
n2, Bp, Bi, Bx = M.lap3d7(22)
want = []
for scale in (1.0, 3.0):
    rc = refd.Ref(); rc.set_matrix(n2, Bp, Bi, Bx * scale, -1); rc.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    rc.factorize(0.0); want.append(rc.Lx().copy())
os.environ["CHOLMOD_USE_GPU"] = "1"
refs = []
for scale in (1.0, 3.0):
    rg = refd.Ref(); rg.seti("useGPU", 1); rg.set_matrix(n2, Bp, Bi, Bx * scale, -1)
    rg.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL); refs.append(rg)
errs = [[], []]
def work(i):
    for rep in range(6):
        st = refs[i].factorize(0.0)
        errs[i].append((st, float(np.max(np.abs(refs[i].Lx() - want[i])) / np.max(np.abs(want[i])))))
th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
[t.start() for t in th]; [t.join() for t in th]
res["threads"] = errs
print("RESULT " + json.dumps(res))
"""


def test_memory_settings_nan_and_threads():
    res = run_child(MISC, {})
    assert res["limited_status"] in (-2 - 100, -2)          # CHOLMOD_OUT_OF_MEMORY (the driver adds -100 when the call failed)
    assert res["unlimited_status"] == 0 and res["unlimited_resid"] <= 1e-12
    for k in ("nan0", "nan2"):
        assert res[k]["status"] in (0, 1) and res[k]["rcond"] == 0
    for per_thread in res["threads"]:
        assert len(per_thread) == 6
        for st, err in per_thread:
            assert st == 0 and err <= 1e-10


NGPUS = r"""
import json, os, sys, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refd
from suitesparse_b200 import matrices as M
res = {}
for kind, size in (("lap3d7", 30), ("lap2d9", 150)):
    n, Ap, Ai, Ax = getattr(M, kind)(size)
    os.environ["CHOLMOD_USE_GPU"] = "0"
    r0 = refd.Ref(); r0.set_matrix(n, Ap, Ai, Ax, -1); r0.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    r0.factorize(0.0); Lcpu = r0.Lx().copy()
    os.environ["CHOLMOD_USE_GPU"] = "1"
    r = refd.Ref(); r.seti("useGPU", 1); r.set_matrix(n, Ap, Ai, Ax, -1); r.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    out = []
    for rep in range(2):
        st = r.factorize(0.0)
        b = M.rhs(n)
        out.append(dict(status=st, err=float(np.max(np.abs(r.Lx() - Lcpu)) / np.max(np.abs(Lcpu))), resid=r.resid(r.solve(b), b)))
    # not positive definite over several GPUs: the truncated re-run
    Ax2 = Ax.copy(); Ax2[Ap[n // 2]] = -1.0
    os.environ["CHOLMOD_USE_GPU"] = "0"
    q0 = refd.Ref(); q0.set_matrix(n, Ap, Ai, Ax2, -1); q0.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    s0 = q0.factorize(0.0)
    os.environ["CHOLMOD_USE_GPU"] = "1"
    q = refd.Ref(); q.seti("useGPU", 1); q.set_matrix(n, Ap, Ai, Ax2, -1); q.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    s1 = q.factorize(0.0)
    bad = dict(status=[s0, s1], minor=[q0.geti("minor"), q.geti("minor")],
               zero_same=bool(np.array_equal(q0.Lx() == 0, q.Lx() == 0)),
               err=float(np.max(np.abs(q0.Lx() - q.Lx())) / max(1.0, np.max(np.abs(q0.Lx())))))
    res[kind] = dict(ok=out, bad=bad)
print("RESULT " + json.dumps(res))
"""


@pytest.mark.parametrize("ngpus", [2, 4, 8])
def test_several_gpus_behind_cholmod_factorize(ngpus):
    """CHOLMOD_B200_NGPUS: an unmodified cholmod_l_factorize caller, ONE process, the factorization spread over
    several devices (subtrees per device, contribution blocks pulled over NVLink, every device copying its
    share of L->x into the caller's one host array)."""
    if b200.device_count() < ngpus:
        pytest.skip("needs %d GPUs" % ngpus)
    res = run_child(NGPUS, {"CHOLMOD_B200_NGPUS": str(ngpus)}, timeout=240)     # normally 15-30 s
    for kind in ("lap3d7", "lap2d9"):
        for ok in res[kind]["ok"]:
            assert ok["status"] == 0 and ok["err"] <= 1e-10 and ok["resid"] <= 1e-12
        bad = res[kind]["bad"]
        assert bad["status"] == [1, 1] and bad["minor"][0] == bad["minor"][1]
        assert bad["zero_same"] and bad["err"] <= 1e-10
