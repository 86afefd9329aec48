"""GPU test of the zero-patch drop-in: the UNMODIFIED reference driver cholmod_l_factorize_p, with
libcholmod_b200_shim.so preloaded, lands in the B200 numeric phase; its cholmod_l_solve then runs on the
factor the GPU wrote into host L->x.  Runs in a child process because LD_PRELOAD must be set at start-up."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "suitesparse_b200", "lib", "libcholmod_b200_shim.so")

CHILD = r"""
import json, sys, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refd
from suitesparse_b200 import matrices as M
out = {}
import os
for use_gpu in (0, 1):
    os.environ["CHOLMOD_USE_GPU"] = str(use_gpu)        # the reference's own switch, read by the shim
    n, Ap, Ai, Ax = M.lap3d7(18)
    r = refd.Ref()
    r.seti("useGPU", use_gpu)
    r.set_matrix(n, Ap, Ai, Ax, -1)
    r.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    st = r.factorize(0.0)                       # cholmod_l_factorize_p -> cholmod_l_super_numeric (interposed)
    b = M.rhs(n)
    x = r.solve(b)                              # the reference's own solve on host L->x
    out[use_gpu] = dict(status=st, minor=r.geti("minor"), resid=r.resid(x, b), Lx=r.Lx().copy())
    # not positive definite through the same door
    Ax2 = Ax.copy(); Ax2[Ap[n // 2]] = -1.0
    r2 = refd.Ref(); r2.seti("useGPU", use_gpu); r2.set_matrix(n, Ap, Ai, Ax2, -1)
    r2.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    out[use_gpu]["np_status"] = r2.factorize(0.0)
    out[use_gpu]["np_minor"] = r2.geti("minor")
    out[use_gpu]["np_Lx"] = r2.Lx().copy()
    # Hermitian, stored zomplex: zd_cholmod_super_numeric_worker's seat (complex L from a zomplex A)
    nz_, Zp, Zi, Zx = M.hermitian(*M.lap3d7(14))
    r3 = refd.Ref(); r3.seti("useGPU", use_gpu); r3.set_matrix(nz_, Zp, Zi, Zx, -1, zomplex=True)
    r3.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    out[use_gpu]["z_status"] = r3.factorize(0.0)
    bz = M.rhs(nz_) * (1 + 0.5j)
    out[use_gpu]["z_resid"] = r3.resid(r3.solve(bz), bz)
    out[use_gpu]["z_Lx"] = r3.Lx().copy()
    # unsymmetric A (stype 0): cholmod_l_factorize_p builds F = A(p,:)' itself and factorizes A*A' + beta*I
    nr, nc, Rp, Ri, Rx = M.aat2d(30)
    r4 = refd.Ref(); r4.seti("useGPU", use_gpu); r4.set_matrix_rect(nr, nc, Rp, Ri, Rx)
    r4.analyze(refd.CHOLMOD_AMD, refd.CHOLMOD_SUPERNODAL)
    out[use_gpu]["u_status"] = r4.factorize(0.5)
    out[use_gpu]["u_Lx"] = r4.Lx().copy()
    xu = r4.solve(M.rhs(nr))
    import scipy.sparse as sp
    Au = sp.csc_matrix((Rx, Ri, Rp), shape=(nr, nc))
    out[use_gpu]["u_resid"] = float(np.max(np.abs(Au @ (Au.T @ xu) + 0.5 * xu - M.rhs(nr))))
import ctypes
shim = ctypes.CDLL(%(shim)r)
res = dict(
    status=[out[0]["status"], out[1]["status"]], minor=[out[0]["minor"], out[1]["minor"]],
    resid=[out[0]["resid"], out[1]["resid"]],
    err=float(np.max(np.abs(out[0]["Lx"] - out[1]["Lx"])) / np.max(np.abs(out[0]["Lx"]))),
    identical=bool(np.array_equal(out[0]["Lx"], out[1]["Lx"])),
    np_status=[out[0]["np_status"], out[1]["np_status"]], np_minor=[out[0]["np_minor"], out[1]["np_minor"]],
    np_err=float(np.max(np.abs(out[0]["np_Lx"] - out[1]["np_Lx"]))),
    np_zero_same=bool(np.array_equal(out[0]["np_Lx"] == 0, out[1]["np_Lx"] == 0)),
    z_status=[out[0]["z_status"], out[1]["z_status"]], z_resid=[out[0]["z_resid"], out[1]["z_resid"]],
    z_err=float(np.max(np.abs(out[0]["z_Lx"] - out[1]["z_Lx"])) / np.max(np.abs(out[0]["z_Lx"]))),
    z_identical=bool(np.array_equal(out[0]["z_Lx"], out[1]["z_Lx"])), z_dtype=str(out[1]["z_Lx"].dtype),
    u_status=[out[0]["u_status"], out[1]["u_status"]], u_resid=[out[0]["u_resid"], out[1]["u_resid"]],
    u_err=float(np.max(np.abs(out[0]["u_Lx"] - out[1]["u_Lx"])) / np.max(np.abs(out[0]["u_Lx"]))),
    u_identical=bool(np.array_equal(out[0]["u_Lx"], out[1]["u_Lx"])),
    dealloc=int(shim.cholmod_l_b200_deallocate(None)))
print("RESULT " + json.dumps(res))
"""


def test_unmodified_cholmod_factorize_runs_on_the_gpu():
    if not os.path.exists(SHIM):
        pytest.skip("libcholmod_b200_shim.so not built (needs the reference headers at build time)")
    env = dict(os.environ, LD_PRELOAD=SHIM)
    r = subprocess.run([sys.executable, "-c", CHILD % dict(root=ROOT, shim=SHIM)], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")][-1]
    res = json.loads(line[len("RESULT "):])
    assert res["status"] == [0, 0] and res["minor"][0] == res["minor"][1]
    assert res["resid"][1] <= 1e-12
    assert res["err"] <= 1e-10
    # useGPU = 1 really took another path than the CPU run (atomics reorder the sums)
    assert not res["identical"]
    assert res["np_status"] == [1, 1] and res["np_minor"][0] == res["np_minor"][1]
    assert res["np_zero_same"] and res["np_err"] <= 1e-10 * 10
    # the Hermitian (zomplex in, complex out) factorization went through the same door
    assert res["z_status"] == [0, 0] and res["z_dtype"] == "complex128"
    assert res["z_err"] <= 1e-10 and res["z_resid"][1] <= 1e-12 and not res["z_identical"]
    # and so did the unsymmetric A*A' factorization
    assert res["u_status"] == [0, 0] and res["u_err"] <= 1e-10 and res["u_resid"][1] <= 1e-10
    assert not res["u_identical"]
    assert res["dealloc"] == 1
