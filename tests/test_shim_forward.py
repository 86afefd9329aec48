"""CPU test of the shim's pass-through: with libcholmod_b200_shim.so preloaded but the GPU switch off
(Common->useGPU == 0 and no CHOLMOD_USE_GPU), the interposed cholmod_l_super_numeric, cholmod_l_free_factor and
cholmod_l_change_factor must hand every call to the reference's own definitions, bit for bit.  Needs the
reference build (oracle/_ref) and the shim, both only present where /root/reference was at build time."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "suitesparse_b200", "lib", "libcholmod_b200_shim.so")

CHILD = r"""
import json, sys, hashlib, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refd
from suitesparse_b200 import matrices as M
n, Ap, Ai, Ax = M.lap2d9(24)
out = []
for rep in range(2):                 # the second Ref frees the first one's factor: cholmod_l_free_factor
    r = refd.Ref()
    r.set_matrix(n, Ap, Ai, Ax, -1)
    r.analyze(refd.CHOLMOD_AMD, refd.CHOLMOD_SUPERNODAL)
    st = r.factorize(0.0)
    st2 = r.factorize(0.0)           # refactorization into the same L->x
    b = M.rhs(n)
    x = r.solve(b)
    out.append(dict(status=[st, st2], minor=r.geti("minor"), resid=r.resid(x, b),
                    sha=hashlib.sha256(r.Lx().tobytes()).hexdigest()))
    r.close()
print("RESULT " + json.dumps(out))
"""


def _run(preload):
    env = dict(os.environ)
    env.pop("CHOLMOD_USE_GPU", None)
    if preload:
        env["LD_PRELOAD"] = SHIM
    r = subprocess.run([sys.executable, "-c", CHILD % dict(root=ROOT)], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_shim_forwards_when_not_activated():
    from oracle import refd
    if not os.path.exists(SHIM) or not refd.have_ref():
        pytest.skip("needs libcholmod_b200_shim.so and oracle/_ref (built where the reference tree is present)")
    plain = _run(False)
    shimmed = _run(True)
    assert plain == shimmed
    assert plain[0]["status"] == [0, 0] and plain[0]["resid"] <= 1e-12
