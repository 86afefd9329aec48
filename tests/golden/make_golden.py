"""Generates tests/golden/*.npz by running the reference CHOLMOD (oracle/_ref, built from /root/reference)
in this container.  Each fixture holds the symbolic factor, the seam matrix S = tril(P A P'), and the
numeric factor L->x, L->minor and Common->status the reference CPU supernodal path produced.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import refd                      # noqa: E402
from suitesparse_b200 import matrices as M   # noqa: E402

REF_DEMO = "/root/reference/CHOLMOD/Demo/Matrix"


def dump(name, ref, beta=0.0):
    sym = ref.symbolic()
    Sp, Si, Sx = ref.seam()
    n, Ap, Ai, Ax, stype = ref.matrix()
    extra = {}
    if stype == 0:          # unsymmetric: S = A(p,:), and F = S^H goes to the numeric phase with it
        Fp, Fi, Fx = ref.seam_F()
        extra = dict(Fp=Fp, Fi=Fi, Fx=Fx, ncol=ref.geti("ncol"))
    status = ref.super_numeric(beta)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), n=sym["n"], nsuper=sym["nsuper"], xsize=sym["xsize"], fl=sym["fl"],
        lnz=sym["lnz"], Perm=sym["Perm"], super=sym["super"], pi=sym["pi"], px=sym["px"], s=sym["s"],
        Ap=Ap, Ai=Ai, Ax=Ax, stype=stype, Sp=Sp, Si=Si, Sx=Sx, beta=beta, status=status,
        minor=ref.geti("minor"), Lx=ref.Lx().copy(), **extra)
    print(name, "n", sym["n"], "nsuper", sym["nsuper"], "xsize", sym["xsize"], "status", status,
          "minor", ref.geti("minor"))


def synthetic(name, gen, size, ordering, beta=0.0, neg=None, hermitian=False):
    n, Ap, Ai, Ax = gen(size)
    if hermitian:
        n, Ap, Ai, Ax = M.hermitian(n, Ap, Ai, Ax)
    if neg is not None:
        Ax = Ax.copy(); Ax[Ap[neg]] = -1.0
    r = refd.Ref(); r.set_matrix(n, Ap, Ai, Ax, -1); r.analyze(ordering, refd.CHOLMOD_SUPERNODAL)
    dump(name, r, beta)


def demo(name, fname, beta=0.0):
    r = refd.Ref(); r.read_matrix(os.path.join(REF_DEMO, fname))
    r.analyze(refd.CHOLMOD_AMD, refd.CHOLMOD_SUPERNODAL)
    dump(name, r, beta)


if __name__ == "__main__":
    only = sys.argv[1:]         # optional: names of the fixtures to (re)generate
    if only:
        _syn, _demo = synthetic, demo
        synthetic = lambda name, *a, **k: _syn(name, *a, **k) if name in only else None     # noqa: E731
        demo = lambda name, *a, **k: _demo(name, *a, **k) if name in only else None         # noqa: E731
    synthetic("lap2d5_24_amd", M.lap2d5, 24, refd.CHOLMOD_AMD)
    synthetic("lap3d7_9_metis", M.lap3d7, 9, refd.CHOLMOD_METIS)
    synthetic("lap2d9_20_metis_beta", M.lap2d9, 20, refd.CHOLMOD_METIS, beta=0.25)
    synthetic("el27_4_metis", M.el27, 4, refd.CHOLMOD_METIS)
    synthetic("lap2d5_16_notposdef", M.lap2d5, 16, refd.CHOLMOD_AMD, neg=100)
    # Hermitian (cd_cholmod_super_numeric_worker): complex entries, complex L
    synthetic("herm_lap3d7_8_metis", M.lap3d7, 8, refd.CHOLMOD_METIS, hermitian=True)
    synthetic("herm_lap2d5_14_notposdef", M.lap2d5, 14, refd.CHOLMOD_AMD, neg=90, hermitian=True)
    # the reference's own ctest matrices (CHOLMOD/CMakeLists.txt:835-878); bcsstk02 is its "supernodal" test
    demo("bcsstk01", "bcsstk01.tri")
    demo("bcsstk02", "bcsstk02.tri")
    # the other three ctest matrices; "demo_" fixtures are checked by the CPU oracle test only.  lp_afiro is
    # unsymmetric (A*A' + beta*I with the demo's beta = 1e-6, cholmod_dl_demo.c:150), c.tri is complex Hermitian
    demo("demo_lp_afiro", "lp_afiro.tri", beta=1e-6)
    demo("demo_c", "c.tri")
    demo("demo_can24", "can___24.mtx")
