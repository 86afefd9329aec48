"""CPU tests of the oracle: the plain-C restatement (oracle/super_numeric_port.c) against the reference
itself (oracle/_ref/libcholmod_ref.so) and against the committed golden fixtures."""
import glob
import os

import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import matrices as M
from cases import Case, CaseAAT, get_case

HERE = os.path.dirname(os.path.abspath(__file__))
needs_ref = pytest.mark.skipif(not refd.have_ref(), reason="oracle/_ref/libref_driver.so not built")


@needs_ref
@pytest.mark.parametrize("kind,size,ordering", [("lap2d5", 100, refd.CHOLMOD_AMD), ("lap3d7", 14, refd.CHOLMOD_METIS),
                                                ("lap2d9", 40, refd.CHOLMOD_METIS), ("el27", 6, refd.CHOLMOD_METIS),
                                                ("lap2d5", 17, refd.CHOLMOD_NATURAL)])
def test_port_matches_reference(kind, size, ordering):
    c = get_case(kind, size, ordering)
    assert c.status == refd.CHOLMOD_OK and c.minor == c.n
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx)
    assert st == 0 and minor == c.n
    assert c.rel_err(Lp) <= 1e-13
    assert c.residual_with(Lp) <= 1e-12


@needs_ref
def test_config1_statistics():
    """the reference analysis run here reproduces the surveyed numbers."""
    c = get_case("lap2d5", 100, refd.CHOLMOD_AMD)
    assert c.sym["nsuper"] == 1282 and c.sym["xsize"] == 403959 and c.sym["ssize"] == 27995
    assert abs(c.sym["fl"] - 1.20883e7) / 1.20883e7 < 1e-4
    assert int(c.sym["lnz"]) == 206332


@needs_ref
def test_port_single_precision():
    c = Case("lap3d7", 10, refd.CHOLMOD_METIS, single=True)
    assert c.Lref.dtype == np.float32 and c.Sx.dtype == np.float32
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx)
    assert st == 0 and Lp.dtype == np.float32
    assert c.rel_err(Lp) <= 1e-5


@needs_ref
@pytest.mark.parametrize("zomplex", [False, True])
@pytest.mark.parametrize("single", [False, True])
def test_port_hermitian(zomplex, single):
    """cd_/zd_/cs_/zs_cholmod_super_numeric_worker: Hermitian A stored complex or zomplex, L complex."""
    c = Case("lap3d7", 10, refd.CHOLMOD_METIS, hermitian=True, zomplex=zomplex, single=single)
    want = np.complex64 if single else np.complex128
    assert c.status == refd.CHOLMOD_OK and c.Lref.dtype == want and c.Sx.dtype == want
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx)
    assert st == 0 and minor == c.n and Lp.dtype == want
    assert c.rel_err(Lp) <= (1e-5 if single else 1e-13)
    if not single:
        assert c.residual_with(Lp) <= 1e-12
        b = M.rhs(c.n) * (1 + 0.5j)
        x = refd.port_solve(c.sym, Lp, b[c.sym["Perm"]])
        xx = np.empty(c.n, dtype=np.complex128); xx[c.sym["Perm"]] = x
        xr = c.ref.solve(b)
        assert np.max(np.abs(xx - xr)) <= 1e-12 * np.max(np.abs(xr))


@needs_ref
@pytest.mark.parametrize("cplx,zomplex,single,beta", [(False, False, False, 0.0), (False, False, True, 0.25),
                                                      (True, False, False, 0.5), (True, True, False, 0.0)])
def test_port_unsymmetric_aat(cplx, zomplex, single, beta):
    """A->stype == 0: S*F + beta*I is factorized, F = S^H given explicitly (SNW:471-506)."""
    c = CaseAAT(12, refd.CHOLMOD_AMD, single=single, cplx=cplx, zomplex=zomplex, beta=beta)
    assert c.status == refd.CHOLMOD_OK and c.minor == c.n
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, beta=beta, F=(c.Fp, c.Fi, c.Fx))
    assert st == 0 and minor == c.n and Lp.dtype == c.Lref.dtype
    assert c.rel_err(Lp) <= (1e-5 if single else 1e-13)
    if not single:
        # L L^H really is P (A A^H + beta I) P': solve through the port and check against the dense product
        b = M.rhs(c.n) * ((1 + 0.5j) if cplx else 1.0)
        perm = c.sym["Perm"]
        x = refd.port_solve(c.sym, Lp, b[perm])
        xx = np.empty_like(x); xx[perm] = x
        assert np.max(np.abs(c.dense_product() @ xx - b)) <= 1e-11 * np.max(np.abs(b))


@needs_ref
def test_port_unsymmetric_not_posdef():
    c = CaseAAT(10, refd.CHOLMOD_AMD, beta=-0.3, factor=False)      # A A' + beta I is indefinite
    c.ref.super_numeric(-0.3)
    assert c.ref.geti("status") == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, beta=-0.3, F=(c.Fp, c.Fi, c.Fx))
    assert st == 1 and minor == minor_ref
    assert np.array_equal(Lp == 0, Lref == 0)
    assert np.max(np.abs(Lp - Lref)) <= 1e-12 * max(1.0, np.max(np.abs(Lref)))


@needs_ref
@pytest.mark.parametrize("col,quick", [(37, False), (37, True), (0, False)])
def test_port_hermitian_not_posdef(col, quick):
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(col, -1.0), factor=False, hermitian=True)
    c.ref.seti("quick_return_if_not_posdef", int(quick))
    c.ref.super_numeric(0.0)
    assert c.ref.geti("status") == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, quick_return=quick)
    assert st == 1 and minor == minor_ref
    assert np.array_equal(Lp == 0, Lref == 0)
    assert np.max(np.abs(Lp - Lref)) <= 1e-13 * max(1.0, np.max(np.abs(Lref)))


@needs_ref
@pytest.mark.parametrize("col,quick", [(0, False), (37, False), (37, True), (399, False), (250, False)])
def test_port_not_posdef_semantics(col, quick):
    """L->minor, zero fill and the repeated supernode of SNW:1084-1146."""
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(col, -1.0), factor=False)
    c.ref.seti("quick_return_if_not_posdef", int(quick))
    c.ref.super_numeric(0.0)
    assert c.ref.geti("status") == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, quick_return=quick)
    assert st == 1 and minor == minor_ref
    assert np.array_equal(Lp == 0, Lref == 0)
    assert np.max(np.abs(Lp - Lref)) <= 1e-13 * max(1.0, np.max(np.abs(Lref)))


@needs_ref
def test_port_beta_and_solve():
    c = Case("lap2d5", 25, refd.CHOLMOD_AMD, beta=0.5)
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, beta=0.5)
    assert st == 0 and c.rel_err(Lp) <= 1e-13
    # port solve agrees with the reference solve on (A + beta I) x = b
    b = M.rhs(c.n)
    x = refd.port_solve(c.sym, Lp, b[c.sym["Perm"]])
    xx = np.empty(c.n); xx[c.sym["Perm"]] = x
    xr = c.ref.solve(b)
    assert np.max(np.abs(xx - xr)) <= 1e-12 * np.max(np.abs(xr))


def test_golden_fixtures():
    """The port reproduces the factors the reference produced when the fixtures were generated
    (tests/golden/make_golden.py); runs without the reference library."""
    files = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))
    assert files, "no golden fixtures"
    for f in files:
        g = np.load(f)
        sym = dict(n=int(g["n"]), nsuper=int(g["nsuper"]), xsize=int(g["xsize"]), super=g["super"], pi=g["pi"],
                   px=g["px"], s=g["s"])
        F = (g["Fp"], g["Fi"], g["Fx"]) if "Fp" in g.files else None     # the unsymmetric fixtures carry F = S^H
        st, minor, Lp = refd.port_factor(sym, g["Sp"], g["Si"], g["Sx"], beta=float(g["beta"]), F=F)
        assert st == int(g["status"]) and minor == int(g["minor"]), f
        Lref = g["Lx"]
        assert np.max(np.abs(Lp - Lref)) <= 1e-13 * max(1.0, np.max(np.abs(Lref))), f
