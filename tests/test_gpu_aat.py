"""GPU parity tests of the unsymmetric case of cholmod_super_numeric (A->stype == 0: A*F + beta*I is
factorized, F = A^H) through b200_super_numeric_aat, against the reference's CPU path on the
same pair (S, F) and against the plain-C restatement."""
import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200, matrices as M
from cases import CaseAAT

pytestmark = pytest.mark.gpu
TOL_L = 1e-10


@pytest.mark.parametrize("size,ordering,cplx,zomplex,single,beta", [
    (40, refd.CHOLMOD_AMD, False, False, False, 0.0),
    (64, refd.CHOLMOD_METIS, False, False, False, 0.5),
    (40, refd.CHOLMOD_METIS, True, False, False, 0.0),
    (24, refd.CHOLMOD_AMD, True, True, False, 0.25),       # zomplex A and F, complex L
    (40, refd.CHOLMOD_AMD, False, False, True, 0.0),
    (24, refd.CHOLMOD_AMD, True, False, True, 0.0),
])
def test_aat_factor_matches_reference(size, ordering, cplx, zomplex, single, beta):
    c = CaseAAT(size, ordering, single=single, cplx=cplx, zomplex=zomplex, beta=beta)
    assert c.status == refd.CHOLMOD_OK
    F = b200.Factor(c.sym, dtype=c.Lref.dtype, device=0)
    st, Lx = F.super_numeric_aat(c.ncol, c.Sp, c.Si, c.Sx, c.Fp, c.Fi, c.Fx, beta=beta, zomplex=zomplex)
    assert st == b200.OK and F.minor == c.n and Lx.dtype == c.Lref.dtype
    assert c.rel_err(Lx) <= (1e-4 if single else TOL_L)
    assert np.count_nonzero((Lx == 0) != (c.Lref == 0)) < 1e-6 * Lx.size + 1
    # device solve: L L^H x = P b  <=>  (A A^H + beta I) x = b
    b = M.rhs(c.n) * ((1 + 0.5j) if cplx else 1.0)
    perm = c.sym["Perm"]
    x = np.empty_like(b); x[perm] = F.solve(b[perm])
    import scipy.sparse as sp
    A = sp.csc_matrix((c.Ax, c.Ai, c.Ap), shape=(c.nrow, c.ncol))
    r = A @ (A.conj().T @ x) + beta * x - b
    assert np.max(np.abs(r)) <= (1e-4 if single else 1e-11) * np.max(np.abs(b)) * (1 + np.max(np.abs(x)))
    # the device-resident three-step form gives the same factor
    if not zomplex:
        F.upload_matrix_aat(c.ncol, c.Sp, c.Si, c.Sx, c.Fp, c.Fi, c.Fx)
        assert F.factor_device(beta=beta) == b200.OK
        assert c.rel_err(F.download_factor()) <= (1e-4 if single else TOL_L)
    # and a symmetric upload afterwards switches the context back to the stype < 0 scatter
    F.close()


def test_aat_not_posdef_and_port():
    c = CaseAAT(20, refd.CHOLMOD_AMD, beta=-0.3, factor=False)
    assert c.ref.super_numeric(-0.3) == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric_aat(c.ncol, c.Sp, c.Si, c.Sx, c.Fp, c.Fi, c.Fx, beta=-0.3)
    assert st == b200.NOT_POSDEF and F.minor == minor_ref
    assert np.array_equal(Lx == 0, Lref == 0)
    assert np.max(np.abs(Lx - Lref)) <= TOL_L * max(1.0, np.max(np.abs(Lref)))
    # the same context then factorizes the positive definite A*A' + 0.5 I, and agrees with the port
    st, Lx = F.super_numeric_aat(c.ncol, c.Sp, c.Si, c.Sx, c.Fp, c.Fi, c.Fx, beta=0.5)
    stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx, beta=0.5, F=(c.Fp, c.Fi, c.Fx))
    assert st == stp == 0 and np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.close()


def test_aat_unpacked_operands():
    """A and F unpacked (A->nz / F->nz non-NULL, slack after every column), as the seam allows (SNW:437-506)."""
    c = CaseAAT(16, refd.CHOLMOD_AMD, beta=0.25)

    def unpack(p, i, x, ncols):
        cnt = np.diff(p)
        up = np.zeros(ncols + 1, dtype=np.int64)
        up[1:] = np.cumsum(cnt + 2)
        ui = np.zeros(up[-1], dtype=np.int64)
        ux = np.full(up[-1], 777.0)                 # slack entries must never be read
        for k in range(ncols):
            ui[up[k]:up[k] + cnt[k]] = i[p[k]:p[k + 1]]
            ux[up[k]:up[k] + cnt[k]] = x[p[k]:p[k + 1]]
        return up, ui, ux, cnt.astype(np.int64)

    Sp, Si, Sx, Snz = unpack(c.Sp, c.Si, c.Sx, c.ncol)
    Fp, Fi, Fx, Fnz = unpack(c.Fp, c.Fi, c.Fx, c.n)
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric_aat(c.ncol, Sp, Si, Sx, Fp, Fi, Fx, beta=0.25, Anz=Snz, Fnz=Fnz)
    assert st == b200.OK and c.rel_err(Lx) <= TOL_L
    F.close()
