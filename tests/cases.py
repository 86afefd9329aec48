"""Shared helpers of the tests: build a reference session for a synthetic matrix, compare factors."""
import functools

import numpy as np

from oracle import refd
from suitesparse_b200 import matrices as M

GEN = {"lap2d5": M.lap2d5, "lap2d9": M.lap2d9, "lap3d7": M.lap3d7, "el27": M.el27}


def make_matrix(kind, size, shift_diag=None, hermitian=False):
    n, Ap, Ai, Ax = GEN[kind](size)
    if hermitian:
        n, Ap, Ai, Ax = M.hermitian(n, Ap, Ai, Ax)
    if shift_diag is not None:
        # make the matrix indefinite: overwrite one diagonal entry (column k's first stored entry is the diagonal)
        k, val = shift_diag
        assert Ai[Ap[k]] == k
        Ax = Ax.copy()
        Ax[Ap[k]] = val
    return n, Ap, Ai, Ax


def _need_reference():
    """The analysis (ordering, supernodes) is the reference's host code: without oracle/_ref there is no case."""
    if not refd.have_ref():
        import pytest
        pytest.skip("oracle/_ref (the compiled reference) is not here; the golden-fixture tests still run")


class Case:
    """Reference analysis + seam matrix + reference CPU factor for one matrix."""

    def __init__(self, kind, size, ordering, single=False, shift_diag=None, beta=0.0, factor=True,
                 hermitian=False, zomplex=False):
        _need_reference()
        self.kind, self.size, self.single, self.beta = kind, size, single, beta
        self.n, self.Ap, self.Ai, self.Ax = make_matrix(kind, size, shift_diag, hermitian)
        self.ref = refd.Ref()
        self.ref.set_matrix(self.n, self.Ap, self.Ai, self.Ax, -1, single=single, zomplex=zomplex)
        self.ref.analyze(ordering, refd.CHOLMOD_SUPERNODAL)
        self.sym = self.ref.symbolic()
        self.Sp, self.Si, self.Sx = self.ref.seam()
        if factor:
            self.status = self.ref.super_numeric(beta)
            self.Lref = self.ref.Lx().copy()
            self.minor = self.ref.geti("minor")

    def rel_err(self, Lx):
        """max |dL| / max |L| over the supernodal storage."""
        wide = np.complex128 if np.iscomplexobj(self.Lref) else np.float64
        Lx = np.asarray(Lx, dtype=wide)
        Lr = self.Lref.astype(wide)
        scale = np.max(np.abs(Lr))
        return float(np.max(np.abs(Lx - Lr)) / (scale if scale > 0 else 1.0))

    def residual_with(self, Lx):
        """Push factor values through the reference's own cholmod_l_solve; the demo's scaled residual."""
        host = self.ref.numeric_Lx()
        host[:] = Lx
        b = M.rhs(self.n)
        x = self.ref.solve(b)
        return self.ref.resid(x, b)


class CaseAAT:
    """The unsymmetric case: reference analysis of A*A^H, the seam pair S = A(p,:), F = S^H, and the
    reference CPU factor of S*F + beta*I."""

    def __init__(self, size, ordering, single=False, cplx=False, zomplex=False, beta=0.0, factor=True):
        _need_reference()
        self.nrow, self.ncol, self.Ap, self.Ai, self.Ax = M.aat2d(size, complex_phases=cplx)
        self.n, self.beta = self.nrow, beta
        self.ref = refd.Ref()
        self.ref.set_matrix_rect(self.nrow, self.ncol, self.Ap, self.Ai, self.Ax, single=single, zomplex=zomplex)
        self.ref.analyze(ordering, refd.CHOLMOD_SUPERNODAL)
        self.sym = self.ref.symbolic()
        self.Sp, self.Si, self.Sx = self.ref.seam()
        self.Fp, self.Fi, self.Fx = self.ref.seam_F()
        if factor:
            self.status = self.ref.super_numeric(beta)
            self.Lref = self.ref.Lx().copy()
            self.minor = self.ref.geti("minor")

    rel_err = Case.rel_err

    def dense_product(self):
        """A*A^H + beta*I as a dense array (small cases only)."""
        import scipy.sparse as sp
        A = sp.csc_matrix((self.Ax, self.Ai, self.Ap), shape=(self.nrow, self.ncol))
        return (A @ A.conj().T).toarray() + self.beta * np.eye(self.nrow)


@functools.lru_cache(maxsize=None)
def get_case(kind, size, ordering, single=False, shift_diag=None, beta=0.0, hermitian=False, zomplex=False):
    return Case(kind, size, ordering, single=single, shift_diag=shift_diag, beta=beta, hermitian=hermitian,
                zomplex=zomplex)


class Forest:
    """kron(I_k, A): k copies of one case side by side, as a case of its own (symbolic factor, seam matrix,
    reference factor are the copies' concatenation).  The supernodal etree is then a forest of k equal trees:
    every level holds at least k supernodes."""

    def __init__(self, c, copies):
        n, ns, ss, xs = c.sym["n"], c.sym["nsuper"], int(c.sym["pi"][-1]), c.sym["xsize"]
        nnz = int(c.Sp[-1])
        cat = np.concatenate
        self.n = n * copies
        self.sym = dict(n=n * copies, nsuper=ns * copies, xsize=xs * copies,
                        super=cat([c.sym["super"][:-1] + k * n for k in range(copies)] + [[n * copies]]),
                        pi=cat([c.sym["pi"][:-1] + k * ss for k in range(copies)] + [[ss * copies]]),
                        px=cat([c.sym["px"][:-1] + k * xs for k in range(copies)] + [[xs * copies]]),
                        s=cat([c.sym["s"] + k * n for k in range(copies)]))
        self.Sp = cat([c.Sp[:-1] + k * nnz for k in range(copies)] + [[nnz * copies]])
        self.Si = cat([c.Si + k * n for k in range(copies)])
        self.Sx = np.tile(c.Sx, copies)
        self.Lref = np.tile(c.Lref, copies)
        self.status, self.minor = c.status, c.minor

    rel_err = Case.rel_err
