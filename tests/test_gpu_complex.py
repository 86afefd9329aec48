"""GPU parity tests of the Hermitian case: complex double / complex single factors against
the reference's cd_/zd_/cs_ workers (oracle/_ref) on the same inputs, through the C ABI
(b200_super_numeric_z / _c, b200_solve_z).  Tolerances as for the real case: max|dL|/max|L| <= 1e-10 and
scaled residual <= 1e-12 in double, 1e-4 / 1e-5 in single."""
import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200, matrices as M
from cases import Case, get_case

pytestmark = pytest.mark.gpu
TOL_L, TOL_RES = 1e-10, 1e-12


def unpermute(sym, xperm):
    x = np.empty_like(xperm)
    x[sym["Perm"]] = xperm
    return x


def diag_positions(sym):
    """Indices into L->x of the diagonal entries of L."""
    out = []
    for s in range(sym["nsuper"]):
        nscol = sym["super"][s + 1] - sym["super"][s]
        nsrow = sym["pi"][s + 1] - sym["pi"][s]
        out.append(sym["px"][s] + np.arange(nscol) * (nsrow + 1))
    return np.concatenate(out) if out else np.zeros(0, dtype=np.int64)


@pytest.mark.parametrize("kind,size,ordering,zomplex", [
    ("lap2d5", 100, refd.CHOLMOD_AMD, False),       # config 1's shape
    ("lap3d7", 16, refd.CHOLMOD_METIS, False),
    ("lap3d7", 16, refd.CHOLMOD_METIS, True),       # CHOLMOD_ZOMPLEX input, complex L
    ("lap3d7", 40, refd.CHOLMOD_METIS, False),      # root ~ 1.6k columns: 128x64 tiles, 64-wide inverse panels
    ("lap2d9", 200, refd.CHOLMOD_METIS, False),
    ("el27", 12, refd.CHOLMOD_METIS, False),
])
def test_hermitian_factor_matches_reference(kind, size, ordering, zomplex):
    c = get_case(kind, size, ordering, hermitian=True, zomplex=zomplex)
    assert c.status == refd.CHOLMOD_OK and c.Lref.dtype == np.complex128
    F = b200.Factor(c.sym, dtype=np.complex128, device=0)
    if zomplex:
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx.real.copy(), Az=c.Sx.imag.copy())
    else:
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK and F.minor == c.n and Lx.dtype == np.complex128
    assert F.geti("kernel_launches") > 0
    assert c.rel_err(Lx) <= TOL_L
    # the diagonal of L is real (zpotrf), the strict upper triangles of the diagonal blocks stay zero
    assert np.all(Lx[diag_positions(c.sym)].imag == 0)
    assert np.count_nonzero((Lx == 0) != (c.Lref == 0)) < 1e-6 * Lx.size + 1
    # the reference's own solve on our factor, and our device solve with a complex right-hand side
    assert c.residual_with(Lx) <= TOL_RES
    b = M.rhs(c.n) * (1 + 0.5j)
    x = unpermute(c.sym, F.solve(b[c.sym["Perm"]]))
    assert c.ref.resid(x, b) <= TOL_RES
    F.close()


def test_hermitian_golden_and_port():
    c = get_case("lap3d7", 16, refd.CHOLMOD_METIS, hermitian=True)
    stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx)
    F = b200.Factor(c.sym, dtype=np.complex128, device=0)
    _, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert stp == 0 and np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.close()


def test_hermitian_single_precision():
    c = Case("lap3d7", 24, refd.CHOLMOD_METIS, hermitian=True, single=True)
    assert c.Lref.dtype == np.complex64
    F = b200.Factor(c.sym, dtype=np.complex64, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == 0 and Lx.dtype == np.complex64
    assert c.rel_err(Lx) <= 1e-4
    b = M.rhs(c.n) * (1 - 0.25j)
    x = unpermute(c.sym, F.solve(b[c.sym["Perm"]]))
    assert c.ref.resid(x, b) <= 1e-5
    F.close()


@pytest.mark.parametrize("col,quick", [(0, False), (37, False), (37, True), (250, False), (399, False)])
def test_hermitian_not_posdef(col, quick):
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(col, -1.0), factor=False, hermitian=True)
    c.ref.seti("quick_return_if_not_posdef", int(quick))
    assert c.ref.super_numeric(0.0) == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    F = b200.Factor(c.sym, dtype=np.complex128, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx, quick_return=quick)
    assert st == b200.NOT_POSDEF and F.minor == minor_ref
    assert np.array_equal(Lx == 0, Lref == 0)
    assert np.max(np.abs(Lx - Lref)) <= TOL_L * max(1.0, np.max(np.abs(Lref)))
    F.close()


def test_hermitian_beta_refactorization_and_multi_rhs():
    c = get_case("lap2d9", 60, refd.CHOLMOD_METIS, hermitian=True)
    F = b200.Factor(c.sym, dtype=np.complex128, device=0)
    for scale, beta in [(1.0, 0.0), (2.5, 0.0), (1.0, 0.75)]:
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx * scale, beta=beta)
        stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx * scale, beta=beta)
        assert st == stp == 0
        assert np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.upload_matrix(c.Sp, c.Si, c.Sx)
    F.upload_values(c.Sx * 3.0)
    assert F.factor_device() == b200.OK
    stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx * 3.0)
    assert np.max(np.abs(F.download_factor() - Lp)) <= TOL_L * np.max(np.abs(Lp))
    # multi-RHS device solve against the plain-C restatement's solve
    rng = np.random.default_rng(5)
    B = np.asfortranarray(rng.standard_normal((c.n, 3)) + 1j * rng.standard_normal((c.n, 3)))
    X = F.solve(B)
    for j in range(3):
        xp = refd.port_solve(c.sym, Lp, B[:, j])
        assert np.max(np.abs(X[:, j] - xp)) <= 1e-11 * np.max(np.abs(xp))
    F.close()


def test_hermitian_dense_supernode():
    """One dense Hermitian supernode wider than two outer blocks, against numpy's Cholesky."""
    n = 600
    rng = np.random.default_rng(11)
    B = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
    A = B @ B.conj().T / n + 2.0 * np.eye(n)
    sym = dict(n=n, nsuper=1, super=np.array([0, n]), pi=np.array([0, n]), px=np.array([0, n * n]), s=np.arange(n))
    Ap = np.zeros(n + 1, dtype=np.int64); Ap[1:] = np.cumsum(n - np.arange(n))
    Ai = np.concatenate([np.arange(k, n) for k in range(n)])
    Ax = np.concatenate([A[k:, k] for k in range(n)])
    F = b200.Factor(sym, dtype=np.complex128, device=0)
    st, Lx = F.super_numeric(Ap, Ai, Ax)
    L = np.tril(Lx.reshape(n, n).T)
    assert st == b200.OK and np.max(np.abs(L - np.linalg.cholesky(A))) <= TOL_L * np.max(np.abs(L))
    assert np.all(np.triu(Lx.reshape(n, n).T, 1) == 0)
    x = F.solve(A @ np.ones(n, dtype=np.complex128))
    assert np.max(np.abs(x - 1.0)) <= 1e-10
    F.close()


def test_hermitian_unpacked_and_diagonal_imaginary_part_ignored():
    """Unpacked, unsorted columns with a strictly-upper extra entry, and garbage in the imaginary part of the
    diagonal of A: zpotrf (and so the reference) only ever reads the real part."""
    c = get_case("lap2d5", 30, refd.CHOLMOD_AMD, hermitian=True)
    n, Sp, Si, Sx = c.n, c.Sp, c.Si, c.Sx.copy()
    Sx[Sp[:-1]] += 0.375j                                   # first entry of a sorted lower column = the diagonal
    assert np.all(Si[Sp[:-1]] == np.arange(n))
    rng = np.random.default_rng(9)
    cnt = np.diff(Sp)
    Up = np.zeros(n + 1, dtype=np.int64)
    Up[1:] = np.cumsum(cnt + 3)
    Ui = np.zeros(Up[-1], dtype=np.int64)
    Ux = np.full(Up[-1], 99.0 + 1j)
    Unz = np.zeros(n, dtype=np.int64)
    for k in range(n):
        rows, vals = Si[Sp[k]:Sp[k + 1]], Sx[Sp[k]:Sp[k + 1]]
        if k > 0:
            rows, vals = np.append(rows, k - 1), np.append(vals, 5.0 - 2j)      # strictly upper: ignored
        perm = rng.permutation(len(rows))
        Ui[Up[k]:Up[k] + len(rows)] = rows[perm]
        Ux[Up[k]:Up[k] + len(rows)] = vals[perm]
        Unz[k] = len(rows)
    F = b200.Factor(c.sym, dtype=np.complex128, device=0)
    st, Lx = F.super_numeric(Up, Ui, Ux, Anz=Unz)
    assert st == b200.OK and c.rel_err(Lx) <= TOL_L
    F.close()
