"""CPU tests of the device solve's schedule: the tables the solve kernels run from (b200_plan_export "solve_*"),
replayed by tests/solve_interp.py on the reference factor, give the solution of L y = b, L' x = y and of both; both
schedules of the wide supernodes (256 and 64 columns per launch)."""
import numpy as np
import pytest
import scipy.sparse as sp
from scipy.sparse.linalg import spsolve_triangular

from oracle import refd
from suitesparse_b200 import b200
from cases import get_case
import solve_interp

needs_ref = pytest.mark.skipif(not refd.have_ref(), reason="oracle/_ref/libref_driver.so not built")


def sparse_L(sym, Lx):
    sup, pi, px, Ls = sym["super"], sym["pi"], sym["px"], sym["s"]
    rows, cols, vals = [], [], []
    for s_ in range(sym["nsuper"]):
        k1, k2 = sup[s_], sup[s_ + 1]
        nsrow = pi[s_ + 1] - pi[s_]
        blk = Lx[px[s_]:px[s_] + nsrow * (k2 - k1)].reshape(k2 - k1, nsrow).T
        r = Ls[pi[s_]:pi[s_ + 1]]
        for j in range(k2 - k1):
            rows.append(r[j:]); cols.append(np.full(nsrow - j, k1 + j)); vals.append(blk[j:, j])
    n = sym["n"]
    return sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n))


def check(sym, Lx, cols, monkeypatch):
    monkeypatch.setenv("CHOLMOD_B200_SOLVE_COLS", str(cols))
    F = b200.Factor(sym, device=-1)
    S = solve_interp.Schedule(F)
    kinds = set(int(k) for k in S.launches[:, 0])
    assert kinds <= ({0, 2} if cols == 64 else {0, 3, 4, 5})
    L = sparse_L(sym, Lx)
    rng = np.random.default_rng(5)
    B = rng.standard_normal((sym["n"], 2))
    Yref = spsolve_triangular(L, B, lower=True)
    Xref = spsolve_triangular(L.T.tocsr(), Yref, lower=False)
    scale = max(1.0, float(np.max(np.abs(Xref))))
    Y = solve_interp.solve(sym, Lx, S, B, which=1)
    assert np.max(np.abs(Y - Yref)) <= 1e-10 * max(1.0, float(np.max(np.abs(Yref))))
    X2 = solve_interp.solve(sym, Lx, S, Y, which=2)
    assert np.max(np.abs(X2 - Xref)) <= 1e-10 * scale
    X = solve_interp.solve(sym, Lx, S, B, which=3)
    assert np.max(np.abs(X - Xref)) <= 1e-10 * scale
    F.close()
    return S


@needs_ref
@pytest.mark.parametrize("cols", [256, 64])
@pytest.mark.parametrize("kind,size,ordering", [
    ("lap3d7", 22, refd.CHOLMOD_METIS),         # root separator of 484 columns: supersteps of 256 + 228
    ("lap3d7", 33, refd.CHOLMOD_METIS),         # 1089 = 4 x 256 + 65: a 65-column last superstep, a 1-column last block
    ("lap2d9", 90, refd.CHOLMOD_METIS),         # wide supernodes with many rows below them on several levels
    ("el27", 7, refd.CHOLMOD_METIS),
    ("lap2d5", 40, refd.CHOLMOD_AMD)])           # narrow supernodes only
def test_solve_schedule_replayed_on_the_host(kind, size, ordering, cols, monkeypatch):
    c = get_case(kind, size, ordering)
    S = check(c.sym, c.Lref, cols, monkeypatch)
    widest = int(np.max(np.diff(c.sym["super"])))
    assert (len(S.blks) > 0) == (widest > 64)
    if cols == 256 and widest > 64:
        # a launch per superstep and sweep (plus one for the first diagonal solves of a level), not one per 64 columns
        n4 = int(np.sum(S.launches[:, 0] == 4))
        assert n4 == int(np.sum(S.launches[:, 0] == 5)) and n4 < len(S.blks) or widest <= 256


@pytest.mark.parametrize("cols", [256, 64])
def test_solve_schedule_of_dense_supernodes(cols, monkeypatch):
    """One dense supernode (no rows below the last superstep: the backward sweep needs its zero-row tile), 700
    columns = 256 + 256 + 188, under a chain of two (a wide supernode with rows below it)."""
    rng = np.random.default_rng(3)
    n1, n2 = 300, 700
    n = n1 + n2
    A = rng.standard_normal((n, n))
    A = A @ A.T + n * np.eye(n)
    Ld = np.linalg.cholesky(A)
    sym = dict(n=n, nsuper=2, super=np.array([0, n1, n]), pi=np.array([0, n, n + n2]),
               px=np.array([0, n * n1, n * n1 + n2 * n2]), s=np.concatenate([np.arange(n), np.arange(n1, n)]),
               xsize=n * n1 + n2 * n2)
    Lx = np.zeros(sym["xsize"])
    Lx[:n * n1] = np.tril(Ld)[:, :n1].T.reshape(-1)              # column-major n x n1, upper part of the block zero
    Lx[n * n1:] = np.tril(Ld[n1:, n1:]).T.reshape(-1)
    S = check(sym, Lx, cols, monkeypatch)
    if cols == 256:
        assert len(S.steps) == 2 + 3 and int(np.sum(S.launches[:, 0] == 5)) == 2 + 3
