"""GPU parity tests: the CUDA numeric phase, called through the C ABI, against the oracle on the same
inputs.  Tolerances are BASELINE.json's: factor values max|dL|/max|L| <= 1e-10 (double), scaled residual
||Ax-b||/(||A|| ||x|| + ||b||) <= 1e-12; symbolic structure is shared with the reference (same host
analysis) and therefore identical by construction.  Single precision uses 1e-4 / 1e-5."""
import glob
import os

import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200, matrices as M
from cases import Case, Forest, get_case

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL_L, TOL_RES = 1e-10, 1e-12


def unpermute(sym, xperm):
    x = np.empty_like(xperm)
    x[sym["Perm"]] = xperm
    return x


@pytest.mark.parametrize("kind,size,ordering", [
    ("lap2d5", 100, refd.CHOLMOD_AMD),          # config 1
    ("lap3d7", 16, refd.CHOLMOD_METIS),
    ("lap3d7", 40, refd.CHOLMOD_METIS),         # the survey's extra case: root ~ 1.6k columns
    ("lap2d9", 200, refd.CHOLMOD_METIS),        # config 5 shape, reduced
    ("el27", 12, refd.CHOLMOD_METIS),           # config 3 shape, reduced
    ("lap2d5", 17, refd.CHOLMOD_NATURAL),
])
def test_factor_matches_reference(kind, size, ordering):
    c = get_case(kind, size, ordering)
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK and F.minor == c.n
    assert F.geti("kernel_launches") > 0
    assert c.rel_err(Lx) <= TOL_L
    # the strict upper triangles of the diagonal blocks stay zero like the reference's (SNW:361-379)
    assert np.array_equal(Lx == 0, c.Lref == 0) or np.count_nonzero((Lx == 0) != (c.Lref == 0)) < 1e-6 * Lx.size
    # reference solve on our factor, and our device solve
    assert c.residual_with(Lx) <= TOL_RES
    b = M.rhs(c.n)
    x = unpermute(c.sym, F.solve(b[c.sym["Perm"]]))
    assert c.ref.resid(x, b) <= TOL_RES
    F.close()


def test_golden_fixtures_on_gpu():
    """Every committed fixture, the reference's five ctest matrices included: bcsstk01/02 and can___24 (real
    symmetric), c.tri (complex Hermitian), lp_afiro (unsymmetric: A*A' + 1e-6 I through the A*F seam)."""
    seen = set()
    for f in sorted(glob.glob(os.path.join(HERE, "golden", "*.npz"))):
        g = np.load(f)
        seen.add(os.path.basename(f)[:-4])
        sym = dict(n=int(g["n"]), nsuper=int(g["nsuper"]), xsize=int(g["xsize"]), super=g["super"], pi=g["pi"],
                   px=g["px"], s=g["s"])
        F = b200.Factor(sym, dtype=g["Sx"].dtype, device=0)      # the Hermitian fixtures are complex128
        if int(g["stype"]) == 0:
            st, Lx = F.super_numeric_aat(int(g["ncol"]), g["Sp"], g["Si"], g["Sx"], g["Fp"], g["Fi"], g["Fx"],
                                         beta=float(g["beta"]))
        else:
            st, Lx = F.super_numeric(g["Sp"], g["Si"], g["Sx"], beta=float(g["beta"]))
        assert st == int(g["status"]) and F.minor == int(g["minor"]), f
        Lref = g["Lx"]
        assert np.max(np.abs(Lx - Lref)) <= TOL_L * max(1.0, np.max(np.abs(Lref))), f
        assert np.array_equal(Lx == 0, Lref == 0), f
        F.close()
    assert {"bcsstk01", "bcsstk02", "demo_lp_afiro", "demo_c", "demo_can24"} <= seen


@pytest.mark.parametrize("kind,size,ordering,copies", [("lap3d7", 12, refd.CHOLMOD_METIS, 8),
                                                       ("lap2d9", 30, refd.CHOLMOD_METIS, 16)])
def test_block_diagonal_forest(kind, size, ordering, copies):
    """kron(I_k, A): the etree is a forest of k equal trees, the subtree-group fork of the plan runs to the
    last level, and the host copy of L->x must still be complete (it is all the drop-in call returns)."""
    f = Forest(get_case(kind, size, ordering), copies)
    F = b200.Factor(f.sym, device=0)
    Lx = np.full(f.sym["xsize"], np.nan)
    st, Lx = F.super_numeric(f.Sp, f.Si, f.Sx, Lx=Lx)
    assert st == b200.OK and np.all(np.isfinite(Lx)) and f.rel_err(Lx) <= TOL_L
    assert f.rel_err(F.download_factor()) <= TOL_L
    F.close()


def test_port_agrees_too():
    c = get_case("lap3d7", 16, refd.CHOLMOD_METIS)
    st, minor, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx)
    F = b200.Factor(c.sym, device=0)
    _, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.close()


def test_beta_and_refactorization():
    """One analysis, several numeric factorizations with new values (cholmod_factorize.c:67-70)."""
    c = get_case("lap2d9", 60, refd.CHOLMOD_METIS)
    F = b200.Factor(c.sym, device=0)
    for scale, beta in [(1.0, 0.0), (2.5, 0.0), (1.0, 0.75)]:
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx * scale, beta=beta)
        stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx * scale, beta=beta)
        assert st == stp == 0
        assert np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    # values-only path: the pattern stays on the device
    F.upload_matrix(c.Sp, c.Si, c.Sx)
    F.upload_values(c.Sx * 3.0)
    assert F.factor_device() == b200.OK
    stp, _, Lp = refd.port_factor(c.sym, c.Sp, c.Si, c.Sx * 3.0)
    assert np.max(np.abs(F.download_factor() - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.close()


def test_unpacked_unsorted_and_extra_entries():
    """The seam accepts unpacked (A->nz) and unsorted columns, ignores the strictly upper part and
    entries outside the pattern of L (SNW:449-468)."""
    c = get_case("lap2d5", 30, refd.CHOLMOD_AMD)
    n, Sp, Si, Sx = c.n, c.Sp, c.Si, c.Sx
    rng = np.random.default_rng(7)
    cnt = np.diff(Sp)
    Up = np.zeros(n + 1, dtype=np.int64)
    Up[1:] = np.cumsum(cnt + 3)                          # 3 slack slots per column
    Ui = np.full(Up[-1], 0, dtype=np.int64)
    Ux = np.full(Up[-1], 123.0)
    Unz = np.zeros(n, dtype=np.int64)
    for k in range(n):
        rows, vals = Si[Sp[k]:Sp[k + 1]], Sx[Sp[k]:Sp[k + 1]]
        extra_r, extra_v = [], []
        if k > 0:
            extra_r.append(k - 1); extra_v.append(55.0)  # strictly upper entry: ignored
        perm = rng.permutation(len(rows) + len(extra_r))
        rr = np.concatenate([rows, extra_r]).astype(np.int64)[perm]
        vv = np.concatenate([vals, extra_v])[perm]
        Ui[Up[k]:Up[k] + len(rr)] = rr
        Ux[Up[k]:Up[k] + len(rr)] = vv
        Unz[k] = len(rr)
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric(Up, Ui, Ux, Anz=Unz)
    assert st == 0 and c.rel_err(Lx) <= TOL_L
    stp, _, Lp = refd.port_factor(c.sym, Up, Ui, Ux, Snz=Unz)
    assert stp == 0 and np.max(np.abs(Lx - Lp)) <= TOL_L * np.max(np.abs(Lp))
    F.close()


@pytest.mark.parametrize("col,quick", [(0, False), (37, False), (37, True), (250, False), (399, False)])
def test_not_posdef_semantics(col, quick):
    """status, L->minor, the zeroed tail and the partially factorized supernode of SNW:1084-1146."""
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(col, -1.0), factor=False)
    c.ref.seti("quick_return_if_not_posdef", int(quick))
    assert c.ref.super_numeric(0.0) == refd.CHOLMOD_NOT_POSDEF
    Lref, minor_ref = c.ref.Lx().copy(), c.ref.geti("minor")
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx, quick_return=quick)
    assert st == b200.NOT_POSDEF and F.minor == minor_ref
    assert np.array_equal(Lx == 0, Lref == 0)
    assert np.max(np.abs(Lx - Lref)) <= TOL_L * max(1.0, np.max(np.abs(Lref)))
    # the context is still usable afterwards
    good = get_case("lap2d5", 20, refd.CHOLMOD_AMD)
    st, Lx = F.super_numeric(good.Sp, good.Si, good.Sx)
    assert st == b200.OK and good.rel_err(Lx) <= TOL_L
    F.close()


def test_edge_cases_empty_scalar_diagonal_dense():
    """Degenerate shapes: n = 0, n = 1, a diagonal matrix (every supernode one column, no update pair),
    one dense supernode wider than an outer block (all of the two-level blocking, no pair)."""
    sym0 = dict(n=0, nsuper=0, super=np.array([0]), pi=np.array([0]), px=np.array([0]), s=np.zeros(0, dtype=np.int64))
    F = b200.Factor(sym0, device=0)
    st, Lx = F.super_numeric(np.array([0]), np.zeros(0, dtype=np.int64), np.zeros(0))
    assert st == b200.OK and Lx.size == 0 and F.minor == 0
    F.close()
    sym1 = dict(n=1, nsuper=1, super=np.array([0, 1]), pi=np.array([0, 1]), px=np.array([0, 1]), s=np.array([0]))
    F = b200.Factor(sym1, device=0)
    st, Lx = F.super_numeric(np.array([0, 1]), np.array([0]), np.array([9.0]))
    assert st == b200.OK and Lx[0] == 3.0
    st, Lx = F.super_numeric(np.array([0, 1]), np.array([0]), np.array([-4.0]))
    assert st == b200.NOT_POSDEF and F.minor == 0 and Lx[0] == 0.0
    F.close()
    n = 50
    symd = dict(n=n, nsuper=n, super=np.arange(n + 1), pi=np.arange(n + 1), px=np.arange(n + 1), s=np.arange(n))
    F = b200.Factor(symd, device=0)
    d = np.arange(1.0, n + 1)
    st, Lx = F.super_numeric(np.arange(n + 1), np.arange(n), d)
    assert st == b200.OK and np.allclose(Lx, np.sqrt(d), rtol=1e-15) and F.geti("npairs") == 0
    assert np.allclose(F.solve(d), np.ones(n), rtol=1e-14)
    F.close()
    n = 700                                                   # one dense supernode, 3 outer blocks
    rng = np.random.default_rng(3)
    B = rng.standard_normal((n, n))
    A = B @ B.T / n + 2.0 * np.eye(n)
    symn = dict(n=n, nsuper=1, super=np.array([0, n]), pi=np.array([0, n]), px=np.array([0, n * n]), s=np.arange(n))
    Ap = np.zeros(n + 1, dtype=np.int64); Ap[1:] = np.cumsum(n - np.arange(n))
    Ai = np.concatenate([np.arange(k, n) for k in range(n)])
    Ax = np.concatenate([A[k:, k] for k in range(n)])
    F = b200.Factor(symn, device=0)
    st, Lx = F.super_numeric(Ap, Ai, Ax)
    L = np.tril(Lx.reshape(n, n).T)
    assert st == b200.OK and np.max(np.abs(L - np.linalg.cholesky(A))) <= TOL_L * np.max(np.abs(L))
    assert np.all(np.triu(Lx.reshape(n, n).T, 1) == 0)
    x = F.solve(A @ np.ones(n))
    assert np.max(np.abs(x - 1.0)) <= 1e-10
    F.close()


def test_single_precision():
    c = Case("lap3d7", 14, refd.CHOLMOD_METIS, single=True)
    F = b200.Factor(c.sym, dtype=np.float32, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == 0 and Lx.dtype == np.float32
    assert c.rel_err(Lx) <= 1e-4
    b = M.rhs(c.n)
    x = unpermute(c.sym, F.solve(b[c.sym["Perm"]]))
    assert c.ref.resid(x, b) <= 1e-5
    F.close()


def test_device_resident_steps_and_multi_rhs_solve():
    c = get_case("lap3d7", 16, refd.CHOLMOD_METIS)
    F = b200.Factor(c.sym, device=0)
    F.upload_matrix(c.Sp, c.Si, c.Sx)
    assert F.factor_device() == b200.OK
    assert F.getd("ms_factor") > 0
    Lx = F.download_factor()
    assert c.rel_err(Lx) <= TOL_L
    B = np.asfortranarray(np.stack([M.rhs(c.n), np.ones(c.n), np.cos(np.arange(c.n))], axis=1))
    Xp = F.solve(B[c.sym["Perm"], :])
    for j in range(B.shape[1]):
        assert c.ref.resid(unpermute(c.sym, Xp[:, j].copy()), B[:, j].copy()) <= TOL_RES
    # solve with a factor computed elsewhere (the reference's), uploaded into the arena
    F.upload_factor(c.Lref)
    x = unpermute(c.sym, F.solve(M.rhs(c.n)[c.sym["Perm"]]))
    assert c.ref.resid(x, M.rhs(c.n)) <= TOL_RES
    F.close()


def test_full_size_properties_config2_shape():
    """A size the plain-C oracle cannot finish quickly (lap3d7 64^3, n = 262k): size-independent
    properties instead -- solve residual with our own factor, and L L' x = A x for a random x."""
    n, Ap, Ai, Ax = M.lap3d7(64)
    ref = refd.Ref()
    ref.set_matrix(n, Ap, Ai, Ax, -1)
    ref.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    sym = ref.symbolic()
    Sp, Si, Sx = ref.seam()
    F = b200.Factor(sym, device=0)
    st, Lx = F.super_numeric(Sp, Si, Sx)
    assert st == b200.OK and np.all(np.isfinite(Lx))
    b = M.rhs(n)
    x = unpermute(sym, F.solve(b[sym["Perm"]]))
    assert ref.resid(x, b) <= TOL_RES
    host = ref.numeric_Lx()
    host[:] = Lx
    assert ref.resid(ref.solve(b), b) <= TOL_RES
    F.close()


def _full_size(kind, size, single=False):
    """Reference analysis + reference CPU factor + our factor at a BASELINE.json size; returns the two figures of
    north_star: max|dL|/max|L| and the scaled residual of our device solve."""
    n, Ap, Ai, Ax = getattr(M, kind)(size)
    ref = refd.Ref(0)
    ref.set_matrix(n, Ap, Ai, Ax, -1, single=single)
    ref.analyze(refd.CHOLMOD_METIS, refd.CHOLMOD_SUPERNODAL)
    sym = ref.symbolic()
    Sp, Si, Sx = ref.seam()
    dt = np.float32 if single else np.float64
    F = b200.Factor(sym, dtype=dt, device=0)
    st, Lx = F.super_numeric(Sp, Si, Sx)
    assert st == b200.OK and F.minor == n
    assert ref.factorize(0.0) == 0
    Lref = ref.Lx()
    scale = float(np.max(np.abs(Lref)))
    err = 0.0
    for a in range(0, Lx.size, 1 << 26):            # chunked: the factors are gigabytes
        err = max(err, float(np.max(np.abs(Lx[a:a + (1 << 26)].astype(np.float64) - Lref[a:a + (1 << 26)]))))
    b = M.rhs(n)
    x = unpermute(sym, F.solve(b[sym["Perm"]]))
    res = ref.resid(x, b)
    F.close()
    ref.close()
    return err / scale, res


@pytest.mark.slow
@pytest.mark.parametrize("kind,size,single,tol_l,tol_r", [
    ("lap3d7", 100, False, TOL_L, TOL_RES),         # config 2 itself (root supernode: 14.9k columns)
    ("el27", 40, False, TOL_L, TOL_RES),            # config 3 shape at 40^3 x 3 dof
    ("lap2d9", 1000, False, TOL_L, TOL_RES),        # config 5 shape at 1000^2, double ...
    ("lap2d9", 1000, True, 1e-4, 1e-5),             # ... and single (config 5b)
])
def test_factor_values_at_benchmark_sizes(kind, size, single, tol_l, tol_r):
    """north_star's value bar where the big fronts stress it: explicit 64 x 64 inverses instead of
    substitution, split-K atomics, thousands of accumulated updates per entry."""
    err, res = _full_size(kind, size, single)
    assert err <= tol_l, err
    assert res <= tol_r, res


SOLVE_MODES = {"default": {}, "cols64": {"CHOLMOD_B200_SOLVE_COLS": "64"}, "no_w256": {"CHOLMOD_B200_SOLVE_W256": "0"},
               "no_pdl": {"CHOLMOD_B200_SOLVE_PDL": "0"}}


@pytest.mark.parametrize("mode", sorted(SOLVE_MODES))
@pytest.mark.parametrize("kind,size,ordering", [("lap3d7", 40, refd.CHOLMOD_METIS), ("lap3d7", 33, refd.CHOLMOD_METIS),
                                                ("lap2d9", 200, refd.CHOLMOD_METIS)])
def test_one_sided_sweeps_compose_to_the_solve(kind, size, ordering, mode, monkeypatch):
    """b200_solve_sweep: the forward sweep alone equals the reference's L\\b on its own factor (the plain-C
    restatement's triangular solve), and forward followed by backward equals the two-sided solve; several
    right-hand sides, factors with supernodes wider than one 256-column superstep (root separators of 1600 =
    6 x 256 + 64 and 1089 = 4 x 256 + 65 columns: ragged last supersteps and a one-column last block), with both
    schedules of the wide supernodes (256 columns per launch, the default, and 64), with and without the
    whole-block inverses and the programmatic dependent launch."""
    for k in ("CHOLMOD_B200_SOLVE_COLS", "CHOLMOD_B200_SOLVE_W256", "CHOLMOD_B200_SOLVE_PDL"):
        monkeypatch.delenv(k, raising=False)
    for k, v in SOLVE_MODES[mode].items():
        monkeypatch.setenv(k, v)
    c = get_case(kind, size, ordering)
    F = b200.Factor(c.sym, device=0)
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK
    rng = np.random.default_rng(11)
    B = np.asfortranarray(rng.standard_normal((c.n, 3)))
    X = F.solve(B)
    Y = F.solve_sweep(1, B)
    Z = F.solve_sweep(2, Y)
    assert np.max(np.abs(Z - X)) <= 1e-12 * np.max(np.abs(X))
    assert F.getd("solve_cols") == (64 if mode == "cols64" else 256)
    has_wide = bool(np.any(np.diff(c.sym["super"]) > 64))
    assert (F.getd("solve_w256_bytes") > 0) == (mode in ("default", "no_pdl") and has_wide)
    # L * Y = B, checked with a sparse matrix built from the supernodal factor
    import scipy.sparse as sp
    sup, pi, px, Ls = c.sym["super"], c.sym["pi"], c.sym["px"], c.sym["s"]
    rows, cols, vals = [], [], []
    for s_ in range(c.sym["nsuper"]):
        k1, k2 = sup[s_], sup[s_ + 1]
        nsrow = pi[s_ + 1] - pi[s_]
        blk = Lx[px[s_]:px[s_ + 1]].reshape(k2 - k1, nsrow).T          # nsrow x nscol
        r = Ls[pi[s_]:pi[s_ + 1]]
        for j in range(k2 - k1):
            rows.append(r[j:]); cols.append(np.full(nsrow - j, k1 + j)); vals.append(blk[j:, j])
    L = sp.csc_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(c.n, c.n))
    assert np.max(np.abs(L @ Y - B)) <= 1e-11 * np.max(np.abs(B))
    assert np.max(np.abs(L.T @ Z - Y)) <= 1e-11 * max(1.0, np.max(np.abs(Y)))
    F.close()


@pytest.mark.parametrize("kind,size,ordering,dtype", [
    ("lap2d9", 120, refd.CHOLMOD_METIS, "double"), ("lap3d7", 16, refd.CHOLMOD_METIS, "double"),
    ("lap2d5", 30, refd.CHOLMOD_AMD, "double"), ("lap2d5", 40, refd.CHOLMOD_AMD, "single")])
def test_graph_replay_matches_plain_launches(kind, size, ordering, dtype):
    """The launch list captured as a CUDA graph (use_graph = 1, CHOLMOD_B200_GRAPH=1; off by default):
    the capture call, its replays with other values, a second graph for the device-resident form, and a
    not-positive-definite matrix through the same graph all give what the plain launches give."""
    single = dtype == "single"
    c = get_case(kind, size, ordering, single=single)
    tol = 2e-4 if single else TOL_L
    F = b200.Factor(c.sym, device=0, dtype=np.float32 if single else np.float64)
    F.seti("use_graph", 1)
    Lbuf = np.empty(F.xsize, dtype=F.dtype)
    for rep, scale in enumerate([1.0, 2.0, 0.5]):
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx * scale, Lx=Lbuf)
        assert st == b200.OK
        assert np.max(np.abs(Lx / np.sqrt(scale) - c.Lref)) <= tol * np.max(np.abs(c.Lref))
        assert F.getd("graph_captures") == 1 and F.getd("graph_replays") == rep + 1
    launches = F.getd("kernel_launches")
    # the factor stays on the device: another destination, another graph
    F.upload_matrix(c.Sp, c.Si, c.Sx)
    for rep in range(2):
        assert F.factor_device() == b200.OK
        assert F.getd("graph_captures") == 2
    # back to the first destination: its graph is still there
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx, Lx=Lbuf)
    assert st == b200.OK and F.getd("graph_captures") == 2 and c.rel_err(Lx.astype(np.float64)) <= tol
    assert F.getd("kernel_launches") == launches
    # a download into another host array moves the context's page-locked registration away from Lbuf: the graph
    # whose copy nodes were bound to that mapping must not be replayed (it is captured again instead)
    assert c.rel_err(F.download_factor().astype(np.float64)) <= tol
    st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx * 2.0, Lx=Lbuf)
    assert st == b200.OK and np.max(np.abs(Lx / np.sqrt(2.0) - c.Lref)) <= tol * np.max(np.abs(c.Lref))
    assert F.getd("graph_captures") == 3
    # plain launches on the same context give the same number of launches and the same factor
    F.seti("use_graph", 0)
    st, Lx2 = F.super_numeric(c.Sp, c.Si, c.Sx)
    assert st == b200.OK and F.getd("kernel_launches") == launches
    assert c.rel_err(Lx2.astype(np.float64)) <= tol
    # not positive definite through the graph: same status, minor and zero pattern as through plain launches
    Sx = c.Sx.copy()
    col = c.n // 2
    Sx[c.Sp[col]] = -1.0                # first entry of the column is its diagonal in the lower-triangular seam
    if c.Si[c.Sp[col]] == col:
        st0, L0 = F.super_numeric(c.Sp, c.Si, Sx)
        m0 = F.minor
        F.seti("use_graph", 1)
        st1, L1 = F.super_numeric(c.Sp, c.Si, Sx, Lx=Lbuf)
        assert st0 == st1 == b200.NOT_POSDEF and F.minor == m0
        assert np.array_equal(L0 == 0, L1 == 0)
        assert np.max(np.abs(L0.astype(np.float64) - L1.astype(np.float64))) <= tol * max(1.0, float(np.max(np.abs(L0))))
        st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx, Lx=Lbuf)
        assert st == b200.OK and c.rel_err(Lx.astype(np.float64)) <= tol
    F.close()
