"""Synthetic SPD test matrices of BASELINE.json's configs.

All generators are deterministic (no RNG), return the LOWER triangle in CSC with int64 indices
(`stype = -1` in CHOLMOD terms): ``(n, Ap, Ai, Ax)``.  Grid node (x,y,z) -> x + nx*(y + ny*z).
"""
import numpy as np
import scipy.sparse as sp


def _lower_csc(rows, cols, vals, n):
    A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsc()
    A.sum_duplicates()
    A.sort_indices()
    return n, A.indptr.astype(np.int64), A.indices.astype(np.int64), A.data.astype(np.float64)


def _stencil(shape, offsets, diag, off):
    """Lower triangle of a constant-coefficient stencil matrix on a regular grid."""
    dims = len(shape)
    n = int(np.prod(shape))
    grids = np.meshgrid(*[np.arange(s) for s in shape], indexing="ij")     # grids[0] = x, fastest
    strides = [int(np.prod(shape[:d])) for d in range(dims)]
    idx = sum(g * st for g, st in zip(grids, strides)).ravel()
    rows, cols, vals = [idx], [idx], [np.full(n, float(diag))]
    for o in offsets:
        ok = np.ones(n, dtype=bool)
        nb = idx.copy()
        for d in range(dims):
            c = grids[d].ravel() + o[d]
            ok &= (c >= 0) & (c < shape[d])
            nb = nb + o[d] * strides[d]
        r, c = idx[ok], nb[ok]
        keep = r > c                                    # strictly lower part
        rows.append(r[keep]); cols.append(c[keep]); vals.append(np.full(int(keep.sum()), float(off)))
    return _lower_csc(np.concatenate(rows), np.concatenate(cols), np.concatenate(vals), n)


def lap2d5(nx, ny=None):
    """2-D 5-point Laplacian: diag 4, -1 to (x±1), (y±1).  (config 1 at 100x100)"""
    ny = nx if ny is None else ny
    offs = [(1, 0), (-1, 0), (0, 1), (0, -1)]
    return _stencil((nx, ny), offs, 4.0, -1.0)


def lap2d9(nx, ny=None):
    """2-D 9-point Laplacian: diag 8, -1 to the 8 neighbours.  (config 5 at 2000x2000)"""
    ny = nx if ny is None else ny
    offs = [(a, b) for a in (-1, 0, 1) for b in (-1, 0, 1) if (a, b) != (0, 0)]
    return _stencil((nx, ny), offs, 8.0, -1.0)


def lap3d7(nx, ny=None, nz=None):
    """3-D 7-point Laplacian: diag 6, -1 to the 6 neighbours.  (configs 2 and 4)"""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    offs = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    return _stencil((nx, ny, nz), offs, 6.0, -1.0)


def el27(nx, ny=None, nz=None, dof=3):
    """3-D 27-point, `dof`-dof elasticity-style SPD block matrix (config 3 at 80^3 x 3).

    Diagonal block: 81 on its diagonal, -0.25 off-diagonal; each of the 26 neighbour blocks: -1 on
    the diagonal, -0.125 off-diagonal (symmetric and strictly diagonally dominant, hence SPD).
    """
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    shape = (nx, ny, nz)
    nn = nx * ny * nz
    n = nn * dof
    gx, gy, gz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    node = (gx + nx * (gy + ny * gz)).ravel()
    rows, cols, vals = [], [], []
    for a in range(dof):
        for b in range(dof):
            if a >= b:
                rows.append(node * dof + a); cols.append(node * dof + b)
                vals.append(np.full(nn, 81.0 if a == b else -0.25))
    for ox in (-1, 0, 1):
        for oy in (-1, 0, 1):
            for oz in (-1, 0, 1):
                if (ox, oy, oz) == (0, 0, 0):
                    continue
                cx, cy, cz = gx.ravel() + ox, gy.ravel() + oy, gz.ravel() + oz
                ok = (cx >= 0) & (cx < nx) & (cy >= 0) & (cy < ny) & (cz >= 0) & (cz < nz)
                nb = cx + nx * (cy + ny * cz)
                r0, c0 = node[ok], nb[ok]
                keep = r0 > c0
                r0, c0 = r0[keep], c0[keep]
                for a in range(dof):
                    for b in range(dof):
                        rows.append(r0 * dof + a); cols.append(c0 * dof + b)
                        vals.append(np.full(r0.size, -1.0 if a == b else -0.125))
    del shape
    return _lower_csc(np.concatenate(rows), np.concatenate(cols), np.concatenate(vals), n)


def hermitian(n, Ap, Ai, Ax):
    """A Hermitian positive definite matrix with the pattern and the moduli of a real symmetric one: every
    stored off-diagonal entry a_ij (i > j) gets the deterministic phase exp(i*theta_ij); the diagonal stays
    real.  (Irreducible diagonal dominance does not see phases, so the result is positive definite whenever
    the real matrix is diagonally dominant.)  Returns (n, Ap, Ai, Ax) with complex128 Ax."""
    col = np.repeat(np.arange(n, dtype=np.int64), np.diff(Ap))
    theta = 0.37 * ((7 * Ai + 13 * col) % 11) + 0.11
    z = np.asarray(Ax, dtype=np.float64) * np.exp(1j * theta)
    z[Ai == col] = np.asarray(Ax, dtype=np.float64)[Ai == col]
    return n, Ap, Ai, z.astype(np.complex128)


def aat2d(nx, ny=None, complex_phases=False):
    """An unsymmetric (CHOLMOD stype 0) test matrix for the A*A' case: A = [D | E], n = nx*ny rows, one column
    per grid node (D = 0.5 I) and one per grid edge (E = the edge-node incidence matrix, +1 / -1), so that
    A*A' = 0.25 I + the 5-point graph Laplacian.  With complex_phases the edge entries get deterministic
    phases and A*A^H is Hermitian positive definite.  Returns (nrow, ncol, Ap, Ai, Ax), CSC, sorted."""
    ny = nx if ny is None else ny
    n = nx * ny
    x, y = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    node = (x + nx * y).ravel()
    xr, yr = x.ravel(), y.ravel()
    rows, cols, vals = [node], [node], [np.full(n, 0.5)]
    ncol = n
    for dx, dy in ((1, 0), (0, 1)):
        ok = (xr + dx < nx) & (yr + dy < ny)
        a, b = node[ok], (xr[ok] + dx) + nx * (yr[ok] + dy)
        e = ncol + np.arange(a.size)
        rows += [a, b]; cols += [e, e]; vals += [np.full(a.size, 1.0), np.full(a.size, -1.0)]
        ncol += a.size
    rows, cols, vals = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    if complex_phases:
        vals = vals * np.exp(1j * (0.29 * ((5 * rows + 3 * cols) % 13) + 0.05))
    A = sp.coo_matrix((vals, (rows, cols)), shape=(n, ncol)).tocsc()
    A.sort_indices()
    return n, ncol, A.indptr.astype(np.int64), A.indices.astype(np.int64), A.data


def rhs(n):
    """The demo's right-hand side b_i = 1 + i/n (CHOLMOD/Demo/cholmod_dl_demo.c:243-249)."""
    return 1.0 + np.arange(n, dtype=np.float64) / n


CONFIGS = {
    # name: (generator, args, ordering)  -- ordering is a CHOLMOD code: 2 = AMD, 3 = METIS
    "cfg1_lap2d5_100": (lap2d5, (100,), 2),
    "cfg2_lap3d7_100": (lap3d7, (100,), 3),
    "cfg3_el27_80": (el27, (80,), 3),
    "cfg4_lap3d7_160": (lap3d7, (160,), 3),
    "cfg5_lap2d9_2000": (lap2d9, (2000,), 3),
    # reduced sizes for parity tests
    "lap3d7_40": (lap3d7, (40,), 3),
    "lap3d7_20": (lap3d7, (20,), 3),
    "lap2d9_200": (lap2d9, (200,), 3),
    "el27_12": (el27, (12,), 3),
}
