"""GPU test of b200_panel_factor_noalloc_d, the entry point for callers that may not allocate.

Some users of this library run under a "no cudaMalloc after start-up" rule (allocation synchronizes the
device and has no latency bound) or have put all of HBM into their own pool.  They hand in a device
buffer they own; the library's scratch is a static __device__ array.  This test plays such a caller: it
owns the panel buffer (a torch tensor), calls the entry point through the C ABI, and checks the result
against what the ordinary entry point b200_super_numeric_d -- which cudaMallocs its scratch in
b200_create -- produces for the same panel: the first supernode of a two-supernode dense factor
(columns 0..jb-1 with all n rows) is exactly "dpotrf of the leading block, dtrsm of the rows below".

The file name sorts last on purpose, so a run that stops at the first failure has already run the rest.
"""
import numpy as np
import pytest

from suitesparse_b200 import b200

pytestmark = pytest.mark.gpu
TOL_L = 1e-10          # the tolerance of tests/test_gpu_parity.py


def dense_spd_lower(n, seed):
    rng = np.random.default_rng(seed)
    B = rng.standard_normal((n, n))
    A = B @ B.T / n + 2.0 * np.eye(n)
    return A


def two_supernode_factor(A, jb):
    """Dense n x n SPD matrix as supernode 0 = columns [0, jb) x all rows, supernode 1 = the rest."""
    n = A.shape[0]
    if jb == n:
        sym = dict(n=n, nsuper=1, super=np.array([0, n]), pi=np.array([0, n]), px=np.array([0, n * n]),
                   s=np.arange(n))
    else:
        sym = dict(n=n, nsuper=2, super=np.array([0, jb, n]), pi=np.array([0, n, n + (n - jb)]),
                   px=np.array([0, n * jb, n * jb + (n - jb) * (n - jb)]),
                   s=np.concatenate([np.arange(n), np.arange(jb, n)]))
    Ap = np.zeros(n + 1, dtype=np.int64)
    Ap[1:] = np.cumsum(n - np.arange(n))
    Ai = np.concatenate([np.arange(k, n) for k in range(n)])
    Ax = np.concatenate([A[k:, k] for k in range(n)])
    return sym, Ap, Ai, Ax


def existing_entry_point_panel(A, jb):
    sym, Ap, Ai, Ax = two_supernode_factor(A, jb)
    F = b200.Factor(sym, device=0)
    st, Lx = F.super_numeric(Ap, Ai, Ax)
    F.close()
    n = A.shape[0]
    return st, F.minor, Lx[:n * jb].reshape(jb, n).T.copy()       # column-major n x jb panel


def noalloc_panel(A, jb, lda_pad=0):
    """The caller's side: a preallocated device buffer, filled with the lower trapezoid of A(:, 0:jb)."""
    import torch
    n = A.shape[0]
    lda = n + lda_pad
    host = np.full((jb, lda), 777.0)                                # row c of `host` = column c of the panel
    for c in range(jb):
        host[c, c:n] = A[c:, c]
        host[c, :c] = 0.0                                           # strict upper part: left untouched
    buf = torch.from_numpy(host).to("cuda:0")                       # owned by the caller, not by the library
    torch.cuda.synchronize()
    st, info = b200.panel_factor_noalloc(buf.data_ptr(), lda, n, jb)
    out = buf.cpu().numpy()
    assert np.all(out[:, n:] == 777.0)                              # padding rows beyond nrows are not touched
    return st, info, out[:, :n].T.copy()


@pytest.mark.parametrize("cases", [[(300, 64, 11, 0), (257, 40, 12, 5)], [(129, 17, 13, 3), (64, 64, 14, 0), (1000, 7, 15, 0)]])
def test_noalloc_matches_existing_entry_point_on_calls_in_a_row(cases):
    """Back-to-back calls share the static scratch; each must stand on its own."""
    for n, jb, seed, pad in cases:
        A = dense_spd_lower(n, seed)
        st_ref, minor_ref, P_ref = existing_entry_point_panel(A, jb)
        st, info, P = noalloc_panel(A, jb, lda_pad=pad)
        assert st_ref == b200.OK and st == b200.OK and info == 0
        scale = np.max(np.abs(P_ref))
        assert np.max(np.abs(P - P_ref)) <= TOL_L * scale, (n, jb)
        # and it is a Cholesky panel: L11 L11' = A11, L21 L11' = A21
        L11 = np.tril(P[:jb, :])
        assert np.max(np.abs(L11 @ L11.T - A[:jb, :jb])) <= 1e-12 * np.max(np.abs(A))
        if n > jb:
            assert np.max(np.abs(P[jb:, :] @ L11.T - A[jb:, :jb])) <= 1e-12 * np.max(np.abs(A))


def test_noalloc_not_posdef_then_good_call():
    """The pivot flag is re-armed by every call: a failed call does not poison the next one."""
    n, jb = 200, 48
    A = dense_spd_lower(n, 21)
    bad = A.copy()
    bad[30, 30] = -5.0
    st, info, _ = noalloc_panel(bad, jb)
    st_ref, minor_ref, _ = existing_entry_point_panel(bad, jb)
    assert st == b200.NOT_POSDEF and st_ref == b200.NOT_POSDEF
    assert info == minor_ref + 1                                    # 1-based column of the first bad pivot
    st, info, P = noalloc_panel(A, jb)
    _, _, P_ref = existing_entry_point_panel(A, jb)
    assert st == b200.OK and info == 0
    assert np.max(np.abs(P - P_ref)) <= TOL_L * np.max(np.abs(P_ref))


def test_noalloc_rejects_what_the_fixed_scratch_cannot_hold():
    import torch
    buf = torch.zeros(1024, dtype=torch.float64, device="cuda:0")
    with pytest.raises(b200.B200Error) as e:
        b200.panel_factor_noalloc(buf.data_ptr(), 200, 200, b200.NOALLOC_MAX_COLS + 1)
    assert e.value.status == b200.TOO_LARGE
    with pytest.raises(b200.B200Error) as e:
        b200.panel_factor_noalloc(buf.data_ptr(), b200.NOALLOC_MAX_ROWS + 1, b200.NOALLOC_MAX_ROWS + 1, 8)
    assert e.value.status == b200.TOO_LARGE
    with pytest.raises(b200.B200Error) as e:
        b200.panel_factor_noalloc(buf.data_ptr(), 2 ** 31, 16, 8)    # lda does not fit the 32-bit descriptors
    assert e.value.status == b200.TOO_LARGE
    with pytest.raises(b200.B200Error) as e:
        b200.panel_factor_noalloc(0, 10, 10, 4)
    assert e.value.status == b200.INVALID
    with pytest.raises(b200.B200Error) as e:
        b200.panel_factor_noalloc(buf.data_ptr(), 5, 10, 4)          # lda < nrows
    assert e.value.status == b200.INVALID
