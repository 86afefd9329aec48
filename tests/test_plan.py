"""CPU tests of the host side of CHOLMOD/B200: the C ABI loads and exports every declared symbol, the
planner's tables reproduce the reference factor when executed by the numpy interpreter
(tests/plan_interp.py), and a context without a device refuses to compute (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200
from cases import Case, CaseAAT, Forest, get_case
import plan_interp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not refd.have_ref(), reason="oracle/_ref/libref_driver.so not built")


def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "cholmod_b200.h")).read()
    declared = set(re.findall(r"\b(b200_[A-Za-z0-9_]+)\s*\(", hdr))
    assert declared == set(b200.SYMBOLS), declared ^ set(b200.SYMBOLS)
    L = C.CDLL(b200.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in b200.lib().b200_version()


def test_library_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", b200.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    if not out.strip():
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out), out


@needs_ref
@pytest.mark.parametrize("kind,size,ordering", [("lap2d5", 40, refd.CHOLMOD_AMD), ("lap3d7", 12, refd.CHOLMOD_METIS),
                                                ("lap2d9", 30, refd.CHOLMOD_METIS), ("el27", 7, refd.CHOLMOD_METIS),
                                                ("lap3d7", 22, refd.CHOLMOD_METIS)])
def test_plan_reproduces_reference_factor(kind, size, ordering):
    c = get_case(kind, size, ordering)
    F = b200.Factor(c.sym, device=-1)
    Lx = np.zeros(c.sym["xsize"])
    plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
    minor = plan_interp.run_plan(F, Lx)
    assert minor is None
    assert c.rel_err(Lx) <= 1e-12
    # the other legal interleaving of the main, look-ahead and copy streams gives the same factor, and the
    # copy-out schedule delivers every value of L->x exactly once and only after it is final
    Lx2 = np.zeros(c.sym["xsize"])
    host = np.full(c.sym["xsize"], np.nan)
    plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx2)
    assert plan_interp.run_plan(F, Lx2, prefer=1, host=host) is None
    assert c.rel_err(Lx2) <= 1e-12
    assert np.array_equal(host, Lx2)
    copies = F.plan_table("copies", b200.COPY_DT)
    assert copies["len"].sum() == c.sym["xsize"] and copies.size <= 1024 + 4 * F.geti("nlaunches")
    # the plan's flop count is the supernodal formula of SURVEY.md 8d: never below Common->fl's leading terms
    assert F.getd("flops_executed") > 0
    F.close()


@needs_ref
@needs_ref
@pytest.mark.parametrize("kind,size,ordering,dtype", [("lap3d7", 12, refd.CHOLMOD_METIS, np.complex128),
                                                      ("lap2d9", 30, refd.CHOLMOD_METIS, np.complex128),
                                                      ("lap3d7", 16, refd.CHOLMOD_METIS, np.complex64)])
def test_plan_reproduces_reference_factor_hermitian(kind, size, ordering, dtype):
    """The same plan run with Hermitian kernel semantics (A*B^H updates, zpotrf, X*L^-H) gives the factor of
    the reference's cd_/cs_ workers; complex128 plans use the inverse-based panel solves, complex64 ones the
    substitution kernels, like their real counterparts."""
    single = dtype == np.complex64
    c = get_case(kind, size, ordering, hermitian=True, single=single)
    assert c.status == refd.CHOLMOD_OK and c.Lref.dtype == dtype
    F = b200.Factor(c.sym, dtype=dtype, device=-1)
    for prefer in (0, 1):
        Lx = np.zeros(c.sym["xsize"], dtype=np.complex128)
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
        assert plan_interp.run_plan(F, Lx, prefer=prefer) is None
        assert c.rel_err(Lx) <= (1e-5 if single else 1e-12)
    launches = F.plan_table("launches", b200.LAUNCH_DT)
    has_inv = np.isin(launches["kind"], (b200.K_TRSM_INV32, b200.K_TRSM_INV64)).any()
    assert has_inv == (not single) or not np.isin(launches["kind"], b200.K_TRSM_KINDS).any()


@needs_ref
@pytest.mark.parametrize("cplx", [False, True])
def test_plan_unsymmetric_aat(cplx):
    """stype 0: only the scatter differs (column k of S*F instead of column k of S); the plan is the same."""
    c = CaseAAT(14, refd.CHOLMOD_AMD, cplx=cplx, beta=0.125)
    dtype = np.complex128 if cplx else np.float64
    F = b200.Factor(c.sym, dtype=dtype, device=-1)
    Lx = np.zeros(c.sym["xsize"], dtype=dtype)
    plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.125, Lx, F=(c.Fp, c.Fi, c.Fx))
    assert plan_interp.run_plan(F, Lx) is None
    assert c.rel_err(Lx) <= 1e-12


def test_plan_structure():
    c = get_case("lap3d7", 22, refd.CHOLMOD_METIS)
    F = b200.Factor(c.sym, device=-1)
    level = F.plan_table("level", np.int32)
    parent = F.plan_table("parent", np.int64)
    assert level.size == c.sym["nsuper"] and F.geti("nlevels") == level.max() + 1
    has_parent = parent >= 0
    assert np.all(level[parent[has_parent]] > level[has_parent])
    assert np.all(parent[has_parent] > np.nonzero(has_parent)[0])          # postordered
    launches = F.plan_table("launches", b200.LAUNCH_DT)
    assert np.all(np.diff(launches["level"]) >= 0)                           # level by level
    assert F.geti("nevents") > 0 and np.any(launches["stream"] == 1)        # look-ahead is in use
    items = F.plan_table("items", b200.ITEM_DT)
    tiles = F.plan_table("tiles", b200.TILE_DT)
    assert tiles["item"].max() < items.size
    # every pair's relative map addresses rows inside its target supernode
    relmap = F.plan_table("relmap", np.int32)
    mapped = items[items["map_off"] >= 0]
    assert mapped.size == F.geti("npairs")
    for it in mapped[:: max(1, mapped.size // 200)]:
        m = relmap[it["map_off"]:it["map_off"] + it["M"]]
        assert np.all(np.diff(m) > 0) and m[-1] < it["ldc"]
    # a wide supernode exercises the two-level blocking and the DMMA tile classes
    assert np.diff(c.sym["super"]).max() > 256
    kinds = set(launches["kind"].tolist())
    assert {b200.K_POTRF16, b200.K_POTRF64, b200.K_TRSM, b200.K_TRSM_INV64, b200.K_UPD_SIMT} <= kinds and (b200.K_UPD_D64 in kinds or b200.K_UPD_D128 in kinds)
    F.close()


def test_plan_rejects_bad_symbolic():
    sym = dict(n=4, nsuper=2, super=np.array([0, 2, 4]), pi=np.array([0, 3, 5]), px=np.array([0, 6, 10]),
               s=np.array([0, 1, 3, 2, 3]))
    F = b200.Factor(sym, device=-1)                 # valid: supernode 0 = cols 0-1 rows {0,1,3}; 1 = cols 2-3
    assert F.geti("npairs") == 1 and F.geti("nlevels") == 2
    F.close()
    bad = dict(sym, px=np.array([0, 5, 9]))
    with pytest.raises(b200.B200Error) as e:
        b200.Factor(bad, device=-1)
    assert e.value.status == b200.INVALID
    bad = dict(sym, s=np.array([0, 1, 1, 2, 3]))   # parent row inside the supernode itself
    with pytest.raises(b200.B200Error):
        b200.Factor(bad, device=-1)


def test_empty_matrix_plan():
    sym = dict(n=0, nsuper=0, super=np.array([0]), pi=np.array([0]), px=np.array([0]), s=np.zeros(0, dtype=np.int64))
    F = b200.Factor(sym, device=-1)
    assert F.geti("nlaunches") == 0 and F.geti("xsize") == 0
    F.close()


def test_no_cpu_fallback():
    """A context without a device must fail loudly instead of computing on the host."""
    sym = dict(n=4, nsuper=2, super=np.array([0, 2, 4]), pi=np.array([0, 3, 5]), px=np.array([0, 6, 10]),
               s=np.array([0, 1, 3, 2, 3]))
    F = b200.Factor(sym, device=-1)
    Ap = np.array([0, 1, 2, 3, 4]); Ai = np.arange(4); Ax = np.ones(4)
    with pytest.raises(b200.B200Error) as e:
        F.super_numeric(Ap, Ai, Ax)
    assert e.value.status == b200.GPU_PROBLEM
    with pytest.raises(b200.B200Error):
        F.solve(np.ones(4))
    F.close()
    if b200.device_count() == 0:
        with pytest.raises(b200.B200Error) as e:
            b200.Factor(sym, device=0)
        assert e.value.status == b200.GPU_PROBLEM


@needs_ref
def test_repair_plan_semantics_match_reference_notposdef():
    """The not-positive-definite path of the executor (zero tail, rebuild supernode s, factor info-1
    columns) restated with the interpreter agrees with the reference's L, zero pattern and L->minor."""
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(250, -1.0))
    assert c.status == refd.CHOLMOD_NOT_POSDEF
    F = b200.Factor(c.sym, device=-1)
    Lx = np.zeros(c.sym["xsize"])
    plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
    minor = plan_interp.run_plan(F, Lx)
    assert minor == c.minor
    F.close()


@needs_ref
def test_subtree_groups_structure_and_equivalence():
    """Single-GPU plans fork the middle levels into subtree groups (plan.cpp: plan_groups): a group is closed
    under taking descendants inside the forked levels, only those levels use the group streams, and the
    lock-step plan (CHOLMOD_B200_GROUPS=1) gives the same factor."""
    c = get_case("lap3d7", 22, refd.CHOLMOD_METIS)
    F8 = b200.Factor(c.sym, device=-1)
    lo, hi, G = F8.geti("fork_lo"), F8.geti("fork_hi"), F8.geti("ngroups")
    assert G > 1 and 0 <= lo < hi <= F8.geti("nlevels") and F8.geti("nstreams") == 5 + 2 * G
    level = F8.plan_table("level", np.int32)
    parent = F8.plan_table("parent", np.int64)
    group = F8.plan_table("group", np.int32)
    below = level < hi
    assert np.all(group[below] >= 0) and np.all(group[below] < G) and np.all(group[~below] == -1)
    kids = np.nonzero(below & (parent >= 0))[0]
    inside = kids[level[parent[kids]] < hi]
    assert np.array_equal(group[inside], group[parent[inside]])
    L = F8.plan_table("launches", b200.LAUNCH_DT)
    on_group = L["stream"] >= 5
    assert on_group.any() and np.all((L["level"][on_group] >= lo) & (L["level"][on_group] < hi))
    work = (L["count"] > 0) & (L["kind"] < len(b200.KIND_NAMES))
    forked = (L["level"] >= lo) & (L["level"] < hi)
    assert np.all(L["stream"][work & forked] >= 5)          # nothing of a forked level runs on the main pair
    os.environ["CHOLMOD_B200_GROUPS"] = "1"
    try:
        F1 = b200.Factor(c.sym, device=-1)
    finally:
        os.environ.pop("CHOLMOD_B200_GROUPS", None)
    assert F1.geti("ngroups") == 1 and F1.plan_table("launches", b200.LAUNCH_DT)["stream"].max() <= 2
    out = []
    for F in (F1, F8):
        Lx = np.zeros(c.sym["xsize"])
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
        assert plan_interp.run_plan(F, Lx, prefer=1) is None
        assert c.rel_err(Lx) <= 1e-12
        out.append(Lx)
    assert np.max(np.abs(out[0] - out[1])) <= 1e-12 * np.max(np.abs(out[0]))
    # same work either way
    for k in range(len(b200.KIND_NAMES)):
        assert F1.getd("flops_kind%d" % k) == pytest.approx(F8.getd("flops_kind%d" % k), rel=1e-12)


def test_fp64_tensor_instructions_are_in_the_library():
    """The update kernels really are on the FP64 tensor pipe (SASS DMMA) and staged with cp.async (LDGSTS)."""
    import subprocess
    out = subprocess.run(["cuobjdump", "-sass", b200.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    if not out.strip():
        pytest.skip("cuobjdump not available")
    seen = {}
    cur = None
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            continue
        if cur and "DMMA" in ln:
            seen.setdefault(cur, set()).add("DMMA")
        if cur and "LDGSTS" in ln:
            seen.setdefault(cur, set()).add("LDGSTS")
    for name in ("k_upd_dmma", "k_upd_zmma", "k_trsm_inv", "k_trsm_zinv"):
        hits = [f for f, v in seen.items() if name in f and "DMMA" in v and "LDGSTS" in v]
        assert hits, name


def test_plan_on_golden_fixtures():
    """The planner + kernel semantics (numpy interpreter) on every committed fixture, including the reference's
    own ctest matrices (complex c.tri, unsymmetric lp_afiro with F): runs without the reference library."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz")))
    assert files
    for f in files:
        g = np.load(f)
        sym = dict(n=int(g["n"]), nsuper=int(g["nsuper"]), xsize=int(g["xsize"]), super=g["super"], pi=g["pi"],
                   px=g["px"], s=g["s"])
        Lref = g["Lx"]
        F = b200.Factor(sym, dtype=Lref.dtype, device=-1)
        Lx = np.zeros(sym["xsize"], dtype=Lref.dtype)
        Fm = (g["Fp"], g["Fi"], g["Fx"]) if "Fp" in g.files else None
        plan_interp.scatter_A(sym, g["Sp"], g["Si"], g["Sx"], float(g["beta"]), Lx, F=Fm)
        minor = plan_interp.run_plan(F, Lx)
        if int(g["status"]) == 0:
            assert minor is None, f
            assert np.max(np.abs(Lx - Lref)) <= 1e-12 * max(1.0, np.max(np.abs(Lref))), f
        else:
            assert minor == int(g["minor"]), f      # the kernels' atomicMin finds the reference's L->minor


@needs_ref
@pytest.mark.parametrize("kind,size,ordering,copies", [("lap3d7", 12, refd.CHOLMOD_METIS, 8),
                                                       ("lap2d9", 30, refd.CHOLMOD_METIS, 16)])
def test_forest_whose_fork_runs_to_the_last_level_is_still_copied_out(kind, size, ordering, copies):
    """kron(I_k, A) has k roots: no level has fewer than 8 supernodes, so the subtree-group fork of the
    single-GPU plan covers everything up to the last level and no checkpoint follows it.  The copy-out
    schedule must still deliver every value of L->x (the drop-in calls return nothing else)."""
    f = Forest(get_case(kind, size, ordering), copies)
    F = b200.Factor(f.sym, device=-1)
    assert F.geti("ngroups") > 1 and F.geti("fork_hi") == F.geti("nlevels")
    copies_tab = F.plan_table("copies", b200.COPY_DT)
    assert copies_tab["len"].sum() == f.sym["xsize"]
    for prefer in (0, 1):
        Lx = np.zeros(f.sym["xsize"])
        host = np.full(f.sym["xsize"], np.nan)
        plan_interp.scatter_A(f.sym, f.Sp, f.Si, f.Sx, 0.0, Lx)
        assert plan_interp.run_plan(F, Lx, prefer=prefer, host=host) is None
        assert f.rel_err(Lx) <= 1e-12
        assert np.array_equal(host, Lx)
    F.close()


def test_tile_indices_are_32_bit():
    """build_plan admits supernodes of up to 2^30 rows, i.e. 2^25 SIMT tile rows: UpdTile.ti/tj must not be
    16-bit.  (A supernode that tall cannot be planned on the CPU box -- its ancestors are dense -- so the
    check is on the exported table layout the kernels and the interpreter share.)"""
    assert b200.TILE_DT["ti"].itemsize == 4 and b200.TILE_DT["tj"].itemsize == 4 and b200.TILE_DT.itemsize == 20
    hdr = open(os.path.join(ROOT, "suitesparse_b200", "csrc", "plan.hpp")).read()
    assert re.search(r"int32_t\s+ti\s*,\s*tj", hdr) and "uint16_t" not in hdr


def dense_case(n, seed=5):
    """One dense supernode: symbolic factor, lower-triangular CSC of a random SPD matrix, numpy's factor."""
    rng = np.random.default_rng(seed)
    B = rng.standard_normal((n, n))
    A = B @ B.T / n + 2.0 * np.eye(n)
    sym = dict(n=n, nsuper=1, super=np.array([0, n]), pi=np.array([0, n]), px=np.array([0, n * n]), s=np.arange(n), xsize=n * n)
    Ap = np.zeros(n + 1, dtype=np.int64); Ap[1:] = np.cumsum(n - np.arange(n))
    Ai = np.concatenate([np.arange(k, n) for k in range(n)])
    Ax = np.concatenate([A[k:, k] for k in range(n)])
    return sym, Ap, Ai, Ax, np.linalg.cholesky(A)


@pytest.mark.parametrize("far_min", [64, 100000])
def test_near_far_split_of_the_panel_chain(far_min):
    """A supernode with rows far below the diagonal block: the far rows' trsm / updates / Ua run on their own stream
    (rectangular items with shift > 0, rotating inverse slots).  Every legal interleaving of the streams the
    interpreter knows must give the factor; with the split off the plan is the old one."""
    n = 1400
    sym, Ap, Ai, Ax, Lnp = dense_case(n)
    os.environ["CHOLMOD_B200_FAR_MIN"] = str(far_min)
    try:
        F = b200.Factor(sym, device=-1)
    finally:
        os.environ.pop("CHOLMOD_B200_FAR_MIN")
    launches = F.plan_table("launches", b200.LAUNCH_DT)
    items = F.plan_table("items", b200.ITEM_DT)
    on_far = (launches["stream"] == 4) & (launches["count"] > 0) & (launches["kind"] < b200.K_JOIN)
    if far_min == 64:
        assert on_far.sum() > 10 and np.any(items["shift"] > 0) and F.geti("winv_slots") >= 8
    else:
        assert on_far.sum() == 0 and not np.any(items["shift"] > 0)
    for prefer in (0, 1, 2, 11, 12, 13, 14, 15, 16):
        Lx = np.zeros(n * n)
        host = np.full(n * n, np.nan)
        plan_interp.scatter_A(sym, Ap, Ai, Ax, 0.0, Lx)
        assert plan_interp.run_plan(F, Lx, prefer=prefer, host=host) is None
        L = np.tril(Lx.reshape(n, n).T)
        assert np.max(np.abs(L - Lnp)) <= 1e-12 * np.max(np.abs(Lnp)), prefer
        assert np.array_equal(host, Lx)
    F.close()


@needs_ref
def test_near_far_split_inside_a_sparse_factor():
    os.environ["CHOLMOD_B200_FAR_MIN"] = "32"
    try:
        c = get_case("lap3d7", 26, refd.CHOLMOD_METIS)
        F = b200.Factor(c.sym, device=-1)
    finally:
        os.environ.pop("CHOLMOD_B200_FAR_MIN")
    launches = F.plan_table("launches", b200.LAUNCH_DT)
    assert np.any((launches["stream"] == 4) & (launches["count"] > 0) & (launches["kind"] < b200.K_JOIN))
    for prefer in (0, 1, 2, 21, 22, 23, 24):
        Lx = np.zeros(c.sym["xsize"])
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
        assert plan_interp.run_plan(F, Lx, prefer=prefer) is None
        assert c.rel_err(Lx) <= 1e-12, prefer
    F.close()
