"""Test-only numpy interpreter of the CHOLMOD/B200 launch plan.

Executes the plan tables exported by `b200_plan_export` (UpdItem/UpdTile/PotrfBlk/TrsmTile/Launch, see
suitesparse_b200/csrc/plan.hpp) with the semantics the CUDA kernels implement, so that the host planner
(levels, pairs, relative maps, tiling, two-level blocking, split-K) can be checked on a machine without
a GPU.  It is NOT a fallback: nothing in the product imports it.
"""
import numpy as np
from numpy.lib.stride_tricks import as_strided

from suitesparse_b200 import b200


def scatter_A(sym, Sp, Si, Sx, beta, Lx, kfirst=0, klast=None, F=None, px=None, colmine=None):
    """What k_scatter_A does (Lx already zero on the touched supernodes); with F = (Fp, Fi, Fx) what
    k_scatter_AF does: column k of S*F, gathered entry by entry.  px: the arena offsets of a distributed plan
    (table "dpx") instead of L->px; colmine: that plan's column ownership mask."""
    n = sym["n"]
    klast = n if klast is None else klast
    sup, pi, Ls = sym["super"], sym["pi"], sym["s"]
    px = sym["px"] if px is None else px
    supermap = np.repeat(np.arange(sym["nsuper"]), np.diff(sup))
    for k in range(kfirst, klast):
        if colmine is not None and not colmine[k]:
            continue
        s = supermap[k]
        k1, k2 = sup[s], sup[s + 1]
        nsrow = pi[s + 1] - pi[s]
        rows = Ls[pi[s]:pi[s + 1]]
        base = px[s] + (k - k1) * nsrow
        if F is None:
            entries = [(Si[p], Sx[p], False) for p in range(Sp[k], Sp[k + 1])]
        else:
            Fp, Fi, Fx = F
            entries = [(Si[p], Sx[p] * Fx[pf], True) for pf in range(Fp[k], Fp[k + 1])
                       for p in range(Sp[Fi[pf]], Sp[Fi[pf] + 1])]
        for i, v, add in entries:
            if i < k:
                continue
            if i < k2:
                pos = i - k1
            else:
                pos = np.searchsorted(rows, i)
                if pos >= nsrow or rows[pos] != i:
                    continue
            Lx[base + pos] = Lx[base + pos] + v if add else v
        if beta != 0:
            Lx[base + (k - k1)] += beta


def _view(Lx, off, rows, cols, ld):
    return as_strided(Lx[off:], shape=(rows, cols), strides=(Lx.itemsize, ld * Lx.itemsize), writeable=True)


def _schedule(launches, prefer):
    """Order the launches the way two CUDA streams could legally execute them: each stream in order, a
    launch with wait_ev only after that event was recorded.  prefer = 0 runs the main stream as far as
    it can before the look-ahead stream gets a turn; prefer = 1 the other way round."""
    ids = sorted(set([0, 1, 2, 3, 4]) | set(int(v) for v in np.unique(launches["stream"])))
    queues = {st: [i for i, L in enumerate(launches) if L["stream"] == st] for st in ids}
    pos, recorded, order = {st: 0 for st in ids}, set(), []
    # stream 2 is the copy stream, 3 the communication stream, 4 the far stream, 5.. the subtree-group streams (odd:
    # panel chain, even: look-ahead): prefer = 0 runs main, look-ahead, the groups in order, communication, copy; prefer = 1
    # the exact reverse
    groups = [st for st in ids if st >= 5]
    # prefer >= 10: a random legal interleaving (the seed)
    rng = np.random.default_rng(prefer) if prefer >= 10 else None
    pref = {0: [0, 4, 1] + groups + [3, 2], 1: [2, 3] + groups[::-1] + [1, 4, 0], 2: [4, 1, 0] + groups + [2, 3]}[min(prefer, 2)]
    # cudaStreamWaitEvent on an event that the host has not recorded yet is a no-op: every wait must come
    # after its record in issue (= list) order
    rec_at = {int(L["rec_ev"]): i for i, L in enumerate(launches) if L["rec_ev"] >= 0}
    for i, L in enumerate(launches):
        if L["wait_ev"] >= 0:
            assert int(L["wait_ev"]) in rec_at and rec_at[int(L["wait_ev"])] < i, "wait before record"

    def ready(st):
        if pos[st] >= len(queues[st]):
            return False
        w = int(launches[queues[st][pos[st]]]["wait_ev"])
        return w < 0 or w in recorded

    while any(pos[st] < len(queues[st]) for st in ids):
        cand = [st for st in pref if ready(st)]
        assert cand, "dead-lock in the launch plan's event graph"
        st = cand[0] if rng is None else cand[int(rng.integers(len(cand)))]
        i = queues[st][pos[st]]
        order.append(i)
        pos[st] += 1
        r = int(launches[i]["rec_ev"])
        if r >= 0:
            recorded.add(r)
    return order


def run_plan(F, Lx, tables=None, prefer=0, host=None, bcast=None, barrier=None, pull=None):
    """Execute every launch of Factor F's plan on the flat factor array Lx.  Returns the smallest failing
    column (what the kernels record through atomicMin) or None.  With `host` (an array like Lx filled with
    NaN) the K_COPYOUT launches copy their segments into it at the moment the schedule reaches them."""
    if tables is None:
        tables = dict(items=F.plan_table("items", b200.ITEM_DT), tiles=F.plan_table("tiles", b200.TILE_DT),
                      potrf=F.plan_table("potrf", b200.POTRF_DT), trsm=F.plan_table("trsm", b200.TRSM_DT),
                      launches=F.plan_table("launches", b200.LAUNCH_DT), relmap=F.plan_table("relmap", np.int32))
    items, tiles, potrf, trsm = tables["items"], tables["tiles"], tables["potrf"], tables["trsm"]
    relmap = tables["relmap"]
    minor = None
    launches = tables["launches"]
    for st1 in launches[(launches["stream"] == 1) | ((launches["stream"] >= 6) & (launches["stream"] % 2 == 0))]:
        assert st1["wait_ev"] >= 0 or st1["kind"] != b200.K_JOIN
    for idx in _schedule(launches, prefer):
        L = launches[idx]
        kind, cnt, first = int(L["kind"]), int(L["count"]), int(L["first"])
        if kind == b200.K_JOIN:
            continue
        if kind == b200.K_BCAST:
            # bcast (off, len, root): the collective; every rank calls it in the same order
            assert L["stream"] == 3 and bcast is not None
            for seg in F.plan_table("bcasts", b200.BCAST_DT)[first:first + cnt]:
                bcast(int(seg["off"]), int(seg["len"]), int(seg["root"]))
            continue
        if kind == b200.K_COPYOUT:
            assert L["stream"] == 2
            if host is not None:
                for seg in F.plan_table("copies", b200.COPY_DT)[first:first + cnt]:
                    o, ln, h = int(seg["off"]), int(seg["len"]), int(seg["hoff"])
                    assert np.all(np.isnan(host[h:h + ln])), "segment copied twice"
                    host[h:h + ln] = Lx[o:o + ln]
            continue
        if kind == b200.K_BARRIER:
            assert L["stream"] == 3 and barrier is not None
            barrier()
            continue
        if kind == b200.K_PULL:
            # pull (seg, r) -> rank r's copy of the block as an (nrows x ncols) array; only the lower part counts
            assert pull is not None
            for seg in F.plan_table("pulls", b200.PULL_DT)[first:first + cnt]:
                nr, nc, ld = int(seg["nrows"]), int(seg["ncols"]), int(seg["ld"])
                mine = _view(Lx, int(seg["off"]), nr, nc, ld)
                low = np.arange(nr)[:, None] >= np.arange(nc)[None, :]
                for r in range(32):
                    if int(seg["mask"]) >> r & 1:
                        theirs = pull(int(seg["off"]), nr, nc, ld, r)
                        mine[low] += theirs[low]
            continue
        if kind in b200.TILE_SHAPE:
            TM, TN = b200.TILE_SHAPE[kind]
            # all tiles of a launch read operands that no tile of the same launch writes: gather the
            # products first, then apply them
            pending = []
            for t in tiles[first:first + cnt]:
                it = items[t["item"]]
                M, N, lda = int(it["M"]), int(it["N"]), int(it["lda"])
                i0, j0 = int(t["ti"]) * TM, int(t["tj"]) * TN
                i1, j1 = min(M, i0 + TM), min(N, j0 + TN)
                sh = int(it["shift"])                                   # B starts `sh` rows above A
                assert i0 < M and j0 < N and i1 - 1 + sh >= j0, "tile without a wanted entry"
                assert sh == 0 or (sh >= N and it["map_off"] < 0)
                k0, k1 = int(t["k0"]), int(t["k1"])
                assert 0 <= k0 < k1 <= it["K"]
                A = _view(Lx, int(it["a_off"]), M, int(it["K"]), lda)
                B = _view(Lx, int(it["a_off"]) - sh, N, int(it["K"]), lda)
                Cpart = A[i0:i1, k0:k1] @ B[j0:j1, k0:k1].conj().T    # A * B^H (B' for real data)
                ii, jj = np.meshgrid(np.arange(i0, i1), np.arange(j0, j1), indexing="ij")
                keep = ii + sh >= jj
                if it["map_off"] >= 0:
                    m = relmap[int(it["map_off"]):int(it["map_off"]) + M].astype(np.int64)
                    dst = int(it["c_off"]) + m[ii[keep]] + m[jj[keep]] * int(it["ldc"])
                else:
                    dst = int(it["c_off"]) + ii[keep] + jj[keep] * int(it["ldc"])
                pending.append((dst, Cpart[keep]))
            for dst, val in pending:
                np.subtract.at(Lx, dst, val)
        elif kind in b200.K_POTRF_KINDS:
            for b in potrf[first:first + cnt]:
                jb, lda = int(b["jb"]), int(b["lda"])
                assert jb <= (16, 32, 64)[kind - b200.K_POTRF16]
                D = _view(Lx, int(b["off"]), jb, jb, lda)
                Dl = np.tril(D)
                Dsym = Dl + np.tril(Dl, -1).conj().T
                # unblocked right-looking Cholesky with the kernel's failure rule
                Lc = Dsym.copy()
                for j in range(jb):
                    d = Lc[j, j].real
                    if not d > 0:
                        col = int(b["col"]) + j
                        minor = col if minor is None else min(minor, col)
                    with np.errstate(all="ignore"):
                        ljj = np.sqrt(d)
                        Lc[j:, j] = Lc[j:, j] / ljj
                        Lc[j, j] = ljj
                        Lc[j + 1:, j + 1:] -= np.outer(Lc[j + 1:, j], Lc[j + 1:, j].conj())
                r, c = np.tril_indices(jb)
                D[r, c] = Lc[r, c]
        elif kind in (b200.K_FACTOR_SMALL, b200.K_FACTOR_TINY):
            for b in potrf[first:first + cnt]:
                jb, lda = int(b["jb"]), int(b["lda"])
                assert (jb <= 16 and lda - jb <= 256) if kind == b200.K_FACTOR_SMALL else (jb <= 4 and lda - jb <= 64)
                P = _view(Lx, int(b["off"]), lda, jb, lda)      # the whole supernode: b.off is its origin
                Dl = np.tril(P[:jb, :])
                Lc = Dl + np.tril(Dl, -1).conj().T
                for j in range(jb):
                    d = Lc[j, j].real
                    if not d > 0:
                        col = int(b["col"]) + j
                        minor = col if minor is None else min(minor, col)
                    with np.errstate(all="ignore"):
                        ljj = np.sqrt(d)
                        Lc[j:, j] = Lc[j:, j] / ljj
                        Lc[j, j] = ljj
                        Lc[j + 1:, j + 1:] -= np.outer(Lc[j + 1:, j], Lc[j + 1:, j].conj())
                r, c = np.tril_indices(jb)
                P[r, c] = Lc[r, c]
                if lda > jb:
                    L11 = np.tril(P[:jb, :])
                    with np.errstate(all="ignore"):
                        P[jb:, :] = np.linalg.solve(L11, P[jb:, :].conj().T).conj().T if np.all(np.isfinite(L11)) and np.all(np.diag(L11) != 0) \
                            else np.nan
        elif kind in b200.K_TRSM_KINDS:
            for t in trsm[first:first + cnt]:
                jb, lda, nr = int(t["jb"]), int(t["lda"]), int(t["nrows"])
                assert (int(t["slot"]) >= 0) == (kind != b200.K_TRSM)
                L11 = np.tril(_view(Lx, int(t["diag_off"]), jb, jb, lda))
                X = _view(Lx, int(t["row_off"]), nr, jb, lda)
                with np.errstate(all="ignore"):
                    X[:, :] = np.linalg.solve(L11, X.conj().T).conj().T if np.all(np.isfinite(L11)) and np.all(np.diag(L11) != 0) \
                        else np.full_like(X, np.nan)
        else:
            raise AssertionError("unknown launch kind %d" % kind)
    return minor
