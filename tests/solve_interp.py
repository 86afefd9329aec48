"""numpy interpreter of the device solve's schedule (b200_plan_export "solve_*"): what k_solve_fwd / k_solve_bwd
(one CTA per narrow supernode), k_solve_fwd_blk / k_solve_bwd_blk (64 columns per launch) and k_solve_diag_fwd256 /
k_solve_fwd256 / k_solve_bwd256 (256 columns per launch) do, launch by launch and tile by tile, on the host.  It
checks what the kernels rely on -- every entry of a wide supernode below a superstep is covered exactly once, the
tiles marked as holders of the next diagonal block are exactly the ones that touch its rows, counters match the
number of tiles -- and produces the solution, so a CPU test can compare it with a plain triangular solve."""
import numpy as np

BLK_DT = np.dtype([("s", "<i4"), ("c0", "<i4"), ("jb", "<i4"), ("pad", "<i4")])
TILE_DT = np.dtype([("s", "<i4"), ("c0", "<i4"), ("jb", "<i4"), ("r0", "<i4"), ("nrows", "<i4"), ("writer", "<i4"),
                    ("ntiles", "<i4"), ("blk", "<i4")])
TILE256_DT = np.dtype([("s", "<i4"), ("c0", "<i4"), ("nb", "<i4"), ("r0", "<i4"), ("nrows", "<i4"), ("ntiles", "<i4"),
                       ("blk0", "<i4"), ("step", "<i4"), ("next_nb", "<i4"), ("pad", "<i4")])


class Schedule:
    def __init__(self, F):
        self.small = F.plan_table("solve_small", np.dtype("<i4"))
        self.blks = F.plan_table("solve_blks", BLK_DT)
        self.steps = F.plan_table("solve_steps", BLK_DT)
        self.tiles = F.plan_table("solve_tiles", TILE_DT)
        self.tiles256 = F.plan_table("solve_tiles256", TILE256_DT)
        self.launches = F.plan_table("solve_launches", np.dtype("<i8")).reshape(-1, 4)     # kind, first, count, maxcol


class Sn:
    """One supernode of the factor: its nsrow x nscol block (column-major in L->x) and the global rows."""

    def __init__(self, sym, Lx, s):
        sup, pi, px = sym["super"], sym["pi"], sym["px"]
        self.k1, self.nscol = int(sup[s]), int(sup[s + 1] - sup[s])
        self.nsrow = int(pi[s + 1] - pi[s])
        self.rows = np.asarray(sym["s"][pi[s]:pi[s + 1]], dtype=np.int64)
        self.L = Lx[px[s]:px[s] + self.nsrow * self.nscol].reshape(self.nscol, self.nsrow).T
        assert np.array_equal(self.rows[:self.nscol], np.arange(self.k1, self.k1 + self.nscol))

    def index(self, r0, nrows):
        return self.rows[r0:r0 + nrows]


def _tri(L, c0, nb):
    return np.tril(L[c0:c0 + nb, c0:c0 + nb])


def _check_blocks(S, sym):
    """the 64 x 64 blocks and the supersteps of a supernode are consecutive, cover its columns, and point at each other"""
    sup = sym["super"]
    for k, st in enumerate(S.steps):
        nscol = int(sup[st["s"] + 1] - sup[st["s"]])
        assert st["c0"] % 256 == 0 and st["jb"] == min(256, nscol - st["c0"])
        for q in range((st["jb"] + 63) // 64):
            b = S.blks[st["pad"] + q]
            assert (b["s"], b["c0"], b["jb"]) == (st["s"], st["c0"] + 64 * q, min(64, nscol - st["c0"] - 64 * q))
        if st["c0"] > 0:
            assert S.steps[k - 1]["s"] == st["s"] and S.steps[k - 1]["c0"] == st["c0"] - 256


def solve(sym, Lx, S, B, which=3):
    """Replay the schedule on the right-hand sides B (n x nrhs); which: bit 0 forward, bit 1 backward."""
    n = sym["n"]
    X = np.array(B, dtype=np.float64).reshape(n, -1).copy()
    Y = np.zeros_like(X)
    wide_cols = np.zeros(n, dtype=bool)
    for b in S.blks:
        k1 = int(sym["super"][b["s"]])
        wide_cols[k1 + b["c0"]:k1 + b["c0"] + b["jb"]] = True
    _check_blocks(S, sym)
    sn = {}

    def node(s):
        if s not in sn:
            sn[s] = Sn(sym, Lx, int(s))
        return sn[s]

    covered = {}                     # (s, superstep or block start, sweep) -> rows seen

    def cover(key, t_r0, t_nrows, lo, hi):
        seen = covered.setdefault(key, np.zeros(hi, dtype=np.int32))
        assert lo <= t_r0 and t_r0 + t_nrows <= hi
        seen[t_r0:t_r0 + t_nrows] += 1

    if which & 1:
        for kind, first, count, maxcol in S.launches:
            if kind == 0:
                for s in S.small[first:first + count]:
                    N = node(s)
                    assert N.nscol <= maxcol
                    x1 = np.linalg.solve(_tri(N.L, 0, N.nscol), X[N.k1:N.k1 + N.nscol])
                    X[N.k1:N.k1 + N.nscol] = x1
                    np.subtract.at(X, N.rows[N.nscol:], N.L[N.nscol:, :] @ x1)
            elif kind == 2:
                tl = S.tiles[first:first + count]
                # every tile of the launch forms y from the block's x before any tile of the launch changes x below it
                ys = {}
                for t in tl:
                    N = node(t["s"])
                    key = (int(t["s"]), int(t["c0"]))
                    if key not in ys:
                        ys[key] = np.linalg.solve(_tri(N.L, t["c0"], t["jb"]), X[N.k1 + t["c0"]:N.k1 + t["c0"] + t["jb"]])
                        assert t["writer"] == 1
                        Y[N.k1 + t["c0"]:N.k1 + t["c0"] + t["jb"]] = ys[key]
                    else:
                        assert t["writer"] == 0
                for t in tl:
                    N = node(t["s"])
                    y = ys[(int(t["s"]), int(t["c0"]))]
                    cover((int(t["s"]), int(t["c0"]), "f64"), t["r0"], t["nrows"], t["c0"] + t["jb"], N.nsrow)
                    rows = slice(t["r0"], t["r0"] + t["nrows"])
                    np.subtract.at(X, N.index(t["r0"], t["nrows"]), N.L[rows, t["c0"]:t["c0"] + t["jb"]] @ y)
            elif kind == 3:
                for t in S.tiles256[first:first + count]:
                    N = node(t["s"])
                    assert t["c0"] == 0 and tuple(S.steps[t["step"]])[:3] == (t["s"], 0, t["nb"]) and S.steps[t["step"]]["pad"] == t["blk0"]
                    Y[N.k1:N.k1 + t["nb"]] = np.linalg.solve(_tri(N.L, 0, t["nb"]), X[N.k1:N.k1 + t["nb"]])
            elif kind == 4:
                tl = S.tiles256[first:first + count]
                # first the tiles that hold rows of the next diagonal block, then that block's solve (which the last
                # of them does), then all the other tiles: nothing else may touch those rows in this launch
                for t in tl:
                    N = node(t["s"])
                    st = S.steps[t["step"]]
                    assert (st["s"], st["c0"], st["jb"], st["pad"]) == (t["s"], t["c0"], t["nb"], t["blk0"])
                    nxt = min(256, N.nscol - t["c0"] - t["nb"])
                    holds_next = t["r0"] < t["c0"] + t["nb"] + nxt
                    assert t["next_nb"] == (nxt if holds_next else 0)
                    cover((int(t["s"]), int(t["c0"]), "f256"), t["r0"], t["nrows"], t["c0"] + t["nb"], N.nsrow)

                def apply(t):
                    N = node(t["s"])
                    y = Y[N.k1 + t["c0"]:N.k1 + t["c0"] + t["nb"]]
                    rows = slice(t["r0"], t["r0"] + t["nrows"])
                    np.subtract.at(X, N.index(t["r0"], t["nrows"]), N.L[rows, t["c0"]:t["c0"] + t["nb"]] @ y)
                ndiag = {}
                for t in tl:
                    if t["next_nb"] > 0:
                        apply(t)
                        ndiag[(int(t["s"]), int(t["c0"]))] = ndiag.get((int(t["s"]), int(t["c0"])), 0) + 1
                done = set()
                for t in tl:
                    key = (int(t["s"]), int(t["c0"]))
                    if t["next_nb"] > 0 and key not in done:
                        done.add(key)
                        assert ndiag[key] == (t["next_nb"] + 63) // 64          # what the kernel's counter waits for
                        N = node(t["s"])
                        c1 = t["c0"] + t["nb"]
                        nstep = S.steps[t["step"] + 1]
                        assert (nstep["s"], nstep["c0"], nstep["jb"]) == (t["s"], c1, t["next_nb"])
                        Y[N.k1 + c1:N.k1 + c1 + t["next_nb"]] = np.linalg.solve(
                            _tri(N.L, c1, t["next_nb"]), X[N.k1 + c1:N.k1 + c1 + t["next_nb"]])
                for t in tl:
                    if t["next_nb"] == 0:
                        apply(t)
        if which == 1:
            X[wide_cols] = Y[wide_cols]              # k_solve_merge
    if which & 2:
        if which == 2:
            Y[:] = X
        for kind, first, count, maxcol in S.launches[::-1]:
            if kind == 0:
                for s in S.small[first:first + count]:
                    N = node(s)
                    r = X[N.k1:N.k1 + N.nscol] - N.L[N.nscol:, :].T @ X[N.rows[N.nscol:]]
                    X[N.k1:N.k1 + N.nscol] = np.linalg.solve(_tri(N.L, 0, N.nscol).T, r)
            elif kind in (2, 5):
                tl = S.tiles[first:first + count] if kind == 2 else S.tiles256[first:first + count]
                width = "jb" if kind == 2 else "nb"
                groups = {}
                for t in tl:
                    groups.setdefault((int(t["s"]), int(t["c0"])), []).append(t)
                for (s, c0), ts in groups.items():
                    N = node(s)
                    nb = int(ts[0][width])
                    assert all(t["ntiles"] == len(ts) for t in ts)             # the counter of the last-finisher
                    for t in ts:
                        cover((s, c0, "b%d" % kind), t["r0"], t["nrows"], c0 + nb, N.nsrow)
                        if t["nrows"] > 0:
                            rows = slice(t["r0"], t["r0"] + t["nrows"])
                            Y[N.k1 + c0:N.k1 + c0 + nb] -= N.L[rows, c0:c0 + nb].T @ X[N.index(t["r0"], t["nrows"])]
                    X[N.k1 + c0:N.k1 + c0 + nb] = np.linalg.solve(_tri(N.L, c0, nb).T, Y[N.k1 + c0:N.k1 + c0 + nb])
    # every row below a superstep / block step was covered exactly once, in each sweep that ran
    for (s, c0, sweep), seen in covered.items():
        N = node(s)
        width = 64 if sweep in ("f64", "b2") else 256
        lo = c0 + min(width, N.nscol - c0)
        assert np.all(seen[lo:] == 1) and np.all(seen[:lo] == 0), (s, c0, sweep)
    return X
