"""CPU tests of the multi-GPU path: the distributed plan -- etree subtrees owned and stored by
single ranks, their updates into the top separators accumulated per rank as contribution blocks and pulled by
the block owners, the top separators split 1-D block-cyclically over all ranks with broadcasts of finished
panels -- executed by the numpy interpreter, once with the ranks as threads of one process and once as two
real processes talking through torch.distributed (gloo).  Every rank copies out exactly what it owns; together
the ranks' copies are the complete factor, equal to the reference's."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from oracle import refd
from suitesparse_b200 import b200
from cases import get_case
import plan_interp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not refd.have_ref(), reason="oracle/_ref/libref_driver.so not built")


def run_ranks_as_threads(c, nranks, prefer=0, dist_run=None, far_min=None):
    """All ranks in one process, each with its own arena; a broadcast is two barriers around a copy from the
    root's arena, a pull reads the other rank's arena directly (the barrier of the plan has made it final)."""
    if dist_run is None:
        os.environ.pop("CHOLMOD_B200_DIST_RUN", None)
    else:
        os.environ["CHOLMOD_B200_DIST_RUN"] = str(dist_run)     # read when the plans are built
    if far_min is not None:
        os.environ["CHOLMOD_B200_FAR_MIN"] = str(far_min)
    try:
        Fs = [b200.Factor(c.sym, device=-1, nranks=nranks, rank=r) for r in range(nranks)]
    finally:
        os.environ.pop("CHOLMOD_B200_FAR_MIN", None)
    Lxs = [np.zeros(F.geti("dxsize")) for F in Fs]
    hosts = [np.full(c.sym["xsize"], np.nan) for _ in range(nranks)]
    for F, Lx in zip(Fs, Lxs):
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx, px=F.plan_table("dpx", np.int64),
                              colmine=F.plan_table("colmine", np.uint8))
    barrier = threading.Barrier(nranks)
    errors, minors = [], [None] * nranks

    def work(r):
        def bcast(off, ln, root):
            barrier.wait()
            if r != root:
                Lxs[r][off:off + ln] = Lxs[root][off:off + ln]
            barrier.wait()

        def pull(off, nr, nc, ld, src):
            return plan_interp._view(Lxs[src], off, nr, nc, ld)
        try:
            minors[r] = plan_interp.run_plan(Fs[r], Lxs[r], prefer=prefer + (r if prefer >= 10 else 0), host=hosts[r], bcast=bcast,
                                             barrier=barrier.wait, pull=pull)
            barrier.wait()              # nobody leaves (and frees its arena) while another rank may still pull
        except BaseException as e:      # noqa: BLE001 - report and release the other ranks
            errors.append((r, repr(e)))
            barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(nranks)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    return Fs, Lxs, hosts, minors


def assemble(hosts):
    """The ranks' host copies put together: every value of L->x must come from exactly one rank."""
    have = np.stack([~np.isnan(h) for h in hosts])
    assert np.all(have.sum(axis=0) == 1), "a value of L->x was copied out by %s ranks" % sorted(set(have.sum(axis=0).tolist()))
    return np.nansum(np.stack(hosts), axis=0)


@needs_ref
@pytest.mark.parametrize("kind,size,ordering,nranks,dist_run", [("lap3d7", 22, refd.CHOLMOD_METIS, 2, 1),
                                                                ("lap3d7", 22, refd.CHOLMOD_METIS, 4, None),
                                                                ("lap3d7", 26, refd.CHOLMOD_METIS, 3, 1),
                                                                ("lap2d9", 60, refd.CHOLMOD_METIS, 3, 1),
                                                                ("lap2d5", 40, refd.CHOLMOD_AMD, 8, 2)])
def test_distributed_plan_ranks_together_hold_the_reference_factor(kind, size, ordering, nranks, dist_run):
    c = get_case(kind, size, ordering)
    Fs, Lxs, hosts, minors = run_ranks_as_threads(c, nranks, prefer=1, dist_run=dist_run)
    os.environ.pop("CHOLMOD_B200_DIST_RUN", None)
    assert all(m is None for m in minors)
    assert c.rel_err(assemble(hosts)) <= 1e-12
    # the plan is one structure seen from different ranks: same broadcasts, same ownership everywhere
    b0 = Fs[0].plan_table("bcasts", b200.BCAST_DT)
    own0 = Fs[0].plan_table("owner", np.int32)
    top0 = Fs[0].plan_table("top", np.int8)
    for F in Fs[1:]:
        assert np.array_equal(F.plan_table("bcasts", b200.BCAST_DT), b0)
        assert np.array_equal(F.plan_table("owner", np.int32), own0)
        assert np.array_equal(F.plan_table("top", np.int8), top0)
    assert b0.size > 0 and set(b0["root"].tolist()) <= set(range(nranks))
    # storage is sharded: a rank's arena = the top separators + its own subtrees, and the subtrees partition the rest
    top_size = Fs[0].geti("top_size")
    assert all(F.geti("top_size") == top_size for F in Fs) and 0 < top_size < c.sym["xsize"]
    assert sum(F.geti("dxsize") - top_size for F in Fs) == c.sym["xsize"] - top_size
    # the fan-in exists and somebody pulls something
    assert sum(F.geti("npulls") for F in Fs) > 0
    launches = Fs[0].plan_table("launches", b200.LAUNCH_DT)
    assert np.count_nonzero(launches["kind"] == b200.K_BARRIER) == 1
    # work is really divided: every rank has update tiles, and no rank does more than ~75 % of the flops
    flops = np.array([sum(F.getd("flops_kind%d" % k) for k in range(len(b200.KIND_NAMES))) for F in Fs])
    assert np.all(flops > 0)
    if nranks >= 2 and kind == "lap3d7":
        assert flops.max() <= 0.75 * flops.sum()
    total = sum(Fs[0].getd("flops_%s" % k) for k in ("update", "potrf", "trsm"))
    assert abs(flops.sum() - total) <= 0.05 * total         # slabs, blocks and contribution blocks cover the work once
    for F in Fs:
        F.close()


@needs_ref
@pytest.mark.parametrize("nranks,dist_run,prefer", [(2, 1, 0), (2, 4, 31), (3, 2, 41), (4, 1, 51)])
def test_distributed_plan_with_near_far_split(nranks, dist_run, prefer):
    """The far stream inside a distributed plan (panels leave when near AND far part are done; the owner of the
    next block applies Ua_near / Ua_mid / Ua_far), under random interleavings of each rank's streams."""
    c = get_case("lap3d7", 26, refd.CHOLMOD_METIS)
    Fs, Lxs, hosts, minors = run_ranks_as_threads(c, nranks, prefer=prefer, dist_run=dist_run, far_min=32)
    os.environ.pop("CHOLMOD_B200_DIST_RUN", None)
    assert all(m is None for m in minors)
    assert c.rel_err(assemble(hosts)) <= 1e-12
    L = [F.plan_table("launches", b200.LAUNCH_DT) for F in Fs]
    assert any(np.any((l["stream"] == 4) & (l["count"] > 0) & (l["kind"] < b200.K_JOIN)) for l in L)
    for F in Fs:
        F.close()


@needs_ref
def test_single_rank_plan_is_unchanged_by_the_distribution_code():
    c = get_case("lap3d7", 22, refd.CHOLMOD_METIS)
    F = b200.Factor(c.sym, device=-1)
    assert F.geti("nbcasts") == 0 and F.geti("ntop") == 0 and F.geti("ngpus") == 1
    launches = F.plan_table("launches", b200.LAUNCH_DT)
    assert not np.any(launches["kind"] == b200.K_BCAST)
    F.close()


@needs_ref
def test_distributed_not_posdef_minor_is_found():
    from cases import Case
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(250, -1.0))
    Fs, Lxs, hosts, minors = run_ranks_as_threads(c, 2)
    found = [m for m in minors if m is not None]
    assert found and min(found) == c.minor                 # the executor all-reduces the minimum
    for F in Fs:
        F.close()


@needs_ref
@pytest.mark.parametrize("col,quick", [(250, False), (250, True), (399, False), (0, False)])
def test_truncated_plan_of_the_not_posdef_case(col, quick):
    """What the executor runs after a failed pivot (repair_dist): the plan truncated at the failing supernode
    leaves, put together over the ranks, exactly the reference's partially factorized L (SNW:1084-1146)."""
    from cases import Case
    import ctypes as C
    c = Case("lap2d5", 20, refd.CHOLMOD_AMD, shift_diag=(col, -1.0), factor=False)
    c.ref.seti("quick_return_if_not_posdef", int(quick))
    assert c.ref.super_numeric(0.0) == refd.CHOLMOD_NOT_POSDEF
    Lref, minor = c.ref.Lx().copy(), c.ref.geti("minor")
    sup = c.sym["super"]
    s = int(np.searchsorted(sup, minor, side="right") - 1)
    info = minor - int(sup[s]) + 1
    ncol_new = 0 if (info == 1 or quick) else info - 1
    nranks = 3
    os.environ["CHOLMOD_B200_TRUNC"] = "%d,%d" % (s, ncol_new)      # test hook of b200_create_dist (device < 0 only)
    try:
        Fs = [b200.Factor(c.sym, device=-1, nranks=nranks, rank=r) for r in range(nranks)]
    finally:
        os.environ.pop("CHOLMOD_B200_TRUNC")
    klast = int(sup[s + 1]) if ncol_new > 0 else int(sup[s])
    Lxs = [np.zeros(max(1, F.geti("dxsize"))) for F in Fs]
    for F, Lx in zip(Fs, Lxs):
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx, klast=klast, px=F.plan_table("dpx", np.int64),
                              colmine=F.plan_table("colmine", np.uint8))
    barrier = threading.Barrier(nranks)
    errors = []

    def work(r):
        def bcast(off, ln, root):
            barrier.wait()
            if r != root:
                Lxs[r][off:off + ln] = Lxs[root][off:off + ln]
            barrier.wait()
        try:
            m = plan_interp.run_plan(Fs[r], Lxs[r], bcast=bcast, barrier=barrier.wait,
                                     pull=lambda off, nr, nc, ld, src: plan_interp._view(Lxs[src], off, nr, nc, ld))
            assert m is None
            barrier.wait()
        except BaseException as e:      # noqa: BLE001
            errors.append((r, repr(e)))
            barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(nranks)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    # zero the columns of s that were not factorized, then every rank sends what it owns
    pi = c.sym["pi"]
    nsrow, nscol = int(pi[s + 1] - pi[s]), int(sup[s + 1] - sup[s])
    host = np.full(c.sym["xsize"], np.nan)
    for F, Lx in zip(Fs, Lxs):
        dpx = F.plan_table("dpx", np.int64)
        if ncol_new > 0 and dpx[s] >= 0:
            Lx[dpx[s] + nsrow * ncol_new:dpx[s] + nsrow * nscol] = 0
        for seg in F.plan_table("owned", b200.COPY_DT):
            o, ln, h = int(seg["off"]), int(seg["len"]), int(seg["hoff"])
            assert np.all(np.isnan(host[h:h + ln]))
            host[h:h + ln] = Lx[o:o + ln]
        F.close()
    assert not np.any(np.isnan(host))
    assert np.array_equal(host == 0, Lref == 0)
    assert np.max(np.abs(host - Lref)) <= 1e-12 * max(1.0, np.max(np.abs(Lref)))


CHILD = r"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
from oracle import refd
from suitesparse_b200 import b200
from cases import get_case
import plan_interp

dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
c = get_case("lap3d7", 18, refd.CHOLMOD_METIS)
F = b200.Factor(c.sym, device=-1, nranks=world, rank=rank)
Lx = np.zeros(F.geti("dxsize"))
plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx, px=F.plan_table("dpx", np.int64),
                      colmine=F.plan_table("colmine", np.uint8))
t = torch.from_numpy(Lx)                          # shares memory with Lx
top = F.geti("top_size")

def bcast(off, ln, root):
    dist.broadcast(t[off:off + ln], src=root)

# the peers' contribution blocks: processes cannot read each other's memory here, so every rank publishes its
# copy of the top separators right after the barrier (on the GPU this is a load over NVLink)
snap = {}
def barrier():
    parts = [torch.empty(top, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(parts, t[:top].clone())
    for r in range(world):
        snap[r] = parts[r].numpy()

def pull(off, nr, nc, ld, src):
    return plan_interp._view(snap[src], off, nr, nc, ld)

host = np.full(c.sym["xsize"], np.nan)
minor = plan_interp.run_plan(F, Lx, bcast=bcast, barrier=barrier, pull=pull, host=host)
hosts = [None] * world
dist.all_gather_object(hosts, (rank, minor, host))
if rank == 0:
    have = np.stack([~np.isnan(h[2]) for h in hosts])
    assert np.all(have.sum(axis=0) == 1)
    full = np.nansum(np.stack([h[2] for h in hosts]), axis=0)
    err = c.rel_err(full)
    print("RESULT", err, [h[1] for h in hosts])
    assert err <= 1e-12 and all(h[1] is None for h in hosts)
dist.destroy_process_group()
"""


@needs_ref
def test_distributed_plan_two_processes_gloo():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29617", WORLD_SIZE="2")
    procs = []
    for r in range(2):
        e = dict(env, RANK=str(r))
        procs.append(subprocess.Popen([sys.executable, "-c", CHILD % dict(root=ROOT)], env=e, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-3000:]
    assert "RESULT" in outs[0][0]
