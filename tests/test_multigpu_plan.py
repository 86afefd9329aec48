"""CPU tests of the multi-GPU path: the distributed plan -- etree subtrees owned by single
ranks, the top separators split 1-D block-cyclically over all ranks, broadcasts of finished subtrees and
panels -- executed by the numpy interpreter, once with the ranks as threads of one process and once as
two real processes talking through torch.distributed (gloo).  Every rank must end with the complete
factor, equal to the reference's."""
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


def run_ranks_as_threads(c, nranks, prefer=0, dist_run=None):
    """All ranks in one process; a broadcast is two barriers around a copy from the root's array."""
    if dist_run is None:
        os.environ.pop("CHOLMOD_B200_DIST_RUN", None)
    else:
        os.environ["CHOLMOD_B200_DIST_RUN"] = str(dist_run)     # read when the plans are built
    Fs = [b200.Factor(c.sym, device=-1, nranks=nranks, rank=r) for r in range(nranks)]
    Lxs = [np.zeros(c.sym["xsize"]) for _ in range(nranks)]
    hosts = [np.full(c.sym["xsize"], np.nan) for _ in range(nranks)]
    for Lx in Lxs:
        plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
    barrier = threading.Barrier(nranks)
    errors, minors = [], [None] * nranks

    def work(r):
        def bcast(off, ln, root):
            barrier.wait()
            if r != root:
                Lxs[r][off:off + ln] = Lxs[root][off:off + ln]
            barrier.wait()
        try:
            minors[r] = plan_interp.run_plan(Fs[r], Lxs[r], prefer=prefer, host=hosts[r], bcast=bcast)
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


@needs_ref
@pytest.mark.parametrize("kind,size,ordering,nranks,dist_run", [("lap3d7", 22, refd.CHOLMOD_METIS, 2, 1),
                                                                ("lap3d7", 22, refd.CHOLMOD_METIS, 4, None),
                                                                ("lap3d7", 26, refd.CHOLMOD_METIS, 3, 1),
                                                                ("lap2d9", 60, refd.CHOLMOD_METIS, 3, 1),
                                                                ("lap2d5", 40, refd.CHOLMOD_AMD, 8, 2)])
def test_distributed_plan_every_rank_ends_with_the_reference_factor(kind, size, ordering, nranks, dist_run):
    c = get_case(kind, size, ordering)
    Fs, Lxs, hosts, minors = run_ranks_as_threads(c, nranks, prefer=1, dist_run=dist_run)
    os.environ.pop("CHOLMOD_B200_DIST_RUN", None)
    for r in range(nranks):
        assert minors[r] is None
        assert c.rel_err(Lxs[r]) <= 1e-12, r
        assert np.array_equal(hosts[r], Lxs[r]), r         # each rank's copy-out schedule is complete and final
    # the plan is one structure seen from different ranks: same broadcasts, same ownership everywhere
    b0 = Fs[0].plan_table("bcasts", b200.BCAST_DT)
    own0 = Fs[0].plan_table("owner", np.int32)
    top0 = Fs[0].plan_table("top", np.int8)
    for F in Fs[1:]:
        assert np.array_equal(F.plan_table("bcasts", b200.BCAST_DT), b0)
        assert np.array_equal(F.plan_table("owner", np.int32), own0)
        assert np.array_equal(F.plan_table("top", np.int8), top0)
    assert b0.size > 0 and set(b0["root"].tolist()) <= set(range(nranks))
    # work is really divided: every rank has update tiles, and no rank does more than ~60 % of the flops
    flops = np.array([sum(F.getd("flops_kind%d" % k) for k in range(len(b200.KIND_NAMES))) for F in Fs])
    assert np.all(flops > 0)
    if nranks >= 2 and kind == "lap3d7":
        assert flops.max() <= 0.75 * flops.sum()
    total = sum(Fs[0].getd("flops_%s" % k) for k in ("update", "potrf", "trsm"))
    assert abs(flops.sum() - total) <= 0.05 * total         # slabs and blocks cover the work once
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
Lx = np.zeros(c.sym["xsize"])
plan_interp.scatter_A(c.sym, c.Sp, c.Si, c.Sx, 0.0, Lx)
t = torch.from_numpy(Lx)                          # shares memory with Lx

def bcast(off, ln, root):
    dist.broadcast(t[off:off + ln], src=root)

minor = plan_interp.run_plan(F, Lx, bcast=bcast)
err = c.rel_err(Lx)
errs = [None] * world
dist.all_gather_object(errs, (rank, float(err), minor))
if rank == 0:
    print("RESULT", errs)
    assert all(e[1] <= 1e-12 and e[2] is None for e in errs), errs
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
