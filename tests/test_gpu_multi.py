"""GPU test of the multi-GPU factorization, one process per GPU: each rank factorizes its subtrees, accumulates
its contribution blocks to the top separators, the block owners pull them over NVLink (arenas attached through
CUDA IPC handles), the top separators are factorized cooperatively (NCCL broadcasts of finished panels), and
every rank copies what it owns into ONE host L->x (a shared mapping).  That L->x must be within 1e-10 of the
reference's; a rank's arena must hold only its share.  Skipped when the box has fewer than 2 GPUs."""
import json
import os
import subprocess
import sys

import pytest

from suitesparse_b200 import b200

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
from oracle import refd
from suitesparse_b200 import b200, matrices as M
from cases import Case
out = []
for ci, (kind, size, shift, herm) in enumerate((("lap3d7", 30, None, False), ("lap2d9", 150, None, False),
                                                ("lap3d7", 14, (1500, -1.0), False), ("lap3d7", 24, None, True))):
    c = Case(kind, size, refd.CHOLMOD_METIS, shift_diag=shift, hermitian=herm)
    # the caller's ONE host L->x: a shared mapping every rank writes its share into
    path = "/dev/shm/b200_multi_%%s_%%d.bin" %% (os.environ["MASTER_PORT"], ci)
    if rank == 0:
        Lx = np.memmap(path, dtype=c.Lref.dtype, mode="w+", shape=(c.sym["xsize"],))
    dist.barrier()
    if rank != 0:
        Lx = np.memmap(path, dtype=c.Lref.dtype, mode="r+", shape=(c.sym["xsize"],))
    F = b200.Factor(c.sym, dtype=c.Lref.dtype, device=rank, nranks=world, rank=rank)
    F.comm_init_torch()
    res = {}
    for rep in range(2):                                    # twice: the context is reusable
        if rank == 0:
            Lx[:] = np.nan
        dist.barrier()
        st, _ = F.super_numeric(c.Sp, c.Si, c.Sx, Lx=Lx)   # every rank: host A in, its share of L->x out
        torch.cuda.synchronize(); dist.barrier()
        res = dict(status=st, minor=F.minor, ref_status=c.status, ref_minor=c.minor, nccl=F.geti("nccl_calls"),
                   ntop=F.geti("ntop"), dxsize=F.geti("dxsize"), top_size=F.geti("top_size"), xsize=F.geti("xsize"),
                   npulls=F.geti("npulls"), nvlink=F.geti("nvlink_in_bytes"))
        if rank == 0:
            full = np.array(Lx)
            res.update(err=c.rel_err(full), complete=bool(np.all(np.isfinite(full))),
                       zero_same=bool(np.array_equal(full == 0, c.Lref == 0)))
    # a distributed factor is not solved on the device: rank 0 puts the assembled L->x into a one-GPU context
    try:
        F.solve(np.ones(c.n, dtype=c.Lref.dtype))
        res["dist_solve_refused"] = False
    except b200.B200Error as e:
        res["dist_solve_refused"] = (e.status == b200.INVALID)
    if shift is None and rank == 0:
        G = b200.Factor(c.sym, dtype=c.Lref.dtype, device=0)
        G.upload_factor(np.array(Lx))
        b = M.rhs(c.n) * ((1 + 0.5j) if herm else 1.0)
        x = np.empty_like(b); x[c.sym["Perm"]] = G.solve(b[c.sym["Perm"]])
        res["resid"] = c.ref.resid(x, b)
        G.close()
    out.append(res)
    F.close()
    dist.barrier()
    del Lx
    if rank == 0:
        os.unlink(path)
allres = [None] * world
dist.all_gather_object(allres, out)
if rank == 0:
    print("RESULT " + json.dumps(allres))
dist.destroy_process_group()
"""


@pytest.mark.parametrize("world", [2, 4, 8])
def test_factorization_over_several_gpus(world):
    if b200.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29640 + world), WORLD_SIZE=str(world))
    procs = [subprocess.Popen([sys.executable, "-c", CHILD % dict(root=ROOT)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = [p.communicate(timeout=900) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-3000:]
    line = [ln for ln in outs[0][0].splitlines() if ln.startswith("RESULT ")][-1]
    allres = json.loads(line[len("RESULT "):])
    assert len(allres) == world
    for r, per_rank in enumerate(allres):
        ok1, ok2, bad, okz = per_rank          # okz: the Hermitian (complex double) case, (re,im) pairs on the wire
        for ok in (ok1, ok2, okz):
            assert ok["status"] == 0 and ok["nccl"] > 0 and ok["dist_solve_refused"]
            if r == 0:
                assert ok["complete"] and ok["err"] <= 1e-10 and ok["resid"] <= 1e-12
        assert ok1["ntop"] > 0
        assert bad["status"] == 1 == bad["ref_status"] and bad["minor"] == bad["ref_minor"]
        if r == 0:
            assert bad["complete"] and bad["zero_same"] and bad["err"] <= 1e-10
    # sharded storage: top separators everywhere, every subtree on exactly one rank; somebody pulled over NVLink
    for k in (0, 1, 3):
        top = allres[0][k]["top_size"]
        assert sum(per_rank[k]["dxsize"] - top for per_rank in allres) == allres[0][k]["xsize"] - top
        assert sum(per_rank[k]["npulls"] for per_rank in allres) > 0 and sum(per_rank[k]["nvlink"] for per_rank in allres) > 0
