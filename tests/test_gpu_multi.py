"""GPU test of the multi-GPU factorization: two (or more) processes, one per GPU, each with its share of
the distributed plan, NCCL broadcasts of subtrees and panels in between.  Every rank must end with the
complete factor, within 1e-10 of the reference's; rank 0's host L->x (copy-out schedule) too.
Skipped when the box has fewer than 2 GPUs."""
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
for kind, size, shift, herm in (("lap3d7", 30, None, False), ("lap2d9", 150, None, False),
                                ("lap3d7", 14, (1500, -1.0), False), ("lap3d7", 24, None, True)):
    c = Case(kind, size, refd.CHOLMOD_METIS, shift_diag=shift, hermitian=herm)
    F = b200.Factor(c.sym, dtype=c.Lref.dtype, device=rank, nranks=world, rank=rank)
    F.comm_init_torch()
    res = {}
    for rep in range(2):                                    # twice: the context is reusable
        if rank == 0:
            st, Lx = F.super_numeric(c.Sp, c.Si, c.Sx)     # host buffer in / out on rank 0
        else:
            F.upload_matrix(c.Sp, c.Si, c.Sx)
            st = F.factor_device()
            Lx = F.download_factor()                        # every rank holds the whole factor
        res = dict(status=st, minor=F.minor, err=c.rel_err(Lx), ref_status=c.status, ref_minor=c.minor,
                   zero_same=bool(np.array_equal(Lx == 0, c.Lref == 0)), nccl=F.geti("nccl_calls"), ntop=F.geti("ntop"))
    if shift is None:
        b = M.rhs(c.n) * ((1 + 0.5j) if herm else 1.0)
        x = np.empty_like(b); x[c.sym["Perm"]] = F.solve(b[c.sym["Perm"]])
        res["resid"] = c.ref.resid(x, b)
    out.append(res)
    F.close()
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
    for per_rank in allres:
        ok1, ok2, bad, okz = per_rank          # okz: the Hermitian (complex double) case, (re,im) pairs on the wire
        for ok in (ok1, ok2, okz):
            assert ok["status"] == 0 and ok["err"] <= 1e-10 and ok["resid"] <= 1e-12
            assert ok["nccl"] > 0
        assert ok1["ntop"] > 0
        assert bad["status"] == 1 == bad["ref_status"] and bad["minor"] == bad["ref_minor"]
        assert bad["zero_same"] and bad["err"] <= 1e-10
