"""One workload, a few device-resident factorizations, nothing else: the target of ncu captures.

    python profiles/run_once.py [--workload cfg2] [--reps 1] [--peak]

--peak: only the register-resident DMMA loop (b200_dmma_peak_tflops), fresh and after ~2 s of load, so that its
tensor-pipe counters can be set beside the update kernel's."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                   # noqa: E402  (workload table + analyze)
from suitesparse_b200 import b200              # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="cfg2")
ap.add_argument("--reps", type=int, default=1)
ap.add_argument("--peak", action="store_true")
ap.add_argument("--single", action="store_true")
args = ap.parse_args()
if args.peak:
    # under `ncu --profile-from-start off` only the launches between cudaProfilerStart / Stop are captured: three
    # measurements on the idle GPU, then three more after ~2 s of continuous DMMA work
    import torch
    rt = torch.cuda.cudart()
    rt.cudaProfilerStart()
    print("fresh", [round(b200.dmma_peak_tflops(0, 40000), 2) for _ in range(3)])
    rt.cudaProfilerStop()
    t0 = time.time()
    while time.time() - t0 < 2.0:
        b200.dmma_peak_tflops(0, 40000)
    rt.cudaProfilerStart()
    print("after 2 s of load", [round(b200.dmma_peak_tflops(0, 40000), 2) for _ in range(3)])
    rt.cudaProfilerStop()
    sys.exit(0)
kind, size, ordering, desc = bench.WORKLOADS[args.workload]
ref, _ = bench.analyze(kind, size, ordering, single=args.single)
sym = ref.symbolic()
Sp, Si, Sx = ref.seam()
F = b200.Factor(sym, dtype=np.float32 if args.single else np.float64, device=0)
F.upload_matrix(Sp, Si, Sx)
for _ in range(args.reps):
    assert F.factor_device() == b200.OK
    print(desc, "ms_factor", F.getd("ms_factor"), "launches", F.geti("kernel_launches"))
F.close()
