#!/usr/bin/env python
"""bench.py -- cholmod_factorize (supernodal numeric phase) FP64 GFLOP/s on a synthetic 3-D Laplacian.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one numeric factorization of the workload matrix (analysis is done once, untimed, by the
reference's own host code as BASELINE.json's north_star prescribes).  Prints ONE JSON line:

  value      Common->fl / t with A and L resident in HBM (t = device time from CUDA events on the
             library's stream: zero + scatter A + every update/potrf/trsm launch), aggregated over ranks.
  e2e        the same metric through the host-buffer C-ABI call b200_super_numeric_d: H2D of A, the
             factorization, D2H of all of L->x, timed on the host clock around the call.
  roofline   the dominant kernel (128x64 DMMA update) against the FP64 tensor peak measured live on
             this device with a register-resident DMMA loop (MEASURED_PEAKS.json has no FP64 figure).
  cpu_baseline  the reference CPU supernodal path (oracle/_ref, all host cores) on the workload itself: one
             factorization, whose factor is also the full-size parity check (`check.max_dL_rel`).

`--impl reference` times the reference's CPU cholmod_l_factorize on the SAME workload (all host cores).  One
config-2 factorization takes the better part of a minute on the box, so when `--steps K --warmup W` would not
fit the time budget (B200_REF_BUDGET_S, default 600 s) it runs 1 warm-up and as many timed factorizations as
fit (at least 3) and prints the `steps` it actually ran.
`--impl reference-gpu` does the same with the reference's own cuBLAS-offload GPU branch (CHOLMOD/GPU built
for sm_100 by oracle/Makefile into oracle/_ref_cuda/), the kernel-level bar of SURVEY.md 2.3.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cholmod_factorize_fp64_gflops"
UNIT = "GFLOP/s"
WORKLOADS = {
    # name: (generator, size, CHOLMOD ordering code, description)
    "cfg2": ("lap3d7", 100, 3, "3-D 7-point Laplacian 100^3 (n=1M), double, METIS"),
    "cfg1": ("lap2d5", 100, 2, "2-D 5-point Laplacian 100x100 (n=10k), double, AMD"),
    "cfg5": ("lap2d9", 2000, 3, "2-D 9-point Laplacian 2000x2000 (n=4M), double, METIS"),
    "cfg3": ("el27", 80, 3, "3-D 27-point 3-dof elasticity-style 80^3 (n=1.5M), double, METIS"),
    "cfg4": ("lap3d7", 160, 3, "3-D 7-point Laplacian 160^3 (n=4.1M), double, METIS"),
    "lap3d7_64": ("lap3d7", 64, 3, "3-D 7-point Laplacian 64^3 (n=262k), double, METIS"),
    "lap3d7_40": ("lap3d7", 40, 3, "3-D 7-point Laplacian 40^3 (n=64k), double, METIS"),
}
def analyze(kind, size, ordering, nthreads=0, single=False, use_gpu=False):
    """Reference host analysis (ordering, etree, supernodes) + the seam matrix."""
    from oracle import refd
    from suitesparse_b200 import matrices
    gen = getattr(matrices, kind)
    n, Ap, Ai, Ax = gen(size)
    ref = refd.Ref(nthreads)
    if use_gpu:
        ref.seti("useGPU", 1)
    ref.set_matrix(n, Ap, Ai, Ax, -1, single=single)
    t = time.time()
    ref.analyze(ordering, refd.CHOLMOD_SUPERNODAL)
    t_an = time.time() - t
    return ref, t_an


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def ncu_traffic(avg_launch_ms):
    """DRAM bytes per launch of the dominant kernel, from the committed `ncu --set full` capture of the benchmarked
    workload (profiles/r02_ncu_full_dmma_root.json: 22 launches of k_upd_dmma<128,64> at the root of config 2, the
    multi-thousand-tile descendant and look-ahead updates that carry the step among them).  The capture holds some
    launches, not all 1355, so the figure is the kernel's DRAM rate over the capture (bytes read + written / time)
    times the average launch duration measured live -- per launch, like `achieved`.  A number taken under a
    profiler, reported only as the traffic figure."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_full_dmma_root.json")))
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
        nbytes = ms = 0.0
        for r in d["launches"]:
            rd, wr, du = r.get("dram_read"), r.get("dram_write"), r.get("duration")
            if not all(isinstance(v, float) and v == v for v in (rd, wr, du)):
                continue
            nbytes += rd * scale.get(r.get("dram_read_unit", "byte"), 1.0) + wr * scale.get(r.get("dram_write_unit", "byte"), 1.0)
            ms += du * tscale.get(r.get("duration_unit", "ms"), 1.0)
        return nbytes / ms * avg_launch_ms if ms > 0 else None
    except Exception:
        return None


def reference_config(desc, sym, esize=8):
    """The `config` both arms print, key for key: the workload, named and sized the same way, and the cache rule of
    the timing (every step rewrites all of L->x, which is far larger than any cache on either side)."""
    return {"workload": desc, "n": int(sym["n"]), "nsuper": int(sym["nsuper"]), "fl": sym["fl"], "lnz": sym["lnz"],
            "xsize": int(sym["xsize"]),
            "l2": "inputs larger than L2 (L->x = %.2f GB is rewritten every step)" % (sym["xsize"] * esize / 1e9)}


def run_reference(args, gpu_branch=False):
    """The reference's own cholmod_l_factorize on the workload: CPU supernodal path (all host cores), or with
    gpu_branch its cuBLAS-offload GPU path (oracle/_ref_cuda, Common->useGPU = 1)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t_start = time.time()
    budget = float(os.environ.get("B200_REF_BUDGET_S", "600"))
    ncores = os.cpu_count() or 1
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(ncores))
    impl = "reference-gpu" if gpu_branch else "reference"
    if gpu_branch:
        os.environ["B200_REFD_FLAVOUR"] = "cuda"        # oracle/refd.py loads oracle/_ref_cuda/ instead of _ref/
        os.environ["CHOLMOD_USE_GPU"] = "1"
        from oracle import refd
        if not refd.have_ref():
            print(json.dumps({"impl": impl, "unavailable": "oracle/_ref_cuda (reference built with CHOLMOD_HAS_CUDA) is not here"}))
            return
    kind, size, ordering, desc = WORKLOADS[args.workload]
    ref, t_an = analyze(kind, size, ordering, ncores, use_gpu=gpu_branch)
    sym = ref.symbolic()
    fl = sym["fl"]
    st = ref.factorize(0.0)                  # warm-up 1 (also the estimate of what fits)
    assert st == 0, st
    t1 = ref.getd("t_factor")
    warm, steps = args.warmup, args.steps
    left = budget - (time.time() - t_start)
    if (max(warm, 1) - 1 + steps) * t1 > left:
        warm = 1
        steps = int(min(args.steps, max(3, left // max(t1, 1e-9))))
    for _ in range(max(warm, 1) - 1):
        ref.factorize(0.0)
    ts = []
    for _ in range(steps):
        st = ref.factorize(0.0)
        assert st == 0, st
        ts.append(ref.getd("t_factor"))
    b = __import__("suitesparse_b200.matrices", fromlist=["rhs"]).rhs(sym["n"])
    resid = ref.resid(ref.solve(b), b)
    cores = ref.geti("nthreads_max")
    t = float(np.sum(ts))
    gf = fl * steps / t / 1e9
    ms = 1e3 * t / steps
    sample = "the workload itself: %d timed cholmod_l_factorize calls after %d warm-up (median %.2f s)" % (
        steps, max(warm, 1), float(np.median(ts)))
    out = {
        "impl": impl, "metric": METRIC, "value": gf, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": max(warm, 1), "steps_requested": args.steps, "warmup_requested": args.warmup,
        "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": reference_config(desc, sym),
        "cpu_baseline": {"value": gf, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": gf, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "analyze_s": t_an, "scaled_residual": resid,
        "blas": os.path.basename(os.environ.get("B200_REF_BLAS", "OpenBLAS 0.3.15 (opencv wheel), %d threads" % ncores)),
    }
    if gpu_branch:
        out["reference_gpu"] = {"useGPU": ref.geti("useGPU"), "nsuper": int(sym["nsuper"]),
                                "note": "CHOLMOD/GPU cuBLAS-offload branch, nvcc -arch=sm_100, one GPU; its analysis "
                                        "limits supernode sizes to the device buffers, so nsuper differs from the CPU build",
                                "gpu_stats": dict({k: ref.geti(k) for k in ("gpu_syrk_calls", "gpu_gemm_calls", "gpu_potrf_calls",
                                                                              "gpu_trsm_calls", "cpu_syrk_calls", "cpu_gemm_calls",
                                                                              "cpu_potrf_calls", "cpu_trsm_calls")},
                                                  **{k: ref.getd(k) for k in ("gpu_syrk_time", "gpu_gemm_time", "gpu_potrf_time",
                                                                              "gpu_trsm_time")})}
    ref.close()
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "reference-gpu"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--sweep", default="", help="diagnostic: ';'-separated planner settings, each a ','-separated list of "
                    "ENV=value (e.g. 'CHOLMOD_B200_DIST_RUN=1;CHOLMOD_B200_DIST_RUN=4,CHOLMOD_B200_FAR_MIN=100000000'); every "
                    "setting gets its own context and is timed like `value`, after the main measurement (key `sweep`)")
    ap.add_argument("--single", action="store_true", help="single precision (config 5b); the roofline stays FP64-tensor "
                    "and is then only informative: single precision runs on the FP32 FMA pipe")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl != "b200":
        return run_reference(args, gpu_branch=(args.impl == "reference-gpu"))

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    import torch.distributed as dist
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from suitesparse_b200 import b200
    if not b200.probe(local):
        raise SystemExit("bench.py: no sm_100 device %d visible (there is no CPU fallback)" % local)
    # FP64 tensor (DMMA) peak of this device, measured live: burst = a 5 ms register-resident DMMA loop on
    # an idle GPU; sustained = the same loop after ~1.5 s of back-to-back runs (the rate a kernel inside a
    # long step can hope for)
    peak_burst = max(b200.dmma_peak_tflops(local, 40000) for _ in range(3)) if rank == 0 else 0.0

    kind, size, ordering, desc = WORKLOADS[args.workload]
    fdtype = np.float32 if args.single else np.float64
    if args.single:
        desc = desc.replace("double", "single")
    ref, t_an = analyze(kind, size, ordering, single=args.single)
    sym = ref.symbolic()
    Sp, Si, Sx = ref.seam()
    fl = sym["fl"]
    t0 = time.time()
    # N > 1: ONE factorization partitioned over the N GPUs: every rank owns (stores, factorizes, copies out) some
    # etree subtrees; what they send into the top separators is pulled, as contribution blocks, by the owners of
    # the separators' 256-column blocks over NVLink; the separators are factorized 1-D block cyclic with their
    # finished panels broadcast.  Strong scaling: value = fl / t.
    F = b200.Factor(sym, dtype=fdtype, device=local, nranks=world, rank=rank)
    if world > 1:
        F.comm_init_torch()
    t_plan = time.time() - t0
    # the caller's ONE host L->x.  With several ranks it is a shared mapping: every rank copies the part of the
    # factor it owns into it over its own PCIe link.
    shm_path = None
    Lx_host = None
    if world == 1:
        Lx_host = np.empty(sym["xsize"], dtype=fdtype)
    elif not args.no_e2e:
        shm_path = "/dev/shm/b200_bench_%s.bin" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            Lx_host = np.memmap(shm_path, dtype=fdtype, mode="w+", shape=(sym["xsize"],))
        dist.barrier()
        if rank != 0:
            Lx_host = np.memmap(shm_path, dtype=fdtype, mode="r+", shape=(sym["xsize"],))
    # what a one-shot cholmod_factorize pays on top of a steady-state step: the context (plan + tables + arena,
    # `plan_s`) and, in the first host-buffer call, page-locking the caller's L->x (cudaHostRegister)
    first_call_ms = None
    if world == 1 and not args.no_e2e:
        t0 = time.time()
        st, _ = F.super_numeric(Sp, Si, Sx, Lx=Lx_host)
        assert st == b200.OK, st
        first_call_ms = 1e3 * (time.time() - t0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident metric: A in HBM -> L in HBM ---------------------------------------------
    F.upload_matrix(Sp, Si, Sx)
    for _ in range(args.warmup):
        st = F.factor_device()
        assert st == b200.OK, st
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    wall0 = time.time()
    ms_dev = 0.0
    launches = 0
    for _ in range(args.steps):
        st = F.factor_device()
        assert st == b200.OK, st
        ms_dev += F.getd("ms_factor")
        launches += F.geti("kernel_launches")
    barrier()
    wall = time.time() - wall0
    clocks = sampler.stop()
    if world > 1:
        tt = torch.tensor([ms_dev], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_dev = float(tt.item())
    ms_step = ms_dev / args.steps
    value = fl / (ms_step * 1e-3) / 1e9

    # ---- end to end through the host-buffer C-ABI call ----------------------------------------------
    e2e = None
    if not args.no_e2e:
        def one_e2e_step():
            # every rank: host A in (its own PCIe link), its share of the factor out into the one host L->x
            st, _ = F.super_numeric(Sp, Si, Sx, Lx=Lx_host)
            assert st == b200.OK
        for _ in range(2):                      # first call pins the caller's L->x (once per factor object)
            one_e2e_step()
        barrier()
        t0 = time.time()
        for _ in range(args.steps):
            one_e2e_step()
        barrier()
        t_e2e = (time.time() - t0) / args.steps
        if world > 1:
            tt = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_e2e = float(tt.item())
        e2e = {"value": fl / t_e2e / 1e9, "unit": UNIT, "ms_per_step": 1e3 * t_e2e,
               "h2d_bytes_per_step": int(Sp.nbytes + Si.nbytes + Sx.nbytes) * world,
               "d2h_bytes_per_step": int(sym["xsize"] * np.dtype(fdtype).itemsize),
               "ms_h2d": F.getd("ms_h2d"), "ms_factor": F.getd("ms_factor"), "ms_d2h": F.getd("ms_d2h"),
               "first_call_ms": first_call_ms, "context_build_ms": 1e3 * t_plan,
               "e2e_first_call_ms": (first_call_ms + 1e3 * t_plan) if first_call_ms is not None else None}

    # ---- diagnostic sweep over planner settings (collective: every rank builds and runs every variant) ----------
    sweep = []
    for setting in [x for x in args.sweep.split(";") if x.strip()]:
        kv = dict(item.split("=", 1) for item in setting.split(","))
        old = {k: os.environ.get(k) for k in kv}
        os.environ.update(kv)
        try:
            F2 = b200.Factor(sym, dtype=fdtype, device=local, nranks=world, rank=rank)
        finally:
            for k, v in old.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
        if world > 1:
            F2.comm_init_torch()
        F2.upload_matrix(Sp, Si, Sx)
        for _ in range(3):
            assert F2.factor_device() == b200.OK
        barrier()
        ms2 = 0.0
        for _ in range(args.steps):
            assert F2.factor_device() == b200.OK
            ms2 += F2.getd("ms_factor")
        barrier()
        if world > 1:
            tt = torch.tensor([ms2], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms2 = float(tt.item())
        sweep.append({"setting": setting, "ms_per_step": ms2 / args.steps})
        F2.close()

    # one profiled (untimed) factorization for the per-kernel breakdown; collective, so every rank runs it
    F.seti("profile", 1)
    F.factor_device()
    F.seti("profile", 0)
    nccl_calls = F.geti("nccl_calls")
    per_rank = None
    if world > 1:
        mine = dict(rank=rank, ms_factor_profiled=F.getd("ms_factor"), host_registered_bytes=F.geti("host_registered"),
                    level_end_ms=[round(F.getd("ms_level_end%d" % k), 2) for k in range(F.geti("nlevels"))],
                    kernel_ms={nm: round(F.getd("ms_kind%d" % k), 2) for k, nm in enumerate(b200.KIND_NAMES)},
                    device_bytes=F.geti("device_bytes"), arena_bytes=F.geti("dxsize") * np.dtype(fdtype).itemsize,
                    nvlink_in_bytes=F.geti("nvlink_in_bytes"), ms_pull=F.getd("ms_pull"), npulls=F.geti("npulls"),
                    d2h_bytes=sum(int(c["len"]) for c in F.plan_table("owned", b200.COPY_DT)) * np.dtype(fdtype).itemsize)
        if os.environ.get("B200_BENCH_TIMELINE") and rank in (0, 3):
            # diagnostic: the profiled pass's device timeline of the upper levels, launch by launch
            lt = F.plan_table("launches", b200.LAUNCH_DT)
            lo = int(os.environ["B200_BENCH_TIMELINE"])
            mine["timeline"] = [(int(L["kind"]), int(L["stream"]), int(L["level"]), int(L["count"]),
                                 round(F.getd("ms_launch_start%d" % i), 3), round(F.getd("ms_launch%d" % i), 3))
                                for i, L in enumerate(lt) if L["level"] >= lo and L["kind"] < b200.K_JOIN and L["count"] > 0]
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
    if rank != 0:
        F.close()
        del Lx_host
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel, from one profiled (untimed) factorization --------------------
    # best of a few: a single short run right after a long step can catch the clocks mid-ramp
    sustained = [b200.dmma_peak_tflops(local, 40000) for _ in range(150)]
    peak_sustained = float(np.median(sustained[-40:]))
    # The dominant kernel is timed inside a long step, so its roofline denominator is the SUSTAINED figure: an idle
    # GPU runs the first burst of DMMAs at 37 TFLOP/s and settles at 29.7 after ~1 s of continuous FP64 tensor work
    # (same reported SM clock, no throttle reason); `peak_burst` is reported next to it.
    peak = peak_sustained if peak_sustained > 0 else peak_burst
    kinds = {}
    for k, nm in enumerate(b200.KIND_NAMES):
        kinds[nm] = {"ms": F.getd("ms_kind%d" % k), "launches": F.geti("launches_kind%d" % k),
                     "gflop": F.getd("flops_kind%d" % k) / 1e9}
    # device timeline of the profiled pass: when (ms after the first launch) each etree level was finished
    timeline = [round(F.getd("ms_level_end%d" % k), 3) for k in range(F.geti("nlevels"))]
    dom = max(("upd_dmma128", "upd_dmma64"), key=lambda nm: kinds[nm]["ms"])
    dk = kinds[dom]
    # Launches of the dominant kernel overlap each other (look-ahead stream, subtree-group streams), so the sum
    # of their durations counts shared time twice.  The kernel's time is the length of the union of its launch
    # intervals on the device timeline; `avg_launch_ms` = that time / launches.
    launches_tab = F.plan_table("launches", b200.LAUNCH_DT)
    dom_kind = b200.KIND_NAMES.index(dom)
    iv = sorted((F.getd("ms_launch_start%d" % i), F.getd("ms_launch_start%d" % i) + F.getd("ms_launch%d" % i))
                for i in np.nonzero((launches_tab["kind"] == dom_kind) & (launches_tab["count"] > 0))[0])
    union_ms, cur_a, cur_b = 0.0, None, None
    for a, b in iv:
        if cur_b is None or a > cur_b:
            union_ms += (cur_b - cur_a) if cur_b is not None else 0.0
            cur_a, cur_b = a, b
        else:
            cur_b = max(cur_b, b)
    union_ms += (cur_b - cur_a) if cur_b is not None else 0.0
    dk["sum_ms"] = dk["ms"]
    dk["ms"] = union_ms if union_ms > 0 else dk["ms"]
    ach = (dk["gflop"] / 1e3) / (dk["ms"] * 1e-3) if dk["ms"] > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "k_upd_dmma<%s>" % ("128,64,32,64" if dom == "upd_dmma128" else "64,64,32,32"),
                "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak > 0 else None,
                "traffic": ncu_traffic(dk["ms"] / max(1, dk["launches"])),
                "peak_burst": peak_burst, "peak_sustained": peak_sustained,
                "frac_of_sustained": ach / peak_sustained if peak_sustained > 0 else None,
                "peak_source": "FP64 DMMA peak measured live by b200_dmma_peak_tflops (MEASURED_PEAKS.json has no FP64 "
                               "figure): `peak` = `peak_sustained` = after 1.5 s of back-to-back runs (the kernel is timed "
                               "inside a long step), `peak_burst` = first run on an idle GPU",
                "flops_per_launch": dk["gflop"] * 1e9 / max(1, dk["launches"]),
                "avg_launch_ms": dk["ms"] / max(1, dk["launches"]), "share_of_step": dk["ms"] / max(1e-9, max(timeline) if timeline else dk["ms"]),
                "timing": "union of the kernel's launch intervals in one profiled pass (CUDA events around every launch); "
                          "sum of launch durations = %.1f ms" % dk["sum_ms"]}
    whole = {"achieved_tflops": fl / (ms_step * 1e-3) / 1e12, "frac_of_fp64_peak": fl / (ms_step * 1e-3) / 1e12 / peak if peak > 0 else None,
             "executed_tflops": F.getd("flops_executed") / (ms_step * 1e-3) / 1e12}
    # the HBM-bound part: zero all of L->x, then scatter A into it.  Algorithmic bytes =
    # 8 per value of L written by the memset + (8 value + 8 row index read, 8 written) per entry of A.
    esz = np.dtype(fdtype).itemsize
    ms_scatter = F.getd("ms_scatter")
    hbm_peak = None
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    zs_bytes = esz * sym["xsize"] + (esz + 8 + esz) * len(Si)
    hbm_phase = {"kernel": "cudaMemsetAsync + k_scatter_A", "bound": "hbm", "bytes": int(zs_bytes), "ms": ms_scatter,
                 "achieved": zs_bytes / (ms_scatter * 1e-3) / 1e9 if ms_scatter > 0 else None,
                 "peak": hbm_peak if hbm_peak else 6650.0, "unit": "GB/s",
                 "peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm_peak else "fallback 6.65 TB/s (B200_PROFILING.md)"}
    if hbm_phase["achieved"]:
        hbm_phase["frac"] = hbm_phase["achieved"] / hbm_phase["peak"]

    # ---- full-size sanity of the factor just computed: device solve + the demo's scaled residual ----------
    from suitesparse_b200 import matrices
    bvec = matrices.rhs(sym["n"])
    xs = np.empty(sym["n"])
    # several ranks: no rank holds the whole factor; the assembled host L->x goes into a one-GPU context
    if world > 1 and Lx_host is None:
        # --no-e2e with several ranks: the factor never left the GPUs, and no rank holds all of it
        check = {"skipped": "no host copy of L->x in this run (--no-e2e): the solve check needs the assembled factor"}
    else:
        G = F if world == 1 else b200.Factor(sym, dtype=fdtype, device=local)
        if world > 1:
            assert np.all(np.isfinite(Lx_host[::4097])), "the ranks' copy-outs left holes in L->x"
            G.upload_factor(np.array(Lx_host))
        xs[sym["Perm"]] = G.solve(bvec[sym["Perm"]])
        solve_first_ms = G.getd("ms_solve")     # builds the solve schedule and inverts the wide supernodes' diagonal blocks
        xs[sym["Perm"]] = G.solve(bvec[sym["Perm"]])
        check = {"solve_ms": G.getd("ms_solve"), "solve_first_ms": solve_first_ms, "scaled_residual": ref.resid(xs, bvec),
                 "tolerance": 1e-5 if args.single else 1e-12}
        if world > 1:
            G.close()

    # ---- CPU baseline: the reference's CPU path on the workload itself, rank 0, N = 1 only; its factor is the
    # full-size parity check of ours (north_star: max|dL|/max|L| <= 1e-10) ----------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        ncores = os.cpu_count() or 1
        ref.seti("nthreads_max", ncores)
        st = ref.factorize(0.0)
        assert st == 0, st
        t_ref = ref.getd("t_factor")
        cpu = {"value": fl / t_ref / 1e9, "unit": UNIT, "cores": ref.geti("nthreads_max"), "kind": "reference",
               "sample": "the workload itself, one cholmod_l_factorize (numeric phase) after the shared analysis",
               "ms_per_factorization": 1e3 * t_ref}
        if not args.no_e2e:
            Lref = ref.Lx()
            scale, err = float(np.max(np.abs(Lref))), 0.0
            for a in range(0, Lref.size, 1 << 26):
                err = max(err, float(np.max(np.abs(Lx_host[a:a + (1 << 26)].astype(np.float64) - Lref[a:a + (1 << 26)]))))
            check["max_dL_rel"] = err / scale
            check["max_dL_tolerance"] = 1e-4 if args.single else 1e-10

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if args.single else "f64", "data": "synthetic",
        "config": reference_config(desc, sym, np.dtype(fdtype).itemsize),
        "plan": {"npairs": F.geti("npairs"), "nlevels": F.geti("nlevels"),
                   "relmap_len": F.geti("relmap_len"), "device_bytes": F.geti("device_bytes"),
                   "parallelism": ("one factorization over %d GPUs: etree subtrees owned per GPU, their contribution blocks "
                                   "pulled over NVLink by the owners of the top separators' blocks, top separators 1-D "
                                   "block cyclic (%d NCCL calls per step, %d top supernodes), every GPU copies its share of "
                                   "L->x to the host" % (world, nccl_calls, F.geti("ntop")))
                   if world > 1 else "1 GPU",
                   "analyze_s": t_an, "plan_s": t_plan, "wall_s_timed_region": wall},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "whole_step": whole,
        "check": check, "hbm_phase": hbm_phase, "per_rank": per_rank, "sweep": sweep,
        "nvlink_bytes_per_step": sum(p["nvlink_in_bytes"] for p in per_rank) if per_rank else 0,
        "kernels": kinds, "level_end_ms": timeline, "cpu_baseline": cpu,
    }
    print(json.dumps(out))
    F.close()
    del Lx_host
    if world > 1:
        dist.barrier()
        if shm_path and os.path.exists(shm_path):
            os.unlink(shm_path)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
