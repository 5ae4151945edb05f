#!/usr/bin/env python
"""bench.py -- fp64 action+gradient evaluations per second (batched) on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W]                 # our arm (CUDA, sm_100a)
    python bench.py --impl reference [--gpus N] [--steps K] [--warmup W] # the reference's CPU implementation

Workload (BASELINE.json configs[1]): system 3B (3 bodies, no symmetry, 3 RFFT loops),
nint = 4096 = 2^12, nparams = 24570, a batch of 4096 random states PER GPU, unit U1 = one fused
(action, gradient) evaluation (== NBodySyst.params_to_action + params_to_action_grad).  A "step" is one pass of
the hot path over the whole batch.  Multi-GPU: the batch of independent states is sharded, one process per GPU,
no data-path collective (scaling = weak: per-GPU batch fixed).

Legs printed on ONE JSON line by rank 0:
  value       U1/s with the parameters already resident in HBM (CUDA events on the launching stream, max over ranks)
  e2e         U1/s through the public host API (NBodySystB200.params_to_action_and_grad on pinned NumPy buffers):
              H2D of the step's parameters and D2H of actions + gradients inside the timed region
  roofline    algorithmic bytes 8*(2*nparams+1) per U1 x batch / launch duration vs measured HBM peak
  cpu_baseline  the reference (oracle/_ref, unmodified choreo) timed on 1 host core on a bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "fp64 action+grad evals/s (batched)"
UNIT = "evals/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--system", default="3B")
    ap.add_argument("--nint", type=int, default=4096)
    ap.add_argument("--batch", type=int, default=4096, help="states per GPU")
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--no-fused", action="store_true")
    ap.add_argument("--fused-ver", type=int, default=4, help="4 = newest fused kernel the layout supports (default), 3 = force version 3")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-sample", type=int, default=0, help="states in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--ref-sample", type=int, default=0, help="states per step of the reference arm (0 = auto)")
    ap.add_argument("--stages", action="store_true", help="also time the staged operators separately")
    ap.add_argument("--no-newton", action="store_true", help="skip the Newton orbit solves/hour leg")
    ap.add_argument("--newton-batch", type=int, default=8192, help="random starts per GPU in the solves/hour leg")
    ap.add_argument("--newton-workers", type=int, default=4, help="chunks of starts in flight per GPU in the solves/hour leg (threads, one stream each)")
    ap.add_argument("--newton-chunk", type=int, default=1024, help="starts per chunk in the solves/hour leg")
    ap.add_argument("--newton-reps", type=int, default=5, help="timed repetitions of the solves/hour leg (median reported)")
    ap.add_argument("--newton-ref-seeds", type=int, default=0, help="seeds of the CPU arm of the solves/hour leg (0 = auto)")
    ap.add_argument("--no-configs", dest="configs", action="store_false", help="skip the extra BASELINE configurations")
    ap.add_argument("--no-pin-numa", dest="pin_numa", action="store_false",
                    help="do not bind a rank to the CPU set of its GPU (nvidia-smi topo -m) when several ranks share the host")
    return ap.parse_args()


def workload_config(args, nparams, world):
    return {
        "workload": f"{args.system} nint={args.nint} nparams={nparams} batch={args.batch}/GPU x {world} GPU: "
                    "fused action+gradient (U1) per state, rng(1234).random params",
        "system": args.system, "nint": args.nint, "nparams": int(nparams), "batch_per_gpu": args.batch,
        "global_batch": args.batch * world, "parallelism": f"batch-shard x{world} (no collective)",
        "l2": "inputs larger than L2 (params+grad = %.0f MB per step per GPU)" % (2 * args.batch * nparams * 8 / 1e6),
    }


def gpu_cpu_affinity(index):
    """CPU set nvidia-smi reports for a GPU (its NUMA node), or None."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        for line in out.splitlines():
            f = line.split()
            if f and f[0] == f"GPU{index}":
                for tok in f[1:]:
                    if tok[0].isdigit() and ("-" in tok or "," in tok):
                        cpus = set()
                        for part in tok.split(","):
                            a, _, b = part.partition("-")
                            cpus.update(range(int(a), int(b or a) + 1))
                        return cpus
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.lines:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.1):
                continue
            f = [s.strip() for s in line.split(",")]
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU baselines
def _ref_or_port(system, nint):
    """(kind, evaluator) : the unmodified reference when oracle/_ref travelled here, else the NumPy oracle port."""
    from oracle import ref_loader
    if ref_loader.reference_available():
        try:
            NBS = ref_loader.build_named(system)
            NBS.nint = nint
            return "reference", NBS
        except Exception as e:  # pragma: no cover
            print(f"[bench] reference unavailable ({e}); using the oracle port", file=sys.stderr)
    from oracle.choreo_oracle import OracleSyst
    from choreographies2_b200 import PlanTables
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{system}.npz"))
    return "port", OracleSyst(T, nint, backend="fft")


def _limit_threads():
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
        os.environ[k] = "1"  # as the reference does (find.py:840, scripts/NewSymsTime.py:8-14)


def cpu_eval_loop(obj, x):
    s = 0.0
    for i in range(x.shape[0]):
        s += obj.params_to_action(x[i])
        s += float(obj.params_to_action_grad(x[i])[0])
    return s


def cpu_baseline(args, nparams):
    _limit_threads()
    try:
        from threadpoolctl import threadpool_limits
        ctx = threadpool_limits(1)
    except Exception:
        ctx = None
    kind, obj = _ref_or_port(args.system, args.nint)
    rng = np.random.default_rng(1234)
    # calibrate so that the sample is ~10-20 s of single-core work
    probe = rng.random((8, nparams))
    t = time.perf_counter(); cpu_eval_loop(obj, probe); per = (time.perf_counter() - t) / 8
    n = args.cpu_sample or int(min(args.batch, max(16, 4.0 / max(per, 1e-6))))
    x = rng.random((n, nparams))
    best = float("inf")
    for _ in range(3):
        t = time.perf_counter(); cpu_eval_loop(obj, x); best = min(best, time.perf_counter() - t)
    # what find.py:267 actually calls per Newton / line-search evaluation is the gradient alone: reported beside U1
    bestg = float("inf")
    for _ in range(3):
        t = time.perf_counter()
        for i in range(n):
            obj.params_to_action_grad(x[i])
        bestg = min(bestg, time.perf_counter() - t)
    if ctx is not None:
        ctx.restore_original_limits() if hasattr(ctx, "restore_original_limits") else None
    return {"value": n / best, "unit": UNIT, "cores": 1, "kind": kind, "grad_only_per_s": n / bestg,
            "note": "U1 on the CPU = params_to_action + params_to_action_grad (two forward transforms; the action goes through the "
                    "function-pointer law, _NBodySyst.pyx:6425) against ONE fused pass on the GPU; grad_only_per_s is the gradient alone",
            "sample": f"{n} states of the same workload ({args.system}, nint {args.nint}), params_to_action + "
                      f"params_to_action_grad per state, best of 3, 1 thread, FFT backend scipy"}


_W = {}


def _ref_worker_init(system, nint, nparams=0, nmax=0):
    _limit_threads()
    try:
        from threadpoolctl import threadpool_limits
        _W["lim"] = threadpool_limits(1)
    except Exception:
        pass
    _W["kind"], _W["obj"] = _ref_or_port(system, nint)
    # every worker owns its shard of the seeded batch (no per-step pickling of inputs)
    _W["x"] = np.random.default_rng(1234 + os.getpid() % 1000).random((nmax, nparams)) if nmax else None


def _ref_worker_run(n):
    return cpu_eval_loop(_W["obj"], _W["x"][:n])


def _ref_worker_kind(_):
    return _W["kind"]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    from choreographies2_b200 import PlanTables
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{args.system}.npz"))
    nparams = T.level(args.nint).nparams
    P = os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    nmax = (args.batch * max(1, args.gpus) + P - 1) // P
    with ctx.Pool(P, initializer=_ref_worker_init, initargs=(args.system, args.nint, nparams, nmax)) as pool:
        kind = pool.map(_ref_worker_kind, range(P), chunksize=1)[0]
        # calibrate: the whole run should take ~90 s at most, every process gets the same number of states
        t = time.perf_counter(); pool.map(_ref_worker_run, [4] * P, chunksize=1); per = (time.perf_counter() - t) / 4
        budget = 90.0 / max(1, args.steps + args.warmup)
        per_proc = (args.ref_sample // P) if args.ref_sample else int(budget / max(per, 1e-6))
        per_proc = max(1, min(nmax, per_proc))
        n = per_proc * P
        for _ in range(args.warmup):
            pool.map(_ref_worker_run, [per_proc] * P, chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_ref_worker_run, [per_proc] * P, chunksize=1)
        dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = (f"{n} states per step of the same workload, split over {P} processes (the reference's own scaling model, "
              f"find.py:836-850), 1 BLAS thread each, FFT backend scipy")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, nparams, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": P, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_newton:  # second half of the metric on the same host cores: the reference's search loop, P processes
        try:
            line["newton"] = search_cpu_arm(3072, P, max(1, (args.newton_ref_seeds or 4 * P) // P))
        except Exception as exc:
            line["newton"] = {"error": f"{type(exc).__name__}: {exc}"}
    print(json.dumps(line))
    return 0



# refinement depth of the solves/hour leg, both arms.  The reference's 3C configuration allows 4 doublings of nint; the rare starts
# that run into a collision then spend minutes per seed in the REFERENCE at nint 24576 / 49152 (1000 Newton iterations per optimiser
# level, 3 ms per gradient), which no bounded CPU sample can afford.  One doubling keeps the loop, the line search and the six optimiser
# levels of the configuration; the full-depth comparison is profiles/r02_find_compare_3C_nintfac600_32seeds.txt.
NEWTON_RECONVERGE = 1
FP64_PEAK_TFLOPS = 37.0  # measured FMA peak of this pool's B200 (tools/fp64_peak.cu, profiles/r01_fp64_pcie_peaks.json)


def alg_flops_per_u1(T, nint, nparams):
    """Algorithmic FP64 flops of one action + gradient evaluation: two transforms (2.5 N log2 N per real
    sequence of N samples) + the pair loop (8 G + 10 per interaction and sample) + the kinetic term."""
    G, nbin = T.geodim, T.nbin
    lvl = T.level(nint)
    nseq = sum((G if int(T.shortcut[il]) == 1 else int(T.nnz[il] * T.nppl[il])) * 2.5 * (2 * int(lvl.npr[il])) *
               np.log2(max(2.0, 2.0 * int(lvl.npr[il]))) for il in range(T.nloop))
    return float(2.0 * nseq + nbin * lvl.segm_store * (8 * G + 10) + 4.0 * nparams)


def config_entry(label, system, nint, batch, dev, hbm_peak):
    """U1 rate of one of the other BASELINE configurations, device-resident, with both rooflines of SURVEY.md 8(d)."""
    import torch
    from choreographies2_b200 import NBodySystB200, PlanTables
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{system}.npz"))
    nbs = NBodySystB200(T, nint, device=dev.index)
    x = torch.rand((batch, nbs.nparams), dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(1234))
    for _ in range(3):
        nbs.params_to_action_and_grad(x)
    l0 = nbs.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 5
    e0.record()
    for _ in range(steps):
        nbs.params_to_action_and_grad(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    rate = batch / (ms * 1e-3)
    # algorithmic work per U1: bytes 8 (2 nparams + 1); flops = two transforms + pair loop
    G, S = T.geodim, T.nsegm
    lvl = T.level(nint)
    flops = alg_flops_per_u1(T, nint, nbs.nparams)
    alg_bytes = 8 * (2 * nbs.nparams + 1)
    staged_bytes = alg_bytes + 32 * S * lvl.segm_store * G
    fused = bool(nbs.fused)
    return {"workload": f"{label}, batch {batch}, U1", "value": rate, "unit": UNIT, "ms_per_step": ms, "fused_version": int(nbs.fused_version),
            "staged_kernels": list(nbs.staged_kernels), "gpu_launches_per_step": int((nbs.kernel_launches - l0) // steps),
            "roofline_hbm": {"alg_bytes_per_unit": alg_bytes, "achieved_gbs": alg_bytes * rate / 1e9, "frac": alg_bytes * rate / 1e9 / hbm_peak,
                             "staged_bytes_per_unit": None if fused else staged_bytes,
                             "frac_on_staged_bytes": None if fused else staged_bytes * rate / 1e9 / hbm_peak},
            "roofline_fp64": {"alg_flops_per_unit": flops, "achieved_tflops": flops * rate / 1e12, "peak_tflops": FP64_PEAK_TFLOPS,
                              "frac": flops * rate / 1e12 / FP64_PEAK_TFLOPS}}


def _search_cpu_worker(task):
    """CPU arm of the solves/hour leg: the reference's search loop (oracle/ref_find.reference_find = unmodified nonlin_solve_pp,
    KrylovJacobian, params_resize ...) on a list of seeds, one process, BLAS threads pinned to 1."""
    seeds, nint0 = task
    _limit_threads()
    from oracle import ref_loader, ref_find
    ref_loader.import_reference()
    NBS = ref_loader.build_named("3C")
    fac0 = nint0 // (2 * int(NBS.nint_min))
    NBS.nint_fac = fac0
    x0 = ref_find.find_starts(NBS, seeds)
    cfg = {k: v for k, v in ref_find.CONFIG_3C.items() if k not in ("nint_fac_init", "coeff_ampl_o", "k_infl", "k_max", "coeff_ampl_min")}
    cfg["n_reconverge_it_max"] = NEWTON_RECONVERGE
    t = time.perf_counter()
    outs = [ref_find.reference_find(NBS, x0[i], fac0, **cfg)[0] for i in range(len(seeds))]
    return time.perf_counter() - t, outs


def search_cpu_arm(nint0, nproc, seeds_per_proc):
    """solves/hour of the reference's search loop on `nproc` host processes (the reference's own scaling model, find.py:836-850)."""
    import multiprocessing as mp
    from oracle import ref_loader
    if not ref_loader.reference_available():
        return None
    tasks = [(list(range(p * seeds_per_proc, (p + 1) * seeds_per_proc)), nint0) for p in range(nproc)]
    t0 = time.perf_counter()
    if nproc == 1:
        res = [_search_cpu_worker(tasks[0])]
    else:
        with mp.get_context("spawn").Pool(nproc) as pool:
            res = pool.map(_search_cpu_worker, tasks, chunksize=1)
    wall = time.perf_counter() - t0
    busy = max(r[0] for r in res)
    outs = [o for r in res for o in r[1]]
    n = len(outs)
    return {"solves_per_hour": n / busy * 3600.0, "seconds_busiest_process": busy, "seconds_wall_incl_startup": wall, "seeds": n, "processes": nproc,
            "saved": int(sum(1 for o in outs if o in (1, 2))), "kind": "reference",
            "sample": f"{n} find.py seeds through the reference's search loop (oracle/ref_find.reference_find: unmodified nonlin_solve_pp + "
                      f"KrylovJacobian(gmres) + params_resize, settings of tests/NewSym_data/3C/choreo_config.json), {nproc} process(es), 1 BLAS thread each"}


def newton_leg(args, rank, world, dev, barrier, max_over_ranks):
    import torch
    from choreographies2_b200 import PlanTables
    from choreographies2_b200 import refine, sweep
    Tn = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nint0, Bn = 3072, int(args.newton_batch)
    W, chunk = max(1, int(args.newton_workers)), max(1, int(args.newton_chunk))
    kw = dict(refine.SEARCH_CONFIG_3C)
    kw["n_reconverge_it_max"] = NEWTON_RECONVERGE
    # warm-up: plans of the levels, workspaces, allocator (every worker thread has its own set of plans)
    warm = sweep.run_search(Tn, nint0, 128 * W, device=dev.index, chunk=128, workers=W, config=kw, seed_offset=rank * Bn)
    plans = warm["plans"]
    nlaunch = lambda: sum(int(pl.kernel_launches) for d in plans for pl in d.values())  # noqa: E731
    times, out_s = [], None
    l0 = nlaunch()
    for _ in range(max(1, args.newton_reps)):
        barrier()
        t0 = time.perf_counter()
        out_s = sweep.run_search(Tn, nint0, Bn, device=dev.index, chunk=chunk, workers=W, config=kw, seed_offset=rank * Bn, plans=plans)
        torch.cuda.synchronize()
        times.append(max_over_ranks(time.perf_counter() - t0))
    rec = out_s["records"]

    class res:  # the fields the summary below reads
        outcome, level = rec["outcome"], rec["level"]
        n_grad_evals, n_hess_evals = out_s["U1"], out_s["U2"]
    distinct = int(np.unique(rec["dup_of_seed"][rec["dup_of_seed"] >= 0]).size)
    launches = (nlaunch() - l0) // max(1, args.newton_reps)
    ts = sorted(times)
    med = ts[len(ts) // 2]
    saved = int(((res.outcome == refine.FOUND) | (res.outcome == refine.APPROX)).sum())
    out = {"solves_per_hour": world * Bn / med * 3600.0, "saved_orbits_per_hour": world * saved / med * 3600.0,
           "seconds_median": med, "seconds_min": ts[0], "seconds_max": ts[-1], "repetitions": len(ts), "batch_per_gpu": Bn,
           "chunk": chunk, "chunks_in_flight_per_gpu": W,
           "system": "3C", "nint0": nint0, "nparams0": int(plans[0][nint0].nparams),
           "settings": "find.py loop with the reference's 3C configuration: GMRES(20), wolfe line search smin 0.1, optimiser levels "
                       "gradtol 1e-1 .. 1e-11 (2-norm) with maxiter 100/1000/1000/1000/500/500, refinement x2 once (n_reconverge_it_max 1, see NEWTON_RECONVERGE), "
                       "Newt_err_norm_max 1e-11 / 1e-8; starts x_avg + x_ptp U[0,1) from Make_params_bounds(0.1, 3, 50, 1e-16); sweep.run_search: chunks of seeds "
                       "through refine.find_batched, several chunks in flight (one stream each), saved orbits hashed and duplicate-filtered per chunk inside the timed region",
           "rank0": {"found": int((res.outcome == refine.FOUND).sum()), "approximate": int((res.outcome == refine.APPROX).sum()),
                     "not_converged": int((res.outcome == refine.NOT_CONVERGED).sum()), "collision": int((res.outcome == refine.COLLISION).sum()),
                     "escaped": int((res.outcome == refine.ESCAPED).sum()), "refined_at_least_once": int((res.level > 0).sum()),
                     "U1_evals": int(res.n_grad_evals), "U2_evals": int(res.n_hess_evals), "gpu_launches": int(launches),
                     "chunk_representatives_after_duplicate_filter": distinct},
           "note": "a solve = one run of the search loop from one random start to termination (saved, abandoned, escaped ...)"}
    if rank == 0 and world == 1 and not args.no_cpu:
        n = args.newton_ref_seeds or 16
        out["cpu_baseline"] = search_cpu_arm(nint0, 1, n)
    return out


# ------------------------------------------------------------------------------------------------ our arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa = None
    if world > 1 and args.pin_numa:
        # several ranks stream pinned buffers through one host: keep each rank (and the first touch of its buffers) on the
        # NUMA node of its GPU
        numa = gpu_cpu_affinity(local)
        if numa:
            try:
                os.sched_setaffinity(0, numa)
            except Exception:
                numa = None
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{args.system}.npz"))
    nbs = NBodySystB200(T, args.nint, device=local)
    if args.chunk:
        nbs.set_chunk(args.chunk)
    nbs.set_fused(0 if args.no_fused else args.fused_ver)
    B, npar = args.batch, nbs.nparams
    rng = np.random.default_rng(1234 + rank)
    x_host = torch.from_numpy(rng.random((B, npar))).pin_memory()
    x = x_host.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value leg: inputs resident in HBM --------------------------------------------------------------------
    for _ in range(args.warmup):
        act, grad = nbs.params_to_action_and_grad(x)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = nbs.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.time()
    e0.record()
    for _ in range(args.steps):
        act, grad = nbs.params_to_action_and_grad(x)
    e1.record()
    barrier()
    tw1 = time.time()
    launches = nbs.kernel_launches - l0
    ms = max_over_ranks(e0.elapsed_time(e1))
    ms_per_step = ms / args.steps
    value = world * B * args.steps / (ms * 1e-3)
    # keep the GPU busy a little longer so that the clock sampler sees the loaded state even for short runs
    t_end = time.time() + max(0.0, 0.6 - (tw1 - tw0))
    while time.time() < t_end:
        nbs.params_to_action_and_grad(x)
        torch.cuda.synchronize()
    clocks = sampler.stop(tw0, time.time()) if rank == 0 else None
    checksum = float(act.sum().item())

    # ---- optional: staged operators timed separately (explains the step) ---------------------------------------
    stages = None
    if args.stages:
        def tm(fn, reps=5):
            fn(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                out = fn()
            b.record(); torch.cuda.synchronize()
            return a.elapsed_time(b) / reps, out
        t_fwd, seg = tm(lambda: nbs.params_to_segmpos(x))
        t_pair, sg = tm(lambda: nbs.segmpos_to_pot_nrg_grad(seg, -1.0))
        t_adj, _ = tm(lambda: nbs.segmpos_to_params_T(sg))
        t_hess, _ = tm(lambda: nbs.segmpos_dparams_to_action_hess(seg, x))
        stages = {"params_to_segmpos_ms": t_fwd, "pot_nrg_grad_ms": t_pair, "segmpos_to_params_T_ms": t_adj,
                  "hess_vec_ms": t_hess, "hess_vec_per_s": B / (t_hess * 1e-3)}
        del seg, sg

    # ---- roofline ------------------------------------------------------------------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    alg_bytes_per_unit = 8 * (2 * npar + 1)
    launches_per_step = max(1, launches // max(1, args.steps))
    achieved = alg_bytes_per_unit * B / (ms_per_step * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(tpath):
        try:
            traffic = json.load(open(tpath)).get("%s_%d_%d_%s" % (args.system, args.nint, B, "fused" if nbs.fused else "staged"))
        except Exception:
            traffic = None
    roofline = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
        "peak_source": peak_src,
        "kernel": (("k_fused4<NSEG,1>" if nbs.fused_version >= 4 else "k_fused<NSEG,1>") +
                   ": 1 launch per step, persistent CTAs (one resident wave), one state at a time per CTA, whole evaluation "
                   "in shared memory, parameter rows stashed in tensor memory" +
                   (", rows moved by cp.async.bulk per warp" if nbs.fused_version >= 4 else "") if nbs.fused else
                   f"staged path: {launches_per_step} launches per step (k_fwd, k_pair<1>, k_adj, k_finish per chunk); "
                   "achieved = whole-step algorithmic bytes / whole-step time"),
        "alg_bytes_per_unit": alg_bytes_per_unit, "units_per_launch": B,
    }
    # second view of the same launch: the FP64 pipe.  B200's FP64 rate is low enough that the fused kernels sit between the two
    # rooflines (for k_fused4 the FP64 issue slots are 24 k of the 50 k cycles per state, the HBM time 17 k)
    try:
        fl = alg_flops_per_u1(T, args.nint, npar)
        roofline["fp64"] = {"alg_flops_per_unit": fl, "achieved_tflops": fl * B / (ms_per_step * 1e-3) / 1e12, "peak_tflops": FP64_PEAK_TFLOPS,
                            "frac": fl * B / (ms_per_step * 1e-3) / 1e12 / FP64_PEAK_TFLOPS,
                            "note": "algorithmic flops (2.5 N log2 N per real sequence, 8 G + 10 per pair sample); executed FP64 instruction "
                                    "counts and pipe utilisation: profiles/r01_fused_v4_ncu_full_summary.md"}
    except Exception:
        pass

    # ---- e2e leg: public host API on pinned NumPy buffers ------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        # pinned host buffers for inputs AND results (allocated once, like a solver would); every step copies the
        # parameters host->device and the actions + gradients device->host inside the timed region
        xh = nbs.pinned_empty((B, npar))
        xh[...] = x_host.numpy()
        a_out, g_out = nbs.pinned_empty((B,)), nbs.pinned_empty((B, npar))
        ksteps = max(1, min(args.steps, 10))
        for _ in range(2):
            nbs.params_to_action_and_grad(xh, out_action=a_out, out_grad=g_out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(ksteps):
            a_h, g_h = nbs.params_to_action_and_grad(xh, out_action=a_out, out_grad=g_out)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        dt = max_over_ranks(dt)
        e2e = {"value": world * B * ksteps / dt, "unit": UNIT, "h2d_bytes_per_step": int(B * npar * 8),
               "d2h_bytes_per_step": int(B * (npar + 1) * 8), "steps": ksteps,
               "numa_pinned_cpus": (len(numa) if numa else None),
               "api": "NBodySystB200.params_to_action_and_grad(pinned numpy in/out) -> chb_run_host(CHB_OP_ACTION_GRAD): "
                      "3-slot stream pipeline of 48 MB slices (H2D, kernels, D2H overlapped)"}
        assert abs(float(a_h.sum()) - checksum) <= 1e-9 * abs(checksum)

    # ---- second half of the metric: Newton orbit solves/hour ----------------------------------------------------------
    # every rank runs the search loop of find.py (refine.find_batched: the reference's 3C settings -- wolfe line search, smin 0.1,
    # six optimiser levels 1e-1 .. 1e-11 in the 2-norm, refinement x2 up to 4 times) on its own shard of find.py-style random
    # starts (seeded by global index) of the 3C system from nint 3072; no collective on the path.  Timed `--newton-reps` times.
    newton = None
    if not args.no_newton:
        try:
            newton = newton_leg(args, rank, world, dev, barrier, max_over_ranks)
        except Exception as exc:  # the first half of the metric stands on its own
            newton = {"error": f"{type(exc).__name__}: {exc}"}

    # ---- the other BASELINE configurations as extra entries (parity cases with a measured rate; not the headline) -----------
    configs = None
    if args.configs and world == 1:
        configs = []
        for label, system, nint, batch in (("C1 3C nint 3072", "3C", 3072, 4096), ("C3 5Dq nint 10240", "5Dq", 10240, 1024),
                                           ("C4 64C nint 16384", "64C", 16384, 256)):
            try:
                configs.append(config_entry(label, system, nint, batch, dev, peak))
            except Exception as exc:
                configs.append({"workload": label, "error": f"{type(exc).__name__}: {exc}"})

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, npar)

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_config(args, npar, world),
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
            "fused": bool(nbs.fused), "fused_version": int(nbs.fused_version),
        }
        if stages:
            line["stages"] = stages
        if newton is not None:
            line["newton"] = newton
        if configs is not None:
            line["configs"] = configs
        print(json.dumps(line))
    return 0


if __name__ == "__main__":
    a = parse_args()
    sys.exit(run_reference(a) if a.impl == "reference" else run_b200(a))
