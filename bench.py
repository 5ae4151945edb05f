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
    return ap.parse_args()


def workload_config(args, nparams, world):
    return {
        "workload": f"{args.system} nint={args.nint} nparams={nparams} batch={args.batch}/GPU x {world} GPU: "
                    "fused action+gradient (U1) per state, rng(1234).random params",
        "system": args.system, "nint": args.nint, "nparams": int(nparams), "batch_per_gpu": args.batch,
        "global_batch": args.batch * world, "parallelism": f"batch-shard x{world} (no collective)",
        "l2": "inputs larger than L2 (params+grad = %.0f MB per step per GPU)" % (2 * args.batch * nparams * 8 / 1e6),
    }


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
    if ctx is not None:
        ctx.restore_original_limits() if hasattr(ctx, "restore_original_limits") else None
    return {"value": n / best, "unit": UNIT, "cores": 1, "kind": kind,
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
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------ our arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
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
               "api": "NBodySystB200.params_to_action_and_grad(pinned numpy in/out) -> chb_run_host(CHB_OP_ACTION_GRAD): "
                      "3-slot stream pipeline of 48 MB slices (H2D, kernels, D2H overlapped)"}
        assert abs(float(a_h.sum()) - checksum) <= 1e-9 * abs(checksum)

    # ---- second half of the metric: Newton orbit solves/hour ----------------------------------------------------------
    # every rank solves its own shard of find.py-style random starts (seeded by global index) of the 3C system at nint 3072
    # with the batched Newton-Krylov driver (GMRES(20) + Armijo, the reference's settings); no collective on the path
    newton = None
    if not args.no_newton:
        try:
            from choreographies2_b200.newton import newton_krylov_batched
            Tn = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
            n_nint, Bn = 3072, int(args.newton_batch)
            nbs_n = NBodySystB200(Tn, n_nint, device=dev.index)
            x_min, x_max = Tn.make_params_bounds(n_nint, 0.1, 3, 50, 1e-16)
            starts = np.empty((Bn, nbs_n.nparams))
            for i in range(Bn):  # find.py:105-108,158: np.random.seed(seed); x = x_min + x_ptp * U[0, 1)
                np.random.seed(rank * Bn + i)
                starts[i] = x_min + (x_max - x_min) * np.random.random(nbs_n.nparams)
            x0n = torch.from_numpy(starts).to(dev)
            newton_krylov_batched(nbs_n, x0n, f_tol=1e-9, maxiter=2)  # warm-up: plan, workspaces, caching allocator
            barrier()
            l0 = nbs_n.kernel_launches
            t0 = time.perf_counter()
            res = newton_krylov_batched(nbs_n, x0n, f_tol=1e-9, maxiter=60)
            torch.cuda.synchronize()
            dtn = max_over_ranks(time.perf_counter() - t0)
            newton = {"solves_per_hour": world * Bn / dtn * 3600.0, "seconds": dtn, "batch_per_gpu": Bn, "system": "3C",
                      "nint": n_nint, "nparams": int(nbs_n.nparams),
                      "settings": "f_tol 1e-9 (max norm), <= 60 Newton iterations, GMRES(20), Armijo smin 1e-2, "
                                  "starts x_min + (x_max - x_min) U[0,1) from Make_params_bounds(0.1, 3, 50, 1e-16)",
                      "rank0": {"converged": int((res.status == 1).sum()), "iteration_limit": int((res.status == 2).sum()),
                                "non_finite": int((res.status == 3).sum()), "U1_evals": int(res.n_grad_evals),
                                "U2_evals": int(res.n_hess_evals), "gpu_launches": int(nbs_n.kernel_launches - l0)},
                      "note": "a solve = one run from one random start to termination; reference solver on one host core: "
                              "profiles/r01_solves_*.json (about 1.0e5 solves/hour per process)"}
            del nbs_n, x0n, res
        except Exception as exc:  # the first half of the metric stands on its own
            newton = {"error": f"{type(exc).__name__}: {exc}"}

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
        print(json.dumps(line))
    return 0


if __name__ == "__main__":
    a = parse_args()
    sys.exit(run_reference(a) if a.impl == "reference" else run_b200(a))
