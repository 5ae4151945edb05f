#!/usr/bin/env python
"""Newton orbit solves per hour (second half of BASELINE.json's metric): batched lock-step Newton-Krylov on one B200 against
the reference's own solver (nonlin_solve_pp + GetKrylovJacobian, one process) on the host, same system, same starts
(find.py:105-108,158: x = x_min + x_ptp U[0,1), np.random.seed(seed)), same settings (gmres, restart 20, armijo, smin 1e-2).

    python tools/solve_bench.py [--system 3C] [--nint 3072] [--batch 1024] [--ref-seeds 8] [--f-tol 1e-9] [--maxiter 60]
Prints one JSON line.  A "solve" = one run from one random start to termination (converged or iteration limit).
"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--system", default="3C"); ap.add_argument("--nint", type=int, default=3072)
    ap.add_argument("--batch", type=int, default=1024); ap.add_argument("--ref-seeds", type=int, default=8)
    ap.add_argument("--f-tol", type=float, default=1e-9); ap.add_argument("--maxiter", type=int, default=60)
    a = ap.parse_args()
    import torch
    from choreographies2_b200 import NBodySystB200, PlanTables
    from choreographies2_b200.newton import newton_krylov_batched
    from oracle import ref_loader
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{a.system}.npz"))
    nbs = NBodySystB200(T, a.nint, device=0)
    have_ref = ref_loader.reference_available()
    if have_ref:
        ref_loader.import_reference()
        from choreo.scipy_plus.nonlin import nonlin_solve_pp
        NBS = ref_loader.build_named(a.system); NBS.nint = a.nint
        x_min, x_max = map(np.asarray, NBS.Make_params_bounds(0.1, 3, 50, 1e-16))
        np.save(os.path.join(ROOT, "gpurun_out", "bounds_%s_%d.npy" % (a.system, a.nint)), np.stack([x_min, x_max])) \
            if os.path.isdir(os.path.join(ROOT, "gpurun_out")) else None
    else:
        raise SystemExit("oracle/_ref missing: the starts come from the reference's Make_params_bounds")
    starts = []
    for seed in range(a.batch):
        np.random.seed(seed)
        starts.append(x_min + (x_max - x_min) * np.random.random(nbs.nparams))
    x0h = np.stack(starts)
    x0 = torch.from_numpy(x0h).cuda()
    newton_krylov_batched(nbs, x0[:8], f_tol=a.f_tol, maxiter=3)  # warm-up (plans, workspaces)
    newton_krylov_batched(nbs, x0, f_tol=a.f_tol, maxiter=2)      # ... and the caching allocator at the full batch size
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res = newton_krylov_batched(nbs, x0, f_tol=a.f_tol, maxiter=a.maxiter)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    xg = res.x.cpu().numpy()
    # reference arm on the first seeds, one process, 1 thread
    from threadpoolctl import threadpool_limits
    tr, agree, nconv_ref = 0.0, 0, 0
    with threadpool_limits(1):
        for i in range(a.ref_seeds):
            jac = NBS.GetKrylovJacobian(True, {"method": "gmres"})
            t = time.perf_counter()
            xr, info = nonlin_solve_pp(NBS.params_to_action_grad, x0h[i].copy(), jacobian=jac, f_tol=a.f_tol, maxiter=a.maxiter,
                                       line_search="armijo", full_output=True, raise_exception=False, smin=1e-2)
            tr += time.perf_counter() - t
            if info["status"] == 1:
                nconv_ref += 1
                if res.status[i] == 1 and abs(NBS.params_to_action(xr) - NBS.params_to_action(xg[i])) < 1e-7 * abs(NBS.params_to_action(xr)):
                    agree += 1
    print(json.dumps({
        "metric": "Newton orbit solves/hour", "system": a.system, "nint": a.nint, "nparams": nbs.nparams, "batch": a.batch,
        "settings": {"f_tol": a.f_tol, "maxiter": a.maxiter, "krylov": "gmres restart 20", "line_search": "armijo smin 1e-2"},
        "b200": {"solves_per_hour": a.batch / dt * 3600, "seconds": dt, "converged": int((res.status == 1).sum()),
                 "iteration_limit": int((res.status == 2).sum()), "non_finite": int((res.status == 3).sum()),
                 "U1_evals": res.n_grad_evals, "U2_evals": res.n_hess_evals, "newton_iterations_mean": float(res.nit.mean())},
        "reference_1_process": {"solves_per_hour": (a.ref_seeds / tr * 3600) if tr > 0 else None, "seconds": tr, "seeds": a.ref_seeds,
                                "converged": nconv_ref, "same_orbit_as_b200": agree},
    }))


if __name__ == "__main__":
    main()
