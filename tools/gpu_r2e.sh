#!/usr/bin/env bash
# round 2, run E: library Newton driver on the GPU: tests, per-seed comparison with the reference search loop, timings
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_newton.py tests/test_refine.py tests/test_gpu_parity.py -m gpu -x -q > $OUT/r2e_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2e_pytest.txt
tail -5 $OUT/r2e_pytest.txt
timeout 900 python tools/find_compare.py --seeds 32 --nint-fac 600 --reconverge 4 > $OUT/r2e_find_compare_3C_fac600.txt 2>&1
tail -3 $OUT/r2e_find_compare_3C_fac600.txt
timeout 600 python - > $OUT/r2e_newton_timing.txt 2>&1 <<'PY'
import sys, time, json, os
sys.path.insert(0, '.')
import numpy as np, torch
from choreographies2_b200 import NBodySystB200, PlanTables
from choreographies2_b200.newton import newton_krylov_batched
T = PlanTables.load('choreographies2_b200/systems/3C.npz')
for nint, B in ((3072, 8192), (3600, 8192)):
    nbs = NBodySystB200(T, nint)
    x_min, x_max = T.make_params_bounds(nint, 0.1, 3, 50, 1e-16)
    starts = np.empty((B, nbs.nparams))
    for i in range(B):
        np.random.seed(i); starts[i] = x_min + (x_max - x_min) * np.random.random(nbs.nparams)
    x0 = torch.from_numpy(starts).cuda()
    for ls, smin, tn, ftol, mit in (("armijo", 1e-2, "max", 1e-9, 60), ("wolfe", 0.1, "l2", 1e-9, 60), ("wolfe", 0.1, "l2", 1e-1, 100)):
        for lib in ("1", "0"):
            if lib == "0" and ls == "wolfe": continue
            os.environ["CHB_NEWTON_LIB"] = lib
            newton_krylov_batched(nbs, x0[:64], f_tol=ftol, maxiter=3, line_search=ls, smin=smin, tol_norm=tn)
            ts = []
            for rep in range(3):
                torch.cuda.synchronize(); t = time.perf_counter()
                res = newton_krylov_batched(nbs, x0, f_tol=ftol, maxiter=mit, line_search=ls, smin=smin, tol_norm=tn)
                torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
            print(json.dumps({"nint": nint, "batch": B, "line_search": ls, "smin": smin, "tol_norm": tn, "f_tol": ftol, "maxiter": mit, "library_driver": lib == "1",
                              "seconds": ts, "solves_per_hour_median": B / sorted(ts)[1] * 3600, "converged": int((res.status == 1).sum()), "limit": int((res.status == 2).sum()),
                              "nonfinite": int((res.status == 3).sum()), "U1": res.n_grad_evals, "U2": res.n_hess_evals, "launches": res.launches, "nit_mean": float(res.nit.mean())}), flush=True)
PY
cat $OUT/r2e_newton_timing.txt
