#!/usr/bin/env python
"""Launch profile of the library Newton driver at a deep refinement level (few states, long rows): 3C at nint 49152, 32 starts,
a few Newton iterations.  Run plain for the wall time per iteration, or under `ncu --metrics gpu__time_duration.sum` for the
per-kernel launch list (tools/gpu_r2j.sh aggregates it by kernel name)."""
import argparse
import sys
import time

sys.path.insert(0, ".")
import torch  # noqa: E402
from choreographies2_b200 import NBodySystB200, PlanTables, refine  # noqa: E402
from choreographies2_b200.newton import newton_krylov_batched  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nint", type=int, default=49152)
ap.add_argument("--states", type=int, default=32)
ap.add_argument("--maxiter", type=int, default=3)
ap.add_argument("--reps", type=int, default=1)
ap.add_argument("--no-fused", action="store_true", help="staged kernels (residue-split transforms) instead of the fused kernel")
ap.add_argument("--split", type=int, default=0)
a = ap.parse_args()
T = PlanTables.load("choreographies2_b200/systems/3C.npz")
nbs = NBodySystB200(T, a.nint)
if a.no_fused:
    nbs.set_fused(False)
    nbs.set_transform_split(a.split)
x0 = torch.from_numpy(refine.find_starts(T, a.nint, range(a.states), **refine.RANDOM_CONFIG_3C)).cuda()
for rep in range(a.reps + 1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = newton_krylov_batched(nbs, x0, f_tol=1e-30, maxiter=a.maxiter, line_search="wolfe", smin=0.1, tol_norm="l2", track_best=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if rep:
        print(f"nint {a.nint} states {a.states} maxiter {a.maxiter}: {dt * 1e3:.2f} ms = {dt * 1e3 / a.maxiter:.2f} ms per Newton iteration, "
              f"launches {r.launches}, U1 {r.n_grad_evals} U2 {r.n_hess_evals}, staged {nbs.staged_kernels}")
