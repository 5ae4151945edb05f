#!/usr/bin/env bash
# round 2, run G: Arnoldi kernel for long rows (GPU test), sweep tool, default bench line (+ timing of the whole run)
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python - > $OUT/r2g_arnoldi_long.txt 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
import numpy as np, torch
from choreographies2_b200 import NBodySystB200, PlanTables
from choreographies2_b200.newton import newton_krylov_batched
T = PlanTables.load('choreographies2_b200/systems/3C.npz')
nbs = NBodySystB200(T, 3072 * 16)     # nparams 98302 > 25600: candidate vector in global memory
print("nparams", nbs.nparams, "fused", nbs.fused_version, nbs.staged_kernels)
from choreographies2_b200 import refine
x0 = torch.from_numpy(refine.find_starts(T, nbs.nint, range(4))).cuda()
r = newton_krylov_batched(nbs, x0, f_tol=1e-9, maxiter=8, line_search="wolfe", smin=0.1, tol_norm="l2")
g = nbs.params_to_action_grad(r.x)
print("status", r.status, "nit", r.nit, "fnorm", r.fnorm, "recomputed", g.abs().amax(1).cpu().numpy())
nb2 = NBodySystB200(T, 3072)
x1 = nbs.tables.params_resize(r.x, nbs.nint, 3072)
print("ok")
PY
tail -4 $OUT/r2g_arnoldi_long.txt
timeout 900 python tools/sweep_bench.py --guesses 8192 --chunk 2048 --cpu-seeds 32 > $OUT/r2g_sweep_1gpu_8192.json 2> $OUT/r2g_sweep_err.txt; tail -c 1800 $OUT/r2g_sweep_1gpu_8192.json; tail -3 $OUT/r2g_sweep_err.txt
( time timeout 900 python bench.py ) > $OUT/r2g_bench_default.json 2> $OUT/r2g_bench_default_err.txt; tail -c 2500 $OUT/r2g_bench_default.json; tail -5 $OUT/r2g_bench_default_err.txt
