#!/usr/bin/env python
"""Where does the batched Newton-Krylov driver spend its time?  torch profiler over one solve of a batch (GPU)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from choreographies2_b200 import NBodySystB200, PlanTables
from choreographies2_b200.newton import newton_krylov_batched
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
nbs = NBodySystB200(T, 3072)
lo, hi = T.make_params_bounds(3072, 0.1, 3, 50, 1e-16)
x0 = np.empty((B, nbs.nparams))
for s in range(B):
    np.random.seed(s); x0[s] = lo + (hi - lo) * np.random.random(nbs.nparams)
x0 = torch.from_numpy(x0).cuda()
newton_krylov_batched(nbs, x0[:8], f_tol=1e-9, maxiter=2)
torch.cuda.synchronize(); t = time.perf_counter()
newton_krylov_batched(nbs, x0, f_tol=1e-9, maxiter=60)
torch.cuda.synchronize(); print("wall", time.perf_counter() - t)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    newton_krylov_batched(nbs, x0, f_tol=1e-9, maxiter=12)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=60))
