"""Bandwidth of the ODE right-hand-side kernels (chb_ode.cuh) on one GPU: python tools/ode_bench.py [system ...]
Algorithmic bytes: forces 16 vs per vector; grad forces (16 ndof + 8) vs per vector (vs = nsegm*geodim)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from choreographies2_b200 import NBodySystB200, PlanTables  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6457.1)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e-3


for name in (sys.argv[1:] or ["3C", "3B", "5Dq", "20B", "64C"]):
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nbs = NBodySystB200(T, 2 * T.nint_min * 4)
    vs = T.nsegm * T.geodim
    nvec = max(1024, (1 << 27) // vs)          # 1 GiB of positions: larger than L2
    pos = torch.randn((nvec, vs), dtype=torch.float64, device="cuda")
    out = {"system": name, "nsegm": int(T.nsegm), "nbin": int(T.nbin), "vs": vs}
    s = timed(lambda: nbs.Compute_forces_vectorized(None, pos))
    out["forces"] = {"nvec": nvec, "ms": s * 1e3, "Mvec_per_s": nvec / s / 1e6, "GBs": 16 * vs * nvec / s / 1e9,
                     "Gpairs_per_s": int(T.nbin) * nvec / s / 1e9,
                     "frac_hbm": 16 * vs * nvec / s / 1e9 / peak}
    ndof = 2 * vs
    nv2 = max(64, (1 << 27) // (vs * ndof))
    pos2 = pos[:nv2].contiguous()
    dpos = torch.randn((nv2, vs, ndof), dtype=torch.float64, device="cuda")
    s = timed(lambda: nbs.Compute_grad_forces_vectorized(None, pos2, dpos))
    by = (16 * ndof + 8) * vs * nv2
    out["grad_forces"] = {"nvec": nv2, "ndof": ndof, "ms": s * 1e3, "Mvec_per_s": nv2 / s / 1e6, "GBs": by / s / 1e9,
                          "Gpair_columns_per_s": int(T.nbin) * nv2 * ndof / s / 1e9,
                          "frac_hbm": by / s / 1e9 / peak}
    print(json.dumps(out))
