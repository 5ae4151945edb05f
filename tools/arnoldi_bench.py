#!/usr/bin/env python
"""Time chb_gmres_step_dev / chb_arnoldi_step_dev alone: us per launch and algorithmic GB/s (8 n (j + 3) bytes per state) for a
few (batch, j); all states live.  python tools/arnoldi_bench.py [--n 6142]"""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=6142)
    ap.add_argument("--batches", default="1024,8192")
    ap.add_argument("--js", default="0,1,2,4,9,19")
    a = ap.parse_args()
    import torch
    from choreographies2_b200 import NBodySystB200, PlanTables
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    dev = torch.device("cuda:0")
    m = 20
    for A in [int(b) for b in a.batches.split(",")]:
        V = torch.randn((m + 1, A, a.n), dtype=torch.float64, device=dev) / np.sqrt(a.n)
        w = torch.randn((A, a.n), dtype=torch.float64, device=dev)
        h = torch.zeros((A, m + 1), dtype=torch.float64, device=dev)
        R = torch.zeros((A, m + 1, m), dtype=torch.float64, device=dev)
        g = torch.ones((A, m + 1), dtype=torch.float64, device=dev)
        cs = torch.zeros((A, m, 2), dtype=torch.float64, device=dev)
        target = torch.zeros((A,), dtype=torch.float64, device=dev)
        kdone = torch.zeros((A,), dtype=torch.int64, device=dev)
        done = torch.zeros((A,), dtype=torch.bool, device=dev)
        for j in [int(v) for v in a.js.split(",")]:
            # fresh random data per j: nearly orthonormal rows, w far from their span -> one Gram-Schmidt pass (a vector
            # left over from an earlier step would be parallel to w and trigger the second pass = twice the traffic)
            V.normal_()
            V /= np.sqrt(a.n)
            w.normal_()
            for name, fn in (("arnoldi", lambda: nbs.arnoldi_step(V, j, w, None, h)),
                             ("gmres_step", lambda: nbs.gmres_step(V, j, w, done, R, g, cs, target, kdone))):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 10
                e0.record()
                for _ in range(reps):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / reps
                gb = 8.0 * a.n * (j + 3) * A / 1e9
                print(f"batch {A:6d} j {j:2d} {name:10s} {us:9.1f} us  {gb / (us * 1e-6):8.1f} GB/s", flush=True)


if __name__ == "__main__":
    main()
