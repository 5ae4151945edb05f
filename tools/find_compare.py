#!/usr/bin/env python
"""Per-seed comparison of the batched search driver (refine.find_batched on the CUDA library) with the reference's search loop
(oracle/ref_find.reference_find: unmodified nonlin_solve_pp / KrylovJacobian / params_resize) on the settings of
tests/NewSym_data/3C/choreo_config.json: wolfe line search, smin 0.1, six optimiser levels 1e-1 .. 1e-11, refinement x2.

    python tools/find_compare.py [--seeds 32] [--nint-fac 600] [--backend cuda|hostsim] [--reconverge 4]
Prints one line per seed and a JSON summary (outcomes equal, hash / action agreement of the orbits both found, timings)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def compare(nseeds, fac0, backend, reconverge, system="3C", verbose=True):
    from oracle import ref_loader, ref_find
    ref_loader.import_reference()
    from choreographies2_b200 import extract_tables, NBodySystB200, refine
    NBS = ref_loader.build_named(system)
    NBS.nint_fac = fac0
    T = extract_tables(NBS)
    cfg = dict(ref_find.CONFIG_3C)
    kw = {k: cfg[k] for k in ("gradtol_list", "maxiter_list", "krylov_method", "line_search", "linesearch_smin", "Newt_err_norm_max",
                              "Newt_err_norm_max_save", "mul_coarse_to_fine")}
    kw["n_reconverge_it_max"] = reconverge
    x0 = ref_find.find_starts(NBS, range(nseeds), cfg["coeff_ampl_o"], cfg["k_infl"], cfg["k_max"], cfg["coeff_ampl_min"])
    if backend == "hostsim":
        from hostsim import build_hostsim
        lib = build_hostsim.load()
        make = lambda nint: NBodySystB200(T, nint, _lib_handle=lib)  # noqa: E731
        xin = x0.copy()
        sync = lambda: None  # noqa: E731
    else:
        import torch
        make = lambda nint: NBodySystB200(T, nint)  # noqa: E731
        xin = torch.from_numpy(x0).cuda()
        sync = torch.cuda.synchronize
        refine.find_batched(make, xin[:2].clone(), 2 * T.nint_min * fac0, **kw)  # warm-up (plans, workspaces)
    sync()
    t = time.perf_counter()
    res = refine.find_batched(make, xin, 2 * T.nint_min * fac0, **kw)
    sync()
    t_b200 = time.perf_counter() - t
    t = time.perf_counter()
    refs = [ref_find.reference_find(NBS, x0[i], fac0, **kw) for i in range(nseeds)]
    t_ref = time.perf_counter() - t
    rows, n_saved_ref, n_same, worst_hash, worst_action = [], 0, 0, 0.0, 0.0
    for i, (out, lvl, iopt, fn, ff, xb, nit) in enumerate(refs):
        same = bool(res.outcome[i] == out and res.level[i] == lvl)
        row = {"seed": i, "ref": [int(out), int(lvl), int(iopt), int(nit)],
               "b200": [int(res.outcome[i]), int(res.level[i]), int(res.optim_level[i]), int(res.newton_iterations[i])], "same": same}
        if out in (ref_find.FOUND, ref_find.APPROX):
            n_saved_ref += 1
            if same:
                NBS.nint_fac = fac0 * 2 ** lvl
                pb = res.params[i]
                pb = pb.detach().cpu().numpy() if hasattr(pb, "detach") else np.asarray(pb)
                sr, sb = NBS.params_to_segmpos(np.ascontiguousarray(xb)), NBS.params_to_segmpos(np.ascontiguousarray(pb))
                hr, hb = np.asarray(NBS.segmpos_to_hash(sr)), np.asarray(NBS.segmpos_to_hash(sb))
                ar, ab = NBS.params_to_action(np.ascontiguousarray(xb)), NBS.params_to_action(np.ascontiguousarray(pb))
                row["hash_rel"] = float(np.abs(hr - hb).max() / np.abs(hr).max())
                row["action_ref"], row["action_b200"] = float(ar), float(ab)
                row["hash_same"] = bool(NBS.TestHashSame(hr, hb, 1e-8))
                worst_hash, worst_action = max(worst_hash, row["hash_rel"]), max(worst_action, abs(ar - ab) / abs(ar))
                n_same += int(row["hash_same"])
        rows.append(row)
        if verbose:
            print(json.dumps(row))
    summary = {"system": system, "nint0": 2 * T.nint_min * fac0, "seeds": nseeds, "backend": backend, "reference_saved": n_saved_ref,
               "same_outcome_level_and_orbit": n_same, "worst_hash_rel": worst_hash, "worst_action_rel": worst_action,
               "seconds_b200_batch": t_b200, "seconds_reference_one_process": t_ref,
               "solves_per_hour_b200": nseeds / t_b200 * 3600, "solves_per_hour_reference_process": nseeds / t_ref * 3600}
    return rows, summary


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=32)
    ap.add_argument("--nint-fac", type=int, default=600)
    ap.add_argument("--backend", default="cuda")
    ap.add_argument("--reconverge", type=int, default=4)
    a = ap.parse_args()
    rows, summary = compare(a.seeds, a.nint_fac, a.backend, a.reconverge)
    print(json.dumps(summary))
