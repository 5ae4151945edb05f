#!/usr/bin/env python
"""Generate the golden fixtures of this directory from the UNMODIFIED reference.

Run in the build container (needs /root/reference for the 74 test configurations and
oracle/_ref = the reference built by oracle/build_ref.sh):

    python tests/golden/make_golden.py            # systems.json + golden_*.npz + sols.npz

Outputs (all committed):
  systems.json     plain description (geodim, nbody, mass, charge, law, symmetries) of every
                   configuration in the reference's tests/NewSym_data/*/choreo_config.json, as parsed
                   by the reference's own loader choreo.find.ChoreoLoadFromDict
                   (tests/test_config/test_config_NBodySyst.py:47-88), plus the synthetic 64C system
                   of BASELINE.json config 4.
  golden_<name>.npz
                   per configuration and per variant v in {base, gen (ForceGeneralSym),
                   gns (ForceGreaterNStore), pow (inter_pow=+1, 'nozerodiv')}: the plan tables
                   ("<v>/T/<field>") and, per nint level L (nint_fac 1, 4, 10 for base; 4 for the other variants), seeded inputs and the reference's outputs
                   ("<v>/<nint>/<key>"): params, dparams, segm (random segment array),
                   segmpos=params_to_segmpos(params), action, grad, hess (params_to_action_hess),
                   hess_cached (segmpos_dparams_to_action_hess), kin, pot, params_T=segmpos_to_params_T(segm),
                   params_inv=segmpos_to_params(segmpos).
  sols.npz         the 7 converged orbits of tests/NewSym_sols: tables, segmpos, stored Action /
                   Grad_Action, and the reference's recomputed action / gradient norm.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from choreographies2_b200.tables import extract_tables  # noqa: E402

REF = os.environ.get("CHOREO_REFERENCE", "/root/reference")
NINT_FACS = (1, 4, 10)


def export_systems():
    choreo = ref_loader.import_reference()
    data_dir = os.path.join(REF, "tests", "NewSym_data")
    systems = {}
    for name in sorted(os.listdir(data_dir)):
        folder = os.path.join(data_dir, name)
        with open(os.path.join(folder, "choreo_config.json")) as f:
            d = json.load(f)
        kw = choreo.find.ChoreoLoadFromDict(
            d, folder, args_list=["geodim", "nbody", "mass", "charge", "inter_pow", "inter_pm", "Sym_list"])
        systems[name] = {
            "geodim": int(kw["geodim"]), "nbody": int(kw["nbody"]),
            "mass": [float(x) for x in kw["mass"]], "charge": [float(x) for x in kw["charge"]],
            "inter_pow": float(kw["inter_pow"]), "inter_pm": int(kw["inter_pm"]),
            "syms": [{
                "BodyPerm": [int(i) for i in S.BodyPerm],
                "SpaceRot": [[float(x) for x in row] for row in np.asarray(S.SpaceRot)],
                "TimeRev": int(S.TimeRev), "TimeShiftNum": int(S.TimeShiftNum), "TimeShiftDen": int(S.TimeShiftDen),
            } for S in kw["Sym_list"]],
        }
    # BASELINE.json config 4: 64-body choreography (one symmetry: cyclic BodyPerm, TimeShift 1/64)
    for n in (16, 64):
        systems[f"{n}C"] = {
            "geodim": 2, "nbody": n, "mass": [1.0] * n, "charge": [1.0] * n, "inter_pow": -1.0, "inter_pm": 1,
            "syms": [{"BodyPerm": [(i + 1) % n for i in range(n)], "SpaceRot": [[1.0, 0.0], [0.0, 1.0]],
                      "TimeRev": 1, "TimeShiftNum": 1, "TimeShiftDen": n}],
        }
    with open(os.path.join(HERE, "systems.json"), "w") as f:
        json.dump(systems, f, indent=None, separators=(",", ":"), sort_keys=True)
    return systems


def export_configs(names=("3C",)):
    """The reference's own search configurations (tests/NewSym_data/<name>/choreo_config.json) that the drop-in test feeds to
    ``choreo.find.ChoreoLoadFromDict`` / ``Find_Choreo``, copied verbatim as fixtures ``<name>_choreo_config.json``."""
    for name in names:
        with open(os.path.join(REF, "tests", "NewSym_data", name, "choreo_config.json")) as f:
            d = json.load(f)
        with open(os.path.join(HERE, f"{name}_choreo_config.json"), "w") as f:
            json.dump(d, f, indent=None, separators=(",", ":"), sort_keys=True)


def golden_for(NBS, rng, out, prefix, facs=NINT_FACS):
    T = extract_tables(NBS)
    for k, v in T.to_dict().items():
        out[f"{prefix}/T/{k}"] = v
    for fac in facs:
        NBS.nint_fac = fac
        n = NBS.nparams
        p = f"{prefix}/{NBS.nint}"
        params = rng.random(n)
        dparams = rng.random(n)
        segm = rng.random((NBS.nsegm, NBS.segm_store, NBS.geodim))
        segmpos = NBS.params_to_segmpos(params)
        out[f"{p}/params"] = params
        out[f"{p}/dparams"] = dparams
        out[f"{p}/segm"] = segm
        out[f"{p}/segmpos"] = segmpos
        out[f"{p}/kin"] = np.float64(NBS.params_to_kin_nrg(params))
        out[f"{p}/pot"] = np.float64(NBS.params_to_pot_nrg(params))
        out[f"{p}/action"] = np.float64(NBS.params_to_action(params))
        out[f"{p}/grad"] = NBS.params_to_action_grad(params)
        out[f"{p}/hess"] = NBS.params_to_action_hess(params, dparams)
        out[f"{p}/hess_cached"] = NBS.segmpos_dparams_to_action_hess(segmpos, dparams)
        out[f"{p}/params_T"] = NBS.segmpos_to_params_T(segm)
        out[f"{p}/params_inv"] = NBS.segmpos_to_params(segmpos)


def main():
    systems = export_systems()
    print(f"systems.json: {len(systems)} systems")
    export_configs()
    for name, desc in systems.items():
        if name == "64C":
            continue  # setup takes seconds and the vectors are large; covered live + by a dedicated fixture below
        rng = np.random.default_rng(sum(map(ord, name)) + 1000)
        out = {}
        variants = {
            "base": {}, "gen": {"ForceGeneralSym": True}, "gns": {"ForceGreaterNStore": True},
            "pow": {"inter_pow": 1.0},
        }
        for v, kw in variants.items():
            NBS = ref_loader.build_system(desc, **kw)
            golden_for(NBS, rng, out, v, NINT_FACS if v == "base" else (4,))
        np.savez_compressed(os.path.join(HERE, f"golden_{name}.npz"), **out)
        print("golden", name, sum(a.nbytes for a in out.values()))

    # converged orbits
    choreo = ref_loader.import_reference()
    sol_dir = os.path.join(REF, "tests", "NewSym_sols")
    out = {}
    names = []
    for fn in sorted(os.listdir(sol_dir)):
        base, ext = os.path.splitext(fn)
        if ext != ".json":
            continue
        NBS, segmpos = choreo.NBodySyst.FromSolutionFile(os.path.join(sol_dir, base))
        with open(os.path.join(sol_dir, fn)) as f:
            info = json.load(f)
        params = NBS.segmpos_to_params(segmpos)
        names.append(base)
        for k, v in extract_tables(NBS).to_dict().items():
            out[f"{base}/T/{k}"] = v
        out[f"{base}/nint"] = np.int64(NBS.nint)
        out[f"{base}/segmpos"] = np.asarray(segmpos)
        out[f"{base}/params"] = params
        out[f"{base}/stored_action"] = np.float64(info["Action"])
        out[f"{base}/stored_grad_action"] = np.float64(info["Grad_Action"])
        out[f"{base}/action"] = np.float64(NBS.params_to_action(params))
        out[f"{base}/grad"] = NBS.params_to_action_grad(params)
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "sols.npz"), **out)
    print("sols", names)


if __name__ == "__main__":
    main()
