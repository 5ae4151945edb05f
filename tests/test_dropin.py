"""The drop-in claim: the reference's OWN search driver ``choreo.find.Find_Choreo`` (find.py:25-444),
unmodified, with ``choreo.NBodySyst`` replaced by ``choreographies2_b200.dropin.NBodySystDropIn`` -- every numeric call of the
action path goes through the C ABI -- on the reference's 3C configuration (wolfe, six optimiser levels, refinement) with
``np.random.seed(3)``: it must save the figure-eight (action 10.4830237879556, |grad action| < 1e-11)."""
import json
import os

import numpy as np
import pytest

from conftest import ROOT, need_reference


def test_find_choreo_runs_on_the_dropin(backend, tmp_path, capsys):
    ref_loader = need_reference()
    choreo = ref_loader.import_reference()
    import choreo.find  # noqa: F401
    from choreographies2_b200 import dropin
    with open(os.path.join(ROOT, "tests", "golden", "3C_choreo_config.json")) as f:
        params_dict = json.load(f)
    params_dict["Solver_CLI"]["SaveImage"] = False
    if backend is not None:  # the emulation is slow: a coarser start level, same code path
        params_dict["Solver_Discr"]["nint_init"] = 64
    store = str(tmp_path / "sols")
    os.makedirs(store)
    extra = {"store_folder": store, "n_opt_max": 1, "n_find_max": 1}
    with dropin.installed(choreo, lib_handle=backend) as cls:
        kwargs = choreo.find.ChoreoLoadFromDict(params_dict, str(tmp_path), callback=choreo.find.Find_Choreo, extra_args_dict=extra)
        np.random.seed(3)
        choreo.find.Find_Choreo(**kwargs)
        calls = dict(cls.calls)
    capsys.readouterr()
    # the search ran on the library: gradient evaluations, Krylov matvecs, transforms, hash, escape test, centre of mass
    for name in ("params_to_action_grad", "segmpos_dparams_to_action_hess", "params_to_segmpos", "segmpos_to_hash", "DetectEscape",
                 "ComputeCenterOfMass", "segmpos_to_params", "params_resize", "GetKrylovJacobian"):
        assert calls.get(name, 0) > 0, (name, calls)
    assert calls["segmpos_dparams_to_action_hess"] > 50
    saved = sorted(fn for fn in os.listdir(store) if fn.endswith(".json"))
    assert len(saved) == 1, saved
    with open(os.path.join(store, saved[0])) as f:
        info = json.load(f)
    assert abs(info["Action"] - 10.4830237879556) < 1e-9, info["Action"]
    assert info["Grad_Action"] < 1e-11
    seg = np.load(os.path.join(store, saved[0][:-5] + ".npy"))
    assert seg.shape[0] == 3 and seg.shape[2] == 2
    # the saved orbit, re-evaluated by the unmodified reference, is a solution
    NBS = ref_loader.build_named("3C")
    NBS.nint = info["nint"]
    x = NBS.segmpos_to_params(seg)
    assert np.linalg.norm(NBS.params_to_action_grad(x)) < 1e-10
    assert abs(NBS.params_to_action(x) - 10.4830237879556) < 1e-9
