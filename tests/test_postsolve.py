"""Post-solve reductions of the search loop (find.py:285-308, 376-380): centre of mass and escape detection, batched, against
the reference's own NBodySyst.ComputeCenterOfMass / DetectEscape (oracle/_ref) on the same segment positions."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from choreographies2_b200 import NBodySystB200, PlanTables


@pytest.fixture(scope="module")
def sim():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    return build_hostsim.load()


def _states(nbs, B, seed):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((B, nbs.nparams)) * np.logspace(0, -3, nbs.nparams)[None, :]
    return x


@pytest.mark.parametrize("name", ["3C", "3D", "3B", "5Dq", "3C5k"])
def test_center_of_mass_and_escape_match_reference(sim, name):
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    ref_loader.import_reference()
    NBS = ref_loader.build_named(name)
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nint = 2 * T.nint_min * 8
    NBS.nint = nint
    nbs = NBodySystB200(T, nint, _lib_handle=sim)
    x = _states(nbs, 6, 3)
    seg = nbs.params_to_segmpos(x)
    # spread some states out so that both answers of the escape test occur
    seg[1] *= 1e-3
    seg[2, 0] += 50.0
    seg[3, -1, :, 0] -= 200.0
    com = nbs.ComputeCenterOfMass(seg)
    esc = nbs.DetectEscape(seg)
    assert com.shape == (6, T.geodim) and esc.shape == (6,)
    for b in range(6):
        ref_com = np.asarray(NBS.ComputeCenterOfMass(np.ascontiguousarray(seg[b])))
        assert np.abs(com[b] - ref_com).max() < 1e-12 * max(1.0, np.abs(ref_com).max())
        assert bool(esc[b]) == bool(NBS.DetectEscape(np.ascontiguousarray(seg[b]))), (name, b)
    assert nbs.DetectEscape(seg[0]) in (True, False)
    for fac in (0.01, 100.0):
        assert bool(nbs.DetectEscape(seg[0], fac)) == bool(NBS.DetectEscape(np.ascontiguousarray(seg[0]), fac))


@pytest.mark.parametrize("name", ["3C", "3D", "3B", "5Dq"])
def test_path_stats_match_reference(sim, name):
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    ref_loader.import_reference()
    NBS = ref_loader.build_named(name)
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nint = 2 * T.nint_min * 8
    NBS.nint = nint
    nbs = NBodySystB200(T, nint, _lib_handle=sim)
    x = _states(nbs, 4, 7)
    seg = nbs.params_to_segmpos(x)
    vel = np.stack([np.asarray(NBS.params_to_segmvel(x[b].copy())) for b in range(4)])
    loop_len, dx_min = nbs.segm_to_path_stats(seg, vel)
    assert loop_len.shape == (4, T.nloop) and dx_min.shape == (4, T.nbin)
    for b in range(4):
        rl, rd = NBS.segm_to_path_stats(np.ascontiguousarray(seg[b]), np.ascontiguousarray(vel[b]))
        assert np.abs(loop_len[b] - np.asarray(rl)).max() < 1e-12 * max(1.0, np.abs(rl).max())
        assert np.abs(dx_min[b] - np.asarray(rd)).max() < 1e-12 * max(1.0, np.abs(rd).max())


def test_center_of_mass_follows_translation(sim):
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3B.npz"))
    nbs = NBodySystB200(T, 32, _lib_handle=sim)
    seg = nbs.params_to_segmpos(_states(nbs, 3, 1))
    shift = np.array([0.7, -1.3])
    assert np.abs(nbs.ComputeCenterOfMass(seg + shift) - nbs.ComputeCenterOfMass(seg) - shift).max() < 1e-13


def test_tables_without_body_symmetries_are_rejected(sim):
    import dataclasses
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    T2 = PlanTables.from_dict({k: v for k, v in T.to_dict().items() if not k.startswith("body_") and k != "bodysegm"})
    assert T2.bodysegm is None
    nbs = NBodySystB200(T2, 48, _lib_handle=sim)
    with pytest.raises(ValueError):
        nbs.ComputeCenterOfMass(np.zeros(nbs.segmpos_shape))


@pytest.mark.gpu
def test_postsolve_reductions_on_device_tensors():
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3D.npz"))
    nbs = NBodySystB200(T, 2 * T.nint_min * 16)
    x = _states(nbs, 32, 5)
    seg_h = nbs.params_to_segmpos(x)
    seg_h[3, 0] += 80.0
    seg_d = torch.from_numpy(seg_h).cuda()
    assert np.abs(nbs.ComputeCenterOfMass(seg_d).cpu().numpy() - nbs.ComputeCenterOfMass(seg_h)).max() < 1e-12
    assert np.array_equal(nbs.DetectEscape(seg_d).cpu().numpy(), nbs.DetectEscape(seg_h))
    assert np.array_equal(nbs.DetectEscape(seg_d, 0.01).cpu().numpy(), nbs.DetectEscape(seg_h, 0.01))
