"""Velocities, change of variables, hash, post-solve reductions and ODE velocity callbacks (SURVEY.md 8(a) row a15, 8(f) ranks
2 and 4) against the UNMODIFIED reference (oracle/_ref) on the same inputs -- on the g++ emulation in the CPU suite and, with
``-m gpu``, on the CUDA library (the reference build travels to the GPU box)."""
import os

import numpy as np
import pytest

from conftest import ROOT, make_nbs, need_reference, relerr
from choreographies2_b200 import extract_tables

SYSTEMS = ["3C", "3B", "3D", "5Dq", "3C5k", "2D2D", "1D1D1D", "uneven_nnpr", "3D5k", "2D1_3dim", "3C_3dim", "5q", "20B", "4C5k",
           "complex_mass_charge", "3C101k"]


def _pair(name, backend, fac=4, **kw):
    ref_loader = need_reference()
    NBS = ref_loader.build_named(name, **kw)
    NBS.nint_fac = fac
    return NBS, make_nbs(extract_tables(NBS), int(NBS.nint), backend)


def _params(NBS, B, seed):
    rng = np.random.default_rng(seed)
    return rng.standard_normal((B, NBS.nparams)) * np.logspace(0, -2, NBS.nparams)[None, :]


@pytest.mark.parametrize("name", SYSTEMS)
def test_segmvel_matches_reference(name, backend):
    """params_to_segmvel (_NBodySyst.pyx:3638-3666) <= 1e-11, base layout and the two forced variants."""
    for kw in ({}, {"ForceGeneralSym": True}, {"ForceGreaterNStore": True}):
        NBS, nbs = _pair(name, backend, **kw)
        x = _params(NBS, 3, 11)
        vel = nbs.params_to_segmvel(x)
        assert vel.shape == (3,) + nbs.segmpos_shape
        for b in range(3):
            assert relerr(vel[b], NBS.params_to_segmvel(x[b].copy())) < 1e-11, (name, kw, b)
        assert relerr(nbs.params_to_segmvel(x[1]), vel[1]) == 0.0


@pytest.mark.parametrize("name", SYSTEMS)
def test_params_changevar_matches_reference(name, backend):
    """params_changevar (_NBodySyst.pyx:2330-2410): the four diagonal maps, exact to rounding."""
    NBS, nbs = _pair(name, backend)
    assert nbs.nparams_incl_o == NBS.nparams_incl_o
    rng = np.random.default_rng(5)
    for inv in (False, True):
        for tr in (False, True):
            to_pos = (not inv and not tr) or (inv and tr)
            n_in = NBS.nparams if to_pos else NBS.nparams_incl_o
            x = rng.random((2, n_in))
            out = nbs.params_changevar(x, inv, tr)
            for b in range(2):
                assert relerr(out[b], NBS.params_changevar(x[b].copy(), inv, tr)) < 1e-15, (name, inv, tr)


@pytest.mark.parametrize("name", SYSTEMS)
def test_hash_and_path_stats_match_reference(name, backend):
    """segmpos_to_hash (:2542-2560) and segm_to_path_stats (:3903-3943) with velocities from the library itself."""
    NBS, nbs = _pair(name, backend)
    x = _params(NBS, 3, 3)
    seg = nbs.params_to_segmpos(x)
    vel = nbs.params_to_segmvel(x)
    H = nbs.segmpos_to_hash(seg)
    ll, dm = nbs.segm_to_path_stats(seg, vel)
    assert H.shape == (3, 9) and ll.shape == (3, NBS.nloop) and dm.shape == (3, NBS.nbin_segm_unique)
    for b in range(3):
        sr = np.asarray(NBS.params_to_segmpos(x[b].copy()))
        assert relerr(H[b], NBS.segmpos_to_hash(sr)) < 1e-11, name
        rl, rd = NBS.segm_to_path_stats(sr, np.asarray(NBS.params_to_segmvel(x[b].copy())))
        assert relerr(ll[b], rl) < 1e-11 and relerr(dm[b], rd) < 1e-11, name


@pytest.mark.parametrize("name", ["3C", "3D", "5Dq", "non_gravity_2dim", "2D1_non_gravity"])
def test_hash_same_matches_reference_near_threshold(name, backend):
    """TestHashSame (:2036-2067) with the reference's Homo_exp (:1692-1718): the test values themselves, for a rescaled orbit
    and for near-duplicates around the threshold."""
    NBS, nbs = _pair(name, backend)
    assert bool(NBS.LawIsHomo) and abs(NBS.Homo_exp - nbs.Homo_exp) < 1e-9
    x = _params(NBS, 1, 7)[0]
    H0 = np.asarray(NBS.segmpos_to_hash(NBS.params_to_segmpos(x)))
    rng = np.random.default_rng(2)
    for eps in (0.0, 1e-7, 3e-6, 1e-5, 3e-5, 1e-3):
        for scale in (1.0, 1.7):
            y = scale * x * (1.0 + eps * rng.standard_normal(x.shape[0]))
            H1 = np.asarray(NBS.segmpos_to_hash(NBS.params_to_segmpos(y)))
            for mult in (True, False):
                assert nbs.TestHashSame(H0, H1, 1e-5, mult) == bool(NBS.TestHashSame(H0, H1, 1e-5, mult)), (name, eps, scale, mult)


@pytest.mark.parametrize("name", ["3C", "3D", "3B", "5Dq", "3C5k", "2D1_3dim", "20B", "2D2D"])
def test_center_of_mass_and_escape_match_reference(name, backend):
    NBS, nbs = _pair(name, backend, fac=8)
    x = _params(NBS, 6, 3)
    seg = np.array(nbs.params_to_segmpos(x))
    seg[1] *= 1e-3
    seg[2, 0] += 50.0
    seg[3, -1, :, 0] -= 200.0
    com, esc = nbs.ComputeCenterOfMass(seg), nbs.DetectEscape(seg)
    assert com.shape == (6, NBS.geodim) and esc.shape == (6,)
    for b in range(6):
        sb = np.ascontiguousarray(seg[b])
        rc = np.asarray(NBS.ComputeCenterOfMass(sb))
        assert np.abs(com[b] - rc).max() < 1e-12 * max(1.0, np.abs(rc).max())
        assert bool(esc[b]) == bool(NBS.DetectEscape(sb)), (name, b)
    for fac in (0.01, 100.0):
        assert bool(nbs.DetectEscape(seg[0], fac)) == bool(NBS.DetectEscape(np.ascontiguousarray(seg[0]), fac))


@pytest.mark.parametrize("name", ["3C", "5Dq", "complex_mass_charge", "2D1_3dim"])
def test_velocity_callbacks_and_ode_def_match_reference(name, backend):
    """Compute_velocities(_vectorized), Compute_grad_velocities(_vectorized) (:4081-4150), Compute_init_pos_mom (:4060-4077)
    and the dict of Get_ODE_def (:4310-4444, Python callables)."""
    NBS, nbs = _pair(name, backend)
    rng = np.random.default_rng(9)
    vs = NBS.nsegm * NBS.geodim
    mom = rng.random((5, vs))
    gmom = rng.random((5, vs, 7))
    t = np.zeros(5)
    assert relerr(nbs.Compute_velocities_vectorized(t, mom), NBS.Compute_velocities_vectorized(t, mom)) < 1e-15
    assert relerr(nbs.Compute_velocities(0., mom[2]), NBS.Compute_velocities(0., mom[2])) < 1e-15
    assert relerr(nbs.Compute_grad_velocities_vectorized(t, mom, gmom), NBS.Compute_grad_velocities_vectorized(t, mom, gmom)) < 1e-15
    assert relerr(nbs.Compute_grad_velocities(0., mom[1], gmom[1]), NBS.Compute_grad_velocities(0., mom[1], gmom[1])) < 1e-15
    x = _params(NBS, 1, 4)[0]
    xo, po = nbs.Compute_init_pos_mom(x)
    rxo, rpo = NBS.Compute_init_pos_mom(x.copy())
    assert relerr(xo, rxo) < 1e-11 and relerr(po, rpo) < 1e-11
    d, rd = nbs.Get_ODE_def(x, vector_calls=True, LowLevel=False, grad=True), NBS.Get_ODE_def(x.copy(), vector_calls=True, LowLevel=False, grad=True)
    assert set(d.keys()) == set(rd.keys()) and d["t_span"] == rd["t_span"]
    pos = rng.standard_normal((3, vs))
    assert relerr(d["gun"](t[:3], pos), np.stack([NBS.Compute_forces(0., pos[i].copy()) for i in range(3)])) < 1e-11
    assert relerr(d["fun"](t[:3], mom[:3]), rd["fun"](t[:3], mom[:3])) < 1e-15
    with pytest.raises(ValueError):
        nbs.Get_ODE_def(x, LowLevel=True)


@pytest.mark.gpu
def test_api2_on_device_tensors():
    """The same operators on torch CUDA tensors (current stream) agree with the host-array path."""
    import torch
    from choreographies2_b200 import NBodySystB200, PlanTables
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "5Dq.npz"))
    nbs = NBodySystB200(T, 2 * T.nint_min * 64)
    x = np.random.default_rng(1).standard_normal((40, nbs.nparams)) * np.logspace(0, -3, nbs.nparams)[None, :]
    xd = torch.from_numpy(x).cuda()
    seg_h, vel_h = nbs.params_to_segmpos(x), nbs.params_to_segmvel(x)
    seg_d, vel_d = nbs.params_to_segmpos(xd), nbs.params_to_segmvel(xd)
    assert relerr(vel_d.cpu().numpy(), vel_h) < 1e-13
    assert relerr(nbs.segmpos_to_hash(seg_d).cpu().numpy(), nbs.segmpos_to_hash(seg_h)) < 1e-13
    assert relerr(nbs.ComputeCenterOfMass(seg_d).cpu().numpy(), nbs.ComputeCenterOfMass(seg_h)) < 1e-13
    assert np.array_equal(nbs.DetectEscape(seg_d).cpu().numpy(), nbs.DetectEscape(seg_h))
    ll_d, dm_d = nbs.segm_to_path_stats(seg_d, vel_d)
    ll_h, dm_h = nbs.segm_to_path_stats(seg_h, vel_h)
    assert relerr(ll_d.cpu().numpy(), ll_h) < 1e-13 and relerr(dm_d.cpu().numpy(), dm_h) < 1e-13
    assert relerr(nbs.params_changevar(xd).cpu().numpy(), nbs.params_changevar(x)) == 0.0


def test_noalloc_hessian_vector_host_emulation():
    """The entry point without a workspace argument (scratch built into the library; tests/test_zz_no_alloc.py is the GPU test): same
    numbers as the workspace variant on two calls in a row, oversize batches rejected with CHB_ERR_WORKSPACE before any work."""
    import ctypes as C
    from choreographies2_b200 import NBodySystB200, PlanTables
    from hostsim import build_hostsim
    sim = build_hostsim.load()
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    for fused in (True, False):
        nbs = NBodySystB200(T, 768, _lib_handle=sim)
        nbs.set_fused(fused)
        rng = np.random.default_rng(2)
        for _ in range(2):
            x, dx = rng.random((4, nbs.nparams)), rng.random((4, nbs.nparams))
            out = np.full_like(x, np.nan)
            rc = sim.chb_params_to_action_hess_noalloc_dev(nbs._plan, 4, C.c_void_p(x.ctypes.data), C.c_void_p(dx.ctypes.data),
                                                           C.c_void_p(out.ctypes.data), None)
            assert rc == 0
            assert np.array_equal(out, nbs.params_to_action_hess(x, dx))
    nbs = NBodySystB200(T, 3072, _lib_handle=sim)
    nbs.set_fused(False)
    nbs.set_chunk(2048)
    x = np.random.default_rng(3).random((2048, nbs.nparams))
    out = np.full_like(x, 7.0)
    rc = sim.chb_params_to_action_hess_noalloc_dev(nbs._plan, 2048, C.c_void_p(x.ctypes.data), C.c_void_p(x.ctypes.data), C.c_void_p(out.ctypes.data), None)
    assert rc == 6 and b"CHB_NOALLOC_SCRATCH_BYTES" in sim.chb_last_error_string()
    assert np.all(out == 7.0)
