"""GPU parity tests (run on the B200 box: pytest -m gpu).  Everything goes through the C ABI of
libchoreo_b200.so via the Python host layer; the oracle / reference is only the checker.

Tolerances (BASELINE.json north_star): action & gradient rel. err <= 1e-11, Hessian-vector <= 1e-10."""
import os

import numpy as np
import pytest

from conftest import ROOT, TOL_ACTION, TOL_GRAD, TOL_HESS, TOL_TRANSFORM, golden_names, relerr

pytestmark = pytest.mark.gpu

NAMES = golden_names()
SYSTEMS = os.path.join(ROOT, "choreographies2_b200", "systems")


@pytest.fixture(scope="module")
def chb():
    import choreographies2_b200 as pkg
    from choreographies2_b200 import _lib
    lib = _lib.load()
    assert lib.chb_device_count() >= 1, "no sm_100 device visible: GPU tests need a B200"
    return pkg


def _tables(name):
    from choreographies2_b200 import PlanTables
    return PlanTables.load(os.path.join(SYSTEMS, f"{name}.npz"))


@pytest.mark.parametrize("name", NAMES)
def test_cuda_matches_reference_golden(name, chb, golden_cache):
    """All 74 reference test configurations (+16C) x {base, ForceGeneralSym, ForceGreaterNStore, inter_pow=+1}."""
    g = golden_cache(name)
    for variant, nint in g.cases():
        nbs = chb.NBodySystB200(chb.PlanTables.from_dict(g.tables_dict(variant)), nint)
        v = lambda k: g.vec(variant, nint, k)  # noqa: E731
        tag = f"{name}/{variant}/{nint}"
        assert relerr(nbs.params_to_segmpos(v("params")), v("segmpos")) < TOL_TRANSFORM, tag
        assert relerr(nbs.params_to_kin_nrg(v("params")), v("kin")) < TOL_ACTION, tag
        assert relerr(nbs.params_to_pot_nrg(v("params")), v("pot")) < TOL_ACTION, tag
        assert relerr(nbs.params_to_action(v("params")), v("action")) < TOL_ACTION, tag
        for fused in (True, False):
            nbs.set_fused(fused)
            act, grad = nbs.params_to_action_and_grad(v("params"))
            assert relerr(act, v("action")) < TOL_ACTION, (tag, fused)
            assert relerr(grad, v("grad")) < TOL_GRAD, (tag, fused)
        assert relerr(nbs.params_to_action_hess(v("params"), v("dparams")), v("hess")) < TOL_HESS, tag
        assert relerr(nbs.segmpos_dparams_to_action_hess(v("segmpos"), v("dparams")), v("hess_cached")) < TOL_HESS, tag
        assert relerr(nbs.segmpos_params_to_action(v("segmpos"), v("params")), v("action")) < TOL_ACTION, tag
        assert relerr(nbs.segmpos_params_to_action_grad(v("segmpos"), v("params")), v("grad")) < TOL_GRAD, tag
        assert relerr(nbs.segmpos_to_params_T(v("segm")), v("params_T")) < TOL_TRANSFORM, tag
        assert relerr(nbs.segmpos_to_params(v("segmpos")), v("params_inv")) < TOL_TRANSFORM, tag


def test_converged_orbits(chb):
    z = np.load(os.path.join(ROOT, "tests", "golden", "sols.npz"))
    for name in z["names"]:
        pre = f"{name}/T/"
        T = chb.PlanTables.from_dict({k[len(pre):]: z[k] for k in z.files if k.startswith(pre)})
        nbs = chb.NBodySystB200(T, int(z[f"{name}/nint"]))
        params = nbs.segmpos_to_params(z[f"{name}/segmpos"])
        assert relerr(params, z[f"{name}/params"]) < TOL_TRANSFORM
        act, grad = nbs.params_to_action_and_grad(params)
        assert abs(act - float(z[f"{name}/stored_action"])) < 1e-9 * max(1.0, abs(act)), name
        assert relerr(act, z[f"{name}/action"]) < TOL_ACTION
        assert np.max(np.abs(grad - z[f"{name}/grad"])) < 1e-9, name


@pytest.mark.parametrize("name,nint", [("3C", 3072), ("3B", 4096), ("5Dq", 10240), ("5Dq", 1280), ("64C", 16384),
                                        ("20B", 2 * 600), ("3C5k", 30 * 2 * 48), ("3D", 12 * 77)])
def test_cuda_matches_oracle_at_baseline_sizes(name, nint, chb):
    """BASELINE.json shapes, a few seeded states each, against the oracle (and the live reference
    when oracle/_ref travelled to this box)."""
    from oracle.choreo_oracle import OracleSyst
    from oracle import ref_loader
    T = _tables(name)
    nbs = chb.NBodySystB200(T, nint)
    O = OracleSyst(T, nint, backend="fft")
    NBS = None
    if ref_loader.reference_available():
        NBS = ref_loader.build_named(name)
        NBS.nint = nint
    rng = np.random.default_rng(1234)
    B = 3
    x, dx = rng.random((B, nbs.nparams)), rng.random((B, nbs.nparams))
    seg = nbs.params_to_segmpos(x)
    hess = nbs.segmpos_dparams_to_action_hess(seg, dx)
    for fused in (True, False):
        nbs.set_fused(fused)
        act, grad = nbs.params_to_action_and_grad(x)
        for b in range(B):
            for chk in ([O] if NBS is None else [O, NBS]):
                assert relerr(act[b], chk.params_to_action(x[b])) < TOL_ACTION, (name, b, fused)
                assert relerr(grad[b], chk.params_to_action_grad(x[b])) < TOL_GRAD, (name, b, fused)
    for b in range(B):
        for chk in ([O] if NBS is None else [O, NBS]):
            assert relerr(seg[b], chk.params_to_segmpos(x[b])) < TOL_TRANSFORM
            assert relerr(hess[b], chk.segmpos_dparams_to_action_hess(seg[b], dx[b])) < TOL_HESS


def test_full_size_batch_properties(chb):
    """BASELINE config 2 at full size (3B, nint 4096, batch 4096) through size-independent properties."""
    import torch
    nbs = chb.NBodySystB200(_tables("3B"), 4096)
    B = 4096
    gen = torch.Generator(device="cuda").manual_seed(7)
    x = torch.rand((B, nbs.nparams), dtype=torch.float64, device="cuda", generator=gen)
    y = torch.rand((B,) + nbs.segmpos_shape, dtype=torch.float64, device="cuda", generator=gen)
    seg = nbs.params_to_segmpos(x)
    # adjointness <Tx, y> = <x, T^t y> per state (tests/test_NBodySyst.py:312-329)
    lhs = (seg * y).sum(dim=(1, 2, 3))
    rhs = (x * nbs.segmpos_to_params_T(y)).sum(dim=1)
    assert float(((lhs - rhs).abs() / lhs.abs()).max()) < 1e-11
    # round trip params -> segmpos -> params (tests/test_NBodySyst.py:178-259)
    assert float((nbs.segmpos_to_params(seg) - x).abs().max()) < 1e-10
    # fused == staged, batched == single, device path == host path
    nbs.set_fused(True)
    a1, g1 = nbs.params_to_action_and_grad(x)
    nbs.set_fused(False)
    a2, g2 = nbs.params_to_action_and_grad(x)
    assert float(((a1 - a2).abs() / a2.abs()).max()) < TOL_ACTION
    assert float((g1 - g2).abs().max() / g2.abs().max()) < TOL_GRAD
    idx = [0, 1, B // 2, B - 1]
    xh = x[idx].cpu().numpy()
    ah, gh = nbs.params_to_action_and_grad(xh)
    assert relerr(ah, a2[idx].cpu().numpy()) < TOL_ACTION and relerr(gh, g2[idx].cpu().numpy()) < TOL_GRAD
    # cached-segmpos variants agree with the params variants (tests/test_NBodySyst.py:902-926)
    g3 = nbs.segmpos_params_to_action_grad(seg, x)
    assert float((g3 - g2).abs().max() / g2.abs().max()) < TOL_GRAD
    dx = torch.rand((B, nbs.nparams), dtype=torch.float64, device="cuda", generator=gen)
    h1 = nbs.params_to_action_hess(x, dx)
    h2 = nbs.segmpos_dparams_to_action_hess(seg, dx)
    assert float((h1 - h2).abs().max() / h2.abs().max()) < TOL_HESS
    # Hessian symmetry <dy, H dx> = <dx, H dy>
    dy = torch.rand((B, nbs.nparams), dtype=torch.float64, device="cuda", generator=gen)
    h3 = nbs.segmpos_dparams_to_action_hess(seg, dy)
    s1, s2 = (dy * h2).sum(1), (dx * h3).sum(1)
    assert float(((s1 - s2).abs() / s1.abs().clamp_min(1e-300)).max()) < 1e-9


def test_fused_v4_matches_v3_and_oracle(chb):
    """Version 4 of the fused kernel (bulk asynchronous row copies, natural-order in-place passes) against version 3 and the
    oracle, with several states per persistent CTA (load barriers change phase, buffers are refilled while stores drain)."""
    import torch
    from oracle.choreo_oracle import OracleSyst
    T = _tables("3B")
    nbs = chb.NBodySystB200(T, 4096)
    assert nbs.fused_version == 4
    B = 148 * 3 + 17
    gen = torch.Generator(device="cuda").manual_seed(11)
    x = torch.rand((B, nbs.nparams), dtype=torch.float64, device="cuda", generator=gen)
    dx = torch.rand((B, nbs.nparams), dtype=torch.float64, device="cuda", generator=gen)
    a4, g4 = nbs.params_to_action_and_grad(x)
    seg = nbs.params_to_segmpos(x)
    h4 = nbs.segmpos_dparams_to_action_hess(seg, dx)
    nbs.set_fused(3)
    assert nbs.fused_version == 3
    a3, g3 = nbs.params_to_action_and_grad(x)
    h3 = nbs.segmpos_dparams_to_action_hess(seg, dx)
    nbs.set_fused(True)
    assert float(((a4 - a3).abs() / a3.abs()).max()) < TOL_ACTION
    assert float(((g4 - g3).abs().amax(1) / g3.abs().amax(1)).max()) < TOL_GRAD
    assert float(((h4 - h3).abs().amax(1) / h3.abs().amax(1)).max()) < TOL_HESS
    O = OracleSyst(T, 4096, backend="fft")
    for b in (0, 147, 148, B - 1):
        xb, dxb = x[b].cpu().numpy(), dx[b].cpu().numpy()
        assert relerr(a4[b].item(), O.params_to_action(xb)) < TOL_ACTION
        assert relerr(g4[b].cpu().numpy(), O.params_to_action_grad(xb)) < TOL_GRAD
        assert relerr(h4[b].cpu().numpy(), O.segmpos_dparams_to_action_hess(seg[b].cpu().numpy(), dxb)) < TOL_HESS
    # the host-pointer path (pipelined slices) runs the same kernel
    ah, gh = nbs.params_to_action_and_grad(x[:5].cpu().numpy())
    assert relerr(ah, a4[:5].cpu().numpy()) < TOL_ACTION and relerr(gh, g4[:5].cpu().numpy()) < TOL_GRAD


def test_gradient_is_derivative_of_action(chb):
    """Finite differences with the non-singular law (tests/test_NBodySyst.py:377-455)."""
    import dataclasses
    T = _tables("5Dq")
    T = dataclasses.replace(T, inter_law=1, inter_law_params=np.array([-0.5, 0.25, -1.5, 1.0]))  # inter_pow=+1
    nbs = chb.NBodySystB200(T, 20 * 24)
    rng = np.random.default_rng(3)
    x, dx = rng.random(nbs.nparams), rng.random(nbs.nparams)
    eps = 1e-6
    fd = (nbs.params_to_action(x + eps * dx) - nbs.params_to_action(x - eps * dx)) / (2 * eps)
    assert abs(fd - float(np.dot(nbs.params_to_action_grad(x), dx))) < 1e-6 * max(1.0, abs(fd))
    fdh = (nbs.params_to_action_grad(x + eps * dx) - nbs.params_to_action_grad(x - eps * dx)) / (2 * eps)
    assert relerr(nbs.params_to_action_hess(x, dx), fdh) < 1e-6


def test_edge_cases(chb):
    nbs = chb.NBodySystB200(_tables("3D"), 12)
    assert nbs.params_to_action_and_grad(np.zeros((0, nbs.nparams)))[1].shape == (0, nbs.nparams)
    with pytest.raises(ValueError, match="should be divisible by"):
        nbs.nint = 13
    with pytest.raises(ValueError):
        nbs.params_to_action_grad(np.zeros(nbs.nparams - 1))
    rng = np.random.default_rng(0)
    x = rng.random((5, nbs.nparams))
    ref = np.stack([nbs.params_to_action_grad(x[i]) for i in range(5)])
    for chunk in (1, 2, 0):
        nbs.set_chunk(chunk)
        assert np.array_equal(nbs.params_to_action_grad(x), ref)
    import torch
    xt = torch.from_numpy(x).cuda()
    assert np.array_equal(nbs.params_to_action_grad(xt).cpu().numpy(), ref)
    with pytest.raises(ValueError):
        nbs.params_to_action_grad(xt.float())


def test_host_solver_on_gpu_evaluations(chb):
    """The reference's own Newton-Krylov machinery (scipy KrylovJacobian, wired as _NBodySyst.pyx:2004-2032) driven
    by GPU evaluations converges to the figure-eight from the fixture's neighbourhood (config 1)."""
    import scipy.optimize
    z = np.load(os.path.join(ROOT, "tests", "golden", "sols.npz"))
    name = "3C-Figure_eight"
    pre = f"{name}/T/"
    T = chb.PlanTables.from_dict({k[len(pre):]: z[k] for k in z.files if k.startswith(pre)})
    nbs = chb.NBodySystB200(T, int(z[f"{name}/nint"]))
    x0 = z[f"{name}/params"] + 1e-4 * np.random.default_rng(0).standard_normal(nbs.nparams)
    jac = nbs.GetKrylovJacobian(True, {"method": "gmres"})
    sol = scipy.optimize.root(nbs.params_to_action_grad, x0, method="krylov",
                              options={"jac_options": {"method": "gmres"}, "fatol": 1e-11, "maxiter": 30})
    x = sol.x
    assert np.linalg.norm(nbs.params_to_action_grad(x), np.inf) < 1e-9
    assert abs(nbs.params_to_action(x) - float(z[f"{name}/stored_action"])) < 1e-8
    assert jac.NBS is nbs


def test_batched_newton_on_gpu(chb):
    """Batched Newton-Krylov on device tensors: converges to the stored figure-eight from perturbed starts, and the
    torch-tensor path agrees with the NumPy host path of the same driver."""
    import torch
    from choreographies2_b200.newton import newton_krylov_batched
    z = np.load(os.path.join(ROOT, "tests", "golden", "sols.npz"))
    name = "3C-Figure_eight"
    pre = f"{name}/T/"
    T = chb.PlanTables.from_dict({k[len(pre):]: z[k] for k in z.files if k.startswith(pre)})
    nbs = chb.NBodySystB200(T, int(z[f"{name}/nint"]))
    rng = np.random.default_rng(0)
    x0 = np.stack([z[f"{name}/params"] + a * rng.standard_normal(nbs.nparams) for a in (1e-4, 1e-3, 1e-2, 0.0, 3e-3)])
    rd = newton_krylov_batched(nbs, torch.from_numpy(x0).cuda(), f_tol=1e-11, maxiter=30)
    rh = newton_krylov_batched(nbs, x0, f_tol=1e-11, maxiter=30)
    assert (rd.status == 1).all() and (rh.status == 1).all() and rd.nit[3] == 0
    assert np.array_equal(rd.nit, rh.nit)
    xd = rd.x.cpu().numpy()
    assert np.abs(xd - rh.x).max() < 1e-6  # |dx| ~ f_tol / smallest Hessian eigenvalue (translation-like soft modes)
    for i in range(x0.shape[0]):
        assert abs(nbs.params_to_action(xd[i]) - float(z[f"{name}/stored_action"])) < 1e-9


@pytest.mark.gpu
def test_find_batched_on_gpu(chb):
    """Coarse-to-fine refinement driver (refine.py) on device tensors (library GMRES kernels, one plan per nint level) against
    the same driver on NumPy arrays through the host-pointer operators (array-level GMRES): same outcome, level and orbit."""
    import torch
    from choreographies2_b200 import refine
    T = chb.PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nint0 = 2 * T.nint_min * 16
    x_min, x_max = T.make_params_bounds(nint0, 0.1, 3, 50, 1e-16)
    starts = []
    for seed in range(12):
        np.random.seed(seed)
        starts.append(x_min + (x_max - x_min) * np.random.random(x_min.shape[0]))
    x0 = np.stack(starts)
    kw = dict(gradtol=1e-9, maxiter=40, n_reconverge_it_max=2, Newt_err_norm_max=1e-8, Newt_err_norm_max_save=1e-5)
    plans = {}

    def make(nint):
        plans[nint] = chb.NBodySystB200(T, nint)
        return plans[nint]

    rd = refine.find_batched(make, torch.from_numpy(x0).cuda(), nint0, **kw)
    rh = refine.find_batched(lambda nint: plans[nint], x0, nint0, **kw)
    good = np.isin(rh.outcome, (refine.FOUND, refine.APPROX))
    assert good.sum() >= 2
    assert np.array_equal(rd.outcome[good], rh.outcome[good]) and np.array_equal(rd.level[good], rh.level[good])
    for i in np.nonzero(good)[0]:
        nbs = plans[int(rh.nint[i])]
        a_d, a_h = nbs.params_to_action(rd.params[i].cpu().numpy()), nbs.params_to_action(rh.params[i])
        assert abs(a_d - a_h) < 1e-7 * abs(a_h)
