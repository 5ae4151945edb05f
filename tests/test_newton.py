"""Batched Newton-Krylov driver (choreographies2_b200/newton.py) against scipy's GMRES and the reference's own solver.

CPU tests run the driver over the host emulation of the kernels (tests/hostsim); the reference arm is the UNMODIFIED
``choreo.scipy_plus.nonlin.nonlin_solve_pp`` + ``NBodySyst.GetKrylovJacobian`` from oracle/_ref (skipped when it was not
built), started from the same vectors with the same settings."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from choreographies2_b200 import NBodySystB200, PlanTables
from choreographies2_b200.newton import _XP, gmres_batched, newton_krylov_batched


def test_gmres_batched_matches_scipy():
    import scipy.sparse.linalg as sla
    rng = np.random.default_rng(0)
    A, n = 6, 48
    Ms = [np.eye(n) * 3 + 0.3 * rng.standard_normal((n, n)) for _ in range(A)]
    rhs = rng.standard_normal((A, n))
    rhs[3] = 0.0
    rtol = np.array([1e-3, 1e-8, 1e-1, 1e-3, 1e-12, 0.5])

    def mv(idx, v):
        ids = range(A) if idx is None else idx
        return np.stack([Ms[i] @ v[k] for k, i in enumerate(ids)])

    sol, k, nmv = gmres_batched(_XP(rhs), mv, rhs, rtol, 20)
    assert nmv == int(k.sum())
    for b in range(A):
        ref, _ = sla.gmres(Ms[b], rhs[b], rtol=rtol[b], atol=0, restart=20, maxiter=1)
        assert np.abs(sol[b] - ref).max() < 1e-12
    assert k[3] == 0 and k[5] < k[0] <= 20


@pytest.mark.parametrize("mode", ["arnoldi", "step"])
def test_gmres_batched_library_steps_match_scipy(mode):
    """The same check with the library's Arnoldi step (chb_arnoldi_step_dev) and whole inner iteration
    (chb_gmres_step_dev) in place of the array-level Gram-Schmidt / Givens code; host emulation build."""
    import scipy.sparse.linalg as sla
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96, _lib_handle=build_hostsim.load())
    rng = np.random.default_rng(1)
    A, n = 7, 61
    Ms = [np.eye(n) + 0.2 * rng.standard_normal((n, n)) for _ in range(A)]
    rhs = rng.standard_normal((A, n))
    rhs[2] = 0.0
    rtol = np.array([1e-3, 1e-8, 1e-1, 1e-3, 1e-12, 0.5, 1e-6])

    def mv(idx, v):
        ids = range(A) if idx is None else idx
        return np.stack([Ms[i] @ v[k] for k, i in enumerate(ids)])

    l0 = nbs.kernel_launches
    sol, k, nmv = gmres_batched(_XP(rhs), mv, rhs, rtol, 20, nbs.arnoldi_step, nbs.gmres_step if mode == "step" else None)
    assert nbs.kernel_launches - l0 == int(k.max())  # one launch per inner iteration
    sol0, k0, _ = gmres_batched(_XP(rhs), mv, rhs, rtol, 20)
    assert np.array_equal(k, k0) and nmv == int(k.sum())
    for b in range(A):
        ref, _ = sla.gmres(Ms[b], rhs[b], rtol=rtol[b], atol=0, restart=20, maxiter=1)
        assert np.abs(sol[b] - ref).max() < 1e-11
    assert k[2] == 0


def _figure_eight(lib):
    z = np.load(os.path.join(ROOT, "tests", "golden", "sols.npz"))
    name = "3C-Figure_eight"
    pre = f"{name}/T/"
    T = PlanTables.from_dict({k[len(pre):]: z[k] for k in z.files if k.startswith(pre)})
    return NBodySystB200(T, int(z[f"{name}/nint"]), _lib_handle=lib), z[f"{name}/params"], float(z[f"{name}/stored_action"])


@pytest.fixture(scope="module")
def sim():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    return build_hostsim.load()


def test_batched_newton_converges_to_fixture(sim):
    nbs, xs, action = _figure_eight(sim)
    rng = np.random.default_rng(0)
    x0 = np.stack([xs + a * rng.standard_normal(nbs.nparams) for a in (1e-4, 1e-3, 1e-2, 0.0)])
    res = newton_krylov_batched(nbs, x0, f_tol=1e-11, maxiter=30)
    assert (res.status == 1).all() and res.nit[3] == 0 and (res.fnorm <= 1e-11).all()
    for i in range(4):
        assert abs(nbs.params_to_action(res.x[i]) - action) < 1e-9
    # states do not influence each other: a batch equals the same states solved alone
    one = newton_krylov_batched(nbs, x0[2:3], f_tol=1e-11, maxiter=30)
    assert np.array_equal(one.x[0], res.x[2]) and one.nit[0] == res.nit[2]


def test_batched_newton_matches_reference_solver(sim):
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    choreo = ref_loader.import_reference()
    from choreo.scipy_plus.nonlin import nonlin_solve_pp
    NBS = ref_loader.build_named("3C")
    nint = 3 * 2 * 32
    NBS.nint = nint
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, nint, _lib_handle=sim)
    assert nbs.nparams == NBS.nparams
    # starts like find.py:105-108,158: x = x_avg + x_ptp * U[0,1)
    x_min, x_max = NBS.Make_params_bounds(0.1, 3, 50, 1e-16)
    x_min, x_max = np.asarray(x_min), np.asarray(x_max)
    x_avg, x_ptp = (x_min + x_min) / 2, x_max - x_min  # sic: find.py:106-107
    starts = []
    for seed in range(6):
        np.random.seed(seed)
        starts.append(x_avg + x_ptp * np.random.random(NBS.nparams))
    x0 = np.stack(starts)
    res = newton_krylov_batched(nbs, x0, f_tol=1e-9, maxiter=60)
    nconv = 0
    for i in range(x0.shape[0]):
        jac = NBS.GetKrylovJacobian(True, {"method": "gmres"})
        try:
            xr, info = nonlin_solve_pp(NBS.params_to_action_grad, x0[i].copy(), jacobian=jac, f_tol=1e-9, maxiter=60,
                                       line_search="armijo", full_output=True, raise_exception=False, smin=1e-2)
        except Exception:
            continue
        if info["status"] != 1:
            continue
        # the reference converged from this start: so must the batched driver, to the same orbit
        assert res.status[i] == 1, i
        nconv += 1
        ar, ab = NBS.params_to_action(xr), nbs.params_to_action(res.x[i])
        assert abs(ar - ab) < 1e-7 * max(1.0, abs(ar)), (i, ar, ab)
        assert abs(int(res.nit[i]) - int(info["nit"])) <= 3, (i, res.nit[i], info["nit"])
    assert nconv >= 1


def test_make_params_bounds_matches_reference():
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    ref_loader.import_reference()
    from choreographies2_b200 import extract_tables
    for name, nint in (("3C", 192), ("5Dq", 80), ("3B", 64), ("3D", 48)):
        NBS = ref_loader.build_named(name)
        NBS.nint = nint
        lo, hi = map(np.asarray, NBS.Make_params_bounds(0.1, 3, 50, 1e-16))
        mlo, mhi = extract_tables(NBS).make_params_bounds(nint, 0.1, 3, 50, 1e-16)
        assert np.abs(lo - mlo).max() < 1e-300 + 1e-15 * np.abs(lo).max() and np.abs(hi - mhi).max() < 1e-15 * np.abs(hi).max()


@pytest.mark.parametrize("line_search,smin", [("armijo", 1e-2), ("wolfe", 0.1), (None, 0.1)])
def test_library_driver_tracks_reference_iterations(line_search, smin, backend):
    """chb_newton_krylov_dev (whole solve in the library: GMRES(20), forcing terms, Armijo / Wolfe line search on the device)
    against the unmodified nonlin_solve_pp + KrylovJacobian iteration by iteration, from find.py-style starts: the iterates
    agree to rounding for the first iterations (the search is chaotic far from a solution, so rounding grows slowly)."""
    from conftest import make_nbs, need_reference
    ref_loader = need_reference()
    from oracle import ref_find
    from choreo.scipy_plus.nonlin import nonlin_solve_pp
    from choreographies2_b200 import extract_tables
    import warnings
    NBS = ref_loader.build_named("3C")
    NBS.nint_fac = 16
    nbs = make_nbs(extract_tables(NBS), int(NBS.nint), backend)
    x0 = ref_find.find_starts(NBS, range(4))
    # the wolfe search differentiates phi by a forward difference of step ~1e-8 (nonlin.py:211-213): its step lengths, and so the
    # iterates, carry the rounding of phi amplified by 1e8 -- in the reference as much as here
    tols = ((1, 1e-6), (2, 1e-6), (4, 1e-4)) if line_search == "wolfe" else ((1, 1e-10), (2, 1e-9), (4, 1e-6))
    for mit, tol in tols:
        if backend is None:
            import torch
            x = torch.from_numpy(x0).cuda()
        else:
            x = x0.copy()
        r = nbs.newton_krylov(x, 1e-13, mit, 20, smin, line_search, "l2", False)
        xb = x.cpu().numpy() if backend is None else x
        nit = r["nit"].cpu().numpy() if backend is None else r["nit"]
        for i in range(4):
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                jac = NBS.GetKrylovJacobian(True, {"method": "gmres"})
                xr, info = nonlin_solve_pp(NBS.params_to_action_grad, x0[i].copy(), jacobian=jac, f_tol=1e-13, maxiter=mit,
                                           line_search=line_search, full_output=True, raise_exception=False, smin=smin,
                                           tol_norm=np.linalg.norm)
            assert int(nit[i]) == int(info["nit"]) == mit
            assert np.abs(xr - xb[i]).max() < tol * max(1.0, np.abs(xr).max()), (line_search, mit, i)


def test_line_search_argument_is_validated(sim):
    nbs, xs, _ = _figure_eight(sim)
    with pytest.raises(ValueError, match="line_search"):
        newton_krylov_batched(nbs, xs[None, :], line_search="wolf")
    with pytest.raises(ValueError, match="line_search"):
        nbs.newton_krylov(xs[None, :].copy(), 1e-9, 5, line_search="strong")
    res = newton_krylov_batched(nbs, xs[None, :] + 1e-3, f_tol=1e-11, maxiter=30, line_search="wolfe", smin=0.1, tol_norm="l2")
    assert res.status[0] == 1


@pytest.mark.parametrize("split", [0, 8])
def test_library_driver_staged_path_equals_fused(split, backend):
    """The library Newton driver on the staged kernels (residue-split transforms, CTAs of stopped states return at once: byte masks
    of the driver) against the same solve on the fused kernel: same iteration counts, same iterates to rounding.  States stop at
    different iterations, so the masks are exercised."""
    import os
    from conftest import ROOT, make_nbs
    from choreographies2_b200 import PlanTables, refine
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nint = 768
    x0 = refine.find_starts(T, nint, range(6), **refine.RANDOM_CONFIG_3C)
    out = []
    for fused in (True, False):
        nbs = make_nbs(T, nint, backend)
        nbs.set_fused(fused)
        if not fused:
            assert nbs.staged_kernels[0] & 8
            nbs.set_transform_split(split)
        if backend is None:
            import torch
            x = torch.from_numpy(x0).cuda()
        else:
            x = x0.copy()
        r = nbs.newton_krylov(x, 1e-3, 12, 20, 0.1, "wolfe", "l2", True)
        get = (lambda t: t.cpu().numpy()) if backend is None else (lambda t: np.asarray(t))
        out.append((get(x), get(r["nit"]), get(r["status"]), get(r["best_x"])))
    (xf, nf, sf, bf), (xs, ns, ss, bs) = out
    assert len(set(nf.tolist())) > 1, "the starts should stop at different iterations"
    assert np.array_equal(nf, ns) and np.array_equal(sf, ss)
    # the wolfe search differentiates phi by a forward difference (step ~1e-8): rounding differences between the two kernel sets
    # are amplified by every line search, as in test_library_driver_tracks_reference_iterations
    assert np.abs(xf - xs).max() < 1e-3 * max(1.0, np.abs(xf).max())
    assert np.abs(bf - bs).max() < 1e-3 * max(1.0, np.abs(bf).max())
