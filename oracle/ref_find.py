"""TEST INFRASTRUCTURE ONLY -- the search loop of the reference, replayed with the UNMODIFIED reference functions.

``Find_Choreo`` (choreo/find.py:25-444) mixes the numerics with file output, plotting and a store folder.  ``reference_find``
below is the same control flow for ONE start vector -- optimiser levels, best-iterate tracking, escape test, refinement
decisions (find.py:225-440) -- built from the reference's own ``nonlin_solve_pp``, ``current_best``, ``GetKrylovJacobian``,
``params_resize`` and the ``nint_fac`` setter, without the I/O.  The tests compare the batched device driver with it per seed,
and ``bench.py --impl reference`` / ``cpu_baseline`` time it as the CPU arm of the solves/hour metric.
"""
from __future__ import annotations

import numpy as np

FOUND, APPROX, NOT_CONVERGED, COLLISION, TOO_BIG_ON_ENTRY, ESCAPED = 1, 2, 3, 4, 5, 6

# Solver_* sections of tests/NewSym_data/3C/choreo_config.json as ChoreoLoadFromDict maps them (find.py:700-750)
CONFIG_3C = dict(
    gradtol_list=[1e-1, 1e-3, 1e-5, 1e-7, 1e-9, 1e-11], maxiter_list=[100, 1000, 1000, 1000, 500, 500],
    inner_maxiter_list=[30, 30, 50, 60, 70, 80], outer_k_list=[5, 5, 5, 5, 5, 7], krylov_method="gmres", line_search="wolfe",
    linesearch_smin=0.1, Newt_err_norm_max=1e-11, Newt_err_norm_max_save=1e-8, n_reconverge_it_max=4, mul_coarse_to_fine=3.0,
    nint_fac_init=600, coeff_ampl_o=0.1, k_infl=3, k_max=50, coeff_ampl_min=1e-16, max_norm_on_entry=1e20,
)


def find_starts(NBS, seeds, coeff_ampl_o=0.1, k_infl=3, k_max=50, coeff_ampl_min=1e-16):
    """Random starts exactly as find.py:105-108,158 draws them: ``np.random.seed(seed); x = x_avg + x_ptp * U[0,1)`` with
    ``x_avg = (x_min + x_min) / 2`` (sic) and ``x_ptp = x_max - x_min`` from ``Make_params_bounds``."""
    x_min, x_max = map(np.asarray, NBS.Make_params_bounds(coeff_ampl_o, k_infl, k_max, coeff_ampl_min))
    x_avg, x_ptp = (x_min + x_min) / 2, x_max - x_min
    out = []
    for s in seeds:
        np.random.seed(int(s))
        out.append(x_avg + x_ptp * np.random.random(NBS.nparams))
    return np.stack(out)


def reference_find(NBS, x0, nint_fac0, *, gradtol_list, maxiter_list, krylov_method="gmres", line_search="armijo",
                   linesearch_smin=1e-2, Newt_err_norm_max=1e-10, Newt_err_norm_max_save=1e-7, n_reconverge_it_max=4,
                   mul_coarse_to_fine=3.0, max_norm_on_entry=1e20, Check_Escape=True, outer_k_list=None, **_unused):
    """find.py:199-440 for one start vector.  Returns (outcome, refinement level, optimiser level, |grad|, |grad| at twice the
    resolution, parameters of the last best iterate (centred when saved), Newton iterations)."""
    from choreo.scipy_plus.nonlin import nonlin_solve_pp, current_best
    import warnings
    NBS.nint_fac = nint_fac0
    best = current_best(np.array(x0, dtype=np.float64), NBS.params_to_action_grad(np.array(x0, dtype=np.float64)))
    if not (best.f_norm < max_norm_on_entry):
        return TOO_BIG_ON_ENTRY, 0, 0, best.f_norm, np.nan, None, 0
    n_optim = len(gradtol_list)
    lvl = iopt = nit = 0
    while True:
        outer_k = 5 if outer_k_list is None else outer_k_list[iopt]
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            jac = NBS.GetKrylovJacobian(True, {"method": krylov_method, "rdiff": None, "outer_k": outer_k, "inner_tol": 0., "inner_M": None})

        def cb(x, f, f_norm):
            best.update(x, f, f_norm)
            return False
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            _, info = nonlin_solve_pp(F=NBS.params_to_action_grad, x0=np.copy(best.x), jacobian=jac, verbose=False,
                                      maxiter=maxiter_list[iopt], f_tol=gradtol_list[iopt], line_search=line_search, callback=cb,
                                      raise_exception=False, smin=linesearch_smin, full_output=True, tol_norm=np.linalg.norm)
        nit += int(info["nit"])
        segmpos = NBS.params_to_segmpos(best.x)
        if Check_Escape and NBS.DetectEscape(segmpos):
            return ESCAPED, lvl, iopt, best.f_norm, np.nan, best.x, nit
        fac = NBS.nint_fac
        x_fine = NBS.params_resize(best.x, 2 * fac)
        NBS.nint_fac = 2 * fac
        f_fine = NBS.params_to_action_grad(x_fine)
        ff = np.linalg.norm(f_fine)
        NBS.nint_fac = fac
        precise, precise_save = ff < Newt_err_norm_max, ff < Newt_err_norm_max_save
        can_change = iopt < n_optim - 1
        can_refine = lvl < n_reconverge_it_max
        needs_ref = (ff > mul_coarse_to_fine * best.f_norm) and (ff > precise_save)
        collision = (best.f_norm < 1e3 * Newt_err_norm_max) and (ff > 1e6 * best.f_norm)
        found = precise and not needs_ref
        approx = (not found) and precise_save and not can_change and (not can_refine or not needs_ref)
        if found or approx:
            xo = NBS.ComputeCenterOfMass(segmpos)  # find.py:376-380
            for idim in range(NBS.geodim):
                segmpos[:, :, idim] -= xo[idim]
            return (FOUND if found else APPROX), lvl, iopt, best.f_norm, ff, NBS.segmpos_to_params(segmpos), nit
        if (not can_refine or not needs_ref) and not can_change:
            return NOT_CONVERGED, lvl, iopt, best.f_norm, ff, best.x, nit
        if collision:
            return COLLISION, lvl, iopt, best.f_norm, ff, best.x, nit
        if needs_ref and can_refine:
            best = current_best(x_fine, f_fine)
            NBS.nint_fac = 2 * fac
            lvl += 1
        else:
            iopt += 1
