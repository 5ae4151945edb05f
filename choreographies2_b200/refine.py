"""Coarse-to-fine refinement of a batch of orbit searches.

Restates the outer loop of the reference's ``Find_Choreo`` (``choreo/find.py:225-432``) for many states at once: solve at the
current ``nint`` level (batched Newton-Krylov, 2-norm termination and best-iterate tracking exactly as find.py:204,249-264
wires them), re-express the best iterate at twice the resolution (``params_resize``), evaluate the gradient there, and decide
per state -- with the reference's own tests -- whether it is a solution, an approximate solution, a failure, a likely
collision, or has to be re-converged at the finer level.  States that continue are compacted and solved together at the next
level; every level has its own plan (``NBodySystB200``), built once and cached.

Escape detection (find.py:290-297) and the re-centering of saved solutions on their centre of mass (find.py:376-380) are
included.  Not restated here (host-side bookkeeping of the reference that does not touch the numerics): the duplicate search
against the store folder (``segmpos_to_hash`` + ``TestHashSame`` are the pieces), descriptors / images written per solution,
the list of optimiser parameter sets (this driver runs the single-set case, ``n_optim_param = 1``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional

import numpy as np

from .newton import _XP, newton_krylov_batched

# per-state outcome codes
FOUND, APPROX, NOT_CONVERGED, COLLISION, TOO_BIG_ON_ENTRY, ESCAPED = 1, 2, 3, 4, 5, 6
OUTCOME = {FOUND: "found solution", APPROX: "found approximate solution",
           NOT_CONVERGED: "could not converge within prescribed optimizer and refinement parameters",
           COLLISION: "likely narrowing in on a collision solution", TOO_BIG_ON_ENTRY: "norm on entry too big",
           ESCAPED: "one loop escaped"}


@dataclass
class RefineResult:
    outcome: np.ndarray          # [B] one of the codes above
    level: np.ndarray            # [B] refinement level at which the state stopped (0 = the initial nint)
    nint: np.ndarray             # [B] nint of that level
    f_norm: np.ndarray           # [B] 2-norm of the action gradient of the best iterate at that level
    f_fine_norm: np.ndarray      # [B] 2-norm of the gradient of the same iterate re-expressed at twice the resolution
    params: List[Any]            # [B] the best iterate of each state (length nparams of ITS level; None for TOO_BIG_ON_ENTRY)
    newton_iterations: np.ndarray  # [B] Newton iterations summed over the levels
    n_grad_evals: int
    n_hess_evals: int


def find_batched(make_nbs: Callable[[int], Any], x0, nint0: int, *, gradtol: float = 1e-11, maxiter: int = 50,
                 n_reconverge_it_max: int = 4, Newt_err_norm_max: float = 1e-10, Newt_err_norm_max_save: float = 1e-7,
                 mul_coarse_to_fine: float = 3.0, max_norm_on_entry: float = 1e20, linesearch_smin: float = 1e-2,
                 line_search: Optional[str] = "armijo", restart: int = 20, Check_Escape: bool = True,
                 recenter: bool = True) -> RefineResult:
    """``make_nbs(nint)`` returns the evaluator of one level (called once per level); ``x0`` ``[B, nparams(nint0)]`` NumPy
    array or torch CUDA tensor.  Keyword names and defaults follow the reference's ``Find_Choreo`` arguments."""
    xp = _XP(x0)
    B = int(x0.shape[0])
    cache: Dict[int, Any] = {}

    def nbs_at(nint):
        if nint not in cache:
            cache[nint] = make_nbs(int(nint))
        return cache[nint]

    outcome = np.zeros(B, dtype=np.int64)
    level = np.zeros(B, dtype=np.int64)
    nint_out = np.zeros(B, dtype=np.int64)
    f_norm = np.full(B, np.nan)
    f_fine_norm = np.full(B, np.nan)
    nit_tot = np.zeros(B, dtype=np.int64)
    params: List[Any] = [None] * B
    n_grad = n_hess = 0

    ids = np.arange(B)
    x = x0
    nint = int(nint0)
    # norm on entry (find.py:201-209)
    F0 = nbs_at(nint).params_to_action_grad(x)
    n_grad += B
    n0 = xp.to_numpy(xp.norm(F0))
    ok = n0 < max_norm_on_entry
    outcome[~ok] = TOO_BIG_ON_ENTRY
    nint_out[~ok] = nint
    f_norm[~ok] = n0[~ok]
    if not ok.all():
        keep = np.nonzero(ok)[0]
        ids, x = ids[keep], x[xp.from_numpy_index(keep)]
    lvl = 0
    while ids.size > 0:
        nbs = nbs_at(nint)
        res = newton_krylov_batched(nbs, x, f_tol=gradtol, maxiter=maxiter, restart=restart, smin=linesearch_smin,
                                    line_search=line_search, tol_norm="l2", track_best=True)
        n_grad += res.n_grad_evals
        n_hess += res.n_hess_evals
        nit_tot[ids] += res.nit
        best_x = res.best_x
        fn = xp.to_numpy(res.best_fnorm)
        # the best iterate at twice the resolution (find.py:312-320)
        nbs_f = nbs_at(2 * nint)
        x_fine = nbs.tables.params_resize(best_x, nint, 2 * nint)
        ff = xp.to_numpy(xp.norm(nbs_f.params_to_action_grad(x_fine)))
        n_grad += int(ids.size)
        # decisions, find.py:322-352 with n_optim_param = 1 (CanChangeOptimParams = False)
        precise = ff < Newt_err_norm_max
        precise_save = ff < Newt_err_norm_max_save
        can_refine = lvl < n_reconverge_it_max
        # the reference compares f_fine_norm with the BOOLEAN ParamPreciseEnoughSave here (find.py:326): kept as is
        needs_ref = (ff > mul_coarse_to_fine * fn) & (ff > precise_save.astype(np.float64))
        collision = (fn < 1e3 * Newt_err_norm_max) & (ff > 1e6 * fn)
        out = np.zeros(ids.size, dtype=np.int64)
        go = np.ones(ids.size, dtype=bool)
        body_tables = getattr(nbs.tables, "bodysegm", None) is not None
        if Check_Escape and body_tables:  # find.py:290-297: an escaped loop ends the attempt before anything else is looked at
            esc = np.asarray(xp.to_numpy(nbs.DetectEscape(nbs.params_to_segmpos(best_x)))).reshape(-1)
            out[esc], go = ESCAPED, go & ~esc
        hit = go & precise & ~needs_ref
        out[hit], go = FOUND, go & ~hit
        hit = go & precise_save & (~needs_ref | (not can_refine))
        out[hit], go = APPROX, go & ~hit
        hit = go & (~needs_ref | (not can_refine))
        out[hit], go = NOT_CONVERGED, go & ~hit
        hit = go & collision
        out[hit], go = COLLISION, go & ~hit
        stop = ~go
        sid = ids[stop]
        outcome[sid], level[sid], nint_out[sid] = out[stop], lvl, nint
        f_norm[sid], f_fine_norm[sid] = fn[stop], ff[stop]
        saved = np.nonzero(stop & ((out == FOUND) | (out == APPROX)))[0]
        if recenter and body_tables and saved.size > 0:  # find.py:376-380: saved solutions are centred on their centre of mass
            sel = xp.from_numpy_index(saved)
            seg = nbs.params_to_segmpos(best_x[sel])
            seg = seg - nbs.ComputeCenterOfMass(seg)[:, None, None, :]
            best_x[sel] = nbs.segmpos_to_params(seg)
        for k in np.nonzero(stop)[0]:
            params[ids[k]] = best_x[int(k)]
        # re-converge the others at the finer level, starting from the resized best iterate (find.py:426-432)
        cont = np.nonzero(go)[0]
        ids = ids[cont]
        x = x_fine[xp.from_numpy_index(cont)] if cont.size < x_fine.shape[0] else x_fine
        nint *= 2
        lvl += 1
    return RefineResult(outcome=outcome, level=level, nint=nint_out, f_norm=f_norm, f_fine_norm=f_fine_norm, params=params,
                        newton_iterations=nit_tot, n_grad_evals=n_grad, n_hess_evals=n_hess)


def group_duplicates(nbs, hashes, rtol: float = 1e-5, detect_multiples: bool = True) -> np.ndarray:
    """Duplicate filter of the search loop over a batch (``Check_Duplicates``, find.py:905-935, minus the store folder): orbit
    ``b`` joins the first earlier orbit whose hash passes ``TestHashSame`` (_NBodySyst.pyx:2036-2067), else it starts a new
    group.  ``hashes``: ``[B, nexp]`` from ``nbs.segmpos_to_hash`` (NumPy array or device tensor).  Returns the group
    representative (index of the first member) of every orbit, ``[B]``."""
    H = hashes.detach().cpu().numpy() if type(hashes).__module__.startswith("torch") else np.asarray(hashes)
    rep = np.arange(H.shape[0])
    firsts: List[int] = []
    for b in range(H.shape[0]):
        for f in firsts:
            if nbs.TestHashSame(H[f], H[b], rtol, detect_multiples):
                rep[b] = f
                break
        else:
            firsts.append(b)
    return rep
