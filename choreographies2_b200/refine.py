"""Coarse-to-fine refinement of a batch of orbit searches.

Restates the outer loop of the reference's ``Find_Choreo`` (``choreo/find.py:225-432``) for many states at once: solve at the
current ``nint`` level (batched Newton-Krylov, 2-norm termination and best-iterate tracking exactly as find.py:204,249-264
wires them), re-express the best iterate at twice the resolution (``params_resize``), evaluate the gradient there, and decide
per state -- with the reference's own tests -- whether it is a solution, an approximate solution, a failure, a likely
collision, or has to be re-converged at the finer level.  States that continue are compacted and solved together at the next
level; every level has its own plan (``NBodySystB200``), built once and cached.

Escape detection (find.py:290-297), the re-centering of saved solutions on their centre of mass (find.py:376-380) and the list
of optimiser parameter sets (``gradtol_list`` / ``maxiter_list``, find.py:225-271, 424-440) are included.  Not restated here
(host-side bookkeeping of the reference that does not touch the numerics): descriptors / images written per solution; the
duplicate search is ``group_duplicates`` / ``DuplicateStore`` below (hash + ``TestHashSame`` against the orbits found so far).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Sequence

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
    optim_level: Any = None      # [B] optimiser level (index into gradtol_list) at which the state stopped


def find_batched(make_nbs: Callable[[int], Any], x0, nint0: int, *, gradtol: float = 1e-11, maxiter: int = 50,
                 gradtol_list: Optional[Sequence[float]] = None, maxiter_list: Optional[Sequence[int]] = None,
                 inner_maxiter_list: Optional[Sequence[int]] = None, krylov_method: str = "gmres",
                 n_reconverge_it_max: int = 4, Newt_err_norm_max: float = 1e-10, Newt_err_norm_max_save: float = 1e-7,
                 mul_coarse_to_fine: float = 3.0, max_norm_on_entry: float = 1e20, linesearch_smin: float = 1e-2,
                 line_search: Optional[str] = "armijo", restart: int = 20, Check_Escape: bool = True,
                 recenter: bool = True) -> RefineResult:
    """``make_nbs(nint)`` returns the evaluator of one level (called once per level); ``x0`` ``[B, nparams(nint0)]`` NumPy
    array or torch CUDA tensor.  Keyword names and defaults follow the reference's ``Find_Choreo`` arguments.

    Optimiser schedule (find.py:225-271, 424-440): ``gradtol_list`` / ``maxiter_list`` give the termination tolerance and the
    Newton iteration limit of every optimiser level (``Solver_Loop`` of a ``choreo_config.json``; the scalars ``gradtol`` /
    ``maxiter`` are the single-level case).  Each state carries its own (refinement level, optimiser level): after a solve it
    is saved, abandoned, re-converged at twice the resolution (same optimiser level) or moved to the next optimiser level,
    exactly as the reference decides; states that share (refinement level, optimiser level) are solved together.
    ``krylov_method``: the reference wires ``inner_maxiter_list`` into the linear solver only for ``'lgmres'``
    (find.py:246-249); with ``'gmres'`` scipy's ``KrylovJacobian`` default restart (20, ``restart`` here) is used at every
    level.  ``'lgmres'`` is not built on the device and is rejected."""
    if krylov_method != "gmres":
        raise ValueError(f"krylov_method {krylov_method!r}: the batched driver implements 'gmres' (scipy KrylovJacobian's default "
                         "for the reference's configurations); 'lgmres' is not built")
    gradtols = [float(gradtol)] if gradtol_list is None else [float(v) for v in gradtol_list]
    maxiters = [int(maxiter)] * len(gradtols) if maxiter_list is None else [int(v) for v in maxiter_list]
    if len(maxiters) != len(gradtols) or (inner_maxiter_list is not None and len(inner_maxiter_list) != len(gradtols)):
        raise ValueError("gradtol_list, maxiter_list and inner_maxiter_list must have the same length")
    n_optim = len(gradtols)
    xp = _XP(x0)
    B = int(x0.shape[0])
    cache: Dict[int, Any] = {}

    def nbs_at(nint):
        if nint not in cache:
            cache[nint] = make_nbs(int(nint))
        return cache[nint]

    outcome = np.zeros(B, dtype=np.int64)
    level = np.zeros(B, dtype=np.int64)
    optim_level = np.zeros(B, dtype=np.int64)
    nint_out = np.zeros(B, dtype=np.int64)
    f_norm = np.full(B, np.nan)
    f_fine_norm = np.full(B, np.nan)
    nit_tot = np.zeros(B, dtype=np.int64)
    params: List[Any] = [None] * B
    n_grad = n_hess = 0

    def cat(parts):
        if len(parts) == 1:
            return parts[0]
        return xp.t.cat(parts, 0) if xp.torch else np.concatenate(parts, 0)

    # norm on entry (find.py:201-209)
    F0 = nbs_at(int(nint0)).params_to_action_grad(x0)
    n_grad += B
    n0 = xp.to_numpy(xp.norm(F0))
    ok = n0 < max_norm_on_entry
    outcome[~ok] = TOO_BIG_ON_ENTRY
    nint_out[~ok] = nint0
    f_norm[~ok] = n0[~ok]
    # work queue over (refinement level, optimiser level): every transition goes to a later key, one sweep in sorted order
    queue: Dict[Any, List[Any]] = {}
    keep = np.nonzero(ok)[0]
    if keep.size:
        queue[(0, 0)] = [(keep, x0 if keep.size == B else x0[xp.from_numpy_index(keep)])]
    while queue:
        key = min(queue.keys())
        parts = queue.pop(key)
        lvl, iopt = key
        ids = np.concatenate([p[0] for p in parts])
        x = cat([p[1] for p in parts])
        nint = int(nint0) * 2 ** lvl
        nbs = nbs_at(nint)
        res = newton_krylov_batched(nbs, x, f_tol=gradtols[iopt], maxiter=maxiters[iopt], restart=restart, smin=linesearch_smin,
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
        # decisions, find.py:322-352
        precise = ff < Newt_err_norm_max
        precise_save = ff < Newt_err_norm_max_save
        can_change = iopt < n_optim - 1
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
        hit = go & precise_save & (not can_change) & (~needs_ref | (not can_refine))
        out[hit], go = APPROX, go & ~hit
        hit = go & (~needs_ref | (not can_refine)) & (not can_change)
        out[hit], go = NOT_CONVERGED, go & ~hit
        hit = go & collision
        out[hit], go = COLLISION, go & ~hit
        stop = ~go
        sid = ids[stop]
        outcome[sid], level[sid], optim_level[sid], nint_out[sid] = out[stop], lvl, iopt, nint
        f_norm[sid], f_fine_norm[sid] = fn[stop], ff[stop]
        saved = np.nonzero(stop & ((out == FOUND) | (out == APPROX)))[0]
        if recenter and body_tables and saved.size > 0:  # find.py:376-380: saved solutions are centred on their centre of mass
            sel = xp.from_numpy_index(saved)
            seg = nbs.params_to_segmpos(best_x[sel])
            seg = seg - nbs.ComputeCenterOfMass(seg)[:, None, None, :]
            best_x[sel] = nbs.segmpos_to_params(seg)
        for k in np.nonzero(stop)[0]:
            params[ids[k]] = best_x[int(k)]
        # find.py:424-440: re-converge at the finer level from the resized best iterate, else the next optimiser level
        ref = np.nonzero(go & needs_ref & can_refine)[0]
        if ref.size:
            queue.setdefault((lvl + 1, iopt), []).append((ids[ref], x_fine[xp.from_numpy_index(ref)]))
        chg = np.nonzero(go & ~(needs_ref & can_refine) & can_change)[0]
        if chg.size:
            queue.setdefault((lvl, iopt + 1), []).append((ids[chg], best_x[xp.from_numpy_index(chg)]))
    return RefineResult(outcome=outcome, level=level, nint=nint_out, f_norm=f_norm, f_fine_norm=f_fine_norm, params=params,
                        newton_iterations=nit_tot, n_grad_evals=n_grad, n_hess_evals=n_hess, optim_level=optim_level)


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
