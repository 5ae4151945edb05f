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

import time
from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from .newton import _XP, newton_krylov_batched

# Solver settings of the reference's 3C search configuration (tests/NewSym_data/3C/choreo_config.json, sections Solver_Optim /
# Solver_Loop / Solver_Discr / Phys_Random as choreo.find.ChoreoLoadFromDict maps them, find.py:700-750): the keyword arguments of
# find_batched for BASELINE config 1 and the sweep of config 5
SEARCH_CONFIG_3C = dict(
    gradtol_list=[1e-1, 1e-3, 1e-5, 1e-7, 1e-9, 1e-11], maxiter_list=[100, 1000, 1000, 1000, 500, 500], krylov_method="gmres",
    line_search="wolfe", linesearch_smin=0.1, Newt_err_norm_max=1e-11, Newt_err_norm_max_save=1e-8, n_reconverge_it_max=4,
    mul_coarse_to_fine=3.0)
RANDOM_CONFIG_3C = dict(coeff_ampl_o=0.1, k_infl=3, k_max=50, coeff_ampl_min=1e-16)  # Make_params_bounds arguments (Phys_Random)


def find_starts(tables, nint: int, seeds, coeff_ampl_o=0.1, k_infl=3, k_max=50, coeff_ampl_min=1e-16) -> np.ndarray:
    """Random starts exactly as find.py:105-108,158 draws them: ``np.random.seed(seed); x = x_avg + x_ptp * U[0, 1)`` with
    ``x_avg = (x_min + x_min) / 2`` (sic) and ``x_ptp = x_max - x_min`` from ``Make_params_bounds`` (one seed per start)."""
    x_min, x_max = tables.make_params_bounds(nint, coeff_ampl_o, k_infl, k_max, coeff_ampl_min)
    x_avg, x_ptp = (x_min + x_min) / 2, x_max - x_min
    out = np.empty((len(seeds), x_min.shape[0]))
    for r, s in enumerate(seeds):
        # the stream of np.random.seed(s); np.random.random(n), drawn from a private generator: safe from several threads
        out[r] = x_avg + x_ptp * np.random.RandomState(int(s)).random_sample(x_min.shape[0])
    return out


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
                 recenter: bool = True, trace: Optional[list] = None) -> RefineResult:
    """``make_nbs(nint)`` returns the evaluator of one level (called once per level); ``x0`` ``[B, nparams(nint0)]`` NumPy
    array or torch CUDA tensor.  Keyword names and defaults follow the reference's ``Find_Choreo`` arguments.

    Optimiser schedule (find.py:225-271, 424-440): ``gradtol_list`` / ``maxiter_list`` give the termination tolerance and the
    Newton iteration limit of every optimiser level (``Solver_Loop`` of a ``choreo_config.json``; the scalars ``gradtol`` /
    ``maxiter`` are the single-level case).  Each state carries its own (refinement level, optimiser level): after a solve it
    is saved, abandoned, re-converged at twice the resolution (same optimiser level) or moved to the next optimiser level,
    exactly as the reference decides; states that share (refinement level, optimiser level) are solved together.
    ``krylov_method``: the reference wires ``inner_maxiter_list`` into the linear solver only for ``'lgmres'``
    (find.py:246-249); with ``'gmres'`` scipy's ``KrylovJacobian`` default restart (20, ``restart`` here) is used at every
    level.  ``'lgmres'`` is not built on the device and is rejected.  ``trace``: a list that receives one record per solved
    (refinement level, optimiser level) group: states, nint, seconds (device-synchronised), Newton iterations."""
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
        if trace is not None:
            xp.sync()
            t_key = time.perf_counter()
        res = newton_krylov_batched(nbs, x, f_tol=gradtols[iopt], maxiter=maxiters[iopt], restart=restart, smin=linesearch_smin,
                                    line_search=line_search, tol_norm="l2", track_best=True)
        n_grad += res.n_grad_evals
        n_hess += res.n_hess_evals
        nit_tot[ids] += res.nit
        if trace is not None:
            xp.sync()
            trace.append(dict(level=int(lvl), optim_level=int(iopt), nint=nint, states=int(ids.size), seconds=time.perf_counter() - t_key,
                              nit_max=int(np.max(res.nit)), nit_sum=int(np.sum(res.nit)), grad_evals=int(res.n_grad_evals),
                              hess_evals=int(res.n_hess_evals)))
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


def hash_test_values(nbs, Hash_a, Hash_b, detect_multiples: bool = True) -> np.ndarray:
    """The test values of ``TestHashSame`` (_NBodySyst.pyx:2036-2067), vectorised: ``Hash_a`` ``[nexp]`` against every row of
    ``Hash_b`` ``[M, nexp]``; two orbits are "the same" when the infinity norm of a row is below ``rtol``."""
    Ha, Hb = np.asarray(Hash_a, dtype=np.float64), np.atleast_2d(np.asarray(Hash_b, dtype=np.float64))
    exps = np.asarray(nbs.Hash_exp, dtype=np.float64)
    with np.errstate(all="ignore"):
        if detect_multiples and nbs.LawIsHomo:
            pw = (nbs.Homo_exp - 1.0) / (2.0 * exps)
            ref = (Ha[0] / Hb[:, 0]) ** pw[0]
            test = (Ha[None, 1:] / Hb[:, 1:]) ** pw[None, 1:] - ref[:, None]
        else:
            test = (Hb - Ha[None, :]) / (Hb + Ha[None, :])
    return test


def group_duplicates(nbs, hashes, rtol: float = 1e-5, detect_multiples: bool = True) -> np.ndarray:
    """Duplicate filter of the search loop over a batch (``Check_Duplicates``, find.py:905-935, minus the store folder): orbit
    ``b`` joins the first earlier orbit whose hash passes ``TestHashSame`` (_NBodySyst.pyx:2036-2067), else it starts a new
    group.  ``hashes``: ``[B, nexp]`` from ``nbs.segmpos_to_hash`` (NumPy array or device tensor).  Returns the group
    representative (index of the first member) of every orbit, ``[B]``.  One vectorised comparison per group."""
    H = hashes.detach().cpu().numpy() if type(hashes).__module__.startswith("torch") else np.asarray(hashes)
    B = H.shape[0]
    rep = np.full(B, -1, dtype=np.int64)
    for b in range(B):
        if rep[b] >= 0:
            continue
        rep[b] = b
        rest = np.nonzero(rep[b + 1:] < 0)[0] + b + 1
        if rest.size:
            t = hash_test_values(nbs, H[b], H[rest], detect_multiples)
            same = np.abs(np.nan_to_num(t, nan=np.inf)).max(axis=1) < rtol
            rep[rest[same]] = b
    return rep


class DuplicateStore:
    """The orbits found so far (hash + action), the role of the store folder in ``Check_Duplicates`` (find.py:865-935): ``add``
    returns, per candidate, the index of the stored orbit it duplicates (or -1 and stores it).  ``load_folder`` reads the
    ``Hash`` / ``Action`` fields of the reference's solution descriptors (``*.json`` written by ``Write_Descriptor``)."""

    def __init__(self, nbs, rtol: float = 1e-5, detect_multiples: bool = True):
        self.nbs, self.rtol, self.detect_multiples = nbs, float(rtol), bool(detect_multiples)
        self.hashes: List[np.ndarray] = []
        self.actions: List[float] = []

    def load_folder(self, folder: str) -> int:
        import json
        import os
        n = 0
        for fn in sorted(os.listdir(folder)):
            if fn.endswith(".json"):
                with open(os.path.join(folder, fn)) as f:
                    info = json.load(f)
                if info.get("Hash") is not None:
                    self.hashes.append(np.asarray(info["Hash"], dtype=np.float64))
                    self.actions.append(float(info.get("Action", np.nan)))
                    n += 1
        return n

    def add(self, hashes, actions=None) -> np.ndarray:
        H = hashes.detach().cpu().numpy() if type(hashes).__module__.startswith("torch") else np.asarray(hashes, dtype=np.float64)
        H = np.atleast_2d(H)
        out = np.full(H.shape[0], -1, dtype=np.int64)
        for b in range(H.shape[0]):
            if self.hashes:
                t = hash_test_values(self.nbs, H[b], np.stack(self.hashes), self.detect_multiples)
                # TestHashSame(Hash_a = candidate, Hash_b = stored)
                same = np.nonzero(np.abs(np.nan_to_num(t, nan=np.inf)).max(axis=1) < self.rtol)[0]
                if same.size:
                    out[b] = int(same[0])
                    continue
            self.hashes.append(H[b].copy())
            self.actions.append(float(actions[b]) if actions is not None else float("nan"))
        return out
