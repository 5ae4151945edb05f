"""Batched lock-step Newton-Krylov root solve of ``grad Action(x) = 0`` over many independent states.

Restates, for a batch, the solver the reference drives one state at a time:
  * outer iteration, inexact-Newton forcing terms, termination: ``choreo/scipy_plus/nonlin.py:54-171`` (``nonlin_solve_pp``);
  * line search: ``nonlin.py:190-234`` (``_nonlin_line_search_pp``) with scipy's ``scalar_search_armijo`` (c1 = 1e-4,
    quadratic then cubic interpolation, ``amin = smin``; fall back to ``s = smin``);
  * linear solves: scipy ``KrylovJacobian(method='gmres')`` as wired by ``NBodySyst.GetKrylovJacobian``
    (``_NBodySyst.pyx:2004-2032``): GMRES(restart = inner_maxiter = 20, maxiter = 1, atol = 0, rtol = tol), zero initial guess,
    matvec = ``segmpos_dparams_to_action_hess(segmpos, v)`` with ``segmpos = params_to_segmpos(x)`` cached per Newton step.

All states advance in lock step; a state that has converged (or failed) is masked out, and every evaluation runs only on the
compacted set of states that still need it.  The heavy operations are the library's batched operators (U1 / U2 of SURVEY
8(d)); the vector algebra between them (dot products, axpys, Givens rotations on [B, 21] arrays) is array-carrier work:
NumPy for host arrays, torch for CUDA tensors.  Summation order differs from scipy's one-state GMRES, so iterates agree to
rounding and converged orbits to the solver tolerance, not bit for bit.
"""
from __future__ import annotations

from dataclasses import dataclass
import os
from typing import Any, Optional

import numpy as np


class _XP:
    """The few array operations the driver needs, for NumPy arrays or torch tensors."""

    def __init__(self, like: Any):
        self.torch = type(like).__module__.startswith("torch")
        if self.torch:
            import torch
            self.t = torch
            self.dev = like.device

    def zeros(self, shape, dtype="f8"):
        if self.torch:
            dt = {"f8": self.t.float64, "i8": self.t.int64, "i4": self.t.int32, "b": self.t.bool}[dtype]
            return self.t.zeros(shape, dtype=dt, device=self.dev)
        return np.zeros(shape, dtype={"f8": np.float64, "i8": np.int64, "i4": np.int32, "b": bool}[dtype])

    def empty(self, shape):
        return self.t.empty(shape, dtype=self.t.float64, device=self.dev) if self.torch else np.empty(shape)

    def full(self, shape, v):
        return self.zeros(shape) + v

    def copy(self, a):
        return a.clone() if self.torch else a.copy()

    def dot(self, a, b):  # row-wise
        return (a * b).sum(-1)

    def norm(self, a):
        return self.sqrt(self.dot(a, a))

    def maxabs(self, a):
        return a.abs().amax(-1) if self.torch else np.abs(a).max(-1)

    def sqrt(self, a):
        return self.t.sqrt(a) if self.torch else np.sqrt(a)

    def abs(self, a):
        return a.abs() if self.torch else np.abs(a)

    def where(self, c, a, b):
        return self.t.where(c, a, b) if self.torch else np.where(c, a, b)

    def minimum(self, a, b):
        return self.t.minimum(a, b) if self.torch else np.minimum(a, b)

    def maximum(self, a, b):
        return self.t.maximum(a, b) if self.torch else np.maximum(a, b)

    def einsum(self, spec, *ops):
        return self.t.einsum(spec, *ops) if self.torch else np.einsum(spec, *ops)

    def nonzero(self, m):
        return self.t.nonzero(m).reshape(-1) if self.torch else np.nonzero(m)[0]

    def any(self, m) -> bool:
        return bool(m.any())

    def isfinite(self, a):
        return self.t.isfinite(a) if self.torch else np.isfinite(a)

    def axpy(self, x, alpha, d):  # x + alpha[:, None] * d in one pass
        return self.t.addcmul(x, alpha[:, None], d) if self.torch else x + alpha[:, None] * d

    def isfinite_rows(self, a):
        return self.t.isfinite(a).all(-1) if self.torch else np.isfinite(a).all(-1)

    def from_numpy_index(self, idx):
        return self.t.as_tensor(idx, device=self.dev) if self.torch else idx

    def to_numpy(self, a):
        return a.detach().cpu().numpy() if self.torch else np.asarray(a)

    def sync(self):
        if self.torch and self.dev.type == "cuda":
            self.t.cuda.synchronize(self.dev)


@dataclass
class NewtonResult:
    x: Any                 # [B, nparams] final iterates
    status: np.ndarray     # [B] 1 converged, 2 iteration limit, 3 non-finite residual
    nit: np.ndarray        # [B] Newton iterations used
    fnorm: np.ndarray      # [B] max-norm of the final gradient
    n_grad_evals: int      # U1 evaluations (states x calls)
    n_hess_evals: int      # U2 evaluations
    best_x: Any = None     # [B, nparams] iterate with the smallest 2-norm of the gradient seen (track_best; nonlin.py:13-39)
    best_fnorm: Any = None  # [B] that 2-norm (array of the carrier)
    launches: int = 0      # kernel launches of the solve (library driver)


def gmres_batched(xp: _XP, matvec, rhs, rtol, restart: int = 20, arnoldi=None, step=None):
    """One GMRES(restart) cycle per row of ``rhs`` ([A, n]) from a zero initial guess; row b stops at the first inner
    iteration whose residual estimate is <= rtol[b] * |rhs[b]| (scipy.sparse.linalg.gmres with atol = 0, maxiter = 1).
    ``matvec(idx, v)`` applies the operator of rows ``idx`` (None = all rows) to ``v``.  Arnoldi on the shifted operator with
    batched classical Gram-Schmidt (see the comment in the loop) + Givens rotations.  ``arnoldi(V, j, w, skip, h) -> bool``, when
    given, is the library's one-launch Arnoldi step (``NBodySystB200.arnoldi_step``); it returning False selects the array-level
    Gram-Schmidt below; ``step(V, j, w, done, R, g, cs, target, kdone) -> bool`` is the library's whole inner iteration
    (``NBodySystB200.gmres_step``: Arnoldi step + Givens rotations + stopping test in one launch, everything in place).  Returns (solution [A, n], inner iterations per row, number of row-matvecs)."""
    A = rhs.shape[0]
    m = int(restart)
    beta = xp.norm(rhs)
    safe_beta = xp.where(beta > 0, beta, beta + 1.0)
    V = xp.zeros((m + 1,) + tuple(rhs.shape))  # rows that stop early multiply the unused vectors by 0: they must be finite
    V[0] = rhs / safe_beta[:, None]
    R = xp.zeros((A, m + 1, m))
    Q = xp.zeros((A, m + 1, m + 1))
    for i in range(m + 1):
        Q[:, i, i] = 1.0
    g = xp.zeros((A, m + 1))
    g[:, 0] = beta
    target = rtol * beta
    done = ~(beta > 0)
    kdone = xp.zeros((A,), "i8")
    nmv = 0
    cs = xp.zeros((A, m, 2)) if step is not None else None
    for j in range(m):
        idx = xp.nonzero(~done)
        na = int(idx.shape[0])
        if na == 0:
            break
        if na == A:
            w = matvec(None, V[j])
        else:
            w = xp.zeros(rhs.shape)
            w[idx] = matvec(idx, V[j][idx])
        nmv += na
        if step is not None:
            if step(V, j, w, done, R, g, cs, target, kdone):
                continue
            step = None  # not addressable by the library (decided at j = 0): array-level path from here on
        h = xp.zeros((A, m + 1))
        if arnoldi is not None and arnoldi(V, j, w, done if na < A else None, h):
            Vj = None
        else:
            Vj = V[:j + 1]
        # Arnoldi on A - I (same Krylov space, H_A = H_(A-I) + I): the action Hessian is the identity (kinetic term) plus the
        # potential part, so A v_j is dominated by v_j itself and orthogonalising it would cancel most of its digits; without that
        # component one classical Gram-Schmidt pass (two batched contractions) is enough, a second one runs only if some row
        # still lost more than a factor 10
        if Vj is not None:
            w = w - V[j]
            n0 = xp.norm(w)
            h1 = xp.einsum("jan,an->ja", Vj, w)
            w = w - xp.einsum("ja,jan->an", h1, Vj)
            weak = xp.norm(w) < 0.1 * n0
            if xp.any(weak):  # second pass, applied only to the rows that need it (rows never influence each other)
                h2 = xp.einsum("jan,an->ja", Vj, w) * weak[None, :]
                w = w - xp.einsum("ja,jan->an", h2, Vj)
                h1 = h1 + h2
            if xp.torch:
                h[:, :j + 1] = h1.transpose(0, 1)
            else:
                h[:, :j + 1] = h1.T
            h[:, j] = h[:, j] + 1.0
            hn = xp.norm(w)
            h[:, j + 1] = hn
            V[j + 1] = w / xp.where(hn > 0, hn, hn + 1.0)[:, None]
        # previous rotations on the new column in one small batched product: Q accumulates them ([A, m+1, m+1])
        h = xp.einsum("aik,ak->ai", Q, h)
        den = xp.sqrt(h[:, j] * h[:, j] + h[:, j + 1] * h[:, j + 1])
        den = xp.where(den > 0, den, den + 1.0)
        c, s = h[:, j] / den, h[:, j + 1] / den
        qj, qj1 = xp.copy(Q[:, j, :]), xp.copy(Q[:, j + 1, :])
        Q[:, j, :] = c[:, None] * qj + s[:, None] * qj1
        Q[:, j + 1, :] = -s[:, None] * qj + c[:, None] * qj1
        h[:, j] = c * h[:, j] + s * h[:, j + 1]
        h[:, j + 1] = 0.0
        gj = g[:, j]
        gnew = -s * gj
        keep = done  # rows that stopped earlier keep their factorisation
        R[:, :, j] = xp.where(keep[:, None], R[:, :, j], h)
        g[:, j + 1] = xp.where(keep, g[:, j + 1], gnew)
        g[:, j] = xp.where(keep, gj, c * gj)
        kdone = xp.where(keep, kdone, kdone * 0 + (j + 1))
        done = done | (xp.abs(g[:, j + 1]) <= target)
    # back substitution on each row's own k x k triangle (columns >= k are the identity with a zero right-hand side)
    y = xp.zeros((A, m))
    for j in range(m - 1, -1, -1):
        inside = kdone > j
        acc = g[:, j] - (R[:, j, j + 1:] * y[:, j + 1:]).sum(-1) if j + 1 < m else g[:, j]
        djj = xp.where(inside & (R[:, j, j] != 0), R[:, j, j], R[:, j, j] * 0 + 1.0)
        y[:, j] = xp.where(inside, acc / djj, acc * 0)
    sol = xp.einsum("aj,jan->an", y, V[:m])
    return sol, kdone, nmv


class _Buffers:
    """Flat scratch arrays of the library path, reused across the Newton iterations (the active set only shrinks)."""

    def __init__(self, xp: _XP):
        self.xp, self.flat = xp, {}

    def arange(self, n):
        a = self.flat.get("arange")
        if a is None or a.shape[0] < n:
            a = self.flat["arange"] = self.xp.t.arange(n, dtype=self.xp.t.int32, device=self.xp.dev) if self.xp.torch \
                else np.arange(n, dtype=np.int32)
        return a[:n]

    def get(self, name, shape):
        size = 1
        for d in shape:
            size *= int(d)
        buf = self.flat.get(name)
        if buf is None or buf.shape[0] < size:
            buf = self.flat[name] = self.xp.empty((size,))
        return buf[:size].reshape(shape)


def gmres_batched_lib(xp: _XP, nbs, seg, rhs, rtol, restart: int, bufs: _Buffers, check_every: int = 4):
    """``gmres_batched`` with every array-sized step inside the library: per inner iteration one Hessian-vector launch over the
    list of running states (``hess_listed``: no gather / scatter of the active rows, the persistent CTAs share the listed states
    evenly) and one ``gmres_step`` launch (Arnoldi + Givens + stopping test + the list for the next iteration); one
    ``gmres_solution`` launch at the end.  The host looks at
    the stop flags only every ``check_every`` iterations (a device synchronisation).  Returns (solution, inner iterations)."""
    A, n = int(rhs.shape[0]), int(rhs.shape[1])
    m = int(restart)
    V = bufs.get("V", (m + 1, A, n))
    w = bufs.get("w", (A, n))
    beta = xp.norm(rhs)
    V[0] = rhs / xp.where(beta > 0, beta, beta + 1.0)[:, None]
    R = xp.zeros((A, m + 1, m))
    g = xp.zeros((A, m + 1))
    g[:, 0] = beta
    cs = xp.zeros((A, m, 2))
    target = xp.zeros((A,)) + rtol * beta
    done = ~(beta > 0)
    kdone = xp.zeros((A,), "i8")
    lists = xp.zeros((2, A), "i4")
    counts = xp.zeros((m + 1,), "i4")
    lists[0] += bufs.arange(A)  # states with a zero right-hand side included: their matvec is 0 and the step skips them
    counts[0] += A
    for j in range(m):
        if j > 0 and j % check_every == 0 and bool(done.all()):
            break
        nbs.hess_listed(seg, V[j], lists[j & 1], counts[j:j + 1], w)
        nbs.gmres_step(V, j, w, done, R, g, cs, target, kdone, lists[(j + 1) & 1], counts[j + 1:j + 2])
    sol = xp.empty((A, n))
    nbs.gmres_solution(V, R, g, kdone, 1.0, sol)
    return sol, kdone


def _take(xp: _XP, a, idx):
    return a if idx is None else a[idx]


def newton_krylov_batched(nbs, x0, f_tol: float = 1e-10, maxiter: int = 50, restart: int = 20, smin: float = 1e-2,
                          line_search: Optional[str] = "armijo", tol_norm: str = "max", track_best: bool = False) -> NewtonResult:
    """See ``_newton_krylov_batched``; masked-out rows may divide by zero in the interpolation formulas: silenced.
    ``tol_norm``: "max" (nonlin_solve_pp's default ``tol_norm``) or "l2" (``np.linalg.norm``, what find.py:264 passes);
    ``track_best``: also return, per state, the iterate with the smallest gradient 2-norm (``current_best``, find.py:204,252)."""
    if tol_norm not in ("max", "l2"):
        raise ValueError("tol_norm must be 'max' or 'l2'")
    if line_search not in (None, "armijo", "wolfe"):
        raise ValueError(f"line_search must be None, 'armijo' or 'wolfe' (nonlin.py:129), got {line_search!r}")
    # the whole solve inside the library (per-state control on the device, no array-level glue) when it can address the arrays
    if hasattr(nbs, "newton_krylov") and nbs._direct_ok(x0) and os.environ.get("CHB_NEWTON_LIB", "1") not in ("", "0"):
        xp = _XP(x0)
        x = xp.copy(x0)
        if xp.torch and not x.is_contiguous():
            x = x.contiguous()
        r = nbs.newton_krylov(x, f_tol, maxiter, restart, smin, line_search, tol_norm, track_best)
        if r is not None:
            return NewtonResult(x=x, status=xp.to_numpy(r["status"]).astype(np.int64), nit=xp.to_numpy(r["nit"]).astype(np.int64),
                                fnorm=xp.to_numpy(r["fnorm_max"]), n_grad_evals=r["n_grad_evals"], n_hess_evals=r["n_hess_evals"],
                                best_x=r["best_x"], best_fnorm=r["best_fnorm"], launches=r["launches"])
    if line_search == "wolfe":
        raise ValueError("line_search='wolfe' needs the library driver (chb_newton_krylov_dev): device tensors, or NumPy arrays "
                         "with a build whose entry points take host pointers")
    with np.errstate(all="ignore"):
        return _newton_krylov_batched(nbs, x0, f_tol, maxiter, restart, smin, line_search, tol_norm, track_best)


def _newton_krylov_batched(nbs, x0, f_tol, maxiter, restart, smin, line_search, tol_norm="max", track_best=False) -> NewtonResult:
    """Solve ``nbs.params_to_action_grad(x) = 0`` for every row of ``x0`` ([B, nparams]; NumPy array or torch CUDA tensor).

    Termination like the reference's ``TerminationCondition(f_tol=...)`` with the max norm (``nonlin.py:77-78``): a state is
    converged when ``max |F| <= f_tol``; ``maxiter`` Newton iterations otherwise (status 2)."""
    xp = _XP(x0)
    x = xp.copy(x0)
    B = x.shape[0]
    # CHB_NO_ARNOLDI=1: array-level Gram-Schmidt instead of the library's Arnoldi kernel (A/B measurements)
    # CHB_NO_GMRES_STEP=1: the library's Arnoldi step but the Givens / stopping logic at array level
    arnoldi = None if os.environ.get("CHB_NO_ARNOLDI", "0") not in ("", "0") else getattr(nbs, "arnoldi_step", None)
    step = None if arnoldi is None or os.environ.get("CHB_NO_GMRES_STEP", "0") not in ("", "0") else getattr(nbs, "gmres_step", None)
    # everything in the library (masked matvec + whole GMRES iteration + solution kernels) when it can address the arrays
    lib_gmres = step is not None and hasattr(nbs, "hess_masked") and nbs._direct_ok(x0) \
        and os.environ.get("CHB_NO_GMRES_LIB", "0") in ("", "0")
    bufs = _Buffers(xp)
    n_grad = n_hess = 0
    n_hess_lazy = []  # per-cycle device sums, read once at the end

    def func(idx, xs):
        nonlocal n_grad
        n_grad += int(xs.shape[0])
        return nbs.params_to_action_grad(xs)

    row_stats = getattr(nbs, "row_stats", None) if lib_gmres else None

    def stats(F):
        """(sum of squares, max |.|) per row; a row with a non-finite entry has a non-finite sum (so does a row whose squares
        overflow: such a state is reported as non-finite, status 3)."""
        r = row_stats(F) if row_stats is not None else None
        if r is None:
            ss = xp.dot(F, F)
            return xp.where(xp.isfinite_rows(F), ss, ss * 0 + float("inf")), xp.maxabs(F)
        return r

    Fx = func(None, x)
    ss0, Fmax = stats(Fx)
    Fnorm = xp.sqrt(ss0)
    tnorm = (lambda: Fnorm) if tol_norm == "l2" else (lambda: Fmax)  # both norms of Fx are kept up to date
    if track_best:
        best_x, best_n = xp.copy(x), xp.where(Fnorm == Fnorm, Fnorm, Fnorm * 0 + 1e100)
    eta = xp.full((B,), 1e-3)
    gamma, eta_max, eta_threshold = 0.9, 0.9999, 0.1  # nonlin.py:101-105
    status = xp.zeros((B,), "i8")
    nit = xp.zeros((B,), "i8")
    for _ in range(int(maxiter)):
        bad = ~xp.isfinite(Fnorm)
        status = xp.where((status == 0) & bad, status * 0 + 3, status)
        conv = (status == 0) & (tnorm() <= f_tol)
        status = xp.where(conv, status * 0 + 1, status)
        act = xp.nonzero(status == 0)
        na = int(act.shape[0])
        if na == 0:
            break
        allact = na == B
        ia = None if allact else act
        xa, Fa, Fna, etaa = _take(xp, x, ia), _take(xp, Fx, ia), _take(xp, Fnorm, ia), _take(xp, eta, ia)
        seg = nbs.params_to_segmpos(xa)  # jacobian.setup / update (:2013-2019)

        def matvec(idx, v):
            nonlocal n_hess
            n_hess += int(v.shape[0])
            return nbs.segmpos_dparams_to_action_hess(seg if idx is None else seg[idx], v)

        tol = xp.minimum(etaa, etaa * Fna)  # nonlin.py:113
        if lib_gmres:
            sol, kd = gmres_batched_lib(xp, nbs, seg if seg.is_contiguous() else seg.contiguous(), Fa if allact else Fa.contiguous(),
                                        tol, restart, bufs) if xp.torch else \
                gmres_batched_lib(xp, nbs, np.ascontiguousarray(seg), np.ascontiguousarray(Fa), tol, restart, bufs)
            n_hess_lazy.append(kd.sum())
        else:
            sol, _, _ = gmres_batched(xp, matvec, Fa, tol, restart, arnoldi, step)
        dx = -sol
        # ---- line search on phi(s) = |F(x + s dx)|^2, scalar_search_armijo(phi, phi0, -phi0, amin=smin) per state ----
        phi0 = Fna * Fna
        if line_search is None:
            s_fin = xp.full((na,), smin)
            F_fin = func(None, xp.axpy(xa, s_fin, dx))
        else:
            c1 = 1e-4
            s_fin = xp.full((na,), smin)
            F_fin = xp.zeros(Fa.shape)
            have = xp.zeros((na,), "b")          # F_fin holds F(x + s_fin dx)
            search = xp.zeros((na,), "b") | True  # still searching
            a0, a1 = xp.full((na,), 1.0), xp.full((na,), 1.0)
            p0, p1 = xp.zeros((na,)), xp.zeros((na,))
            for trial in range(40):
                ids = xp.nonzero(search)
                if int(ids.shape[0]) == 0:
                    break
                if trial == 0:
                    alpha = a0
                elif trial == 1:
                    den = 2.0 * (p0 - phi0 + phi0 * a0)
                    alpha = phi0 * a0 * a0 / xp.where(den != 0, den, den + 1e-300)
                    a1 = alpha
                else:
                    factor = a0 * a0 * a1 * a1 * (a1 - a0)
                    factor = xp.where(factor != 0, factor, factor + 1e-300)
                    ca = (a0 * a0 * (p1 - phi0 + phi0 * a1) - a1 * a1 * (p0 - phi0 + phi0 * a0)) / factor
                    cb = (-a0 * a0 * a0 * (p1 - phi0 + phi0 * a1) + a1 * a1 * a1 * (p0 - phi0 + phi0 * a0)) / factor
                    ca = xp.where(ca != 0, ca, ca + 1e-300)
                    alpha = (-cb + xp.sqrt(xp.abs(cb * cb + 3.0 * ca * phi0))) / (3.0 * ca)
                sub_all = int(ids.shape[0]) == na
                al = alpha if sub_all else alpha[ids]
                Ft = func(None, xp.axpy(xa if sub_all else xa[ids], al, dx if sub_all else dx[ids]))
                sst, _ = stats(Ft)
                pt = xp.where(xp.isfinite(sst), sst, al * 0 + float("inf"))  # _safe_norm_pp
                ph0 = phi0 if sub_all else phi0[ids]
                ok = pt <= ph0 - c1 * al * ph0
                if sub_all:
                    ptf, okf = pt, ok
                else:
                    ptf = xp.full((na,), float("inf"))
                    ptf[ids] = pt
                    okf = xp.zeros((na,), "b")
                    okf[ids] = ok
                acc = search & okf
                s_fin = xp.where(acc, alpha, s_fin)
                if sub_all:
                    F_fin = xp.where(acc[:, None], Ft, F_fin)
                else:
                    F_fin[ids[ok]] = Ft[ok]
                have = have | acc
                search = search & ~okf
                if trial == 0:
                    p0 = ptf
                elif trial == 1:
                    p1 = ptf
                else:
                    a2 = alpha
                    halve = ((a1 - a2) > a1 / 2.0) | ((1.0 - a2 / a1) < 0.96)
                    a2 = xp.where(halve, a1 / 2.0, a2)
                    a0, a1, p0, p1 = a1, a2, p1, ptf
                if trial >= 1:
                    search = search & (a1 > smin)  # while alpha1 > amin (the failed ones fall back to s = smin)
            miss = xp.nonzero(~have)
            if int(miss.shape[0]) > 0:
                F_fin[miss] = func(None, xp.axpy(xa[miss], s_fin[miss], dx[miss]))
        xn = xp.axpy(xa, s_fin, dx)
        ssn, Fmax_new = stats(F_fin)
        Fn_new = xp.sqrt(ssn)
        # forcing terms (nonlin.py:146-151)
        eta_A = gamma * (Fn_new / Fna) ** 2
        low = gamma * etaa * etaa < eta_threshold
        eta_new = xp.where(low, xp.minimum(eta_A * 0 + eta_max, eta_A),
                           xp.minimum(eta_A * 0 + eta_max, xp.maximum(eta_A, gamma * etaa * etaa)))
        if allact:
            x, Fx, Fnorm, Fmax, eta = xn, F_fin, Fn_new, Fmax_new, eta_new
            nit = nit + 1
        else:
            x[act], Fx[act], Fnorm[act], Fmax[act], eta[act] = xn, F_fin, Fn_new, Fmax_new, eta_new
            nit[act] = nit[act] + 1
        if track_best:  # current_best.update after every iteration (the callback of find.py:249-257)
            better = Fnorm < best_n
            best_x = xp.where(better[:, None], x, best_x)
            best_n = xp.where(better, Fnorm, best_n)
    else:
        conv = (status == 0) & (tnorm() <= f_tol)
        status = xp.where(conv, status * 0 + 1, status)
    status = xp.where(status == 0, status * 0 + 2, status)
    n_hess += sum(int(v) for v in n_hess_lazy)
    return NewtonResult(x=x, status=xp.to_numpy(status), nit=xp.to_numpy(nit), fnorm=xp.to_numpy(Fmax),
                        n_grad_evals=n_grad, n_hess_evals=n_hess,
                        best_x=best_x if track_best else None, best_fnorm=best_n if track_best else None)
