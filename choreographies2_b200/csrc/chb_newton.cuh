// chb_newton.cuh -- per-state control of the batched Newton-Krylov driver, on the device.
//
// The reference solves one state at a time on the host: choreo/scipy_plus/nonlin.py:54-171 (nonlin_solve_pp: termination test,
// inexact-Newton forcing terms), :194-234 (_nonlin_line_search_pp: phi(s) = |F(x + s dx)|^2, derphi by a forward difference of
// step (|s| + |x|/|dx| + 1) rdiff) on top of scipy's scalar_search_armijo and scalar_search_wolfe1 (MINPACK-2 dcsrch / dcstep,
// scipy/optimize/_linesearch.py and _dcsrch.py; scipy is a dependency of the reference, pinned < 1.14 in requirements.txt, the
// algorithm restated here is the published one).  Here every state of a batch carries its own copy of those scalars in device
// memory (NK_STRIDE doubles per state) and small kernels -- one thread per state -- advance all the state machines at once:
// the host only sequences launches and reads a few counters.
//
// All functions below are host+device so that tests/hostsim runs them on the CPU.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__) && !defined(CHB_HOSTSIM)
#define NK_HD __host__ __device__ __forceinline__
#else
#define NK_HD inline
#endif

#define NK_STRIDE 48
enum {
    NK_STATUS = 0,   // 0 running, 1 converged, 2 iteration limit, 3 non-finite residual, 4 zero Newton step
    NK_NIT, NK_ETA, NK_FNORM, NK_FMAX, NK_BEST, NK_TOL, NK_PHI0,
    NK_LS_TASK,      // 0 idle, 1 wants phi(s_eval) (stored), 2 wants phi(s_eval) (probe of derphi), 3 finished
    NK_S_EVAL, NK_S_STORED, NK_PHI_STORED, NK_FMAX_STORED, NK_S_FINAL, NK_NEED_FINAL,
    NK_A0, NK_A1, NK_PA0, NK_PA1, NK_TRIAL,                       // Armijo backtracking
    NK_STAGE, NK_BRACKT, NK_GINIT, NK_GTEST, NK_GX, NK_GY, NK_FINIT, NK_FX, NK_FY, NK_STX, NK_STY, NK_STMIN, NK_STMAX,
    NK_WIDTH, NK_WIDTH1, NK_STP, NK_F, NK_ITER, NK_SNORM, NK_DS, NK_STARTED,  // dcsrch
    NK_XNORM2, NK_DXNORM2, NK_LIVE,
    NK_NFIELDS
};
static_assert(NK_NFIELDS <= NK_STRIDE, "NK_STRIDE too small");

struct NkOpts {
    double f_tol, smin;
    int maxiter, restart;
    int line_search;  // 0 none (s = smin), 1 armijo, 2 wolfe
    int tol_norm;     // 0 max norm, 1 2-norm
    int track_best;
};

// ---- dcstep / dcsrch (scipy/optimize/_dcsrch.py, MINPACK-2) -------------------------------------------------------------------
NK_HD double nk_sign(double v) { return v > 0.0 ? 1.0 : (v < 0.0 ? -1.0 : 0.0); }
NK_HD double nk_max3(double a, double b, double c) { return fmax(a, fmax(b, c)); }
NK_HD double nk_clip(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

NK_HD void nk_dcstep(double &stx, double &fx, double &dx, double &sty, double &fy, double &dy, double &stp, double fp, double dp,
                     int &brackt, double stpmin, double stpmax) {
    const double sgnd = nk_sign(dp) * nk_sign(dx);
    double stpf;
    if (fp > fx) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = nk_max3(fabs(theta), fabs(dx), fabs(dp));
        double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
        if (stp < stx) gamma = -gamma;
        const double p = (gamma - dx) + theta, q = ((gamma - dx) + gamma) + dp, r = p / q;
        const double stpc = stx + r * (stp - stx);
        const double stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx);
        stpf = fabs(stpc - stx) <= fabs(stpq - stx) ? stpc : stpc + (stpq - stpc) / 2.0;
        brackt = 1;
    } else if (sgnd < 0.0) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = nk_max3(fabs(theta), fabs(dx), fabs(dp));
        double gamma = s * sqrt((theta / s) * (theta / s) - (dx / s) * (dp / s));
        if (stp > stx) gamma = -gamma;
        const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dx, r = p / q;
        const double stpc = stp + r * (stx - stp);
        const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
        stpf = fabs(stpc - stp) > fabs(stpq - stp) ? stpc : stpq;
        brackt = 1;
    } else if (fabs(dp) < fabs(dx)) {
        const double theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp;
        const double s = nk_max3(fabs(theta), fabs(dx), fabs(dp));
        double gamma = s * sqrt(fmax(0.0, (theta / s) * (theta / s) - (dx / s) * (dp / s)));
        if (stp > stx) gamma = -gamma;
        const double p = (gamma - dp) + theta, q = (gamma + (dx - dp)) + gamma, r = p / q;
        double stpc;
        if (r < 0.0 && gamma != 0.0) stpc = stp + r * (stx - stp);
        else if (stp > stx) stpc = stpmax;
        else stpc = stpmin;
        const double stpq = stp + (dp / (dp - dx)) * (stx - stp);
        if (brackt) {
            stpf = fabs(stpc - stp) < fabs(stpq - stp) ? stpc : stpq;
            if (stp > stx) stpf = fmin(stp + 0.66 * (sty - stp), stpf);
            else stpf = fmax(stp + 0.66 * (sty - stp), stpf);
        } else {
            stpf = fabs(stpc - stp) > fabs(stpq - stp) ? stpc : stpq;
            stpf = nk_clip(stpf, stpmin, stpmax);
        }
    } else {
        if (brackt) {
            const double theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp;
            const double s = nk_max3(fabs(theta), fabs(dy), fabs(dp));
            double gamma = s * sqrt((theta / s) * (theta / s) - (dy / s) * (dp / s));
            if (stp > sty) gamma = -gamma;
            const double p = (gamma - dp) + theta, q = ((gamma - dp) + gamma) + dy, r = p / q;
            stpf = stp + r * (sty - stp);
        } else if (stp > stx) {
            stpf = stpmax;
        } else {
            stpf = stpmin;
        }
    }
    if (fp > fx) {
        sty = stp; fy = fp; dy = dp;
    } else {
        if (sgnd < 0.0) { sty = stx; fy = fx; dy = dx; }
        stx = stp; fx = fp; dx = dp;
    }
    stp = stpf;
}

// one call of DCSRCH._iterate; returns the task: 0 "FG" (evaluate at the returned stp), 1 convergence, 2 warning, 3 error
NK_HD int nk_dcsrch_iterate(double *st, double &stp, double f, double g, double ftol, double gtol, double xtol, double stpmin,
                            double stpmax) {
    const double p5 = 0.5, p66 = 0.66, xtrapl = 1.1, xtrapu = 4.0;
    if (st[NK_STARTED] == 0.0) {
        if (stp < stpmin || stp > stpmax || g >= 0.0 || stpmax < stpmin) return 3;
        st[NK_STARTED] = 1.0;
        st[NK_BRACKT] = 0.0;
        st[NK_STAGE] = 1.0;
        st[NK_FINIT] = f;
        st[NK_GINIT] = g;
        st[NK_GTEST] = ftol * g;
        st[NK_WIDTH] = stpmax - stpmin;
        st[NK_WIDTH1] = st[NK_WIDTH] / p5;
        st[NK_STX] = 0.0; st[NK_FX] = f; st[NK_GX] = g;
        st[NK_STY] = 0.0; st[NK_FY] = f; st[NK_GY] = g;
        st[NK_STMIN] = 0.0;
        st[NK_STMAX] = stp + xtrapu * stp;
        return 0;
    }
    int brackt = st[NK_BRACKT] != 0.0;
    const double gtest = st[NK_GTEST], ginit = st[NK_GINIT], finit = st[NK_FINIT];
    const double ftest = finit + stp * gtest;
    if (st[NK_STAGE] == 1.0 && f <= ftest && g >= 0.0) st[NK_STAGE] = 2.0;
    int task = 0;
    if (brackt && (stp <= st[NK_STMIN] || stp >= st[NK_STMAX])) task = 2;
    if (brackt && st[NK_STMAX] - st[NK_STMIN] <= xtol * st[NK_STMAX]) task = 2;
    if (stp == stpmax && f <= ftest && g <= gtest) task = 2;
    if (stp == stpmin && (f > ftest || g >= gtest)) task = 2;
    if (f <= ftest && fabs(g) <= gtol * -ginit) task = 1;
    if (task) return task;
    double stx = st[NK_STX], fx = st[NK_FX], gx = st[NK_GX], sty = st[NK_STY], fy = st[NK_FY], gy = st[NK_GY];
    if (st[NK_STAGE] == 1.0 && f <= fx && f > ftest) {
        const double fm = f - stp * gtest;
        double fxm = fx - stx * gtest, fym = fy - sty * gtest;
        const double gm = g - gtest;
        double gxm = gx - gtest, gym = gy - gtest;
        nk_dcstep(stx, fxm, gxm, sty, fym, gym, stp, fm, gm, brackt, st[NK_STMIN], st[NK_STMAX]);
        fx = fxm + stx * gtest;
        fy = fym + sty * gtest;
        gx = gxm + gtest;
        gy = gym + gtest;
    } else {
        nk_dcstep(stx, fx, gx, sty, fy, gy, stp, f, g, brackt, st[NK_STMIN], st[NK_STMAX]);
    }
    if (brackt) {
        if (fabs(sty - stx) >= p66 * st[NK_WIDTH1]) stp = stx + p5 * (sty - stx);
        st[NK_WIDTH1] = st[NK_WIDTH];
        st[NK_WIDTH] = fabs(sty - stx);
    }
    if (brackt) {
        st[NK_STMIN] = fmin(stx, sty);
        st[NK_STMAX] = fmax(stx, sty);
    } else {
        st[NK_STMIN] = stp + xtrapl * (stp - stx);
        st[NK_STMAX] = stp + xtrapu * (stp - stx);
    }
    stp = nk_clip(stp, stpmin, stpmax);
    if ((brackt && (stp <= st[NK_STMIN] || stp >= st[NK_STMAX])) || (brackt && st[NK_STMAX] - st[NK_STMIN] <= xtol * st[NK_STMAX]))
        stp = stx;
    st[NK_STX] = stx; st[NK_FX] = fx; st[NK_GX] = gx;
    st[NK_STY] = sty; st[NK_FY] = fy; st[NK_GY] = gy;
    st[NK_BRACKT] = brackt ? 1.0 : 0.0;
    return 0;
}

// ---- Newton iteration: begin (termination test, forcing term, GMRES right-hand side) --------------------------------------------
// returns 1 when the state takes part in this iteration
NK_HD int nk_begin_state(double *st, const NkOpts &O) {
    st[NK_LIVE] = 0.0;
    if (st[NK_STATUS] != 0.0) return 0;
    const double fn = st[NK_FNORM];
    if (!(fn == fn) || fn > 1.7e308) { st[NK_STATUS] = 3.0; return 0; }
    const double tn = O.tol_norm == 1 ? fn : st[NK_FMAX];
    if (tn <= O.f_tol) { st[NK_STATUS] = 1.0; return 0; }
    const double eta = st[NK_ETA];
    st[NK_TOL] = fmin(eta, eta * fn);  // nonlin.py:113
    st[NK_LIVE] = 1.0;
    return 1;
}

// ---- line search: start (after dx is known; NK_XNORM2 / NK_DXNORM2 hold |x|^2 and |dx|^2) ---------------------------------------
NK_HD void nk_ls_start(double *st, const NkOpts &O) {
    if (st[NK_LIVE] == 0.0) { st[NK_LS_TASK] = 0.0; return; }
    if (!(st[NK_DXNORM2] > 0.0)) {  // nonlin.py:118-121: "Jacobian inversion yielded zero vector" (a ValueError there)
        if (st[NK_DXNORM2] == 0.0) { st[NK_STATUS] = 4.0; st[NK_LIVE] = 0.0; st[NK_LS_TASK] = 0.0; return; }
    }
    const double phi0 = st[NK_FNORM] * st[NK_FNORM];
    st[NK_PHI0] = phi0;
    st[NK_S_STORED] = 0.0;       // tmp_s = [0]: F(x) itself is the stored evaluation
    st[NK_PHI_STORED] = phi0;
    st[NK_NEED_FINAL] = 0.0;
    st[NK_S_FINAL] = O.smin;
    if (O.line_search == 0) {    // no search: s = smin
        st[NK_S_EVAL] = O.smin;
        st[NK_LS_TASK] = 1.0;
        st[NK_TRIAL] = -1.0;
    } else if (O.line_search == 1) {
        st[NK_A0] = 1.0; st[NK_A1] = 1.0; st[NK_PA0] = 0.0; st[NK_PA1] = 0.0;
        st[NK_TRIAL] = 0.0;
        st[NK_S_EVAL] = 1.0;
        st[NK_LS_TASK] = 1.0;
    } else {
        st[NK_SNORM] = sqrt(st[NK_XNORM2]) / sqrt(st[NK_DXNORM2]);
        st[NK_STARTED] = 0.0;
        st[NK_ITER] = 0.0;
        st[NK_STP] = 0.0;        // derphi(0) first
        st[NK_F] = phi0;
        const double ds = (0.0 + st[NK_SNORM] + 1.0) * 1e-8;
        st[NK_DS] = ds;
        st[NK_S_EVAL] = ds;
        st[NK_LS_TASK] = 2.0;
    }
}

// ---- line search: one evaluation phi_t = |F(x + s_eval dx)|^2 (inf when not finite) has arrived ---------------------------------
// kind: 1 the stored evaluations were run, 2 the probes were run; states whose task is the other kind are left alone
NK_HD void nk_ls_update(double *st, const NkOpts &O, int kind, double phi_t, double fmax_t) {
    if ((int)st[NK_LS_TASK] != kind) return;
    if (!(phi_t == phi_t) || phi_t > 1.7e308) phi_t = INFINITY;  // _safe_norm_pp
    const double phi0 = st[NK_PHI0];
    if (kind == 1) {  // phi(s, store=True)
        st[NK_S_STORED] = st[NK_S_EVAL];
        st[NK_PHI_STORED] = phi_t;
        st[NK_FMAX_STORED] = fmax_t;
    }
    if (st[NK_NEED_FINAL] != 0.0) {  // the closing evaluation F(x + s dx) of _nonlin_line_search_pp
        st[NK_NEED_FINAL] = 0.0;
        st[NK_LS_TASK] = 3.0;
        return;
    }
    if (O.line_search == 0) {
        st[NK_S_FINAL] = O.smin;
        st[NK_LS_TASK] = 3.0;
        return;
    }
    if (O.line_search == 1) {  // scalar_search_armijo(phi, phi0, -phi0, amin = smin), c1 = 1e-4
        const double c1 = 1e-4, derphi0 = -phi0;
        const int trial = (int)st[NK_TRIAL];
        const double alpha = st[NK_S_EVAL];
        if (phi_t <= phi0 + c1 * alpha * derphi0) {
            st[NK_S_FINAL] = alpha;
            st[NK_LS_TASK] = 3.0;
            return;
        }
        double next;
        if (trial == 0) {
            st[NK_PA0] = phi_t;
            next = -(derphi0) * st[NK_A0] * st[NK_A0] / 2.0 / (phi_t - phi0 - derphi0 * st[NK_A0]);
            st[NK_A1] = next;
        } else {
            double a0 = st[NK_A0], a1 = st[NK_A1], pa0 = st[NK_PA0], pa1 = st[NK_PA1];
            if (trial == 1) {
                pa1 = phi_t;
            } else {  // alpha = alpha2 failed: shift the interpolation points
                double a2 = alpha;
                if ((a1 - a2) > a1 / 2.0 || (1.0 - a2 / a1) < 0.96) a2 = a1 / 2.0;
                a0 = a1; a1 = a2; pa0 = pa1; pa1 = phi_t;
            }
            st[NK_A0] = a0; st[NK_A1] = a1; st[NK_PA0] = pa0; st[NK_PA1] = pa1;
            if (!(a1 > O.smin)) {  // while alpha1 > amin: failed, s = smin (needs its own evaluation)
                st[NK_S_FINAL] = O.smin;
                st[NK_S_EVAL] = O.smin;
                st[NK_NEED_FINAL] = 1.0;
                st[NK_LS_TASK] = 1.0;
                return;
            }
            const double factor = a0 * a0 * a1 * a1 * (a1 - a0);
            const double a = (a0 * a0 * (pa1 - phi0 - derphi0 * a1) - a1 * a1 * (pa0 - phi0 - derphi0 * a0)) / factor;
            const double b = (-a0 * a0 * a0 * (pa1 - phi0 - derphi0 * a1) + a1 * a1 * a1 * (pa0 - phi0 - derphi0 * a0)) / factor;
            next = (-b + sqrt(fabs(b * b - 3.0 * a * derphi0))) / (3.0 * a);
        }
        st[NK_TRIAL] = (double)(trial + 1);
        st[NK_S_EVAL] = next;
        st[NK_LS_TASK] = 1.0;
        return;
    }
    // ---- wolfe: scalar_search_wolfe1(phi, derphi, phi0, xtol = 1e-2, amin = smin), c1 = 1e-4, c2 = 0.9, amax = 50 ----
    const double ftol = 1e-4, gtol = 0.9, xtol = 1e-2, stpmin = O.smin, stpmax = 50.0;
    if (kind == 1) {  // phi(stp) arrived: now its derivative probe phi(stp + ds)
        st[NK_F] = phi_t;
        const double ds = (fabs(st[NK_STP]) + st[NK_SNORM] + 1.0) * 1e-8;
        st[NK_DS] = ds;
        st[NK_S_EVAL] = st[NK_STP] + ds;
        st[NK_LS_TASK] = 2.0;
        return;
    }
    // probe arrived: derphi(stp) = (phi(stp + ds) - phi(stp)) / ds, then one dcsrch iteration
    const double g = (phi_t - st[NK_F]) / st[NK_DS];
    double stp = st[NK_STARTED] == 0.0 ? 1.0 : st[NK_STP];  // alpha1 = 1 (old_phi0 is None)
    int task = nk_dcsrch_iterate(st, stp, st[NK_F], g, ftol, gtol, xtol, stpmin, stpmax);
    const int iter = (int)st[NK_ITER] + 1;
    st[NK_ITER] = (double)iter;
    double s = 0.0;
    bool found = false;
    if (!(stp == stp) || fabs(stp) > 1.7e308) {
        task = 2;
    } else if (task == 0) {
        if (iter >= 100) {
            task = 2;  // dcsrch did not converge within max iterations
        } else {
            st[NK_STP] = stp;
            st[NK_S_EVAL] = stp;
            st[NK_LS_TASK] = 1.0;
            return;
        }
    } else if (task == 1) {
        s = stp;
        found = true;
    }
    if (!found) s = O.smin;  // warnings / errors: stp = None -> s = smin
    st[NK_S_FINAL] = s;
    if (s == st[NK_S_STORED]) {
        st[NK_LS_TASK] = 3.0;
    } else {
        st[NK_S_EVAL] = s;
        st[NK_NEED_FINAL] = 1.0;
        st[NK_LS_TASK] = 1.0;
    }
}

// ---- Newton iteration: end (forcing terms nonlin.py:146-151, iteration count); NK_PHI_STORED / NK_FMAX_STORED are those of the accepted step
NK_HD void nk_end_state(double *st, const NkOpts &O) {
    (void)O;
    if (st[NK_LIVE] == 0.0) return;
    const double gamma = 0.9, eta_max = 0.9999, eta_threshold = 0.1;
    const double fn_old = st[NK_FNORM], fn_new = sqrt(st[NK_PHI_STORED]);
    const double eta = st[NK_ETA];
    const double eta_A = gamma * (fn_new / fn_old) * (fn_new / fn_old);
    double eta_new;
    if (gamma * eta * eta < eta_threshold) eta_new = fmin(eta_max, eta_A);
    else eta_new = fmin(eta_max, fmax(eta_A, gamma * eta * eta));
    if (!(eta_new == eta_new)) eta_new = eta_max;
    st[NK_ETA] = eta_new;
    st[NK_FNORM] = fn_new;
    st[NK_FMAX] = st[NK_FMAX_STORED];
    st[NK_NIT] += 1.0;
}
