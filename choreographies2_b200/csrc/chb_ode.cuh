// ODE right-hand sides of the reduced N-body system, batched over many phase-space points.
// Reference (choreo/cython/_NBodySyst.pyx): Compute_forces_vectorized 7531-7638, Compute_grad_forces_vectorized 7641-7790
// (Python entry points 4154-4268).  One time sample per vector: pos is [nvec][nsegm][G]; the tangent version carries
// dpos / grad_forces as [nvec][nsegm][G][grad_ndof].
//   for every binary interaction (src, tgt, R, c):   dx = R x_src - x_tgt,  (V, V', V'') = inter_law(|dx|^2)
//       forces:       df  = -4 c V' dx                               F_src += R^T df,   F_tgt -= df
//       grad forces:  ddx = R dx_src' - dx_tgt',  ddf = -8 c V'' (dx . ddx) dx - 4 c V' ddx   (same scatter)
// A "column" is one vector (forces) or one (vector, degree of freedom) pair (grad forces); columns are independent, so one
// thread owns one column and accumulates it in a shared-memory tile laid out [entry][column] (pitch T + 1: the transposing
// loads and the per-thread accesses are both bank-conflict free), read and written with coalesced global accesses.
// HBM-bound: 16 nsegm G bytes per column.
// Note: in the non-identity branch of the vectorised reference functions the source position is read from vector 0
// (`pos + isegm*geodim`, 7582 / 7705) instead of the current vector; this restates the per-vector definition
// (Compute_forces / Compute_grad_forces called once per vector), which is what the tests compare against.
#pragma once
#include "chb_core.cuh"

template <bool GRAD, class PosF, class DinF, class OutF>
CHB_HD void chb_ode_column(const ChbSys &S, int vs, PosF pos, DinF din, OutF out) {
    const int G = S.G;
    for (int e = 0; e < vs; ++e) out(e) = 0.0;
    for (int ib = 0; ib < S.nbin; ++ib) {
        const ChbBin &B = S.bins[ib];
        const int so = B.src * G, to = B.tgt * G;
        double dx[CHB_MAX_G], ddx[CHB_MAX_G], df[CHB_MAX_G];
        double dx2 = 0.0, dxtddx = 0.0;
        for (int i = 0; i < G; ++i) {
            double a, da = 0.0;
            if (B.isid) {
                a = pos(so + i) - pos(to + i);
                if (GRAD) da = din(so + i) - din(to + i);
            } else {
                a = -pos(to + i);
                if (GRAD) da = -din(to + i);
                for (int j = 0; j < G; ++j) {
                    a += B.rot[i * G + j] * pos(so + j);
                    if (GRAD) da += B.rot[i * G + j] * din(so + j);
                }
            }
            dx[i] = a;
            ddx[i] = da;
            dx2 += a * a;
            dxtddx += a * da;
        }
        double V, V1, V2;
        chb_inter_law(S, dx2, &V, &V1, &V2);
        const double fac = -4.0 * B.fac;
        const double a1 = V1 * fac;
        const double b1 = GRAD ? 2.0 * fac * V2 * dxtddx : 0.0;
        for (int i = 0; i < G; ++i) df[i] = GRAD ? b1 * dx[i] + a1 * ddx[i] : a1 * dx[i];
        if (B.isid) {
            for (int i = 0; i < G; ++i) out(so + i) += df[i];
        } else {
            for (int i = 0; i < G; ++i) {
                double r = 0.0;
                for (int j = 0; j < G; ++j) r += B.rot[j * G + i] * df[j];
                out(so + i) += r;
            }
        }
        for (int i = 0; i < G; ++i) out(to + i) -= df[i];
    }
}

// Address of entry e of column c: forces [c][e]; grad forces [v][e][k] with c = v*ndof + k.
CHB_HD long long chb_ode_addr(bool grad, long long c, int e, int vs, long long ndof) {
    if (!grad) return c * vs + e;
    const long long v = c / ndof, k = c - v * ndof;
    return (v * vs + e) * ndof + k;
}

#if defined(__CUDACC__) && !defined(CHB_HOSTSIM)
// SMEM: 1 = input and output tiles in shared memory, 0 = straight to global memory (vectors too long for a tile).
template <bool GRAD, bool SMEM>
__global__ void k_ode_forces(ChbSys S, long long ncol, long long ndof, const double *__restrict__ pos,
                             const double *__restrict__ din, double *__restrict__ out) {
    extern __shared__ double ode_sm[];
    const int vs = S.nsegm * S.G;
    const int T = blockDim.x, TP = T + 1, tid = threadIdx.x;
    double *in_s = ode_sm;                    // [vs][TP]
    double *out_s = ode_sm + (size_t)vs * TP;  // [vs][TP]
    const long long ntile = (ncol + T - 1) / T;
    for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
        const long long c0 = tile * T;
        const int nc = (int)min((long long)T, ncol - c0);
        const long long c = c0 + tid;
        const long long v = GRAD ? c / ndof : c;
        if (SMEM) {
            if (!GRAD) {
                // the tile is one contiguous block of nc*vs doubles: coalesced read, transposed into [e][column]
                // (loads batched in registers: the shared-memory stores would otherwise order them one latency apart)
                const double *src = pos + c0 * vs;
                const int n = nc * vs;
                for (int base = tid; base < n; base += 8 * T) {
                    double r[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) r[u] = base + u * T < n ? __ldg(src + base + u * T) : 0.0;
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int idx = base + u * T;
                        if (idx < n) {
                            const int cl = idx / vs, e = idx - cl * vs;
                            in_s[e * TP + cl] = r[u];
                        }
                    }
                }
            } else if (tid < nc) {
                // consecutive columns of one vector are consecutive in memory for every entry
                const long long k = c - v * ndof;
                const double *src = din + v * vs * ndof + k;
                for (int e0 = 0; e0 < vs; e0 += 8) {
                    double r[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) r[u] = e0 + u < vs ? __ldg(src + (long long)(e0 + u) * ndof) : 0.0;
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (e0 + u < vs) in_s[(e0 + u) * TP + tid] = r[u];
                }
            }
            __syncthreads();
            if (tid < nc) {
                if (!GRAD) {
                    auto P = [&](int e) { return in_s[e * TP + tid]; };
                    auto O = [&](int e) -> double & { return out_s[e * TP + tid]; };
                    chb_ode_column<false>(S, vs, P, P, O);
                } else {
                    const double *pv = pos + v * vs;
                    auto P = [&](int e) { return __ldg(pv + e); };
                    auto D = [&](int e) { return in_s[e * TP + tid]; };
                    auto O = [&](int e) -> double & { return out_s[e * TP + tid]; };
                    chb_ode_column<true>(S, vs, P, D, O);
                }
            }
            __syncthreads();
            if (!GRAD) {
                double *dst = out + c0 * vs;
                for (int idx = tid; idx < nc * vs; idx += T) {
                    const int cl = idx / vs, e = idx - cl * vs;
                    dst[idx] = out_s[e * TP + cl];
                }
            } else if (tid < nc) {
                const long long k = c - v * ndof;
                double *dst = out + v * vs * ndof + k;
                for (int e = 0; e < vs; ++e) dst[(long long)e * ndof] = out_s[e * TP + tid];
            }
            __syncthreads();
        } else if (tid < nc) {
            const double *pv = pos + v * vs;
            auto P = [&](int e) { return __ldg(pv + e); };
            auto D = [&](int e) { return din[chb_ode_addr(true, c, e, vs, ndof)]; };
            auto O = [&](int e) -> double & { return out[chb_ode_addr(GRAD, c, e, vs, ndof)]; };
            chb_ode_column<GRAD>(S, vs, P, D, O);
        }
    }
}
// Dense variants (systems with S.pair_w: identity BinSpaceRot everywhere, no self-interaction; 20B, 64C ...): every segment
// gathers from all the others through the symmetric weight matrix W[s][t] = sum of BinProdChargeSum over the bins {s, t},
//     F_s = sum_t -4 W[s][t] V'(|x_s - x_t|^2) (x_s - x_t),
// so each pair is evaluated from both sides but nothing is scattered: accumulators live in registers, shared memory holds
// only the positions (and a 32-column slab of tangents), and the SMs run at full occupancy.  FP64-pipe bound.
template <int G>
__global__ void k_ode_forces_dense(ChbSys S, long long nvec, const double *__restrict__ pos, double *__restrict__ out) {
    extern __shared__ double ode_sm[];
    const int ns = S.nsegm, vs = ns * G, spad = S.pair_spad;
    const int nv_tile = blockDim.x / ns, tid = threadIdx.x;
    const int vl = tid / ns, s = tid - vl * ns;
    const long long ntile = (nvec + nv_tile - 1) / nv_tile;
    for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
        const long long v0 = tile * nv_tile;
        const int nv = (int)min((long long)nv_tile, nvec - v0);
        const double *src = pos + v0 * vs;
        for (int idx = tid; idx < nv * vs; idx += blockDim.x) ode_sm[idx] = src[idx];
        __syncthreads();
        if (vl < nv) {
            const double *xv = ode_sm + vl * vs;
            double xs[G], f[G];
#pragma unroll
            for (int i = 0; i < G; ++i) { xs[i] = xv[s * G + i]; f[i] = 0.0; }
            for (int t = 0; t < ns; ++t) {
                const double w = __ldg(S.pair_w + (size_t)t * spad + s);  // symmetric: row t read along s is coalesced
                if (w != 0.0) {
                    double dx[G], dx2 = 0.0;
#pragma unroll
                    for (int i = 0; i < G; ++i) { dx[i] = xs[i] - xv[t * G + i]; dx2 += dx[i] * dx[i]; }
                    double V, V1, V2;
                    chb_inter_law(S, dx2, &V, &V1, &V2);
                    const double a1 = -4.0 * w * V1;
#pragma unroll
                    for (int i = 0; i < G; ++i) f[i] += a1 * dx[i];
                }
            }
            double *dst = out + (v0 + vl) * vs + s * G;
#pragma unroll
            for (int i = 0; i < G; ++i) dst[i] = f[i];
        }
        __syncthreads();
    }
}

// One work item = (vector, slab of 32 tangent columns): lane = column, warps stride over the segments.
#define CHB_ODE_KC 32
template <int G>
__global__ void k_ode_grad_forces_dense(ChbSys S, long long nvec, long long ndof, const double *__restrict__ pos,
                                        const double *__restrict__ din, double *__restrict__ out) {
    extern __shared__ double ode_sm[];
    const int ns = S.nsegm, vs = ns * G, spad = S.pair_spad;
    double *xv = ode_sm;       // [vs]
    double *ds = ode_sm + vs;  // [vs][KC]
    const int tid = threadIdx.x, kl = tid & (CHB_ODE_KC - 1), w0 = tid / CHB_ODE_KC, nw = blockDim.x / CHB_ODE_KC;
    const long long nslab = (ndof + CHB_ODE_KC - 1) / CHB_ODE_KC, nitem = nvec * nslab;
    for (long long item = blockIdx.x; item < nitem; item += gridDim.x) {
        const long long v = item / nslab, k0 = (item - v * nslab) * CHB_ODE_KC;
        const bool live = k0 + kl < ndof;
        const double *dv = din + v * vs * ndof + k0 + kl;
        for (int e = tid; e < vs; e += blockDim.x) xv[e] = pos[v * vs + e];
        for (int e = w0; e < vs; e += nw) ds[e * CHB_ODE_KC + kl] = live ? dv[(long long)e * ndof] : 0.0;
        __syncthreads();
        for (int s = w0; s < ns; s += nw) {
            double xs[G], dsv[G], g[G];
#pragma unroll
            for (int i = 0; i < G; ++i) { xs[i] = xv[s * G + i]; dsv[i] = ds[(s * G + i) * CHB_ODE_KC + kl]; g[i] = 0.0; }
            for (int t = 0; t < ns; ++t) {
                const double w = __ldg(S.pair_w + (size_t)s * spad + t);  // warp-uniform
                if (w != 0.0) {
                    double dx[G], ddx[G], dx2 = 0.0, dxtddx = 0.0;
#pragma unroll
                    for (int i = 0; i < G; ++i) {
                        dx[i] = xs[i] - xv[t * G + i];
                        ddx[i] = dsv[i] - ds[(t * G + i) * CHB_ODE_KC + kl];
                        dx2 += dx[i] * dx[i];
                        dxtddx += dx[i] * ddx[i];
                    }
                    double V, V1, V2;
                    chb_inter_law(S, dx2, &V, &V1, &V2);
                    const double a1 = -4.0 * w * V1, b1 = -8.0 * w * V2 * dxtddx;
#pragma unroll
                    for (int i = 0; i < G; ++i) g[i] += b1 * dx[i] + a1 * ddx[i];
                }
            }
            if (live) {
                double *dst = out + v * vs * ndof + k0 + kl;
#pragma unroll
                for (int i = 0; i < G; ++i) dst[(long long)(s * G + i) * ndof] = g[i];
            }
        }
        __syncthreads();
    }
}
#else
// host emulation: one column after the other
template <bool GRAD>
static void ode_forces_host(const ChbSys &S, long long ncol, long long ndof, const double *pos, const double *din, double *out) {
    const int vs = S.nsegm * S.G;
    for (long long c = 0; c < ncol; ++c) {
        const long long v = GRAD ? c / ndof : c;
        const double *pv = pos + v * vs;
        auto P = [&](int e) { return pv[e]; };
        auto D = [&](int e) { return GRAD ? din[chb_ode_addr(true, c, e, vs, ndof)] : 0.0; };
        auto O = [&](int e) -> double & { return out[chb_ode_addr(GRAD, c, e, vs, ndof)]; };
        chb_ode_column<GRAD>(S, vs, P, D, O);
    }
}
#endif
