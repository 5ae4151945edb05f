// chb_post.cuh -- post-solve reductions and the small diagonal operators next to the action path, one CTA per state.
//
// Reference restated (file:line in choreo/cython/_NBodySyst.pyx @ 8c8dd3f2):
//   segm_pos_to_hash 6271-6369 (all exponents of Hash_exp in ONE pass over the pair distances)
//   ComputeCenterOfMass 3968-4009 | DetectEscape 2125-2192 | segmpos_to_unary_path_stats / _binary_path_stats 7420-7529
//   changevar_mom_pos / _T / _inv / _invT 4489-4827 (params_changevar 2330-2410)
//   Compute_velocities_vectorized / Compute_grad_velocities_vectorized 8106-8160
// CTA programs over a context (ctx.par) like chb_core.cuh, so tests/hostsim runs the same code on the CPU.
#pragma once
#include "chb_core.cuh"

#define CHB_MAX_HASH 16
#define CHB_POST_THREADS 128
#define CHB_MAX_BODY 64

struct ChbBodyTab {  // per (body, interval of the period): stored segment and the symmetry that carries the body onto it
    int nbody, nint_min;
    const int *bodysegm;     // [nbody][nint_min]
    const int *timerev;      // [nbody][nint_min]
    const double *rot;       // [nbody][nint_min][G][G]
    const double *bodymass;  // [nbody]  loopmass[bodyloop[ib]]
    double totmass;          // sum over bodies and intervals
};
struct ChbHashExp {
    int n;
    double e[CHB_MAX_HASH];
};

// part: [K][nthr] partial values of the phase before; thread k < K folds row k with `op` (0 sum, 1 min, 2 max) into out[k]
template <class Ctx>
CHB_D void cta_fold(Ctx &ctx, const double *part, int K, double *out, int op_lo, int op_hi, int split) {
    ctx.par([&](int tid, int nthr) {
        if (tid >= K) return;
        const int op = tid < split ? op_lo : op_hi;
        const double *row = part + (long long)tid * nthr;
        double v = row[0];
        for (int j = 1; j < nthr; ++j) v = op == 0 ? v + row[j] : (op == 1 ? (row[j] < v ? row[j] : v) : (row[j] > v ? row[j] : v));
        out[tid] = v;
    });
}

CHB_HD void chb_bin_dpos(const ChbSys &S, const ChbBin &B, const double *seg_b, int i, double *d) {
    const int G = S.G;
    const double *xs = seg_b + ((long long)B.src * S.segm_store + i) * G;
    const double *xt = seg_b + ((long long)B.tgt * S.segm_store + i) * G;
    for (int a = 0; a < G; ++a) {
        double t;
        if (B.isid) {
            t = xs[a];
        } else {
            t = 0.0;
            for (int b = 0; b < G; ++b) t += B.rot[a * G + b] * xs[b];
        }
        d[a] = t - xt[a];
    }
}

// hash[b][e] = sum_bins BinProdChargeSum / segm_size * sum_i w_i |R x_src - x_tgt|^(2 exp_e)   (6271-6369)
template <class Ctx>
CHB_D void cta_hash_multi(Ctx &ctx, const ChbSys &S, const ChbHashExp &H, const double *seg_b, double *hash_b, double *part) {
    const int store = S.segm_store;
    const double inv_size = 1.0 / (double)S.segm_size;
    ctx.par([&](int tid, int nthr) {
        double h[CHB_MAX_HASH];
        for (int e = 0; e < H.n; ++e) h[e] = 0.0;
        for (int ib = 0; ib < S.nbin; ++ib) {
            const ChbBin &B = S.bins[ib];
            const double fac = B.fac * inv_size;
            for (int i = tid; i < store; i += nthr) {
                double d[CHB_MAX_G], xsq = 0.0;
                chb_bin_dpos(S, B, seg_b, i, d);
                for (int a = 0; a < S.G; ++a) xsq += d[a] * d[a];
                const double w = (store != S.segm_size && (i == 0 || i == store - 1)) ? 0.5 * fac : fac;
                for (int e = 0; e < H.n; ++e) h[e] += w * pow(xsq, H.e[e]);
            }
        }
        for (int e = 0; e < H.n; ++e) part[(long long)e * nthr + tid] = h[e];
    });
    cta_fold(ctx, part, H.n, hash_b, 0, 0, H.n);
}

// centre of mass over all bodies and the whole period (3968-4009): out[G]
template <class Ctx>
CHB_D void cta_center_of_mass(Ctx &ctx, const ChbSys &S, const ChbBodyTab &T, const double *seg_b, double *out_b, double *part,
                               double *ssum) {
    const int G = S.G, size = S.segm_size, store = S.segm_store;
    // ssum[s][0..G): sum over samples 0 .. size-1 ; ssum[s][G..2G): sum over samples 1 .. size (time-reversed bodies)
    for (int s = 0; s < S.nsegm; ++s) {
        ctx.par([&](int tid, int nthr) {
            double a[CHB_MAX_G];
            for (int g = 0; g < G; ++g) a[g] = 0.0;
            for (int i = tid; i < size; i += nthr)
                for (int g = 0; g < G; ++g) a[g] += seg_b[((long long)s * store + i) * G + g];
            for (int g = 0; g < G; ++g) part[(long long)g * nthr + tid] = a[g];
        });
        cta_fold(ctx, part, G, ssum + (long long)s * 2 * G, 0, 0, G);
    }
    ctx.par([&](int tid, int nthr) {
        (void)nthr;
        if (tid != 0) return;
        for (int s = 0; s < S.nsegm; ++s)
            for (int g = 0; g < G; ++g) {
                const double f = ssum[(long long)s * 2 * G + g];
                ssum[(long long)s * 2 * G + G + g] =
                    store > size ? f - seg_b[((long long)s * store) * G + g] + seg_b[((long long)s * store + size) * G + g] : f;
            }
        double tot[CHB_MAX_G];
        for (int g = 0; g < G; ++g) tot[g] = 0.0;
        for (int ib = 0; ib < T.nbody; ++ib)
            for (int ii = 0; ii < T.nint_min; ++ii) {
                const int k = ib * T.nint_min + ii;
                const double *src = ssum + (long long)T.bodysegm[k] * 2 * G + (T.timerev[k] > 0 ? 0 : G);
                const double *R = T.rot + (long long)k * G * G;
                for (int a = 0; a < G; ++a) {
                    double v = 0.0;
                    for (int b = 0; b < G; ++b) v += R[a * G + b] * src[b];
                    tot[a] += T.bodymass[ib] * v;
                }
            }
        for (int g = 0; g < G; ++g) out_b[g] = tot[g] / ((double)size * T.totmass);
    });
}

// escape test (2125-2192): flag_b = 1 when the graph of overlapping body boxes is not connected
template <class Ctx>
CHB_D void cta_detect_escape(Ctx &ctx, const ChbSys &S, const ChbBodyTab &T, const double *seg_b, double fac, int *flag_b,
                              double *part, double *mm, double *box) {
    const int G = S.G, store = S.segm_store;
    // mm[s][0..G) = min over the stored samples, mm[s][G..2G) = max
    for (int s = 0; s < S.nsegm; ++s) {
        ctx.par([&](int tid, int nthr) {
            double lo[CHB_MAX_G], hi[CHB_MAX_G];
            for (int g = 0; g < G; ++g) { lo[g] = 1e300; hi[g] = -1e300; }
            for (int i = tid; i < store; i += nthr)
                for (int g = 0; g < G; ++g) {
                    const double v = seg_b[((long long)s * store + i) * G + g];
                    lo[g] = v < lo[g] ? v : lo[g];
                    hi[g] = v > hi[g] ? v : hi[g];
                }
            for (int g = 0; g < G; ++g) {
                part[(long long)g * nthr + tid] = lo[g];
                part[(long long)(G + g) * nthr + tid] = hi[g];
            }
        });
        cta_fold(ctx, part, 2 * G, mm + (long long)s * 2 * G, 1, 2, G);
    }
    // per body: box of the rotated corners, centre and diagonal: box[ib][0..G) centre, box[ib][G] diagonal
    ctx.par([&](int tid, int nthr) {
        for (int ib = tid; ib < T.nbody; ib += nthr) {
            double lo[CHB_MAX_G], hi[CHB_MAX_G];
            for (int g = 0; g < G; ++g) { lo[g] = 1e300; hi[g] = -1e300; }
            for (int ii = 0; ii < T.nint_min; ++ii) {
                const int k = ib * T.nint_min + ii;
                const double *R = T.rot + (long long)k * G * G;
                for (int c = 0; c < 2; ++c) {
                    const double *x = mm + (long long)T.bodysegm[k] * 2 * G + c * G;
                    for (int a = 0; a < G; ++a) {
                        double v = 0.0;
                        for (int b = 0; b < G; ++b) v += R[a * G + b] * x[b];
                        lo[a] = v < lo[a] ? v : lo[a];
                        hi[a] = v > hi[a] ? v : hi[a];
                    }
                }
            }
            double sz = 0.0;
            for (int g = 0; g < G; ++g) {
                box[(long long)ib * (G + 1) + g] = (hi[g] + lo[g]) / 2;
                sz += (hi[g] - lo[g]) * (hi[g] - lo[g]);
            }
            box[(long long)ib * (G + 1) + G] = sqrt(sz);
        }
    });
    // connectivity from body 0 (breadth-first over the implicit adjacency), one thread
    ctx.par([&](int tid, int nthr) {
        (void)nthr;
        if (tid != 0) return;
        unsigned char seen[CHB_MAX_BODY];
        int queue[CHB_MAX_BODY];
        const int nb = T.nbody;
        for (int i = 0; i < nb; ++i) seen[i] = 0;
        int head = 0, tail = 0, count = 1;
        seen[0] = 1;
        queue[tail++] = 0;
        while (head < tail) {
            const int a = queue[head++];
            for (int b = 0; b < nb; ++b) {
                if (seen[b]) continue;
                double dist = 0.0;
                for (int g = 0; g < G; ++g) {
                    const double dx = box[(long long)a * (G + 1) + g] - box[(long long)b * (G + 1) + g];
                    dist += dx * dx;
                }
                if (sqrt(dist) < fac * (box[(long long)a * (G + 1) + G] + box[(long long)b * (G + 1) + G])) {
                    seen[b] = 1;
                    queue[tail++] = b;
                    ++count;
                }
            }
        }
        *flag_b = count == nb ? 0 : 1;
    });
}

// loop lengths (from segment velocities) and minimum distance per binary interaction (7420-7529, 3903-3943)
template <class Ctx>
CHB_D void cta_path_stats(Ctx &ctx, const ChbSys &S, int nint_min, const double *seg_b, const double *vel_b, double *loop_len_b,
                           double *bin_min_b, double *part, double *seglen) {
    const int G = S.G, size = S.segm_size, store = S.segm_store;
    for (int s = 0; s < S.nsegm; ++s) {
        ctx.par([&](int tid, int nthr) {
            double a = 0.0;
            for (int i = tid; i < store; i += nthr) {
                double vsq = 0.0;
                for (int g = 0; g < G; ++g) {
                    const double v = vel_b[((long long)s * store + i) * G + g];
                    vsq += v * v;
                }
                const double w = (store != size && (i == 0 || i == store - 1)) ? 0.5 : 1.0;
                a += w * sqrt(vsq);
            }
            part[tid] = a;
        });
        cta_fold(ctx, part, 1, seglen + s, 0, 0, 1);
    }
    for (int ib = 0; ib < S.nbin; ++ib) {
        ctx.par([&](int tid, int nthr) {
            double m = 1.7976931348623157e308;
            const ChbBin &B = S.bins[ib];
            for (int i = tid; i < store; i += nthr) {
                double d[CHB_MAX_G], xsq = 0.0;
                chb_bin_dpos(S, B, seg_b, i, d);
                for (int a = 0; a < G; ++a) xsq += d[a] * d[a];
                m = xsq < m ? xsq : m;
            }
            part[tid] = m;
        });
        cta_fold(ctx, part, 1, bin_min_b + S.bins[ib].pad, 1, 1, 1);  // pad = index of the interaction in the caller's order
    }
    ctx.par([&](int tid, int nthr) {
        (void)nthr;
        if (tid != 0) return;
        for (int il = 0; il < S.nloop; ++il) loop_len_b[il] = 0.0;
        for (int s = 0; s < S.nsegm; ++s) loop_len_b[S.segs[s].loop] += (double)nint_min * seglen[s] / (double)size;
        for (int il = 0; il < S.nloop; ++il) loop_len_b[il] /= (double)nint_min;
        for (int ib = 0; ib < S.nbin; ++ib) bin_min_b[ib] = sqrt(bin_min_b[ib]);
    });
}

// params_changevar (2330-2410): mode = inv + 2 * transpose.
//   mode 0: mom [nparams] -> pos [nparams_incl_o]   (x scale, removed zero-frequency entries and the padding half are 0)
//   mode 2: pos -> mom (x scale)     mode 1: pos -> mom (/ scale)     mode 3: mom -> pos (/ scale)
// pos layout: loop il occupies [2 pos_off, 2 pos_off + 2 nmem), nmem = npr * C values then nmem zeros
template <class Ctx>
CHB_D void cta_changevar(Ctx &ctx, const ChbSys &S, const long long *pos_off, int mode, const double *in_b, double *out_b) {
    ctx.par([&](int tid, int nthr) {
        for (int il = 0; il < S.nloop; ++il) {
            const ChbLoop &L = S.loops[il];
            const long long nmem = (long long)L.npr * L.C, base = 2 * pos_off[il];
            const bool to_pos = (mode == 0 || mode == 3);
            for (long long idx = tid; idx < nmem; idx += nthr) {
                const int n = (int)(idx / L.C), c = (int)(idx - (long long)n * L.C);
                int kk;
                double sc = chb_scale(L, n, c, &kk);
                if (mode == 1 || mode == 3) sc = 1.0 / sc;
                const long long im = chb_mom_index(L, n, c);
                if (to_pos) {
                    out_b[base + idx] = im >= 0 ? in_b[im] * sc : 0.0;
                    out_b[base + nmem + idx] = 0.0;
                } else if (im >= 0) {
                    out_b[im] = in_b[base + idx] * sc;
                }
            }
        }
    });
}
