// chb_core.cuh -- CTA programs of the generic (staged) B200 path.
//
// Every kernel body is written ONCE as a "CTA program": a sequence of parallel phases
// `ctx.par([&](int tid, int nthr){...})` separated by block barriers.  On the GPU `par` runs the
// phase with (threadIdx.x, blockDim.x) and ends with __syncthreads(); the CPU-only build container
// has no GPU, so tests/hostsim compiles the very same header with g++ and a context that runs the
// phase for tid = 0..nthr-1 sequentially.  That emulation exists to validate index arithmetic before
// spending GPU minutes; it is test infrastructure and is never loaded by the product.
//
// Reference restated here (file:line in choreo/cython/_NBodySyst.pyx @ 8c8dd3f2):
//   changevar_mom_pos(_T,_inv) 4489-4827 | params_to_pos_slice 5342-5506 | partial_fft_to_pos_slice_{1,2}_sub 5041-5175
//   pos_slice_to_params 5509-5735 | pos_slice_to_partial_fft_{1,2}_sub 5070-5108,5178-5265 | inplace_twiddle 4976-5038
//   Adjust_after_last_gen(_T) 5267-5340 | pos_slice_to_segmpos 5737-5813 | segmpos_to_pos_slice(_T) 5815-5976
//   segm_pos_to_pot_nrg(_grad,_hess) 6372-6533, 6899-6962 (+ inner variants) | params_to_kin_nrg(_grad_daxpy) 4918-4973
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CHB_HD __host__ __device__ __forceinline__
#define CHB_D __device__ __forceinline__
typedef double2 cd;
#else
#define CHB_HD inline
#define CHB_D inline
struct alignas(16) cd {
    double x, y;
};
#endif

#define CHB_MAX_G 4
#define CHB_MAX_C 32
#define CHB_MAX_NNZ 8
#define CHB_MAX_NPPL 16
#define CHB_MAX_FAC 24
#define CHB_MAX_PRIME 31

CHB_HD cd mk(double x, double y) {
    cd r;
    r.x = x;
    r.y = y;
    return r;
}
CHB_HD cd cadd(cd a, cd b) { return mk(a.x + b.x, a.y + b.y); }
CHB_HD cd csub(cd a, cd b) { return mk(a.x - b.x, a.y - b.y); }
CHB_HD cd cmul(cd a, cd b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
CHB_HD cd cconj(cd a) { return mk(a.x, -a.y); }
CHB_HD cd cscale(cd a, double s) { return mk(a.x * s, a.y * s); }
CHB_HD cd cmuli(cd a) { return mk(-a.y, a.x); }   // i * a
CHB_HD cd cmulmi(cd a) { return mk(a.y, -a.x); }  // -i * a

// ------------------------------------------------------------------------------------------------
// device-visible plan structures (built on the host by chb_plan.inl, uploaded once per nint level)
// ------------------------------------------------------------------------------------------------
struct ChbFft {
    int M;                  // complex length (= npr); the real transforms have length N = 2M
    int N;
    int nfac;
    int fac[CHB_MAX_FAC];   // DIF radices, product = M
    const cd *roots;        // exp(-2 pi i j / N), j in [0, N)
    const int *rev;         // rev[k] = storage slot of output index k after the in-place DIF passes
};

struct ChbLoop {
    int shortcut, n_sub_fft, nnz, nppl, C, ncm, has_k0;
    int npr, ninter, ncomp;
    int nseq;    // real sequences transformed: geodim (RFFT) or C = nnz*nppl (GENERAL)
    int stride;  // complex slots per sequence = M + 1
    long long mom_off;  // offset of this loop in the parameter vector
    int nco_in;         // zero-frequency parameters removed from the parameter vector
    int co_prefix[CHB_MAX_NPPL];
    int co_in[CHB_MAX_NPPL];
    int nnz_k[CHB_MAX_NNZ];
    double loopmul;  // 1/sqrt(loopnb*loopmass*4pi^2)
    int vel;         // velocity plan (changevar_mom_vel, 4830-4914): scale velmul for k > 0, 0 for k = 0
    double velmul;   // 1/sqrt(loopnb*loopmass)
    double invmass;  // 1/loopmass (Compute_velocities: momentum -> velocity of the loop's segments)
    double bre[CHB_MAX_G][CHB_MAX_C], bim[CHB_MAX_G][CHB_MAX_C];
    int has_alg, alg_src;  // Adjust_after_last_gen: extra row 2*npr built from row alg_src
    double alg_rot[CHB_MAX_G * CHB_MAX_G];
    cd wmir[CHB_MAX_NNZ];  // exp(+2 pi i nnz_k / ncoeff_min)
    ChbFft fft;
    const cd *gen_tw;  // [npr+1][nnz]  exp(-2 pi i m nnz_k / nint)
    int n_iint;        // number of segment slots along the slice
    const int *iint_off;  // CSR over slots -> segment ids (ascending)
    const int *iint_seg;
    int seg_count;
    const int *segs;        // generating segments of this loop
    long long scratch_off;  // offset (in cd) of this loop inside one state's FFT scratch
    long long smem_cd;      // cd slots of the FFT buffer = nseq*stride
};

struct ChbSeg {
    int loop, iintrel, timerev, isid;
    double rot[CHB_MAX_G * CHB_MAX_G];
};

struct ChbBin {
    int src, tgt, isid, pad;
    double fac;  // BinProdChargeSum
    double rot[CHB_MAX_G * CHB_MAX_G];
};

struct ChbSys {
    int G, nloop, nsegm, nbin;
    int segm_size, segm_store;
    long long nparams, segsz;  // segsz = nsegm*segm_store*G
    int law_kind;
    double law[4];
    const ChbLoop *loops;
    const ChbSeg *segs;
    const ChbBin *bins;  // sorted by colour round
    int nround;
    const int *round_off;  // [nround+1]
    int pair_tile, pair_ntiles;
    // dense pair kernel (cta_pair_dense): symmetric weight matrix W[s][t] = sum of BinProdChargeSum over the bins {s,t},
    // [nsegm][pair_spad], pair_spad = nsegm rounded up to a multiple of 32; pair_dense = 1 when the level uses it
    int pair_dense, pair_spad;
    const double *pair_w;
    int pair_uniform;  // all interacting pairs share the weight pair_wu
    double pair_wu;
    int pair_tilek, tile_nblk, tile_ngrp;  // cta_pair_tile serves the level: segment blocks of 8, groups of 32 samples per CTA
    long long scratch_per_state;  // cd slots
    const unsigned char *skip;    // staged kernels inside the Newton driver: states whose byte is non-zero are left out (NULL: all run)
};

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
// Block-wide deterministic sum: every thread has stored its partial in red[tid] during the previous
// phase; the result lands in red[0].
template <class Ctx>
CHB_D void cta_reduce(Ctx &ctx, double *red) {
    ctx.par([&](int tid, int nthr) {
        if (tid < 32) {
            double s = 0.0;
            for (int j = tid; j < nthr; j += 32) s += red[j];
            red[tid] = s;  // safe: slot tid is only read by thread tid in this phase
        }
    });
    ctx.par([&](int tid, int nthr) {
        if (tid == 0) {
            double s = 0.0;
            int n = nthr < 32 ? nthr : 32;
            for (int j = 0; j < n; ++j) s += red[j];
            red[0] = s;
        }
    });
}

// index of pos-parameter (n, c) of loop L in the user parameter vector, or -1 if removed (co_in False)
CHB_HD long long chb_mom_index(const ChbLoop &L, int n, int c) {
    if (n == 0 && L.has_k0 && c < L.nppl) {
        if (!L.co_in[c]) return -1;
        return L.mom_off + L.co_prefix[c];
    }
    return L.mom_off + (long long)n * L.C + c - L.nco_in;
}

// changevar_mom_pos scale: loopmul/k for k>0, loopmul for k=0  (4512-4566)
CHB_HD double chb_scale(const ChbLoop &L, int n, int c, int *kout) {
    int ik = c / L.nppl;
    int k = n * L.ncm + L.nnz_k[ik];
    *kout = k;
    if (L.vel) return k > 0 ? L.velmul : 0.0;
    return k > 0 ? L.loopmul / (double)k : L.loopmul;
}

// ------------------------------------------------------------------------------------------------
// in-place mixed-radix DIF FFT on `nseq` sequences (natural order in, slot rev[k] holds output k)
// sign = -1: forward exp(-i..), sign = +1: backward (unnormalised)
// ------------------------------------------------------------------------------------------------
CHB_HD cd chb_root(const ChbFft &f, int idx, int sign) {
    cd w = f.roots[idx];
    if (sign > 0) w.y = -w.y;
    return w;
}

template <class Ctx>
CHB_D void cta_fft_dif(Ctx &ctx, cd *buf, int nseq, int stride, const ChbFft &f, int sign) {
    int L = f.M;
    for (int s = 0; s < f.nfac; ++s) {
        const int r = f.fac[s];
        const int Lp = L / r;
        const int nb = f.M / r;
        const int twstep = f.N / L;  // exp(-+2 pi i j t / L) = roots[j*t*twstep]
        const int rstep = f.N / r;   // exp(-+2 pi i q / r)   = roots[q*rstep]
        ctx.par([&](int tid, int nthr) {
            for (int idx = tid; idx < nseq * nb; idx += nthr) {
                int c = idx / nb;
                int b = idx - c * nb;
                int blk = b / Lp;
                int j = b - blk * Lp;
                cd *x = buf + (long long)c * stride + blk * L + j;
                if (r == 2) {
                    cd a0 = x[0], a1 = x[Lp];
                    x[0] = cadd(a0, a1);
                    x[Lp] = cmul(csub(a0, a1), chb_root(f, j * twstep, sign));
                } else if (r == 4) {
                    cd a0 = x[0], a1 = x[Lp], a2 = x[2 * Lp], a3 = x[3 * Lp];
                    cd t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), t3 = csub(a1, a3);
                    t3 = sign < 0 ? cmulmi(t3) : cmuli(t3);
                    x[0] = cadd(t0, t2);
                    x[Lp] = cmul(cadd(t1, t3), chb_root(f, j * twstep, sign));
                    x[2 * Lp] = cmul(csub(t0, t2), chb_root(f, 2 * j * twstep, sign));
                    x[3 * Lp] = cmul(csub(t1, t3), chb_root(f, 3 * j * twstep, sign));
                } else {
                    cd a[CHB_MAX_PRIME];
                    for (int q = 0; q < r; ++q) a[q] = x[q * Lp];
                    for (int t = 0; t < r; ++t) {
                        cd y = a[0];
                        for (int q = 1; q < r; ++q) y = cadd(y, cmul(a[q], chb_root(f, ((q * t) % r) * rstep, sign)));
                        if (t > 0) y = cmul(y, chb_root(f, j * t * twstep, sign));
                        x[t * Lp] = y;
                    }
                }
            }
        });
        L = Lp;
    }
}

// c2r preprocessing (natural order, in place): X_0..X_M  ->  Z_0..Z_{M-1} such that the backward complex FFT
// z = sum_k Z_k exp(+2 pi i k m / M) gives x[2m] = Re z[m], x[2m+1] = Im z[m] of the unnormalised c2r
// x[n] = Re X_0 + 2 Re sum_{0<k<M} X_k exp(2 pi i k n / N) + Re X_M (-1)^n.
template <class Ctx>
CHB_D void cta_c2r_pre(Ctx &ctx, cd *buf, int nseq, int stride, const ChbFft &f) {
    const int M = f.M;
    const int np = M / 2 + 1;  // pairs (k, M-k), k = 0..M/2
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < nseq * np; idx += nthr) {
            int c = idx / np;
            int k = idx - c * np;
            cd *x = buf + (long long)c * stride;
            if (k == 0) {
                double x0 = x[0].x, xm = x[M].x;
                x[0] = mk(x0 + xm, x0 - xm);
            } else {
                int k2 = M - k;
                if (k > k2) continue;
                cd A = x[k], Bc = cconj(x[k2]);
                cd E = cadd(A, Bc), D = csub(A, Bc);
                cd wD = cmul(cconj(f.roots[k]), D);  // exp(+2 pi i k / N) * D
                x[k] = cadd(E, cmuli(wD));
                if (k2 != k) x[k2] = cadd(cconj(E), cmuli(cconj(wD)));
            }
        }
    });
}

// r2c postprocessing: Z (output of the forward DIF FFT of z[m] = x[2m] + i x[2m+1], slot rev[k]) -> X_k in slot
// rev[k] (k < M) and X_M in slot M.
template <class Ctx>
CHB_D void cta_r2c_post(Ctx &ctx, cd *buf, int nseq, int stride, const ChbFft &f) {
    const int M = f.M;
    const int np = M / 2 + 1;
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < nseq * np; idx += nthr) {
            int c = idx / np;
            int k = idx - c * np;
            cd *x = buf + (long long)c * stride;
            if (k == 0) {
                cd z = x[0];
                x[0] = mk(z.x + z.y, 0.0);
                x[M] = mk(z.x - z.y, 0.0);
            } else {
                int k2 = M - k;
                if (k > k2) continue;
                int p1 = f.rev[k], p2 = f.rev[k2];
                cd Zk = x[p1], Zc = cconj(x[p2]);
                cd E = cadd(Zk, Zc), D = csub(Zk, Zc);
                cd wD = cmul(f.roots[k], D);  // exp(-2 pi i k / N) * D
                x[p1] = cscale(csub(E, cmuli(wD)), 0.5);
                if (k2 != k) x[p2] = cscale(csub(cconj(E), cmuli(cconj(wD))), 0.5);
            }
        }
    });
}

// ------------------------------------------------------------------------------------------------
// forward: params -> segment positions, one CTA per (state, loop)
// ------------------------------------------------------------------------------------------------
// value of real sample n of sequence c after the backward transform
CHB_HD double chb_real_at(const cd *buf, int stride, const ChbFft &f, int c, int n) {
    cd v = buf[(long long)c * stride + f.rev[n >> 1]];
    return (n & 1) ? v.y : v.x;
}

// pos_slice row `row` < ncomp of loop L, recomputed on the fly from the transformed buffer
CHB_HD void chb_eval_pos_row_computed(const ChbSys &S, const ChbLoop &L, const cd *buf, int row, double *out) {
    const int G = S.G;
    if (L.shortcut == 1) {
        for (int g = 0; g < G; ++g) out[g] = chb_real_at(buf, L.stride, L.fft, g, row);
        return;
    }
    const int npr = L.npr;
    const bool mirror = row > npr;
    const int mm = mirror ? 2 * npr - row : row;
    for (int g = 0; g < G; ++g) out[g] = 0.0;
    for (int c = 0; c < L.C; ++c) {
        int ik = c / L.nppl;
        cd F = (mm < L.fft.M) ? buf[(long long)c * L.stride + L.fft.rev[mm]] : buf[(long long)c * L.stride + L.fft.M];
        cd W = cmul(F, L.gen_tw[(long long)mm * L.nnz + ik]);
        if (!mirror) {
            for (int g = 0; g < G; ++g) out[g] += 2.0 * (L.bre[g][c] * W.x + L.bim[g][c] * W.y);
        } else {
            cd V = cmul(L.wmir[ik], W);
            for (int g = 0; g < G; ++g) out[g] += 2.0 * (L.bre[g][c] * V.x - L.bim[g][c] * V.y);
        }
    }
}

CHB_HD void chb_eval_pos_row(const ChbSys &S, const ChbLoop &L, const cd *buf, int row, double *out) {
    const int G = S.G;
    if (L.nnz == 0) {
        for (int g = 0; g < G; ++g) out[g] = 0.0;
        return;
    }
    if (row >= L.ncomp) {  // Adjust_after_last_gen row (5290-5302); rows past ncomp only exist when has_alg
        double t[CHB_MAX_G];
        chb_eval_pos_row_computed(S, L, buf, L.alg_src, t);
        for (int a = 0; a < G; ++a) {
            double s = 0.0;
            for (int b = 0; b < G; ++b) s += L.alg_rot[a * G + b] * t[b];
            out[a] = s;
        }
        return;
    }
    chb_eval_pos_row_computed(S, L, buf, row, out);
}

// params_b: this state's parameter vector; segmpos_b: this state's [nsegm][store][G] output;
// buf: nseq*stride cd slots (shared memory or global scratch); red: nthr doubles (shared);
// kin_out: if non-null receives 1/2 sum of squares of this loop's k>0 parameters.
template <class Ctx>
CHB_D void cta_params_to_segmpos(Ctx &ctx, const ChbSys &S, const ChbLoop &L, const double *params_b,
                                  double *segmpos_b, cd *buf, double *red, double *kin_out) {
    const int G = S.G;
    const int M = L.fft.M;
    if (L.nnz > 0) {
        if (L.shortcut == 1) {
            // spectrum X_g[k] = p_pos[k][2g] + i p_pos[k][2g+1]  (5473 / 5499 complex view), X_g[M] = 0
            ctx.par([&](int tid, int nthr) {
                double kin = 0.0;
                for (int idx = tid; idx < (M + 1) * G; idx += nthr) {
                    int k = idx / G, g = idx - k * G;
                    cd X = mk(0.0, 0.0);
                    if (k < M) {
                        int kk;
                        double sc = chb_scale(L, k, 0, &kk);
                        long long i0 = chb_mom_index(L, k, 2 * g), i1 = chb_mom_index(L, k, 2 * g + 1);
                        double p0 = i0 >= 0 ? params_b[i0] : 0.0, p1 = i1 >= 0 ? params_b[i1] : 0.0;
                        X = mk(p0 * sc, p1 * sc);
                        if (kk > 0) kin += p0 * p0 + p1 * p1;
                    }
                    buf[(long long)g * L.stride + k] = X;
                }
                red[tid] = kin;
            });
            cta_c2r_pre(ctx, buf, L.nseq, L.stride, L.fft);
            cta_fft_dif(ctx, buf, L.nseq, L.stride, L.fft, +1);
        } else {
            // real sequences p_c[n] (n < npr), zero-padded to 2*npr, packed as z[m] = p[2m] + i p[2m+1]
            ctx.par([&](int tid, int nthr) {
                double kin = 0.0;
                for (int idx = tid; idx < 2 * L.npr * L.C; idx += nthr) {
                    int n = idx / L.C, c = idx - n * L.C;
                    double v = 0.0;
                    if (n < L.npr) {
                        int kk;
                        double sc = chb_scale(L, n, c, &kk);
                        long long im = chb_mom_index(L, n, c);
                        double p = im >= 0 ? params_b[im] : 0.0;
                        if (kk > 0) kin += p * p;
                        v = p * sc;
                        if (n == 0 && L.has_k0 && c < L.nppl) v *= 0.5;  // 5389-5392
                    }
                    ((double *)(buf + (long long)c * L.stride))[n] = v;
                }
                red[tid] = kin;
            });
            cta_fft_dif(ctx, buf, L.nseq, L.stride, L.fft, -1);
            cta_r2c_post(ctx, buf, L.nseq, L.stride, L.fft);
        }
    } else {
        ctx.par([&](int tid, int nthr) { red[tid] = 0.0; });
    }
    // scatter into the generating segments of this loop (pos_slice_to_segmpos, 5737-5813)
    if (segmpos_b) {
        ctx.par([&](int tid, int nthr) {
            const int store = S.segm_store;
            for (int idx = tid; idx < L.seg_count * store; idx += nthr) {
                int si = idx / store, i = idx - si * store;
                int s = L.segs[si];
                const ChbSeg &sg = S.segs[s];
                int row = sg.iintrel * S.segm_size + (sg.timerev > 0 ? i : store - 1 - i);
                double v[CHB_MAX_G];
                chb_eval_pos_row(S, L, buf, row, v);
                double *dst = segmpos_b + ((long long)s * store + i) * G;
                if (sg.isid) {
                    for (int a = 0; a < G; ++a) dst[a] = v[a];
                } else {
                    for (int a = 0; a < G; ++a) {
                        double t = 0.0;
                        for (int b = 0; b < G; ++b) t += v[b] * sg.rot[b * G + a];  // x @ R
                        dst[a] = t;
                    }
                }
            }
        });
    }
    if (kin_out) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            if (tid == 0) *kin_out = 0.5 * red[0];
        });
    }
}

// ------------------------------------------------------------------------------------------------
// adjoint / inverse: segment array -> params, one CTA per (state, loop)
//   mode 0: segmpos_to_params_T (accumulating gather + Adjust_after_last_gen_T, direction -1)
//   mode 1: segmpos_to_params   (last-writer-wins gather, direction +1)
// out_b[idx] = value (+ kin_src_b[idx] on k>0 entries when kin_src_b != NULL: params_to_kin_nrg_grad_daxpy)
// ------------------------------------------------------------------------------------------------
CHB_HD void chb_seg_contrib(const ChbSys &S, const double *seg_b, int s, int i, double *v) {
    const int G = S.G;
    const ChbSeg &sg = S.segs[s];
    int ii = sg.timerev > 0 ? i : S.segm_store - 1 - i;
    const double *src = seg_b + ((long long)s * S.segm_store + ii) * G;
    if (sg.isid) {
        for (int a = 0; a < G; ++a) v[a] = src[a];
    } else {
        for (int a = 0; a < G; ++a) {
            double t = 0.0;
            for (int b = 0; b < G; ++b) t += sg.rot[a * G + b] * src[b];  // R . x  (transt dgemm, 5863/5945)
            v[a] = t;
        }
    }
}

CHB_HD void chb_gather_row(const ChbSys &S, const ChbLoop &L, const double *seg_b, int row, int mode, double *out) {
    const int G = S.G;
    for (int g = 0; g < G; ++g) out[g] = 0.0;
    int q = row / S.segm_size;
    int i = row - q * S.segm_size;
    int best = -1, best_i = 0;
    double v[CHB_MAX_G];
    if (q < L.n_iint) {
        for (int e = L.iint_off[q]; e < L.iint_off[q + 1]; ++e) {
            int s = L.iint_seg[e];
            if (mode == 0) {
                chb_seg_contrib(S, seg_b, s, i, v);
                for (int g = 0; g < G; ++g) out[g] += v[g];
            } else if (s > best) {
                best = s;
                best_i = i;
            }
        }
    }
    if (S.segm_store > S.segm_size && i == 0 && q >= 1 && q - 1 < L.n_iint) {
        for (int e = L.iint_off[q - 1]; e < L.iint_off[q]; ++e) {
            int s = L.iint_seg[e];
            if (mode == 0) {
                chb_seg_contrib(S, seg_b, s, S.segm_size, v);
                for (int g = 0; g < G; ++g) out[g] += v[g];
            } else if (s > best) {
                best = s;
                best_i = S.segm_size;
            }
        }
    }
    if (mode == 1 && best >= 0) chb_seg_contrib(S, seg_b, best, best_i, out);
}

CHB_HD void chb_eval_P_row(const ChbSys &S, const ChbLoop &L, const double *seg_b, int row, int mode, double *out) {
    chb_gather_row(S, L, seg_b, row, mode, out);
    if (mode == 0 && L.has_alg && row == L.alg_src) {  // Adjust_after_last_gen_T (5323-5340)
        double t[CHB_MAX_G];
        chb_gather_row(S, L, seg_b, 2 * L.npr, 0, t);
        const int G = S.G;
        for (int b = 0; b < G; ++b) {
            double s = 0.0;
            for (int a = 0; a < G; ++a) s += L.alg_rot[a * G + b] * t[a];
            out[b] += s;
        }
    }
}

template <class Ctx>
CHB_D void cta_segm_to_params(Ctx &ctx, const ChbSys &S, const ChbLoop &L, const double *seg_b, double *out_b,
                               const double *kin_src_b, cd *buf, int mode) {
    const int G = S.G;
    const int M = L.fft.M;
    const int N = 2 * M;
    if (L.nnz == 0) return;
    if (L.shortcut == 1) {
        ctx.par([&](int tid, int nthr) {
            for (int n = tid; n < N; n += nthr) {
                double v[CHB_MAX_G];
                chb_eval_P_row(S, L, seg_b, n, mode, v);
                for (int g = 0; g < G; ++g) ((double *)(buf + (long long)g * L.stride))[n] = v[g];
            }
        });
        cta_fft_dif(ctx, buf, L.nseq, L.stride, L.fft, -1);
        cta_r2c_post(ctx, buf, L.nseq, L.stride, L.fft);
        ctx.par([&](int tid, int nthr) {
            for (int idx = tid; idx < M * G; idx += nthr) {
                int k = idx / G, g = idx - k * G;
                cd Y = buf[(long long)g * L.stride + L.fft.rev[k]];
                double f = mode == 0 ? (k > 0 ? 2.0 : 1.0) : 1.0 / (double)N;  // 5728-5735
                int kk;
                double sc = chb_scale(L, k, 0, &kk);
                double mul = mode == 0 ? f * sc : f / sc;
                long long i0 = chb_mom_index(L, k, 2 * g), i1 = chb_mom_index(L, k, 2 * g + 1);
                if (i0 >= 0) out_b[i0] = Y.x * mul + ((kin_src_b && kk > 0) ? kin_src_b[i0] : 0.0);
                if (i1 >= 0) out_b[i1] = Y.y * mul + ((kin_src_b && kk > 0) ? kin_src_b[i1] : 0.0);
            }
        });
        return;
    }
    // GENERAL_SYM: X_m = alpha_m * exp(+2 pi i m nnz_k/nint) * ( B[m] (+ conj(w) conj(B[2npr-m])) ),  B[m] = basis^T P[m]
    const int npr = L.npr;
    ctx.par([&](int tid, int nthr) {
        for (int m = tid; m <= npr; m += nthr) {
            double P1[CHB_MAX_G], P2[CHB_MAX_G];
            chb_eval_P_row(S, L, seg_b, m, mode, P1);
            const bool mir = (L.n_sub_fft == 1) && m > 0 && m < npr;
            if (mir) chb_eval_P_row(S, L, seg_b, 2 * npr - m, mode, P2);
            const bool end = (m == 0 || m == npr);
            double alpha;
            if (mode == 0)
                alpha = end ? 2.0 : 1.0;  // 5102-5106 / 5260-5263
            else
                alpha = (L.n_sub_fft == 1 && !end) ? 0.5 : 1.0;  // dfac 5094-5095 / 5210-5211 (+ end doubling 5260-5263)
            for (int c = 0; c < L.C; ++c) {
                int ik = c / L.nppl;
                cd B = mk(0.0, 0.0);
                for (int g = 0; g < G; ++g) {
                    B.x += L.bre[g][c] * P1[g];
                    B.y += L.bim[g][c] * P1[g];
                }
                if (mir) {
                    cd B2 = mk(0.0, 0.0);
                    for (int g = 0; g < G; ++g) {
                        B2.x += L.bre[g][c] * P2[g];
                        B2.y += L.bim[g][c] * P2[g];
                    }
                    B = cadd(B, cmul(cconj(L.wmir[ik]), cconj(B2)));
                }
                cd X = cscale(cmul(cconj(L.gen_tw[(long long)m * L.nnz + ik]), B), alpha);
                buf[(long long)c * L.stride + m] = X;
            }
        }
    });
    cta_c2r_pre(ctx, buf, L.nseq, L.stride, L.fft);
    cta_fft_dif(ctx, buf, L.nseq, L.stride, L.fft, +1);
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < npr * L.C; idx += nthr) {
            int n = idx / L.C, c = idx - n * L.C;
            long long im = chb_mom_index(L, n, c);
            if (im < 0) continue;
            double x = chb_real_at(buf, L.stride, L.fft, c, n);
            int kk;
            double sc = chb_scale(L, n, c, &kk);
            double val;
            if (mode == 0) {
                if (n == 0 && L.has_k0 && c < L.nppl) x *= 0.5;  // 5634-5641
                val = x * sc;
            } else {
                val = x / ((double)N * sc);
            }
            out_b[im] = val + ((kin_src_b && kk > 0) ? kin_src_b[im] : 0.0);
        }
    });
}

// ------------------------------------------------------------------------------------------------
// pair interactions over one tile of samples of one state.
//   MODE 0: potential energy only; MODE 1: energy + gradient; MODE 2: Hessian-vector product
// Bins are processed in conflict-free colour rounds: inside a round no two bins touch the same segment,
// so (bin, sample) work items update disjoint shared-memory accumulators.
// ------------------------------------------------------------------------------------------------
CHB_HD void chb_inter_law(const ChbSys &S, double xsq, double *V, double *V1, double *V2) {
    double a, b;
    if (S.law_kind == 0) {
#if defined(__CUDA_ARCH__)
        double r = rsqrt(xsq);
#else
        double r = 1.0 / sqrt(xsq);
#endif
        double r2 = r * r;
        a = r2 * r2 * r;  // xsq^-2.5   (inline_gravity_pot, 111-118)
        b = xsq * a;
        *V = -xsq * b;
        *V1 = 0.5 * b;
        *V2 = -0.75 * a;
    } else {
        a = S.law[3] * pow(xsq, S.law[2]);  // power_law_pot, 125-133
        b = xsq * a;
        *V = -xsq * b;
        *V1 = S.law[0] * b;
        *V2 = S.law[1] * a;
    }
}

template <int MODE, class Ctx>
CHB_D void cta_pair(Ctx &ctx, const ChbSys &S, const double *seg_b, const double *dseg_b, double *out_b,
                     double *pot_out, int tile, double globalmul, double *spos, double *sdpos, double *sout,
                     double *red) {
    const int G = S.G, T = S.pair_tile, store = S.segm_store;
    const int i0 = tile * T;
    const int n = (store - i0) < T ? (store - i0) : T;
    const int nS = S.nsegm;
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < nS * n * G; idx += nthr) {
            int s = idx / (n * G);
            int r = idx - s * (n * G);
            long long src = ((long long)s * store + i0) * G + r;
            int dst = s * T * G + r;
            spos[dst] = seg_b[src];
            if (MODE == 2) sdpos[dst] = dseg_b[src];
            if (MODE >= 1) sout[dst] = 0.0;
        }
        red[tid] = 0.0;
    });
    const double inv_size = 1.0 / (double)S.segm_size;
    for (int rd = 0; rd < S.nround; ++rd) {
        const int b0 = S.round_off[rd], nb = S.round_off[rd + 1] - b0;
        ctx.par([&](int tid, int nthr) {
            double e = 0.0;
            for (int item = tid; item < nb * n; item += nthr) {
                int kb = item / n, i = item - kb * n;
                const ChbBin &B = S.bins[b0 + kb];
                const double *xs = spos + (B.src * T + i) * G;
                const double *xt = spos + (B.tgt * T + i) * G;
                double d[CHB_MAX_G], dd[CHB_MAX_G], g[CHB_MAX_G];
                double xsq = 0.0, dot = 0.0;
                for (int a = 0; a < G; ++a) {
                    double t;
                    if (B.isid) {
                        t = xs[a];
                    } else {
                        t = 0.0;
                        for (int b = 0; b < G; ++b) t += B.rot[a * G + b] * xs[b];
                    }
                    d[a] = t - xt[a];
                    xsq += d[a] * d[a];
                }
                if (MODE == 2) {
                    const double *ys = sdpos + (B.src * T + i) * G;
                    const double *yt = sdpos + (B.tgt * T + i) * G;
                    for (int a = 0; a < G; ++a) {
                        double t;
                        if (B.isid) {
                            t = ys[a];
                        } else {
                            t = 0.0;
                            for (int b = 0; b < G; ++b) t += B.rot[a * G + b] * ys[b];
                        }
                        dd[a] = t - yt[a];
                        dot += d[a] * dd[a];
                    }
                }
                double V, V1, V2;
                chb_inter_law(S, xsq, &V, &V1, &V2);
                int gi = i0 + i;
                double w = (store != S.segm_size && (gi == 0 || gi == store - 1)) ? 0.5 : 1.0;
                double binfac = B.fac * inv_size;
                if (MODE <= 1) e += w * V * binfac;
                if (MODE >= 1) {
                    double f = 2.0 * binfac * globalmul * w;
                    if (MODE == 1) {
                        for (int a = 0; a < G; ++a) g[a] = f * V1 * d[a];
                    } else {
                        double bb = 2.0 * V2 * dot;
                        for (int a = 0; a < G; ++a) g[a] = f * (bb * d[a] + V1 * dd[a]);
                    }
                    double *os = sout + (B.src * T + i) * G;
                    double *ot = sout + (B.tgt * T + i) * G;
                    if (B.isid) {
                        for (int a = 0; a < G; ++a) os[a] += g[a];
                    } else {
                        for (int a = 0; a < G; ++a) {
                            double t = 0.0;
                            for (int b = 0; b < G; ++b) t += B.rot[b * G + a] * g[b];  // R^T g
                            os[a] += t;
                        }
                    }
                    for (int a = 0; a < G; ++a) ot[a] -= g[a];
                }
            }
            if (MODE <= 1) red[tid] += e;
        });
    }
    if (MODE >= 1) {
        ctx.par([&](int tid, int nthr) {
            for (int idx = tid; idx < nS * n * G; idx += nthr) {
                int s = idx / (n * G);
                int r = idx - s * (n * G);
                out_b[((long long)s * store + i0) * G + r] = sout[s * T * G + r];
            }
        });
    }
    if (MODE <= 1 && pot_out) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            if (tid == 0) *pot_out = red[0];
        });
    }
}

// kinetic energy of one state: 1/2 sum p^2 over k>0 entries (4918-4944); one CTA per state
template <class Ctx>
CHB_D void cta_kin(Ctx &ctx, const ChbSys &S, const double *params_b, double *kin_out, double *red) {
    ctx.par([&](int tid, int nthr) {
        double kin = 0.0;
        for (int il = 0; il < S.nloop; ++il) {
            const ChbLoop &L = S.loops[il];
            for (int idx = tid; idx < L.npr * L.C; idx += nthr) {
                int n = idx / L.C, c = idx - n * L.C;
                long long im = chb_mom_index(L, n, c);
                int kk;
                chb_scale(L, n, c, &kk);
                if (im >= 0 && kk > 0) {
                    double p = params_b[im];
                    kin += p * p;
                }
            }
        }
        red[tid] = kin;
    });
    cta_reduce(ctx, red);
    ctx.par([&](int tid, int nthr) {
        if (tid == 0) *kin_out = 0.5 * red[0];
    });
}

// out[b] = sum_il kin_part[b][il] + psign * sum_t pot_part[b][t]   (fixed order => deterministic)
CHB_HD double chb_finish_action(const double *kin_part, int nk, const double *pot_part, int np, double psign) {
    double k = 0.0, p = 0.0;
    for (int i = 0; i < nk; ++i) k += kin_part[i];
    for (int i = 0; i < np; ++i) p += pot_part[i];
    return k + psign * p;
}

// ------------------------------------------------------------------------------------------------
// Dense pair kernel for the compute-bound regime (64 segments, 2016 interactions per sample).
// Conditions (checked at plan time): planar, every BinSpaceRot the identity.  Then the scatter of segm_pos_to_pot_nrg_grad /
// _hess (:6502-6529, :6927-6958) is symmetric: segment t receives  -f_st [V' d]  resp.  -f_st [2 V'' (d.dd) d + V' dd]
// from every partner s, with d = x_s - x_t, dd = dx_s - dx_t, f_st = 2 w_i W[s][t] / segm_size * globalmul, whether t is the
// bin's source or its target.  One thread owns (sample i, segment t) and accumulates in REGISTERS over all s: x_s is a
// broadcast read (the lanes of a warp share i), W[s][t] a conflict-free one; no shared-memory read-modify-write, no colour
// rounds, no barriers inside the loop.  Every unordered pair is evaluated twice (once per side): that doubles the FP64 work of
// the pair sum but removes the scatter, which is what bounded the generic kernel (cta_pair) on this shape.
// Gravity uses the closed forms V = -r, V' = r^3/2, V'' = -3 r^5/4 with r = (x^2)^(-1/2) from the hardware seed and one
// third-order correction step (relative error of the seed 2^-20 -> below 2^-55).
// ------------------------------------------------------------------------------------------------
CHB_HD double chb_rsqrt_fast(double d) {
#if defined(__CUDA_ARCH__)
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double e = fma(-d * y, y, 1.0);
    return fma(y * e, fma(e, 0.375, 0.5), y);
#else
    return 1.0 / sqrt(d);
#endif
}

// UNI: every interacting pair has the same weight S.pair_wu (equal masses and charges -- every nC / nB test system): no weight
// matrix in shared memory (more resident CTAs), no weight loads or multiplies inside the loop
// LAW0: the plan's law is the fast power law (law_kind 0, V = -1/r): a compile-time fact inside the kernel, so the unrolled pair
// evaluations are straight-line code the compiler can interleave (with the run-time branch every evaluation was a basic block of
// its own, and the generic branch's pow() made the tiled kernel 15 k instructions long)
template <int MODE, bool UNI, bool LAW0, class Ctx>
CHB_D void cta_pair_dense(Ctx &ctx, const ChbSys &S, const double *seg_b, const double *dseg_b, double *out_b,
                           double *pot_out, int tile, double globalmul, int nthreads, cd *spos, cd *sdpos, cd *sout,
                           double *sw, double *red) {
    const int T = S.pair_tile, store = S.segm_store, nS = S.nsegm, SP = S.pair_spad;
    const int i0 = tile * T;
    const int n = (store - i0) < T ? (store - i0) : T;
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < nS * n; idx += nthr) {
            const int s = idx / n, i = idx - s * n;
            const long long src = (long long)s * store + i0 + i;
            spos[i * SP + s] = ((const cd *)seg_b)[src];
            if (MODE == 2) sdpos[i * SP + s] = ((const cd *)dseg_b)[src];
        }
        if (!UNI)
            for (int idx = tid; idx < nS * SP; idx += nthr) sw[idx] = S.pair_w[idx];
        red[tid] = 0.0;
    });
    const double inv_size = 1.0 / (double)S.segm_size;
    const int nI = nthreads / SP;  // samples in flight
    ctx.par([&](int tid, int nthr) {
        (void)nthr;
        const int t = tid % SP, il = tid / SP;
        if (t >= nS || il >= nI) return;
        double esum = 0.0;
        for (int i = il; i < n; i += nI) {
            const int gi = i0 + i;
            const double wi = (store != S.segm_size && (gi == 0 || gi == store - 1)) ? 0.5 : 1.0;
            const cd *xrow = spos + i * SP;
            const cd *drow = sdpos + i * SP;
            const cd xt = xrow[t];
            cd dxt = mk(0.0, 0.0);
            if (MODE == 2) dxt = drow[t];
            double gx = 0.0, gy = 0.0, e = 0.0;
#pragma unroll 4
            for (int s = 0; s < nS; ++s) {
                const double w = UNI ? 1.0 : sw[s * SP + t];  // UNI: the s == t term (d = 0, x^2 forced to 1) only adds to e, removed below
                const cd xs = xrow[s];
                const double dx = xs.x - xt.x, dy = xs.y - xt.y;
                double xsq = fma(dx, dx, dy * dy);
                xsq = (s == t) ? 1.0 : xsq;  // W[t][t] = 0: the term vanishes, keep it finite
                if (LAW0 || S.law_kind == 0) {
                    const double r = chb_rsqrt_fast(xsq);
                    const double r2 = r * r;
                    const double c = w * (r2 * r);    // 2 V' w
                    if (MODE <= 1) e = fma(w, r, e);  // -V w
                    if (MODE == 1) {
                        gx = fma(c, dx, gx);
                        gy = fma(c, dy, gy);
                    } else if (MODE == 2) {  // 2 (2 V'' (d.dd) d + V' dd) = r^3 (dd - 3 r^2 (d.dd) d)
                        const cd ds = drow[s];
                        const double ddx = ds.x - dxt.x, ddy = ds.y - dxt.y;
                        const double k3 = 3.0 * r2 * fma(dx, ddx, dy * ddy);
                        gx = fma(c, fma(-k3, dx, ddx), gx);
                        gy = fma(c, fma(-k3, dy, ddy), gy);
                    }
                } else {
                    double V, V1, V2;
                    chb_inter_law(S, xsq, &V, &V1, &V2);
                    if (MODE <= 1) e = fma(-w, V, e);
                    if (MODE == 1) {
                        const double c = 2.0 * w * V1;
                        gx = fma(c, dx, gx);
                        gy = fma(c, dy, gy);
                    } else if (MODE == 2) {
                        const cd ds = drow[s];
                        const double ddx = ds.x - dxt.x, ddy = ds.y - dxt.y;
                        const double bb = 2.0 * V2 * fma(dx, ddx, dy * ddy);
                        gx = fma(2.0 * w, fma(bb, dx, V1 * ddx), gx);
                        gy = fma(2.0 * w, fma(bb, dy, V1 * ddy), gy);
                    }
                }
            }
            if (UNI && MODE <= 1) {  // take the self term out of the energy sum
                if (LAW0 || S.law_kind == 0) {
                    e -= chb_rsqrt_fast(1.0);
                } else {
                    double V, V1, V2;
                    chb_inter_law(S, 1.0, &V, &V1, &V2);
                    e += V;
                }
            }
            // g_t = -w_i / segm_size * globalmul * sum_s [...] (the factor 2 and W[s][t] are inside the sums)
            const double wu = UNI ? S.pair_wu : 1.0;
            const double f = -wi * inv_size * globalmul * wu;
            if (MODE >= 1) sout[i * SP + t] = mk(f * gx, f * gy);
            esum = fma(-0.5 * wi * inv_size * wu, e, esum);  // ordered pairs count every bin twice; e holds -sum V w
        }
        if (MODE <= 1) red[tid] += esum;
    });
    if (MODE >= 1) {
        ctx.par([&](int tid, int nthr) {
            for (int idx = tid; idx < nS * n; idx += nthr) {
                const int s = idx / n, i = idx - s * n;
                ((cd *)out_b)[(long long)s * store + i0 + i] = sout[i * SP + s];
            }
        });
    }
    if (MODE <= 1 && pot_out) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            if (tid == 0) *pot_out = red[0];
        });
    }
}

// ------------------------------------------------------------------------------------------------
// Tiled pair kernel: every unordered pair evaluated ONCE, accumulators in registers.
// Same conditions as cta_pair_dense (planar, identity BinSpaceRot, no self-interaction; weights W[s][t] in S.pair_w).
// A warp owns (32 samples, block A of 8 segments): positions and gradients of block A stay in registers while the warp sweeps a
// partner block B in sub-blocks of SBB segments -- 8 x SBB pair evaluations between register operands, no shared-memory traffic
// inside.  Partner blocks are taken in rounds B = (A + r) mod nblk, r = 0 (pairs inside A) .. nblk / 2, so that within a round
// every block is "A" for exactly one warp and "B" for exactly one warp: the A-side sums go to accA, the B-side sums to accB,
// plain read-modify-write, no atomics, fixed order (deterministic).  Compared with cta_pair_dense the FP64 work per sample drops
// from 15 instructions per ORDERED pair to 17 per UNORDERED pair.
// Layout: xs / dxs / accA / accB [SP][TS] cd (SP = nsegm padded to a multiple of 8, TS = samples per CTA = 32 * groups).
// ------------------------------------------------------------------------------------------------
#define CHB_TILE_SB 8
template <int MODE, bool LAW0>
CHB_HD void chb_tile_eval(int law_kind, const double *law, double w, const cd &xa, const cd &xb, const cd &dxa, const cd &dxb, cd &ga,
                          cd &gb, double &e) {
    const double dx = xa.x - xb.x, dy = xa.y - xb.y;
    const double xsq = fma(dx, dx, dy * dy);
    if (LAW0 || law_kind == 0) {
        const double r = chb_rsqrt_fast(xsq);
        const double r2 = r * r;
        const double c = w * (r2 * r);  // 2 V' w
        if (MODE <= 1) e = fma(w, r, e);  // -V w
        if (MODE == 1) {
            ga.x = fma(-c, dx, ga.x); ga.y = fma(-c, dy, ga.y);
            gb.x = fma(c, dx, gb.x);  gb.y = fma(c, dy, gb.y);
        } else if (MODE == 2) {  // h = c (dd - 3 r^2 (d.dd) d), odd under a <-> b
            const double ddx = dxa.x - dxb.x, ddy = dxa.y - dxb.y;
            const double k3 = 3.0 * r2 * fma(dx, ddx, dy * ddy);
            const double hx = c * fma(-k3, dx, ddx), hy = c * fma(-k3, dy, ddy);
            ga.x -= hx; ga.y -= hy;
            gb.x += hx; gb.y += hy;
        }
    } else {
        double a = law[3] * pow(xsq, law[2]);  // power_law_pot, 125-133
        double b = xsq * a;
        const double V = -xsq * b, V1 = law[0] * b, V2 = law[1] * a;
        if (MODE <= 1) e = fma(-w, V, e);
        if (MODE == 1) {
            const double c = 2.0 * w * V1;
            ga.x = fma(-c, dx, ga.x); ga.y = fma(-c, dy, ga.y);
            gb.x = fma(c, dx, gb.x);  gb.y = fma(c, dy, gb.y);
        } else if (MODE == 2) {
            const double ddx = dxa.x - dxb.x, ddy = dxa.y - dxb.y;
            const double bb = 2.0 * V2 * fma(dx, ddx, dy * ddy);
            const double hx = 2.0 * w * fma(bb, dx, V1 * ddx), hy = 2.0 * w * fma(bb, dy, V1 * ddy);
            ga.x -= hx; ga.y -= hy;
            gb.x += hx; gb.y += hy;
        }
    }
}

template <int MODE, bool UNI, bool LAW0, int SBBV = 0, class Ctx>
CHB_D void cta_pair_tile(Ctx &ctx, const ChbSys &S, const double *seg_b, const double *dseg_b, double *out_b, double *pot_out,
                          int tile, double globalmul, int nblk, int ngrp, cd *xs, cd *dxs, cd *accA, cd *accB, double *red) {
    // partner sub-blocks of 2 segments: fewest live registers, which is what the two resident CTAs per SM need (A/B of sub-blocks of
    // 2 / 4 / 8 and of one CTA per SM with up to 255 registers: profiles/r02_pair_tile_variants_ab.txt); SBBV: the A/B variants
    constexpr int SB = CHB_TILE_SB, SBB = SBBV ? SBBV : 2;
    const int store = S.segm_store, nS = S.nsegm, SP = nblk * SB, TS = 32 * ngrp, WSP = S.pair_spad;
    const int i0 = tile * TS;
    const int n = (store - i0) < TS ? (store - i0) : TS;
    ctx.par([&](int tid, int nthr) {
        for (int idx = tid; idx < SP * TS; idx += nthr) {
            const int s = idx / TS, j = idx - s * TS;
            // padding segments / samples: far apart from everything and from each other, weight 0
            cd v = mk(1.0e8 * (double)(s + 1), 1.0e8 * (double)(j + 1));
            cd dv = mk(0.0, 0.0);
            if (s < nS && j < n) {
                const long long src = (long long)s * store + i0 + j;
                v = ((const cd *)seg_b)[src];
                if (MODE == 2) dv = ((const cd *)dseg_b)[src];
            }
            xs[idx] = v;
            if (MODE == 2) dxs[idx] = dv;
            if (MODE >= 1) { accA[idx] = mk(0.0, 0.0); accB[idx] = mk(0.0, 0.0); }
        }
        red[tid] = 0.0;
    });
    // Rounds.  Every unordered block pair {A, B} once: warp A takes B = (A + r) mod nblk in round r.  For an even number of blocks
    // the middle round r = nblk / 2 needs only half of the warps (each pair would come up twice); the other half does the work of
    // the "diagonal round" there instead -- the pairs inside its own block and inside block A - nblk / 2, whose B-side accumulator
    // nobody else writes in that round -- so there is no separate diagonal round and no idle half: nblk / 2 rounds of (almost)
    // equal work per warp instead of nblk / 2 + 1 unequal ones (the round barrier was 19 % of the stall samples,
    // profiles/r02_pair_tile_ncu_summary.md).  Odd nblk: a diagonal round (r = 0) and (nblk - 1) / 2 full rounds.
    const bool even = (nblk % 2) == 0;
    for (int r = even ? 1 : 0; r <= nblk / 2; ++r) {
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            const int warp = tid >> 5, lane = tid & 31;
            const int grp = warp / nblk, A = warp - grp * nblk;
            if (grp >= ngrp) return;
            const int j = grp * 32 + lane;
            double e = 0.0;
            // the pairs inside block D, sums added to acc (accA of the block's own warp, or the free accB of the middle round)
            auto diagonal = [&](int D, cd *acc) {
                cd xd[SB], dxd[SB], gd[SB];
#pragma unroll
                for (int a = 0; a < SB; ++a) {
                    xd[a] = xs[(D * SB + a) * TS + j];
                    dxd[a] = (MODE == 2) ? dxs[(D * SB + a) * TS + j] : mk(0.0, 0.0);
                    gd[a] = mk(0.0, 0.0);
                }
#pragma unroll
                for (int a = 0; a < SB; ++a)
#pragma unroll
                    for (int b = a + 1; b < SB; ++b) {
                        const int sa = D * SB + a, sb = D * SB + b;
                        const double w = UNI ? 1.0 : ((sa < nS && sb < nS) ? S.pair_w[sa * WSP + sb] : 0.0);
                        chb_tile_eval<MODE, LAW0>(S.law_kind, S.law, w, xd[a], xd[b], dxd[a], dxd[b], gd[a], gd[b], e);
                    }
                if (MODE >= 1) {
#pragma unroll
                    for (int a = 0; a < SB; ++a) {
                        const int o = (D * SB + a) * TS + j;
                        acc[o] = cadd(acc[o], gd[a]);
                    }
                }
            };
            const bool middle = even && 2 * r == nblk;
            if (r == 0) {
                diagonal(A, accA);
            } else if (middle && A >= nblk / 2) {
                diagonal(A, accA);
                diagonal(A - nblk / 2, accB);
            } else {
                const int B = (A + r) % nblk;
                cd xa[SB], dxa[SB], ga[SB];
#pragma unroll
                for (int a = 0; a < SB; ++a) {
                    xa[a] = xs[(A * SB + a) * TS + j];
                    dxa[a] = (MODE == 2) ? dxs[(A * SB + a) * TS + j] : mk(0.0, 0.0);
                    ga[a] = mk(0.0, 0.0);
                }
#pragma unroll 1
                for (int jb = 0; jb < SB / SBB; ++jb) {
                    cd xb[SBB], dxb[SBB], gb[SBB];
#pragma unroll
                    for (int b = 0; b < SBB; ++b) {
                        xb[b] = xs[(B * SB + jb * SBB + b) * TS + j];
                        dxb[b] = (MODE == 2) ? dxs[(B * SB + jb * SBB + b) * TS + j] : mk(0.0, 0.0);
                        gb[b] = mk(0.0, 0.0);
                    }
#pragma unroll
                    for (int a = 0; a < SB; ++a)
#pragma unroll
                        for (int b = 0; b < SBB; ++b) {
                            const int sa = A * SB + a, sb = B * SB + jb * SBB + b;
                            const double w = UNI ? 1.0 : ((sa < nS && sb < nS) ? S.pair_w[sa * WSP + sb] : 0.0);
                            chb_tile_eval<MODE, LAW0>(S.law_kind, S.law, w, xa[a], xb[b], dxa[a], dxb[b], ga[a], gb[b], e);
                        }
                    if (MODE >= 1) {
#pragma unroll
                        for (int b = 0; b < SBB; ++b) {
                            const int o = (B * SB + jb * SBB + b) * TS + j;
                            accB[o] = cadd(accB[o], gb[b]);
                        }
                    }
                }
                if (MODE >= 1) {
#pragma unroll
                    for (int a = 0; a < SB; ++a) {
                        const int o = (A * SB + a) * TS + j;
                        accA[o] = cadd(accA[o], ga[a]);
                    }
                }
            }
            if (MODE <= 1) {
                const int gi = i0 + j;
                const double wi = (store != S.segm_size && (gi == 0 || gi == store - 1)) ? 0.5 : 1.0;
                if (j < n) red[tid] += -wi * (UNI ? S.pair_wu : 1.0) * e / (double)S.segm_size;
            }
        });
    }
    if (MODE >= 1) {
        ctx.par([&](int tid, int nthr) {
            const double inv_size = 1.0 / (double)S.segm_size;
            for (int idx = tid; idx < nS * TS; idx += nthr) {
                const int s = idx / TS, j = idx - s * TS;
                if (j >= n) continue;
                const int gi = i0 + j;
                const double wi = (store != S.segm_size && (gi == 0 || gi == store - 1)) ? 0.5 : 1.0;
                const double f = -wi * inv_size * globalmul * (UNI ? S.pair_wu : 1.0);
                const cd g = cadd(accA[idx], accB[idx]);
                ((cd *)out_b)[(long long)s * store + gi] = mk(f * g.x, f * g.y);
            }
        });
    }
    if (MODE <= 1 && pot_out) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            if (tid == 0) *pot_out = red[0];
        });
    }
}
