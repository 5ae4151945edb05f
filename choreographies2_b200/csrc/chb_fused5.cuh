// chb_fused5.cuh -- fused one-CTA-per-state evaluation for planar systems with ANY mix of loop layouts.
//
// chb_fused.cuh / chb_fused4.cuh serve systems whose loops all take the RFFT shortcut and whose segments store segm_size
// samples.  This header generalises the same idea (parameters read once, result written once, transforms + pair interactions
// in shared memory) to the other layout classes of SURVEY.md Appendix A:
//   * GENERAL_SYM loops (params_to_pos_slice :5342-5437, partial_fft_to_pos_slice_{1,2}_sub :5041-5175 and their adjoints
//     pos_slice_to_partial_fft_{1,2}_sub :5070-5108 / :5178-5265, pos_slice_to_params :5509-5641), n_sub_fft 1 and 2;
//   * segm_store = segm_size + 1 (trapezoid end weights :6430-6460, :6631-6679, :7067-7133) and
//     Adjust_after_last_gen(_T) (:5267-5340);
//   * segments with a space rotation and / or time reversal (pos_slice_to_segmpos :5737-5813, segmpos_to_pos_slice_T :5896-5976).
//
// GENERAL loop, forward.  The reference transforms C = nnz*nppl REAL sequences p_c[n] (n < npr, zero padded to N = 2 npr) with an
// r2c of length N and contracts the result with the complex basis.  Here two real sequences share one complex sequence,
//     z_P[n] = p_2P[n] + i p_2P+1[n],   Z_P = DFT_N(z_P),   F_2P[m] = (Z_P[m] + conj Z_P[N-m]) / 2,   F_2P+1[m] = (Z_P[m] - conj Z_P[N-m]) / 2i
// (C is even for every planar layout), so a loop costs C/2 in-place complex FFTs of length N in shared memory -- the register
// butterflies of chb_fused.cuh -- followed by a "combine" pass: one thread per pair (m, N-m) forms F_c[m], applies the twiddle
// exp(-2 pi i m nnz_k / nint) and the basis, and writes the position of row m (and of the mirrored row 2 npr - m when
// n_sub_fft = 1) as ONE complex number x + i y into the slots Z_0[m] / Z_0[N-m] just consumed.  The pair interactions then run in
// place on those slots exactly as in chb_fused.cuh, and the adjoint is the mirror image: combine^T (basis^T, conjugate twiddle,
// Hermitian packing Y_P[m] = X_2P[m] + i X_2P+1[m], Y_P[N-m] = conj X_2P[m] + i conj X_2P+1[m]), inverse complex FFT, real / imaginary
// parts = the gradient with respect to p_2P / p_2P+1.
//
// segm_store = segm_size + 1: the first and last sample of consecutive segments share a row of the loop's slice.  Interior
// samples 1 .. segm_size - 1 touch every row exactly once (plain stores); the two end samples of all segments are evaluated by
// two threads into a small scratch and accumulated by one thread after the block barrier (deterministic, no atomics).
//
// Written as CTA programs over a context like the other headers, so tests/hostsim runs the same index arithmetic on the CPU.
#pragma once
#include "chb_fused.cuh"

#define CH5_MAX_SEGS 8
#define CH5_MAX_C 8     // real sequences per GENERAL loop (nnz * nppl), even
#define CH5_MAX_NNZ 4

struct Ch5Loop {
    ChfLoop f;        // FFT plan of length N (all sequences of the loop) + the RFFT fields of chb_fused.cuh
    int kind;         // CHB_GENERAL_SYM (0) or CHB_RFFT (1)
    int n_sub_fft, C, Cp, nnz, nppl, has_k0, tw_trivial;
    int nrows;        // rows of the slice produced by the transform: npr + 1 (n_sub_fft 2) or 2 npr
    int has_alg, alg_src;  // Adjust_after_last_gen: extra row 2 npr = alg_rot . row alg_src
    int seq_stride;   // cd slots between the sequences of the loop
    int co_in[CHB_MAX_NPPL], co_prefix[CHB_MAX_NPPL];
    int nnz_k[CH5_MAX_NNZ];
    int ikc[CH5_MAX_C];  // c / nppl: index into nnz_k of real sequence c (a table: no integer division in the kernels)
    double bre[2][CH5_MAX_C], bim[2][CH5_MAX_C];
    double alg_rot[4];
    cd wmir[CH5_MAX_NNZ];  // exp(+2 pi i nnz_k / ncoeff_min)
    const cd *gen_tw;      // [npr + 1][nnz] exp(-2 pi i m nnz_k / nint)
};
struct Ch5Sys {
    int nloop, nseg, T, segm_size, segm_store, law_kind;
    int tw_smem_off, tw_smem_len;  // twiddle region (cd slots behind the loop buffers), filled once per persistent CTA
    long long nparams, segsz;
    double law[4];
    Ch5Loop loops[CHF_MAX_LOOPS];
    ChfSeg segs[CH5_MAX_SEGS];
    ChfPair pairs[CH5_MAX_SEGS * CH5_MAX_SEGS];  // [src * CH5_MAX_SEGS + tgt]
};

// slot (cd index inside the loop's buffer) of row `row` of the loop's slice; row 2 npr (Adjust_after_last_gen) sits behind the sequences
CHB_HD int ch5_row_slot(const Ch5Loop &L, int row) {
    if (row >= L.f.N) return L.Cp * L.seq_stride;
    return chf_swz(L.kind == 1 ? row : L.f.rev[row], L.f.swm);
}
CHB_HD long long ch5_mom_index(const Ch5Loop &L, int n, int c) {
    if (n == 0 && L.has_k0 && c < L.nppl) {
        if (!L.co_in[c]) return -1;
        return L.f.mom_off + L.co_prefix[c];
    }
    return L.f.mom_off + (long long)n * L.C + c - L.f.nco_in;
}
CHB_HD double ch5_scale(const Ch5Loop &L, int n, int c, int *kout) {  // changevar_mom_pos (:4512-4566)
    const int k = n * L.f.ncm + L.nnz_k[L.ikc[c]];
    *kout = k;
    return k > 0 ? L.f.loopmul / (double)k : L.f.loopmul;
}


// ------------------------------------------------------------------------------------------------
// RFFT loop inside a mixed system: parameters -> spectrum Z[k] = X0[k] + i X1[k] in digit-reversed slots and back
// (same arithmetic as chf_load_loop / chf_store_loop of chb_fused.cuh, scalar accesses: any alignment)
// ------------------------------------------------------------------------------------------------
template <class Grp>
CHB_D void ch5_rfft_load(Grp &grp, const ChfLoop &L, const double *vec_b, cd *buf, double *red, bool want_kin) {
    grp.par([&](int tid, int T) {
        double kin = 0.0;
        const double *base = vec_b + L.mom_off - L.nco_in;
        for (int k = tid; k < L.npr; k += T) {
            if (k == 0) {
                const double p0 = L.co0 >= 0 ? vec_b[L.mom_off + L.co0] : 0.0;
                const double p2 = L.co2 >= 0 ? vec_b[L.mom_off + L.co2] : 0.0;
                buf[chf_swz(L.rev[0], L.swm)] = mk(p0 * L.loopmul, p2 * L.loopmul);  // the c2r ignores Im of the DC term
                buf[chf_swz(L.rev[L.npr], L.swm)] = mk(0.0, 0.0);                    // Nyquist row = zero padding
            } else {
                const double *p = base + 4LL * k;
                kin += (p[0] * p[0] + p[1] * p[1]) + (p[2] * p[2] + p[3] * p[3]);
                const double sc = L.loopmul / (double)((long long)k * L.ncm);
                buf[chf_swz(L.rev[k], L.swm)] = mk((p[0] - p[3]) * sc, (p[1] + p[2]) * sc);
                buf[chf_swz(L.rev[L.N - k], L.swm)] = mk((p[0] + p[3]) * sc, (p[2] - p[1]) * sc);
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}
template <class Grp>
CHB_D void ch5_rfft_store(Grp &grp, const ChfLoop &L, const cd *buf, const double *kin_b, double *out_b, double *red,
                           bool want_kin) {
    grp.par([&](int tid, int T) {
        double kin = 0.0;
        const long long off = L.mom_off - L.nco_in;
        for (int k = tid; k < L.npr; k += T) {
            const cd A = buf[chf_swz(L.rev[k], L.swm)];
            if (k == 0) {
                for (int c = 0; c < 4 - L.nco_in; ++c) out_b[L.mom_off + c] = 0.0;
                if (L.co0 >= 0) out_b[L.mom_off + L.co0] = A.x * L.loopmul;
                if (L.co2 >= 0) out_b[L.mom_off + L.co2] = A.y * L.loopmul;
            } else {
                const cd Bm = buf[chf_swz(L.rev[L.N - k], L.swm)];
                const double sc = L.loopmul / (double)((long long)k * L.ncm);
                double v[4] = {(A.x + Bm.x) * sc, (A.y - Bm.y) * sc, (A.y + Bm.y) * sc, (Bm.x - A.x) * sc};
                if (kin_b) {
                    const double *p = kin_b + off + 4LL * k;
                    kin += (p[0] * p[0] + p[1] * p[1]) + (p[2] * p[2] + p[3] * p[3]);
                    v[0] += p[0]; v[1] += p[1]; v[2] += p[2]; v[3] += p[3];
                }
                double *o = out_b + off + 4LL * k;
                o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}

// ------------------------------------------------------------------------------------------------
// GENERAL loop: parameters -> packed real sequences (natural order), forward FFTs, combine -> rows
// CP = C / 2 complex sequences at compile time (1: dihedral classes, 2: k-fold / forced general layouts; 0: run-time count)
// ------------------------------------------------------------------------------------------------
template <int CP>
struct Ch5Basis {  // the loop's basis and mirror factors in registers for the combine passes
    double bre[2][CP > 0 ? 2 * CP : CH5_MAX_C], bim[2][CP > 0 ? 2 * CP : CH5_MAX_C];
    cd wmir[CP > 0 ? 2 * CP : CH5_MAX_C];
    int ik[CP > 0 ? 2 * CP : CH5_MAX_C];
    CHB_HD void init(const Ch5Loop &L) {
        const int C = CP > 0 ? 2 * CP : L.C;
#pragma unroll
        for (int c = 0; c < (CP > 0 ? 2 * CP : CH5_MAX_C); ++c) {
            if (c >= C) break;
            ik[c] = L.ikc[c];
            wmir[c] = L.wmir[ik[c]];
#pragma unroll
            for (int g = 0; g < 2; ++g) { bre[g][c] = L.bre[g][c]; bim[g][c] = L.bim[g][c]; }
        }
    }
};

// one row n >= 1 of the parameter array: C values at rows[n C ..], harmonic k_c = n ncm + nnz_k[ikc[c]] > 0
template <int CP, class Grp>
CHB_D void ch5_general_load(Grp &grp, const Ch5Loop &L, const double *vec_b, cd *buf, double *red, bool want_kin) {
    const int npr = L.f.npr, C = CP > 0 ? 2 * CP : L.C, Cp = CP > 0 ? CP : L.Cp;
    grp.par([&](int tid, int T) {
        double kin = 0.0;
        const double *rows = vec_b + L.f.mom_off - L.f.nco_in;
        const double lm = L.f.loopmul;
        const int ncm = L.f.ncm, nppl = L.nppl, swm = L.f.swm, stride = L.seq_stride;
        for (int n = tid; n < npr; n += T) {
            const int sl = chf_swz(n, swm), sz = chf_swz(npr + n, swm);
            if (n == 0) {  // row 0: removed zero-frequency parameters, halved zero-frequency terms (:5389-5392)
                for (int P = 0; P < Cp; ++P) {
                    double v[2];
                    for (int h = 0; h < 2; ++h) {
                        const int c = 2 * P + h;
                        int kk;
                        const double sc = ch5_scale(L, 0, c, &kk);
                        const long long im = ch5_mom_index(L, 0, c);
                        const double p = im >= 0 ? vec_b[im] : 0.0;
                        if (kk > 0) kin += p * p;
                        v[h] = p * sc;
                        if (L.has_k0 && c < nppl) v[h] *= 0.5;
                    }
                    buf[P * stride + sl] = mk(v[0], v[1]);
                    buf[P * stride + sz] = mk(0.0, 0.0);
                }
                continue;
            }
            const double *r = rows + (long long)n * C;
            double sc_prev = 0.0;
            int ik_prev = -1;
#pragma unroll
            for (int P = 0; P < (CP > 0 ? CP : CH5_MAX_C / 2); ++P) {
                if (P >= Cp) break;
                double v[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = 2 * P + h, ik = L.ikc[c];
                    if (ik != ik_prev) {
                        sc_prev = lm * chf_rcp((double)(n * ncm + L.nnz_k[ik]));
                        ik_prev = ik;
                    }
                    const double p = r[c];
                    kin = fma(p, p, kin);
                    v[h] = p * sc_prev;
                }
                buf[P * stride + sl] = mk(v[0], v[1]);
                buf[P * stride + sz] = mk(0.0, 0.0);  // zero padding of the second half
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}

template <int CP, class Grp>
CHB_D void ch5_general_combine(Grp &grp, const Ch5Loop &L, cd *buf) {
    const int N = L.f.N, npr = L.f.npr, Cp = CP > 0 ? CP : L.Cp;
    grp.par([&](int tid, int T) {
        Ch5Basis<CP> Bs;
        Bs.init(L);
        const int swm = L.f.swm, stride = L.seq_stride, nnz = L.nnz;
        const int *rev = L.f.rev;
        const bool sub1 = L.n_sub_fft == 1, twt = L.tw_trivial != 0;
        for (int m = tid; m <= npr; m += T) {
            const int s1 = chf_swz(rev[m], swm);
            const int s2 = chf_swz(rev[m == 0 ? 0 : N - m], swm);
            cd x = mk(0.0, 0.0), xm = mk(0.0, 0.0);
            const bool mir = sub1 && m > 0 && m < npr;
#pragma unroll
            for (int P = 0; P < (CP > 0 ? CP : CH5_MAX_C / 2); ++P) {
                if (P >= Cp) break;
                const cd A = buf[P * stride + s1], Bq = buf[P * stride + s2];
                // F_2P = (A + conj B) / 2, F_2P+1 = (A - conj B) / 2i ; the factor 2 of x = 2 Re(...) cancels the halves
                cd F[2];
                F[0] = mk(A.x + Bq.x, A.y - Bq.y);
                F[1] = mk(A.y + Bq.y, Bq.x - A.x);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = 2 * P + h;
                    const cd W = twt ? F[h] : cmul(F[h], L.gen_tw[(long long)m * nnz + Bs.ik[c]]);
                    x.x = fma(Bs.bre[0][c], W.x, fma(Bs.bim[0][c], W.y, x.x));
                    x.y = fma(Bs.bre[1][c], W.x, fma(Bs.bim[1][c], W.y, x.y));
                    if (mir) {
                        const cd V = cmul(Bs.wmir[c], W);  // :5147-5175
                        xm.x = fma(Bs.bre[0][c], V.x, fma(-Bs.bim[0][c], V.y, xm.x));
                        xm.y = fma(Bs.bre[1][c], V.x, fma(-Bs.bim[1][c], V.y, xm.y));
                    }
                }
            }
            buf[s1] = x;
            if (mir) buf[s2] = xm;
        }
    });
}

template <int CP, class Grp>
CHB_D void ch5_general_combine_T(Grp &grp, const Ch5Loop &L, cd *buf) {
    const int N = L.f.N, npr = L.f.npr, Cp = CP > 0 ? CP : L.Cp;
    grp.par([&](int tid, int T) {
        Ch5Basis<CP> Bs;
        Bs.init(L);
        const int swm = L.f.swm, stride = L.seq_stride, nnz = L.nnz;
        const int *rev = L.f.rev;
        const bool sub1 = L.n_sub_fft == 1, twt = L.tw_trivial != 0;
        for (int m = tid; m <= npr; m += T) {
            const int s1 = chf_swz(rev[m], swm);
            const int s2 = chf_swz(rev[m == 0 ? 0 : N - m], swm);
            const bool end = (m == 0 || m == npr);
            const bool mir = sub1 && !end;
            const cd P1 = buf[s1];
            const cd P2 = mir ? buf[s2] : mk(0.0, 0.0);
            const double alpha = end ? 2.0 : 1.0;  // :5102-5106 / :5260-5263
#pragma unroll
            for (int P = 0; P < (CP > 0 ? CP : CH5_MAX_C / 2); ++P) {
                if (P >= Cp) break;
                cd X[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = 2 * P + h;
                    cd B = mk(Bs.bre[0][c] * P1.x + Bs.bre[1][c] * P1.y, Bs.bim[0][c] * P1.x + Bs.bim[1][c] * P1.y);
                    if (mir) {
                        const cd B2 = mk(Bs.bre[0][c] * P2.x + Bs.bre[1][c] * P2.y, Bs.bim[0][c] * P2.x + Bs.bim[1][c] * P2.y);
                        B = cadd(B, cmul(cconj(Bs.wmir[c]), cconj(B2)));
                    }
                    if (!twt) B = cmul(cconj(L.gen_tw[(long long)m * nnz + Bs.ik[c]]), B);
                    X[h] = cscale(B, alpha);
                }
                if (end) {  // the c2r ignores the imaginary parts of the first and the last coefficient
                    buf[P * stride + s1] = mk(X[0].x, X[1].x);
                } else {
                    buf[P * stride + s1] = mk(X[0].x - X[1].y, X[0].y + X[1].x);  // X_2P + i X_2P+1
                    buf[P * stride + s2] = mk(X[0].x + X[1].y, X[1].x - X[0].y);  // conj X_2P + i conj X_2P+1
                }
            }
        }
    });
}

template <int CP, class Grp>
CHB_D void ch5_general_store(Grp &grp, const Ch5Loop &L, const cd *buf, const double *kin_b, double *out_b, double *red,
                              bool want_kin) {
    const int npr = L.f.npr, C = CP > 0 ? 2 * CP : L.C, Cp = CP > 0 ? CP : L.Cp;
    grp.par([&](int tid, int T) {
        double kin = 0.0;
        const long long off = L.f.mom_off - L.f.nco_in;
        const double lm = L.f.loopmul;
        const int ncm = L.f.ncm, nppl = L.nppl, swm = L.f.swm, stride = L.seq_stride;
        for (int n = tid; n < npr; n += T) {
            const int sl = chf_swz(n, swm);
            if (n == 0) {
                for (int c = 0; c < C; ++c) {
                    const long long im = ch5_mom_index(L, 0, c);
                    if (im < 0) continue;
                    int kk;
                    const double sc = ch5_scale(L, 0, c, &kk);
                    double x = ((const double *)(buf + (c >> 1) * stride + sl))[c & 1];
                    if (L.has_k0 && c < nppl) x *= 0.5;  // :5634-5641
                    double val = x * sc;
                    if (kin_b && kk > 0) {
                        const double p = kin_b[im];
                        kin += p * p;
                        val += p;
                    }
                    out_b[im] = val;
                }
                continue;
            }
            double sc_prev = 0.0;
            int ik_prev = -1;
#pragma unroll
            for (int P = 0; P < (CP > 0 ? CP : CH5_MAX_C / 2); ++P) {
                if (P >= Cp) break;
                const cd y = buf[P * stride + sl];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = 2 * P + h, ik = L.ikc[c];
                    if (ik != ik_prev) {
                        sc_prev = lm * chf_rcp((double)(n * ncm + L.nnz_k[ik]));
                        ik_prev = ik;
                    }
                    double val = (h == 0 ? y.x : y.y) * sc_prev;
                    const long long im = off + (long long)n * C + c;
                    if (kin_b) {
                        const double p = kin_b[im];
                        kin = fma(p, p, kin);
                        val += p;
                    }
                    out_b[im] = val;
                }
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}

template <int CP, class Grp>
CHB_D void ch5_general_forward_t(Grp &grp, const Ch5Loop &L, const double *vec_b, cd *buf, double *red, bool want_kin, const cd *twsm) {
    ch5_general_load<CP>(grp, L, vec_b, buf, red, want_kin);
    for (int P = 0; P < L.Cp; ++P) chf_fft_time_to_spec(grp, buf + P * L.seq_stride, L.f, false, twsm, 0);
    ch5_general_combine<CP>(grp, L, buf);
}
template <class Grp>
CHB_D void ch5_general_forward(Grp &grp, const Ch5Loop &L, const double *vec_b, cd *buf, double *red, bool want_kin, const cd *twsm) {
    if (L.Cp == 1) ch5_general_forward_t<1>(grp, L, vec_b, buf, red, want_kin, twsm);
    else if (L.Cp == 2) ch5_general_forward_t<2>(grp, L, vec_b, buf, red, want_kin, twsm);
    else ch5_general_forward_t<0>(grp, L, vec_b, buf, red, want_kin, twsm);
}

// rows (gradient with respect to the slice, x + i y per row) -> Hermitian-packed spectra, inverse FFTs, parameter-space result
template <int CP, class Grp>
CHB_D void ch5_general_adjoint_t(Grp &grp, const Ch5Loop &L, cd *buf, const double *kin_b, double *out_b, double *red,
                                  bool want_kin, const cd *twsm) {
    ch5_general_combine_T<CP>(grp, L, buf);
    for (int P = 0; P < L.Cp; ++P) chf_fft_spec_to_time(grp, buf + P * L.seq_stride, L.f, false, twsm, 0);
    ch5_general_store<CP>(grp, L, buf, kin_b, out_b, red, want_kin);
}
template <class Grp>
CHB_D void ch5_general_adjoint(Grp &grp, const Ch5Loop &L, cd *buf, const double *kin_b, double *out_b, double *red,
                                bool want_kin, const cd *twsm) {
    if (L.Cp == 1) ch5_general_adjoint_t<1>(grp, L, buf, kin_b, out_b, red, want_kin, twsm);
    else if (L.Cp == 2) ch5_general_adjoint_t<2>(grp, L, buf, kin_b, out_b, red, want_kin, twsm);
    else ch5_general_adjoint_t<0>(grp, L, buf, kin_b, out_b, red, want_kin, twsm);
}

// ------------------------------------------------------------------------------------------------
// pair phase in place on the rows; PMODE 0 energy, 1 energy + gradient, 2 Hessian-vector, 3 segment positions only
// ------------------------------------------------------------------------------------------------
template <int NSEG, int PMODE>
struct Ch5Samp {
    cd g[NSEG];
    double e;
};
// evaluate sample i of every segment: positions from the rows (or seg_in), gradient / Hessian-vector per segment in out.g
template <int NSEG, int PMODE>
CHB_HD void ch5_eval_sample(const Ch5Sys &F, const cd *sm, bool have_fwd, const double *seg_in_b, double *seg_out_b, int i,
                            double wi, int *slot, Ch5Samp<NSEG, PMODE> &out) {
    const int size = F.segm_size, store = F.segm_store;
    const double inv_size = 1.0 / (double)size;
    cd x[NSEG], dx[NSEG];
#pragma unroll
    for (int s = 0; s < NSEG; ++s) {
        const ChfSeg &sg = F.segs[s];
        const Ch5Loop &L = F.loops[sg.loop];
        const int row = sg.row0 + (sg.timerev > 0 ? i : store - 1 - i);
        slot[s] = L.f.smem_off + ch5_row_slot(L, row);
        cd v = mk(0.0, 0.0);
        if (have_fwd) {
            const cd u = sm[slot[s]];
            v = sg.isid ? u : mk(u.x * sg.rot[0] + u.y * sg.rot[2], u.x * sg.rot[1] + u.y * sg.rot[3]);  // x @ R
        }
        if (PMODE == 2) {
            dx[s] = v;
            x[s] = ((const cd *)seg_in_b)[(long long)s * store + i];
        } else {
            x[s] = seg_in_b ? ((const cd *)seg_in_b)[(long long)s * store + i] : v;
            dx[s] = mk(0.0, 0.0);
        }
        if (seg_out_b) ((cd *)seg_out_b)[(long long)s * store + i] = v;
        out.g[s] = mk(0.0, 0.0);
    }
    double e = 0.0;
    if (PMODE != 3) {
#pragma unroll
        for (int s = 0; s < NSEG; ++s) {
#pragma unroll
            for (int t = 0; t < NSEG; ++t) {
                const ChfPair &P = F.pairs[s * CH5_MAX_SEGS + t];
                if (!P.on) continue;
                cd xs = x[s], ds = dx[s];
                if (!P.isid) {
                    xs = mk(P.rot[0] * x[s].x + P.rot[1] * x[s].y, P.rot[2] * x[s].x + P.rot[3] * x[s].y);
                    if (PMODE == 2) ds = mk(P.rot[0] * dx[s].x + P.rot[1] * dx[s].y, P.rot[2] * dx[s].x + P.rot[3] * dx[s].y);
                }
                const cd d = csub(xs, x[t]);
                const double xsq = fma(d.x, d.x, d.y * d.y);
                const double binfac = P.fac * inv_size * wi;
                cd gv = mk(0.0, 0.0);
                if (F.law_kind == 0) {  // gravity in closed form: V = -r, V' = r^3 / 2, V'' = -3 r^5 / 4, r = (x^2)^(-1/2)
                    const double r = chb_rsqrt_fast(xsq);
                    const double r2 = r * r;
                    const double c3 = binfac * (r2 * r);
                    if (PMODE <= 1) e = fma(-binfac, r, e);
                    if (PMODE == 1) {  // -2 binfac V' d
                        gv = mk(-c3 * d.x, -c3 * d.y);
                    } else if (PMODE == 2) {  // -2 binfac (2 V'' (d.dd) d + V' dd) = binfac r^3 (3 r^2 (d.dd) d - dd)
                        const cd dd = csub(ds, dx[t]);
                        const double k3 = 3.0 * r2 * fma(d.x, dd.x, d.y * dd.y);
                        gv = mk(c3 * fma(k3, d.x, -dd.x), c3 * fma(k3, d.y, -dd.y));
                    }
                } else {
                    double V, V1, V2;
                    chf_inter_law(F.law_kind, F.law, xsq, &V, &V1, &V2);
                    if (PMODE <= 1) e += V * binfac;
                    const double f = -2.0 * binfac;  // globalmul = -1 (:2843)
                    if (PMODE == 1) {
                        gv = mk(f * V1 * d.x, f * V1 * d.y);
                    } else if (PMODE == 2) {
                        const cd dd = csub(ds, dx[t]);
                        const double bb = 2.0 * V2 * (d.x * dd.x + d.y * dd.y);
                        gv = mk(f * (bb * d.x + V1 * dd.x), f * (bb * d.y + V1 * dd.y));
                    }
                }
                if (PMODE == 1 || PMODE == 2) {
                    if (P.isid) {
                        out.g[s] = cadd(out.g[s], gv);
                    } else {  // R^T g
                        out.g[s].x += P.rot[0] * gv.x + P.rot[2] * gv.y;
                        out.g[s].y += P.rot[1] * gv.x + P.rot[3] * gv.y;
                    }
                    out.g[t] = csub(out.g[t], gv);
                }
            }
        }
    }
    if (PMODE == 1 || PMODE == 2) {  // segmpos_to_pos_slice_T (:5896-5976): the row receives R . g
#pragma unroll
        for (int s = 0; s < NSEG; ++s) {
            const ChfSeg &sg = F.segs[s];
            if (!sg.isid)
                out.g[s] = mk(sg.rot[0] * out.g[s].x + sg.rot[1] * out.g[s].y, sg.rot[2] * out.g[s].x + sg.rot[3] * out.g[s].y);
        }
    }
    out.e = e;
}

// bnd: scratch of 2 * NSEG cd (gradients of the two end samples) + 2 * NSEG ints (their slots)
template <int NSEG, int PMODE, class Ctx>
CHB_D void ch5_pair(Ctx &ctx, const Ch5Sys &F, cd *sm, bool have_fwd, const double *seg_in_b, double *seg_out_b, double *red,
                     cd *bnd) {
    const bool gns = F.segm_store != F.segm_size;
    int *bslot = (int *)(bnd + 2 * NSEG);
    ctx.par([&](int tid, int nthr) {
        double e = 0.0;
        const int i0 = gns ? 1 : 0, i1 = gns ? F.segm_size - 1 : F.segm_size - 1;  // interior samples i0 .. i1
        for (int i = i0 + tid; i <= i1; i += nthr) {
            Ch5Samp<NSEG, PMODE> r;
            int slot[NSEG];
            ch5_eval_sample<NSEG, PMODE>(F, sm, have_fwd, seg_in_b, seg_out_b, i, 1.0, slot, r);
            e += r.e;
            if (PMODE == 1 || PMODE == 2) {
#pragma unroll
                for (int s = 0; s < NSEG; ++s) sm[slot[s]] = r.g[s];
            }
        }
        if (gns) {
            // end samples (trapezoid weight 1/2): the last threads of the last two warps take them (they own the fewest
            // interior samples), i = 0 and i = segm_size
            const int t1 = nthr - 1, t2 = nthr > 32 ? nthr - 33 : nthr - 1;
            for (int which = 0; which < 2; ++which) {
                if (tid != (which == 0 ? t1 : t2)) continue;
                const int i = which == 0 ? 0 : F.segm_size;
                Ch5Samp<NSEG, PMODE> r;
                int slot[NSEG];
                ch5_eval_sample<NSEG, PMODE>(F, sm, have_fwd, seg_in_b, seg_out_b, i, 0.5, slot, r);
                e += r.e;
#pragma unroll
                for (int s = 0; s < NSEG; ++s) {
                    bnd[which * NSEG + s] = r.g[s];
                    bslot[which * NSEG + s] = slot[s];
                }
            }
        }
        if (PMODE <= 1) red[tid] -= e;  // e accumulates +V: the action is kinetic - potential
    });
    if (PMODE == 1 || PMODE == 2) {
        bool any_alg = false;
        for (int il = 0; il < F.nloop; ++il) any_alg = any_alg || F.loops[il].has_alg;
        if (gns || any_alg) {
            ctx.par([&](int tid, int nthr) {
                (void)nthr;
                if (tid != 0) return;
                if (gns) {
                    // the rows shared by two consecutive segments: zero, then both contributions in a fixed order
                    for (int k = 0; k < 2 * NSEG; ++k) sm[bslot[k]] = mk(0.0, 0.0);
                    for (int k = 0; k < 2 * NSEG; ++k) sm[bslot[k]] = cadd(sm[bslot[k]], bnd[k]);
                }
                for (int il = 0; il < F.nloop; ++il) {  // Adjust_after_last_gen_T (:5323-5340): row alg_src += alg_rot^T . row 2 npr
                    const Ch5Loop &L = F.loops[il];
                    if (!L.has_alg) continue;
                    const cd t = sm[L.f.smem_off + ch5_row_slot(L, L.f.N)];
                    const int sa = L.f.smem_off + ch5_row_slot(L, L.alg_src);
                    sm[sa].x += L.alg_rot[0] * t.x + L.alg_rot[2] * t.y;
                    sm[sa].y += L.alg_rot[1] * t.x + L.alg_rot[3] * t.y;
                }
            });
        }
    }
}

// once per persistent CTA: the stage tables that fit behind the loop buffers are copied into shared memory
template <class Ctx>
CHB_D void cta_fused5_prologue(Ctx &ctx, const Ch5Sys &F, cd *sm) {
    if (F.tw_smem_len <= 0) return;
    ctx.par([&](int tid, int nthr) {
        for (int il = 0; il < F.nloop; ++il) {
            const ChfLoop &L = F.loops[il].f;
            for (int s = 0; s < L.nfac; ++s) {
                if (L.tw_soff[s] < 0) continue;
                cd *dst = sm + F.tw_smem_off + L.tw_soff[s];
                for (int i = tid; i < L.tw_len[s]; i += nthr) dst[i] = L.tw[s][i];
            }
        }
    });
}

// ------------------------------------------------------------------------------------------------
// the CTA program for one state
// ------------------------------------------------------------------------------------------------
template <int NSEG, int PMODE, class Ctx>
CHB_D void cta_fused5(Ctx &ctx, const Ch5Sys &F, const ChfArgs &A, long long b, long long b_pf, double *red, cd *sm, cd *bnd) {
    const double *vec_b = A.vec_in ? A.vec_in + b * F.nparams : (const double *)0;
    const double *kin_b = A.kin_src ? A.kin_src + b * F.nparams : (const double *)0;
    const double *seg_in_b = A.seg_in ? A.seg_in + b * F.segsz : (const double *)0;
    double *seg_out_b = A.seg_out ? A.seg_out + b * F.segsz : (double *)0;
    double *out_b = A.vec_out ? A.vec_out + b * F.nparams : (double *)0;
    const bool have_fwd = vec_b != (const double *)0;
    const cd *twsm = F.tw_smem_len > 0 ? sm + F.tw_smem_off : (const cd *)0;
    ctx.groups(F.nloop, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) { red[grp.g * T + tid] = 0.0; });
        const Ch5Loop &L = F.loops[grp.g];
        if (!have_fwd) return;
        cd *buf = sm + L.f.smem_off;
        if (L.kind == 1) {
            ch5_rfft_load(grp, L.f, vec_b, buf, red, A.kin_from == 1);
            chf_fft_spec_to_time(grp, buf, L.f, false, twsm, 0);
        } else {
            ch5_general_forward(grp, L, vec_b, buf, red, A.kin_from == 1, twsm);
        }
        if (b_pf < A.nstates)  // ask L2 for the rows of the state this CTA takes next (they arrive while this one is computed)
            grp.one([&]() {
                chf_prefetch_l2(A.vec_in + b_pf * F.nparams, L.f.mom_off - L.f.nco_in, (long long)L.f.npr * L.C,
                                A.vec_in + A.nstates * F.nparams);
            });
        if (L.has_alg)  // Adjust_after_last_gen (:5290-5302): row 2 npr = alg_rot . row alg_src
            grp.par([&](int tid, int T) {
                (void)T;
                if (tid != 0) return;
                const cd t = buf[ch5_row_slot(L, L.alg_src)];
                buf[ch5_row_slot(L, L.f.N)] = mk(L.alg_rot[0] * t.x + L.alg_rot[1] * t.y, L.alg_rot[2] * t.x + L.alg_rot[3] * t.y);
            });
    });
    ch5_pair<NSEG, PMODE>(ctx, F, sm, have_fwd, seg_in_b, seg_out_b, red, bnd);
    if (out_b) {
        ctx.groups(F.nloop, F.T, [&](auto &grp) {
            const Ch5Loop &L = F.loops[grp.g];
            cd *buf = sm + L.f.smem_off;
            if (L.kind == 1) {
                chf_fft_time_to_spec(grp, buf, L.f, false, twsm, 0);
                ch5_rfft_store(grp, L.f, buf, kin_b, out_b, red, A.kin_from == 2);
            } else {
                ch5_general_adjoint(grp, L, buf, kin_b, out_b, red, A.kin_from == 2, twsm);
            }
        });
    }
    if (A.scalar_out) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            if (tid == 0) A.scalar_out[b] = red[0];
        });
    }
}
