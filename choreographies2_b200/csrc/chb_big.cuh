// chb_big.cuh -- staged transforms for planar RFFT systems that do not fit the one-CTA-per-state kernels: many loops / segments
// (20B, 16C, 2C3C5C ...) or loops too long for one CTA's shared memory (64C at nint 16384: 262 KB per state).
//
// One CTA per (state, loop, parity).  A loop of N = 2 npr samples is the complex sequence z[n] = x[n] + i y[n]
// (chb_fused.cuh); its transform is split by the PARITY of the time index,
//     z[2m + par] = sum_{k < N/2} (Z[k] + (-1)^par Z[k + N/2]) w^(k par) exp(+2 pi i k m / (N/2)),      w = exp(+2 pi i / N),
// so each CTA runs a complex FFT of length N/2 in shared memory (the register butterflies of chb_fused.cuh, stage tables in
// shared memory) and writes every other sample of the loop's segments; both CTAs of a pair read all parameter rows (the second
// read comes from L2).  The adjoint mirrors it on the spectral side: the CTA of parity par folds the segment gradient
// g[n] + (-1)^par g[n + N/2], multiplies by conj(w)^(n par), transforms, and owns the result rows k = 2 k' + par -- the
// Hermitian partner Y[N - k] has the same parity, so the split needs no exchange between the two CTAs.
// Twiddles w^k come from a two-level table (k = 128 hi + lo), exact entries, one complex product.
//
// Reference restated: params_to_pos_slice (RFFT branch) :5463-5506, pos_slice_to_segmpos :5737-5813, segmpos_to_pos_slice_T
// :5896-5976, pos_slice_to_params (RFFT branch, direction -1) :5645-5735, changevar_mom_pos(_T) :4489-4573 / :4746-4827,
// params_to_kin_nrg(_grad_daxpy) :4918-4973 of choreo/cython/_NBodySyst.pyx.
#pragma once
#include "chb_fused.cuh"

#define CHB_BIG_MAX_LOOPS 24

struct ChbBigLoop {
    ChfLoop h;              // FFT plan of the half length Nh = N / 2 = npr (+ mom_off, nco_in, ncm, co0, co2, loopmul of the loop)
    const cd *t_hi, *t_lo;  // exp(-2 pi i 128 j / N), j < ceil(Nh / 128);  exp(-2 pi i j / N), j < 128
    const int *slot_seg;    // [N / segm_size] generating segment stored in each slot of the slice
};
struct ChbBigSys {
    int nloop, T, segm_size, pad;
    int tw_smem_len;        // cd slots of stage tables copied to shared memory (behind the buffer)
    long long nparams, segsz;
    ChbBigLoop loops[CHB_BIG_MAX_LOOPS];
};

CHB_HD cd chb_big_root(const ChbBigLoop &L, int k) {  // exp(-2 pi i k / N)
    return cmul(L.t_hi[k >> 7], L.t_lo[k & 127]);
}

// row k >= 1 of the loop's parameter array -> Z[k] and its Hermitian mirror Z[N - k]
CHB_HD void chb_big_row(const double *rows, int k, double sc, cd *Z, cd *Zm, double *kin) {
    const double *p = rows + 4LL * k;
    const double p0 = p[0], p1 = p[1], p2 = p[2], p3 = p[3];
    *kin += (p0 * p0 + p1 * p1) + (p2 * p2 + p3 * p3);
    *Z = mk((p0 - p3) * sc, (p1 + p2) * sc);
    *Zm = mk((p0 + p3) * sc, (p2 - p1) * sc);
}

template <class Ctx>
CHB_D void cta_big_fwd(Ctx &ctx, const ChbBigSys &F, const ChbSys &S, const ChbBigLoop &L, int par, const double *vec_b,
                        double *segmpos_b, double *kin_out, cd *buf, const cd *twsm, double *red) {
    const ChfLoop &h = L.h;
    const int Nh = h.N;
    const double sgn = par ? -1.0 : 1.0;
    ctx.groups(1, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) {
            double kin = 0.0;
            const double *rows = vec_b + h.mom_off - h.nco_in;
            const double lm = h.loopmul / (double)h.ncm;
            for (int q = tid; q <= Nh / 2; q += T) {
                const int k1 = q, k2 = Nh - q;
                if (q == 0) {
                    const double c0 = h.co0 >= 0 ? vec_b[h.mom_off + h.co0] : 0.0;
                    const double c2 = h.co2 >= 0 ? vec_b[h.mom_off + h.co2] : 0.0;
                    buf[chf_swz(h.rev[0], h.swm)] = mk(c0 * h.loopmul, c2 * h.loopmul);  // Z[0] + (-1)^par Z[N/2], Nyquist row = 0
                    continue;
                }
                cd Z1, Zm1, Z2, Zm2;
                chb_big_row(rows, k1, lm * chf_rcp((double)k1), &Z1, &Zm1, &kin);
                if (k2 != k1) {
                    chb_big_row(rows, k2, lm * chf_rcp((double)k2), &Z2, &Zm2, &kin);
                } else {
                    Z2 = Z1;
                    Zm2 = Zm1;
                }
                // Z[k + Nh] = mirror of row Nh - k
                cd A1 = mk(Z1.x + sgn * Zm2.x, Z1.y + sgn * Zm2.y), A2 = mk(Z2.x + sgn * Zm1.x, Z2.y + sgn * Zm1.y);
                if (par) {
                    A1 = cmul(A1, cconj(chb_big_root(L, k1)));
                    A2 = cmul(A2, cconj(chb_big_root(L, k2)));
                }
                buf[chf_swz(h.rev[k1], h.swm)] = A1;
                if (k2 != k1) buf[chf_swz(h.rev[k2], h.swm)] = A2;
            }
            red[tid] = 0.5 * kin;
        });
        chf_fft_spec_to_time(grp, buf, h, false, twsm, 0);
        if (segmpos_b) {
            grp.par([&](int tid, int T) {
                const int size = F.segm_size;
                for (int m = tid; m < Nh; m += T) {
                    const int n = 2 * m + par;
                    const int q = n / size, i = n - q * size;
                    const int s = L.slot_seg[q];
                    const ChbSeg &sg = S.segs[s];
                    const cd u = buf[chf_swz(m, h.swm)];
                    const cd v = sg.isid ? u : mk(u.x * sg.rot[0] + u.y * sg.rot[2], u.x * sg.rot[1] + u.y * sg.rot[3]);  // x @ R
                    const int ii = sg.timerev > 0 ? i : size - 1 - i;
                    ((cd *)segmpos_b)[(long long)s * size + ii] = v;
                }
            });
        }
    });
    if (kin_out && par == 0) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            if (tid == 0) *kin_out = red[0];
        });
    }
}

// segment gradient -> result rows k = 2 k' + par of the loop (+ kin_src rows on k > 0: params_to_kin_nrg_grad_daxpy)
template <class Ctx>
CHB_D void cta_big_adj(Ctx &ctx, const ChbBigSys &F, const ChbSys &S, const ChbBigLoop &L, int par, const double *seg_b,
                        double *out_b, const double *kin_b, cd *buf, const cd *twsm) {
    const ChfLoop &h = L.h;
    const int Nh = h.N;
    const double sgn = par ? -1.0 : 1.0;
    ctx.groups(1, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) {
            const int size = F.segm_size;
            for (int n1 = tid; n1 < Nh; n1 += T) {
                cd g[2];
                for (int hh = 0; hh < 2; ++hh) {
                    const int n = n1 + hh * Nh;
                    const int q = n / size, i = n - q * size;
                    const int s = L.slot_seg[q];
                    const ChbSeg &sg = S.segs[s];
                    const int ii = sg.timerev > 0 ? i : size - 1 - i;
                    const cd u = ((const cd *)seg_b)[(long long)s * size + ii];
                    g[hh] = sg.isid ? u : mk(sg.rot[0] * u.x + sg.rot[1] * u.y, sg.rot[2] * u.x + sg.rot[3] * u.y);  // R . g
                }
                cd Bv = mk(g[0].x + sgn * g[1].x, g[0].y + sgn * g[1].y);
                if (par) Bv = cmul(Bv, chb_big_root(L, n1));
                buf[chf_swz(n1, h.swm)] = Bv;
            }
        });
        chf_fft_time_to_spec(grp, buf, h, false, twsm, 0);
        grp.par([&](int tid, int T) {
            const long long off = h.mom_off - h.nco_in;
            const double lm = h.loopmul / (double)h.ncm;
            for (int kp = tid; 2 * kp + par < Nh; kp += T) {
                const int k = 2 * kp + par;
                const cd A = buf[chf_swz(h.rev[kp], h.swm)];
                if (k == 0) {
                    // f = 1, Y_0[0] = Re, the imaginary parts of the zero-frequency term get no gradient
                    for (int c = 0; c < 4 - h.nco_in; ++c) out_b[h.mom_off + c] = 0.0;
                    if (h.co0 >= 0) out_b[h.mom_off + h.co0] = A.x * h.loopmul;
                    if (h.co2 >= 0) out_b[h.mom_off + h.co2] = A.y * h.loopmul;
                    continue;
                }
                const int kq = par ? Nh - 1 - kp : Nh - kp;  // Y[N - k] = Y'[kq]
                const cd Bm = buf[chf_swz(h.rev[kq], h.swm)];
                const double sc = lm * chf_rcp((double)k);
                // Y_0 = (A + conj(Bm)) / 2, Y_1 = (A - conj(Bm)) / (2i); the adjoint carries a factor 2 for k > 0
                double v[4] = {(A.x + Bm.x) * sc, (A.y - Bm.y) * sc, (A.y + Bm.y) * sc, (Bm.x - A.x) * sc};
                if (kin_b) {
                    const double *p = kin_b + off + 4LL * k;
                    v[0] += p[0]; v[1] += p[1]; v[2] += p[2]; v[3] += p[3];
                }
                double *o = out_b + off + 4LL * k;
                o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
            }
        });
    });
}

template <class Ctx>
CHB_D void cta_big_prologue(Ctx &ctx, const ChfLoop &h, cd *twsm_dst) {
    ctx.par([&](int tid, int nthr) {
        for (int s = 0; s < h.nfac; ++s) {
            if (h.tw_soff[s] < 0) continue;
            cd *dst = twsm_dst + h.tw_soff[s];
            for (int i = tid; i < h.tw_len[s]; i += nthr) dst[i] = h.tw[s][i];
        }
    });
}
