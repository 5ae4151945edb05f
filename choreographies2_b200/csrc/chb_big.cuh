// chb_big.cuh -- staged transforms for planar RFFT systems that do not fit the one-CTA-per-state kernels: many loops / segments
// (20B, 16C, 2C3C5C ...) or loops too long for one CTA's shared memory (64C at nint 16384: 262 KB per state; the deep refinement
// levels of a search, 3C at nint 49152: 786 KB).
//
// One CTA per (state, loop, residue q of the time index modulo R), R = 2, 4 or 8.  A loop of N = 2 npr samples is the complex
// sequence z[n] = x[n] + i y[n] (chb_fused.cuh); with M = N / R, w = exp(+2 pi i / N), omega = exp(+2 pi i / R),
//     z[R m + q] = sum_{k < M} ( w^(k q) sum_{r < R} Z[k + r M] omega^(r q) ) exp(+2 pi i k m / M),
// so each CTA folds the spectrum to M terms, runs a complex FFT of length M in shared memory (the register butterflies of
// chb_fused.cuh, stage tables in shared memory) and writes every R-th sample of the loop's segments; all R CTAs of a loop read
// all parameter rows (the repeats come from L2).  R is chosen per launch: the smallest that fits when the batch fills the GPU,
// a larger one when few states are live (tail of a search: parallelism inside the state instead of across states).
//
// The adjoint mirrors it on the spectral side: the CTA of residue q folds the segment gradient sum_r g[n + r M] omega^(-q r),
// multiplies by conj(w)^(n q), transforms, and holds Y[R k' + q].  A result row k needs Y[k] and Y[N - k]: for q = 0 and
// q = R / 2 the Hermitian partner has the same residue (packed x + i y transform, nothing exchanged).  The other residues come
// in pairs (q, R - q); there the two CTAs split by COORDINATE instead: CTA q transforms the x components for residue q, CTA
// R - q the y components for residue q.  A real sequence has Y[N - k] = conj(Y[k]), so one transform serves the rows k = q and
// k = R - q (mod R) of its coordinate: every row element is written by exactly one CTA (no atomics, deterministic).
// Twiddles w^j come from a two-level table (j = 128 hi + lo), exact entries, one complex product.
//
// Reference restated: params_to_pos_slice (RFFT branch) :5463-5506, pos_slice_to_segmpos :5737-5813, segmpos_to_pos_slice_T
// :5896-5976, pos_slice_to_params (RFFT branch, direction -1) :5645-5735, changevar_mom_pos(_T) :4489-4573 / :4746-4827,
// params_to_kin_nrg(_grad_daxpy) :4918-4973 of choreo/cython/_NBodySyst.pyx.
#pragma once
#include "chb_fused.cuh"

#define CHB_BIG_MAX_LOOPS 24
#define CHB_BIG_MAX_THREADS 512  // CHF_RED_BYTES holds 512 partial sums

struct ChbBigLoop {
    ChfLoop h;              // FFT plan of the length M = N / R (+ mom_off, nco_in, ncm, co0, co2, loopmul of the loop)
    const cd *t_hi, *t_lo;  // exp(-2 pi i 128 j / N), j <= N / 128;  exp(-2 pi i j / N), j < 128
    const int *slot_seg;    // [N / segm_size] generating segment stored in each slot of the slice
};
struct ChbBigSlot {         // one slot of a loop's slice, staged in shared memory by the prologue (the loops below then have no
    int s, timerev, isid, pad;  // dependent global loads: slot -> segment -> flags / rotation)
    double rot[4];
};
struct ChbBigSys {
    int nloop, T, segm_size, R;
    int tw_smem_len;        // cd slots of stage tables copied to shared memory (behind the buffer)
    int max_slots;          // ChbBigSlot entries behind the stage tables
    int stagger_ns, first_wave;  // > 0: CTAs of the first wave on odd SMs start this much later (see k_big_fwd)
    long long nparams, segsz;
    ChbBigLoop loops[CHB_BIG_MAX_LOOPS];
};

CHB_HD cd chb_big_root(const ChbBigLoop &L, int k) {  // exp(-2 pi i k / N), 0 <= k < N
    return cmul(L.t_hi[k >> 7], L.t_lo[k & 127]);
}
CHB_HD cd chb_root8(int j) {  // exp(+2 pi i j / 8)
    const double s = 0.70710678118654752440;
    j &= 7;
    const double c = (j == 0) ? 1.0 : (j == 4) ? -1.0 : (j == 2 || j == 6) ? 0.0 : (j == 1 || j == 7) ? s : -s;
    const double sn = (j == 2) ? 1.0 : (j == 6) ? -1.0 : (j == 0 || j == 4) ? 0.0 : (j < 4) ? s : -s;
    return mk(c, sn);
}
CHB_HD cd chb_big_omega(int R, int j) {  // exp(+2 pi i j / R), R in {2, 4, 8}
    return chb_root8(j * (8 / R));
}

// row k >= 1 of the loop's parameter array -> Z[k] and its Hermitian mirror Z[N - k]
CHB_HD void chb_big_row(const double *rows, int k, double sc, cd *Z, cd *Zm, double *kin) {
    const double *p = rows + 4LL * k;
    const double p0 = p[0], p1 = p[1], p2 = p[2], p3 = p[3];
    *kin += (p0 * p0 + p1 * p1) + (p2 * p2 + p3 * p3);
    *Z = mk((p0 - p3) * sc, (p1 + p2) * sc);
    *Zm = mk((p0 + p3) * sc, (p2 - p1) * sc);
}
CHB_HD cd chb_cfma(cd acc, cd a, cd b) {  // acc + a * b
    return mk(acc.x + (a.x * b.x - a.y * b.y), acc.y + (a.x * b.y + a.y * b.x));
}

// (quotient, remainder) of n / size kept incrementally: n advances by a fixed step whose quotient / remainder are dq / dr
struct ChbDivStep {
    int q, r;
};
CHB_HD void chb_div_advance(ChbDivStep &d, int dq, int dr, int size) {
    d.q += dq;
    d.r += dr;
    if (d.r >= size) {
        d.r -= size;
        ++d.q;
    }
}

// row (already in registers) -> Z[k] and its Hermitian mirror Z[N - k]
CHB_HD void chb_big_row_regs(const double *p, double sc, cd *Z, cd *Zm, double *kin) {
    *kin += (p[0] * p[0] + p[1] * p[1]) + (p[2] * p[2] + p[3] * p[3]);
    *Z = mk((p[0] - p[3]) * sc, (p[1] + p[2]) * sc);
    *Zm = mk((p[0] + p[3]) * sc, (p[2] - p[1]) * sc);
}

// The loops below have a handful of iterations per thread and one CTA per SM, so the latency of their global loads is what they
// cost (profiles/r02_big_transforms_ncu_summary.md: half of all stall samples were long-scoreboard waits).  Each loop therefore
// works in batches of CHB_BIG_U(R) independent iterations: all global loads of a batch are issued first, into registers, then
// the arithmetic and the shared-memory traffic follow (written out by hand: buf and the global arrays are generic pointers, the
// compiler may not move the loads of one iteration above the shared-memory stores of the previous one).
#define CHB_BIG_U(R) ((R) == 2 ? 4 : ((R) == 4 ? 2 : 1))

// params -> the samples n = R m + q of the loop's segments (+ kinetic energy of the loop from the CTA of residue 0)
template <int R, class Ctx>
CHB_D void cta_big_fwd(Ctx &ctx, const ChbBigSys &F, const ChbBigLoop &L, int q, const double *vec_b, double *segmpos_b,
                        double *kin_out, cd *buf, const cd *twsm, const ChbBigSlot *slots, double *red) {
    const ChfLoop &h = L.h;
    const int M = h.N;
    constexpr int H = R / 2, U = CHB_BIG_U(R);
    cd om[R];  // omega^(r q)
    for (int r = 0; r < R; ++r) om[r] = chb_big_omega(R, r * q);
    ctx.groups(1, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) {
            double kin = 0.0;
            const double *rows = vec_b + h.mom_off - h.nco_in;
            const double lm = h.loopmul / (double)h.ncm;
            if (tid == 0) {
                // k = 0: Z[0] (zero frequency), Z[r M] for 0 < r < R / 2, the Nyquist term is zero, mirrors above it
                const double c0 = h.co0 >= 0 ? vec_b[h.mom_off + h.co0] : 0.0;
                const double c2 = h.co2 >= 0 ? vec_b[h.mom_off + h.co2] : 0.0;
                cd acc = mk(c0 * h.loopmul, c2 * h.loopmul), Z, Zm;
#pragma unroll
                for (int r = 1; r < H; ++r) {
                    chb_big_row(rows, r * M, lm * chf_rcp((double)(r * M)), &Z, &Zm, &kin);
                    acc = chb_cfma(acc, Z, om[r]);
                    acc = chb_cfma(acc, Zm, om[R - r]);
                }
                buf[chf_swz(h.rev[0], h.swm)] = acc;
            }
            for (int kk0 = tid; kk0 <= M / 2; kk0 += U * T) {
                double pr[U][2 * H][4];
                int rv[U][2];
#pragma unroll
                for (int u = 0; u < U; ++u) {  // loads of the batch
                    const int kk = kk0 + u * T;
                    const bool on = kk >= 1 && kk <= M / 2;
                    const int k1 = on ? kk : 1, k2 = on ? M - kk : 1;
                    rv[u][0] = h.rev[k1];
                    rv[u][1] = h.rev[k2];
#pragma unroll
                    for (int r = 0; r < H; ++r) {
                        const double *p1 = rows + 4LL * (k1 + r * M), *p2 = rows + 4LL * (k2 + r * M);
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            pr[u][2 * r][c] = p1[c];
                            pr[u][2 * r + 1][c] = p2[c];
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int kk = kk0 + u * T;
                    if (kk < 1 || kk > M / 2) continue;
                    const int k1 = kk, k2 = M - kk;
                    // A[k] = w^(k q) ( sum_{r < R/2} Z[k + r M] omega^(r q) + sum_{r < R/2} mirror(row (M - k) + r M) omega^((R - 1 - r) q) )
                    cd a1 = mk(0.0, 0.0), a2 = mk(0.0, 0.0), Z, Zm;
#pragma unroll
                    for (int r = 0; r < H; ++r) {
                        chb_big_row_regs(pr[u][2 * r], lm * chf_rcp((double)(k1 + r * M)), &Z, &Zm, &kin);
                        a1 = chb_cfma(a1, Z, om[r]);
                        if (k2 != k1) {
                            a2 = chb_cfma(a2, Zm, om[R - 1 - r]);
                            chb_big_row_regs(pr[u][2 * r + 1], lm * chf_rcp((double)(k2 + r * M)), &Z, &Zm, &kin);
                            a2 = chb_cfma(a2, Z, om[r]);
                        }
                        a1 = chb_cfma(a1, Zm, om[R - 1 - r]);
                    }
                    if (q) a1 = cmul(a1, cconj(chb_big_root(L, q * k1)));
                    buf[chf_swz(rv[u][0], h.swm)] = a1;
                    if (k2 != k1) {
                        if (q) a2 = cmul(a2, cconj(chb_big_root(L, q * k2)));
                        buf[chf_swz(rv[u][1], h.swm)] = a2;
                    }
                }
            }
            red[tid] = 0.5 * kin;
        });
        chf_fft_spec_to_time(grp, buf, h, false, twsm, 0);
        if (segmpos_b) {
            grp.par([&](int tid, int T) {
                const int size = F.segm_size, step = R * T;
                const int dq = step / size, dr = step - dq * size;
                const int n0 = R * tid + q;
                ChbDivStep d = {n0 / size, n0 % size};
                for (int m = tid; m < M; m += T) {
                    const ChbBigSlot &sl = slots[d.q];
                    const cd u = buf[chf_swz(m, h.swm)];
                    const cd v = sl.isid ? u : mk(u.x * sl.rot[0] + u.y * sl.rot[2], u.x * sl.rot[1] + u.y * sl.rot[3]);  // x @ R
                    const int ii = sl.timerev > 0 ? d.r : size - 1 - d.r;
                    ((cd *)segmpos_b)[(long long)sl.s * size + ii] = v;
                    chb_div_advance(d, dq, dr, size);
                }
            });
        }
    });
    if (kin_out && q == 0) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            (void)nthr;
            if (tid == 0) *kin_out = red[0];
        });
    }
}

// segment gradient -> result rows of the loop (+ kin_src rows on k > 0: params_to_kin_nrg_grad_daxpy).  q = 0, R / 2: packed
// transform, rows k = q (mod R), all four entries; other q: one coordinate of the rows k = +-rho (mod R), rho = min(q, R - q)
template <int R, class Ctx>
CHB_D void cta_big_adj(Ctx &ctx, const ChbBigSys &F, const ChbBigLoop &L, int q, const double *seg_b, double *out_b,
                        const double *kin_b, cd *buf, const cd *twsm, const ChbBigSlot *slots) {
    const ChfLoop &h = L.h;
    const int M = h.N, N = R * M;
    constexpr int U = CHB_BIG_U(R), UO = 4;
    const bool packed = (q == 0) || (2 * q == R);
    const int rho = packed ? q : (q < R - q ? q : R - q);
    const int dim = (2 * q > R) ? 1 : 0;  // coordinate of the unpacked CTAs
    cd omc[R];  // omega^(-rho r)
    for (int r = 0; r < R; ++r) omc[r] = cconj(chb_big_omega(R, r * rho));
    ctx.groups(1, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) {
            const int size = F.segm_size;
            const int Tq = T / size, Tr = T - Tq * size, Mq = M / size, Mr = M - Mq * size;
            ChbDivStep d0 = {tid / size, tid % size};  // of n1, advanced by T per iteration
            for (int n10 = tid; n10 < M; n10 += U * T) {
                cd g[U][R];
                int sq[U][R];
#pragma unroll
                for (int u = 0; u < U; ++u) {  // loads of the batch
                    const bool on = n10 + u * T < M;
                    ChbDivStep d = d0;
#pragma unroll
                    for (int r = 0; r < R; ++r) {  // n = n1 + r M
                        const int qs = on ? d.q : 0;
                        const ChbBigSlot &sl = slots[qs];
                        const int ii = sl.timerev > 0 ? d.r : size - 1 - d.r;
                        sq[u][r] = qs;
                        g[u][r] = ((const cd *)seg_b)[(long long)sl.s * size + (on ? ii : 0)];
                        chb_div_advance(d, Mq, Mr, size);
                    }
                    chb_div_advance(d0, Tq, Tr, size);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int n1 = n10 + u * T;
                    if (n1 >= M) continue;
                    cd acc = mk(0.0, 0.0);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const ChbBigSlot &sl = slots[sq[u][r]];
                        const cd w = g[u][r];
                        const cd gg = sl.isid ? w : mk(sl.rot[0] * w.x + sl.rot[1] * w.y, sl.rot[2] * w.x + sl.rot[3] * w.y);  // R . g
                        if (packed) {
                            acc = chb_cfma(acc, gg, omc[r]);
                        } else {
                            const double gd = dim ? gg.y : gg.x;
                            acc = mk(acc.x + gd * omc[r].x, acc.y + gd * omc[r].y);
                        }
                    }
                    if (rho) acc = cmul(acc, chb_big_root(L, rho * n1));
                    buf[chf_swz(n1, h.swm)] = acc;
                }
            }
        });
        chf_fft_time_to_spec(grp, buf, h, false, twsm, 0);
        grp.par([&](int tid, int T) {
            const long long off = h.mom_off - h.nco_in;
            const double lm = h.loopmul / (double)h.ncm;
            if (packed) {
                for (int kp0 = tid; R * kp0 + q < N / 2; kp0 += UO * T) {
                    int ra[UO], rb[UO];
                    double kn[UO][4];
#pragma unroll
                    for (int u = 0; u < UO; ++u) {  // loads of the batch
                        const int kp = kp0 + u * T, k = R * kp + q;
                        const bool on = k < N / 2, row = on && k > 0;
                        ra[u] = h.rev[on ? kp : 0];
                        rb[u] = h.rev[row ? (q ? M - 1 - kp : M - kp) : 0];  // Y[N - k] = Y'[kq]
#pragma unroll
                        for (int c = 0; c < 4; ++c) kn[u][c] = (kin_b && row) ? kin_b[off + 4LL * k + c] : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < UO; ++u) {
                        const int kp = kp0 + u * T, k = R * kp + q;
                        if (k >= N / 2) continue;
                        const cd A = buf[chf_swz(ra[u], h.swm)];
                        if (k == 0) {
                            // f = 1, Y_0[0] = Re, the imaginary parts of the zero-frequency term get no gradient
                            for (int c = 0; c < 4 - h.nco_in; ++c) out_b[h.mom_off + c] = 0.0;
                            if (h.co0 >= 0) out_b[h.mom_off + h.co0] = A.x * h.loopmul;
                            if (h.co2 >= 0) out_b[h.mom_off + h.co2] = A.y * h.loopmul;
                            continue;
                        }
                        const cd Bm = buf[chf_swz(rb[u], h.swm)];
                        const double sc = lm * chf_rcp((double)k);
                        // Y_0 = (A + conj(Bm)) / 2, Y_1 = (A - conj(Bm)) / (2i); the adjoint carries a factor 2 for k > 0
                        double *o = out_b + off + 4LL * k;
                        o[0] = (A.x + Bm.x) * sc + kn[u][0];
                        o[1] = (A.y - Bm.y) * sc + kn[u][1];
                        o[2] = (A.y + Bm.y) * sc + kn[u][2];
                        o[3] = (Bm.x - A.x) * sc + kn[u][3];
                    }
                }
            } else {
                for (int kp0 = tid; kp0 < M; kp0 += UO * T) {
                    int ra[UO];
                    double kn[UO][2];
#pragma unroll
                    for (int u = 0; u < UO; ++u) {  // loads of the batch
                        const int kp = kp0 + u * T;
                        const bool on = kp < M;
                        const int k = R * (on ? kp : 0) + rho;  // never 0 or N / 2
                        const int row = k < N / 2 ? k : N - k;
                        ra[u] = h.rev[on ? kp : 0];
                        const long long at = off + 4LL * row + 2 * dim;
                        kn[u][0] = (kin_b && on) ? kin_b[at] : 0.0;
                        kn[u][1] = (kin_b && on) ? kin_b[at + 1] : 0.0;
                    }
#pragma unroll
                    for (int u = 0; u < UO; ++u) {
                        const int kp = kp0 + u * T;
                        if (kp >= M) continue;
                        const int k = R * kp + rho;
                        const cd A = buf[chf_swz(ra[u], h.swm)];
                        const int row = k < N / 2 ? k : N - k;           // Y_d[N - k] = conj(Y_d[k]) for the real sequence of coordinate d
                        const double im = k < N / 2 ? A.y : -A.y;
                        const double sc = 2.0 * lm * chf_rcp((double)row);
                        const long long at = off + 4LL * row + 2 * dim;
                        out_b[at] = A.x * sc + kn[u][0];
                        out_b[at + 1] = im * sc + kn[u][1];
                    }
                }
            }
        });
    });
}

// stage tables of the FFT and the slot table of the loop's slice -> shared memory
template <class Ctx>
CHB_D void cta_big_prologue(Ctx &ctx, const ChbBigSys &F, const ChbSys &S, const ChbBigLoop &L, cd *twsm_dst, ChbBigSlot *slots) {
    const ChfLoop &h = L.h;
    ctx.par([&](int tid, int nthr) {
        for (int s = 0; s < h.nfac; ++s) {
            if (h.tw_soff[s] < 0) continue;
            cd *dst = twsm_dst + h.tw_soff[s];
            for (int i = tid; i < h.tw_len[s]; i += nthr) dst[i] = h.tw[s][i];
        }
        const int nslot = (F.R * h.N) / F.segm_size;
        for (int j = tid; j < nslot; j += nthr) {
            const int s = L.slot_seg[j];
            const ChbSeg &sg = S.segs[s];
            ChbBigSlot sl;
            sl.s = s;
            sl.timerev = sg.timerev;
            sl.isid = sg.isid;
            sl.pad = 0;
            for (int c = 0; c < 4; ++c) sl.rot[c] = sg.rot[c];
            slots[j] = sl;
        }
    });
}
