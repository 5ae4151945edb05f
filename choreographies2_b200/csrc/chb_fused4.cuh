// chb_fused4.cuh -- fused kernel, version 4: natural-order in-place transform fed by bulk asynchronous copies (TMA).
//
// Specialisation of the fused path of chb_fused.cuh for the layout of the headline workload:
// planar system, every loop an RFFT loop of N = 2 npr = 4096 = 16^3 samples covered by exactly one identity segment,
// gravity, one interaction per unordered pair.  Same arithmetic as version 3 (same butterflies, same Hermitian
// merge / split, same tensor-memory stash of the parameter rows for the kinetic term); what changes is where data lives:
//
//   * the parameter rows of a loop are copied global -> shared by cp.async.bulk (no per-thread global loads, no
//     lane-pair redistribution) and land in NATURAL order: row k = (p0 p1 | p2 p3) occupies complex slots 2k, 2k+1;
//   * every pass works IN PLACE on that order.  With k = k0 + 16 k1 + 256 q and n = t + 16 u + 256 v (t = t_lo + 2 t_hi):
//         load pass    radix 16 over q  -> t    slot(t_hi, k1, k0, t_lo) = 512 t_hi + 32 k1 + 2 k0 + t_lo   (own rows)
//         pass "k1"    radix 16 over k1 -> u    reads natural, writes slot(t_hi, u, k0 ^ (u & 3), t_lo)
//         pass "k0"    radix 16 over k0 -> v    in place on the swizzled slots
//     the XOR of the k0 / v digit with u & 3 makes the stride-2 accesses of pass "k0" bank-conflict free; pass "k1" is the
//     only place where a thread writes slots that another thread (of the same warp) reads, hence its __syncwarp;
//   * the adjoint runs the mirror image and leaves the result ROWS in natural order in the same buffer, from where
//     cp.async.bulk writes them to global memory while the threads already work on the second half of the butterflies;
//   * copies are issued per half of the butterflies ("chunk" A = pairs 0..63, B = pairs 64..127), each half of the
//     next state's rows has its own mbarrier, so the load pass starts on chunk A while chunk B is still in flight.
//
// Reference restated: the same functions as chb_fused.cuh (params_to_pos_slice :5463-5506, segm_pos_to_pot_nrg_grad /
// _hess :6472-6533 / :6899-6962, segmpos_to_params_T :6179-6251, changevar_mom_pos(_T) :4489-4573 / :4746-4827,
// params_to_kin_nrg(_grad_daxpy) :4918-4973 of choreo/cython/_NBodySyst.pyx).
#pragma once
#include "chb_fused.cuh"

#define CH4_N 4096
#define CH4_M 256
#define CH4_T 128  // threads per loop group

// The 16 lane pairs of warp w of a loop group own, in pass `it`, the butterflies klo in [a, a + 16) and their Hermitian partners
// M - klo, a = 64 it + 16 w (a = 0: butterflies 0..15, 128 and 241..255).  Their rows klo + 256 q form nruns(a) contiguous
// runs; run idx = (first row, number of rows).  Each run is one bulk copy, issued by lane idx of the warp.
CHB_HD int ch4_nruns(int a) { return a == 0 ? 24 : 16; }
CHB_HD void ch4_run(int a, int idx, int *r0, int *nr) {
    if (a == 0) {
        const int q = (idx * 11) >> 5, kind = idx - 3 * q;  // idx / 3 for idx < 24
        *r0 = 256 * q + (kind == 0 ? 0 : (kind == 1 ? 241 : 128));
        *nr = kind == 0 ? 16 : (kind == 1 ? 15 : 1);
    } else {
        const int q = idx >> 1;
        *r0 = 256 * q + ((idx & 1) ? CH4_M - a - 15 : a);
        *nr = 16;
    }
}
// doubles [d0, d0 + nd) of the loop's row array covered by a run (the removed zero-frequency parameters of row 0 are absent)
CHB_HD void ch4_run_doubles(const ChfLoop &L, int a, int idx, int *d0, int *nd) {
    int r0, nr;
    ch4_run(a, idx, &r0, &nr);
    *d0 = 4 * r0;
    *nd = 4 * nr;
    if (r0 == 0) { *d0 = L.nco_in; *nd -= L.nco_in; }
}

// global -> shared copies of the rows of (warp, it) of loop L of the state whose vector starts at vec_b; called by every lane
template <class Grp>
CHB_D void ch4_issue_loads(Grp &grp, const ChfLoop &L, const double *vec_b, cd *buf, int warp, int lane, int it) {
    const int a = 64 * it + 16 * warp;
    const double *rows = vec_b + L.mom_off - L.nco_in;
    double *dst = (double *)buf;
    if (lane == 0) grp.tma_expect(warp, it, 8192u - (a == 0 ? 8u * (unsigned)L.nco_in : 0u));
    if (lane < ch4_nruns(a)) {
        int d0, nd;
        ch4_run_doubles(L, a, lane, &d0, &nd);
        grp.tma_load(dst + d0, rows + d0, (unsigned)nd * 8u, warp, it);
    }
}
template <class Grp>
CHB_D void ch4_issue_stores(Grp &grp, const ChfLoop &L, double *out_b, const cd *buf, int warp, int lane, int it) {
    const int a = 64 * it + 16 * warp;
    double *rows = out_b + L.mom_off - L.nco_in;
    const double *src = (const double *)buf;
    if (lane < ch4_nruns(a)) {
        int d0, nd;
        ch4_run_doubles(L, a, lane, &d0, &nd);
        grp.tma_store(rows + d0, src + d0, (unsigned)nd * 8u);
    }
    grp.tma_store_commit();  // every lane commits one (possibly empty) group per pass: wait_group counts are uniform
}

// ------------------------------------------------------------------------------------------------
// load pass of one chunk: rows (shared, natural) -> Hermitian merge -> radix-16 butterfly over q, in place
// ------------------------------------------------------------------------------------------------
template <class Grp>
CHB_D void ch4_load_pass(Grp &grp, const ChfLoop &L, cd *buf, double *red, bool want_kin, int it) {
    constexpr int R = 16, H = 8, M = CH4_M;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    grp.template par_xchg<ChfXS<R>, H>(
        [&](int tid, int T, ChfXS<R> &s) {
            (void)T;
            const int pr = it * 64 + (tid >> 1);
            const int klo = chf_pair_klo(pr, tid & 1, M, M / 2);
            s.klo = klo;
            s.xm = pr == 0 ? 0 : 1;
            s.base = 2 * klo;
            double p[H][4];
            // a row is two 16-byte halves and the lanes of a pair sit 32 bytes apart (klo ascending / M - klo descending): the
            // even lane takes the low half first, the odd lane the high half first, so the two lanes of every pair hit different
            // bank groups in both requests (no 2-way conflict); a select puts the halves back in order
            const int hsw = tid & 1;
#pragma unroll
            for (int q = 0; q < H; ++q) {
                const cd u = buf[512 * q + 2 * klo + hsw], v = buf[512 * q + 2 * klo + (hsw ^ 1)];
                const cd lo = hsw ? v : u, hi = hsw ? u : v;
                p[q][0] = lo.x; p[q][1] = lo.y; p[q][2] = hi.x; p[q][3] = hi.y;
            }
            double c0 = 0.0, c2 = 0.0;
            if (klo == 0) {  // row 0 holds the zero-frequency parameters that are present, packed behind the removed ones
                const double *r0 = (const double *)buf;
                c0 = L.co0 >= 0 ? r0[L.nco_in + L.co0] : 0.0;
                c2 = L.co2 >= 0 ? r0[L.nco_in + L.co2] : 0.0;
                p[0][0] = p[0][1] = p[0][2] = p[0][3] = 0.0;
            }
            grp.template tmem_st<4 * H>(it * 8 * H, &p[0][0]);
            double kin = 0.0;
            const double dklo = (double)klo;
#pragma unroll
            for (int q = 0; q < H; ++q) {
                if (klo == 0 && q == 0) {
                    s.a[q] = mk(c0 * L.loopmul, c2 * L.loopmul);  // the c2r ignores Im of the zero-frequency term
                    s.x[q] = mk(0.0, 0.0);
                } else {
                    const double sc = lm * chf_rcp(fma((double)q, dM, dklo));
                    kin += (p[q][0] * p[q][0] + p[q][1] * p[q][1]) + (p[q][2] * p[q][2] + p[q][3] * p[q][3]);
                    // X0 = (p0 + i p1) sc, X1 = (p2 + i p3) sc ; Z[k] = X0 + i X1 ; Z[N-k] = conj(X0) + i conj(X1)
                    s.a[q] = mk((p[q][0] - p[q][3]) * sc, (p[q][1] + p[q][2]) * sc);
                    s.x[q] = mk((p[q][0] + p[q][3]) * sc, (p[q][2] - p[q][1]) * sc);
                }
            }
            if (want_kin) red[grp.g * grp.T + tid] += 0.5 * kin;
        },
        [&](int tid, int T, ChfXS<R> &s) {
            (void)tid; (void)T;
            const bool dc = s.klo == 0;
#pragma unroll
            for (int q = H; q < R; ++q) {
                // mirror of row (R-1-q) of the partner butterfly; butterfly 0 pairs q with R - q and holds the Nyquist zero
                const cd gen = s.x[R - 1 - q];
                const cd spc = q == H ? mk(0.0, 0.0) : s.x[(R - q) < H ? (R - q) : 0];
                s.a[q] = dc ? spc : gen;
            }
            ChfDft<R, +1>::run(s.a);
            cd *p = buf + s.base;
            const int hsw = tid & 1;  // same half-swap as the row reads: conflict-free stores
#pragma unroll
            for (int t = 0; t < R; t += 2) {
                p[512 * (t >> 1) + hsw] = hsw ? s.a[t + 1] : s.a[t];
                p[512 * (t >> 1) + (hsw ^ 1)] = hsw ? s.a[t] : s.a[t + 1];
            }
        });
}

// ------------------------------------------------------------------------------------------------
// store pass of one chunk: radix-16 butterfly over t -> Hermitian split + changevar_mom_pos_T + kinetic term -> rows, in place
// ------------------------------------------------------------------------------------------------
template <class Grp>
CHB_D void ch4_store_pass(Grp &grp, const ChfLoop &L, cd *buf, double *red, bool want_kin, bool add_kin, int it, int dbg) {
    constexpr int R = 16, H = 8, M = CH4_M;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    grp.template par_xchg<ChfXS<R>, H>(
        [&](int tid, int T, ChfXS<R> &s) {
            (void)T;
            const int pr = it * 64 + (tid >> 1);
            const int klo = chf_pair_klo(pr, tid & 1, M, M / 2);
            s.klo = klo;
            s.xm = pr == 0 ? 0 : 1;
            s.base = 2 * klo;
            const cd *p = buf + s.base;
            const int hsw = tid & 1;  // half-swap between the lanes of a pair: conflict-free reads (see ch4_load_pass)
#pragma unroll
            for (int t = 0; t < R; t += 2) {
                const cd u = p[512 * (t >> 1) + hsw], v = p[512 * (t >> 1) + (hsw ^ 1)];
                s.a[t] = hsw ? v : u;
                s.a[t + 1] = hsw ? u : v;
            }
            ChfDft<R, -1>::run(s.a);  // a[q] = Y[klo + q M]
            const bool dc = klo == 0;
#pragma unroll
            for (int t = 0; t < H; ++t) {
                // Y[N - (klo + t M)] is output R-1-t of butterfly M - klo (output R - t of butterfly 0)
                const cd gen = s.a[R - 1 - t];
                const cd spc = t == 0 ? mk(0.0, 0.0) : s.a[R - t];
                s.x[t] = dc ? spc : gen;
            }
        },
        [&](int tid, int T, ChfXS<R> &s) {
            (void)T;
            const int klo = s.klo;
            double kin = 0.0;
            const double dklo = (double)klo;
            cd *rows = buf + s.base;
            constexpr int HC = 4;  // rows per tensor-memory read: keeps the live registers below the cap
#pragma unroll
            for (int t0 = 0; t0 < H; t0 += HC) {
                double p[HC][4];
                if (!(dbg & 2048)) grp.template tmem_ld<4 * HC>(it * 8 * H + 8 * t0, &p[0][0]);
                else for (int u = 0; u < HC; ++u) p[u][0] = p[u][1] = p[u][2] = p[u][3] = 1.0;
#pragma unroll
                for (int u = 0; u < HC; ++u) {
                    const int t = t0 + u;
                    const cd A = s.a[t], Bm = s.x[t];
                    if (klo == 0 && t == 0) {
                        // f = 1, Y_0[0] = Re, imaginary parts of the zero-frequency term get no gradient
                        double *r0 = (double *)buf;
                        for (int c = 0; c < 4 - L.nco_in; ++c) r0[L.nco_in + c] = 0.0;
                        if (L.co0 >= 0) r0[L.nco_in + L.co0] = A.x * L.loopmul;
                        if (L.co2 >= 0) r0[L.nco_in + L.co2] = A.y * L.loopmul;
                    } else {
                        const double sc = lm * chf_rcp(fma((double)t, dM, dklo));
                        // Y_0 = (A + conj(Bm))/2, Y_1 = (A - conj(Bm))/(2i); the adjoint carries a factor 2 for k > 0
                        double v[4] = {(A.x + Bm.x) * sc, (A.y - Bm.y) * sc, (A.y + Bm.y) * sc, (Bm.x - A.x) * sc};
                        if (add_kin) {
                            if (want_kin) kin += (p[u][0] * p[u][0] + p[u][1] * p[u][1]) + (p[u][2] * p[u][2] + p[u][3] * p[u][3]);
                            v[0] += p[u][0]; v[1] += p[u][1]; v[2] += p[u][2]; v[3] += p[u][3];
                        }
                        const int hsw = tid & 1;
                        rows[512 * t + hsw] = hsw ? mk(v[2], v[3]) : mk(v[0], v[1]);
                        rows[512 * t + (hsw ^ 1)] = hsw ? mk(v[0], v[1]) : mk(v[2], v[3]);
                    }
                }
            }
            if (want_kin) red[grp.g * grp.T + tid] += 0.5 * kin;
            if (!(dbg & 1024)) grp.fence_async();  // the rows are read by the bulk copy engine next
        });
    grp.warp_sync();
}

// ------------------------------------------------------------------------------------------------
// middle passes
// ------------------------------------------------------------------------------------------------
struct Ch4S {
    cd a[16];
};
// radix 16 over k1 <-> u (length-256 stage, twiddles w256^(k1 t)): SIGN +1 reads natural / writes swizzled, -1 the reverse
template <int SIGN, class Grp>
CHB_D void ch4_pass_k1(Grp &grp, cd *buf, const cd *tw4) {
    for (int i = 0; i < 2; ++i) {
        grp.template par2<Ch4S>(
            [&](int tid, int T, Ch4S &s) {
                (void)T;
                const int t_lo = tid & 1, k0 = ((tid >> 1) & 7) + 8 * i, t_hi = tid >> 4;
                const cd *p = buf + ((t_hi << 9) | t_lo);
                const cd *wt = tw4 + (2 * t_hi + t_lo);  // w256^(q t) at wt[16 q], conjugated for the backward transform
                if (SIGN > 0) {
#pragma unroll
                    for (int q = 0; q < 16; ++q) s.a[q] = p[32 * q + 2 * k0];
#pragma unroll
                    for (int q = 1; q < 16; ++q) s.a[q] = cmul(s.a[q], cconj(wt[16 * q]));
                    ChfDft<16, SIGN>::run(s.a);
                } else {
#pragma unroll
                    for (int q = 0; q < 16; ++q) s.a[q] = p[32 * q + 2 * (k0 ^ (q & 3))];
                    ChfDft<16, SIGN>::run(s.a);
#pragma unroll
                    for (int q = 1; q < 16; ++q) s.a[q] = cmul(s.a[q], wt[16 * q]);
                }
            },
            [&](int tid, int T, Ch4S &s) {
                (void)T;
                const int t_lo = tid & 1, k0 = ((tid >> 1) & 7) + 8 * i, t_hi = tid >> 4;
                cd *p = buf + ((t_hi << 9) | t_lo);
                if (SIGN > 0) {
#pragma unroll
                    for (int u = 0; u < 16; ++u) p[32 * u + 2 * (k0 ^ (u & 3))] = s.a[u];
                } else {
#pragma unroll
                    for (int q = 0; q < 16; ++q) p[32 * q + 2 * k0] = s.a[q];
                }
            });
    }
    grp.sync();
}
// radix 16 over k0 <-> v (length-4096 stage, twiddles w4096^(k0 m), m = t + 16 u), in place on the swizzled slots.
// tw0p is the stage table permuted so that butterfly b = t_lo + 2 u + 32 t_hi reads entry b.
template <int SIGN, class Grp>
CHB_D void ch4_pass_k0(Grp &grp, cd *buf, const cd *tw0p) {
    grp.par([&](int tid, int T) {
        (void)T;  // T = 128: butterflies tid and tid + 128 (t_hi and t_hi + 4: 2048 slots apart, same u and t_lo)
        const int t_lo = tid & 1, u = (tid >> 1) & 15, t_hi = tid >> 5;
        cd *p = buf + ((t_hi << 9) | (u << 5) | t_lo);
        const int x = u & 3;
        // both butterflies' inputs are requested before the first one is computed: the second set of shared-memory reads
        // overlaps the first butterfly's FP64 work (12 warps per SM run the same phase, nothing else hides that latency)
        cd a[16], c[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) a[q] = p[2 * (q ^ x)];
#pragma unroll
        for (int q = 0; q < 16; ++q) c[q] = p[2048 + 2 * (q ^ x)];
        if (SIGN > 0) chf_twiddle<16, SIGN>(a, tw0p + tid, 256);
        ChfDft<16, SIGN>::run(a);
        if (SIGN < 0) chf_twiddle<16, SIGN>(a, tw0p + tid, 256);
#pragma unroll
        for (int t = 0; t < 16; ++t) p[2 * (t ^ x)] = a[t];
        if (SIGN > 0) chf_twiddle<16, SIGN>(c, tw0p + tid + 128, 256);
        ChfDft<16, SIGN>::run(c);
        if (SIGN < 0) chf_twiddle<16, SIGN>(c, tw0p + tid + 128, 256);
#pragma unroll
        for (int t = 0; t < 16; ++t) p[2048 + 2 * (t ^ x)] = c[t];
    });
}

// time index of slot s of the swizzled layout
CHB_HD int ch4_slot_to_n(int s) {
    const int t_lo = s & 1, u = (s >> 5) & 15, v = ((s >> 1) & 15) ^ (u & 3), t_hi = s >> 9;
    return t_lo + 2 * t_hi + 16 * u + 256 * v;
}

// ------------------------------------------------------------------------------------------------
// pair phase in place (gravity closed form, see chf_pair_fast): PMODE 1 energy + gradient, 2 Hessian-vector
// ------------------------------------------------------------------------------------------------
template <int NSEG, int PMODE, class Ctx>
CHB_D void ch4_pair(Ctx &ctx, const ChfSys &F, cd *sm, const double *seg_in_b, double *seg_out_b, double *red) {
    ctx.par([&](int tid, int nthr) {
        const double inv_size = 1.0 / (double)CH4_N;
        double w[NSEG][NSEG];
        int off[NSEG];
#pragma unroll
        for (int s = 0; s < NSEG; ++s) {
            off[s] = F.loops[F.segs[s].loop].smem_off;
#pragma unroll
            for (int t = 0; t < NSEG; ++t) w[s][t] = t > s ? F.pairs[s * CHF_MAX_SEGS + t].fac * inv_size : 0.0;
        }
        double e = 0.0;
        for (int sl = tid; sl < CH4_N; sl += nthr) {
            const int n = ch4_slot_to_n(sl);
            cd x[NSEG], dx[NSEG], g[NSEG];
#pragma unroll
            for (int s = 0; s < NSEG; ++s) {
                const cd v = sm[off[s] + sl];
                if (PMODE == 2) {
                    dx[s] = v;
                    x[s] = ((const cd *)seg_in_b)[(long long)s * CH4_N + n];
                } else {
                    x[s] = v;
                    dx[s] = mk(0.0, 0.0);
                }
                if (seg_out_b) ((cd *)seg_out_b)[(long long)s * CH4_N + n] = v;
                g[s] = mk(0.0, 0.0);
            }
#pragma unroll
            for (int s = 0; s < NSEG; ++s) {
#pragma unroll
                for (int t = s + 1; t < NSEG; ++t) {
                    const cd d = csub(x[s], x[t]);
                    const double r = chb_rsqrt_fast(fma(d.x, d.x, d.y * d.y));  // seed + one third-order step: < 2^-55 relative
                    if (PMODE == 1) e = fma(w[s][t], r, e);
                    const double r2 = r * r;
                    const double c = w[s][t] * (r2 * r);
                    if (PMODE == 1) {  // -2 w V' d = -w r^3 d on the source, + on the target
                        g[s].x = fma(-c, d.x, g[s].x);
                        g[s].y = fma(-c, d.y, g[s].y);
                        g[t].x = fma(c, d.x, g[t].x);
                        g[t].y = fma(c, d.y, g[t].y);
                    } else {  // -2 w (2 V'' (d.dd) d + V' dd) = w r^3 (3 r^2 (d.dd) d - dd)
                        const cd dd = csub(dx[s], dx[t]);
                        const double k3 = 3.0 * r2 * fma(d.x, dd.x, d.y * dd.y);
                        const double hx = c * fma(k3, d.x, -dd.x), hy = c * fma(k3, d.y, -dd.y);
                        g[s].x += hx;
                        g[s].y += hy;
                        g[t].x -= hx;
                        g[t].y -= hy;
                    }
                }
            }
#pragma unroll
            for (int s = 0; s < NSEG; ++s) sm[off[s] + sl] = g[s];
        }
        if (PMODE == 1) red[tid] += e;  // - potential
    });
}

// once per persistent CTA: twiddle tables into shared memory (stage 0 permuted for ch4_pass_k0, stage 1 as is)
template <class Ctx>
CHB_D void cta_fused4_prologue(Ctx &ctx, const ChfSys &F, cd *sm) {
    const ChfLoop &L = F.loops[F.tw_src_loop];
    ctx.par([&](int tid, int nthr) {
        cd *d0 = sm + F.tw_smem_off + L.tw_soff[0];
        for (int i = tid; i < 4 * 256; i += nthr) {
            const int e = i >> 8, b = i & 255;
            const int m = (b & 1) + 2 * (b >> 5) + 16 * ((b >> 1) & 15);
            d0[i] = L.tw[0][e * 256 + m];
        }
        cd *d1 = sm + F.tw_smem_off + L.tw_soff[1];
        for (int i = tid; i < 4 * 16; i += nthr) d1[i] = L.tw[1][i];
        // pass "k1": all 15 twiddles of a butterfly are read, not generated: w256^(q t) at [q][t] (exact table entries
        // exp(-2 pi i j 2^e / 256) of stage 1 combined: q t = sum of at most four powers of two times j <= 15 ... simpler and
        // as accurate: one correctly rounded sincospi per entry, the argument q t / 128 is exact)
        cd *d4 = sm + F.tw4_off;
        for (int i = tid; i < 256; i += nthr) {
            const int q = i >> 4, t = i & 15;
            double sn, cs;
#if defined(__CUDA_ARCH__)
            sincospi((double)(q * t) / 128.0, &sn, &cs);
#else
            sn = sin(3.14159265358979323846 * (double)(q * t) / 128.0);
            cs = cos(3.14159265358979323846 * (double)(q * t) / 128.0);
#endif
            d4[i] = mk(cs, -sn);
        }
    });
}

// first loads of a CTA (state b0); later states are requested at the end of the previous state's store pass
template <class Ctx>
CHB_D void cta_fused4_first_loads(Ctx &ctx, const ChfSys &F, const ChfArgs &A, long long b0, cd *sm) {
    if (b0 >= A.nstates) return;
    ctx.groups(F.nloop, F.T, [&](auto &grp) {
        const ChfLoop &L = F.loops[grp.g];
        grp.lanes([&](int warp, int lane) {
            ch4_issue_loads(grp, L, A.vec_in + b0 * F.nparams, sm + L.smem_off, warp, lane, 0);
            ch4_issue_loads(grp, L, A.vec_in + b0 * F.nparams, sm + L.smem_off, warp, lane, 1);
        });
    });
}

// ------------------------------------------------------------------------------------------------
// the CTA program for one state; parity = phase bit of the load barriers (number of states this CTA has done, mod 2)
// ------------------------------------------------------------------------------------------------
template <int NSEG, int PMODE, class Ctx>
CHB_D void cta_fused4(Ctx &ctx, const ChfSys &F, const ChfArgs &A, long long b, long long b_next, unsigned parity,
                       double *red, cd *sm) {
    const double *seg_in_b = A.seg_in ? A.seg_in + b * F.segsz : (const double *)0;
    double *seg_out_b = A.seg_out ? A.seg_out + b * F.segsz : (double *)0;
    double *out_b = A.vec_out + b * F.nparams;
    const cd *tw0p = sm + F.tw_smem_off + F.loops[F.tw_src_loop].tw_soff[0];
    const cd *tw1 = sm + F.tw4_off;
    ctx.groups(F.nloop, F.T, [&](auto &grp) {
        const ChfLoop &L = F.loops[grp.g];
        cd *buf = sm + L.smem_off;
        grp.par([&](int tid, int T) { red[grp.g * T + tid] = 0.0; });
        for (int it = 0; it < 2; ++it) {
            grp.tma_wait(it, parity);  // the rows of this warp's butterflies of pass `it`
            if (!(F.dbg & 1)) ch4_load_pass(grp, L, buf, red, A.kin_from == 1, it);
        }
        grp.sync();
        if (b_next < A.nstates && !(F.dbg & 512))
            grp.one([&]() {
                chf_prefetch_l2(A.vec_in + b_next * F.nparams, L.mom_off - L.nco_in, 4LL * L.npr,
                                A.vec_in + A.nstates * F.nparams);
            });
        if (!(F.dbg & 2)) {
            ch4_pass_k1<+1>(grp, buf, tw1);
            ch4_pass_k0<+1>(grp, buf, tw0p);
        }
    });
    if (!(F.dbg & 4)) ch4_pair<NSEG, PMODE>(ctx, F, sm, seg_in_b, seg_out_b, red);
    ctx.groups(F.nloop, F.T, [&](auto &grp) {
        const ChfLoop &L = F.loops[grp.g];
        cd *buf = sm + L.smem_off;
        if (!(F.dbg & 8)) {
            ch4_pass_k0<-1>(grp, buf, tw0p);
            ch4_pass_k1<-1>(grp, buf, tw1);
        }
        // every warp is on its own from here to the end of the next state's load pass: it writes the result rows of its
        // butterflies, hands them to the copy engine, and asks for the next state's rows as soon as the engine has read them
        for (int it = 0; it < 2; ++it) {
            if (!(F.dbg & 16)) ch4_store_pass(grp, L, buf, red, A.kin_from == 2, A.kin_src != (const double *)0, it, F.dbg);
            grp.lanes([&](int warp, int lane) { ch4_issue_stores(grp, L, out_b, buf, warp, lane, it); });
        }
        // the rows of pass 0 were handed over one store pass ago: refill them first, the load pass of the next state starts
        // on them while the rows of pass 1 are still draining
        for (int it = 0; it < 2; ++it) {
            grp.lanes([&](int warp, int lane) {
                (void)warp; (void)lane;
                if (it == 0) grp.tma_store_wait_read1(); else grp.tma_store_wait_read();  // the engine has read this lane's rows ...
            });
            grp.warp_sync();  // ... and those of the other lanes of the warp: the warp's rows of this pass may be refilled
            if (b_next < A.nstates)
                grp.lanes([&](int warp, int lane) { ch4_issue_loads(grp, L, A.vec_in + b_next * F.nparams, buf, warp, lane, it); });
        }
    });
    // the partial sums live in one of two buffers (state parity): the first warp adds them up and stores the action while
    // the other warps already start the next state -- no block-wide barrier behind the reduction
    if (A.scalar_out && !(F.dbg & 32)) ctx.reduce_store(red, A.scalar_out + b);
}
