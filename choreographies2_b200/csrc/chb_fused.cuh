// chb_fused.cuh -- the fused one-CTA-per-state path for planar systems whose loops all use the RFFT shortcut.
//
// One CTA owns one state.  Every loop's sampled path lives in shared memory as ONE complex sequence
// z[n] = x[n] + i y[n] (geodim == 2, so a segment position IS a complex number), which turns the reference's
// "c2r of length 2*npr, howmany geodim" (params_to_pos_slice, _NBodySyst.pyx:5463-5506) into a single in-place complex
// FFT of length N = 2*npr whose spectrum is Z[k] = X0[k] + i X1[k], Z[N-k] = conj(X0[k]) + i conj(X1[k]).
// The pair interactions (segm_pos_to_pot_nrg_grad / _hess, :6472-6533, :6899-6962) then run IN PLACE on the same
// buffers (one thread per time sample reads every segment's position and overwrites it with the segment gradient),
// and the adjoint transform (segmpos_to_params_T, :6179-6251) is the forward complex FFT of the same buffers followed
// by a Hermitian split.  Parameters are read from HBM once, the result is written once; segment positions, segment
// gradients and all FFT passes never leave the SM.
//
// Work split inside the CTA: one thread GROUP per loop (named barrier per group) for the load / FFT / store phases,
// so the loops' memory phases and FFT passes overlap instead of draining the whole SM at every pass; the pair phase
// uses all threads.  FFT passes are register butterflies of radix 16/8/4/2/3/5/7 on a XOR-swizzled buffer
// (conflict-free for every power-of-two stride); per-pass twiddles come from a compact table of base powers
// (w, w^2, w^4, w^8), the other powers are products.
//
// Like chb_core.cuh this header is written as CTA programs over a context (ctx.par / ctx.groups) so that the
// GPU-less build container can run the very same index arithmetic under tests/hostsim; the product is the nvcc build.
#pragma once
#include "chb_core.cuh"

#define CHF_MAX_LOOPS 6
#define CHF_MAX_SEGS 6
#define CHF_MAX_FAC 8
#define CHF_MAX_THREADS 384
#define CHF_RED_BYTES 4096  // CHF_MAX_THREADS doubles, rounded up

struct ChfLoop {
    int N, npr, nfac, smem_off;  // smem_off in cd slots
    int swm;                     // swizzle mask (15 when N % 16 == 0, else 0)
    int nco_in;                  // zero-frequency parameters removed from the parameter vector
    int ncm;                     // ncoeff_min: row k of the loop is harmonic k * ncm
    int co0, co2;                // index (relative to mom_off) of the DC parameters of x and y, or -1 when removed
    int fac[CHF_MAX_FAC];        // DIF radices, product N
    const int *rev;              // [N] slot of spectral index k after the DIF passes
    const cd *tw[CHF_MAX_FAC];   // DIF stage s: base twiddles exp(-2 pi i j 2^e / L_s), [Lp_s][nbase(r_s)]
    int tw_soff[CHF_MAX_FAC];    // >= 0: the stage's table is also kept in shared memory at this offset (cd) of the twiddle region
    int tw_len[CHF_MAX_FAC];     // table length in cd
    long long mom_off;
    double loopmul;  // 1/sqrt(loopnb*loopmass*4pi^2)
};
struct ChfSeg {
    int loop, row0, timerev, isid;
    double rot[4];
};
struct ChfPair {
    int on, isid;
    double fac;
    double rot[4];
};
struct ChfSys {
    int nloop, nseg, T, segm_size;
    int aligned16, law_kind;
    int tmem_cols;  // > 0: tensor-memory columns each CTA allocates for the parameter-row stash of ver 3 (power of two >= 32)
    int tmem_thread_cols;  // columns per thread (pair iterations * 4 R), threads of warp w use block w / 4 of their lane quarter
    int ver;        // 3: ver 2 + half-row coalescing of every global access + rows stashed in tensor memory for the kinetic term;
                    // 2: parameter load / result store fused into the first / last FFT pass (lane-pair exchange), 1: separate phases
    int tw_smem_off, tw_smem_len;  // twiddle region (cd slots after the loop buffers), filled once per persistent CTA
    int tw_src_loop;               // loop whose tables the region holds (loops of equal length share them)
    int dbg;        // profiling only (env CHB_FUSED_DBG): bit 0..5 skip load pass / inverse passes / pair / forward passes / store pass / reduce
    int pair_fast;  // gravity, every pair s < t interacts once with identity rotation, segments are plain slices
    int tw4_off;    // version 4: cd offset (from the loop buffers' base) of the full table w256^(q t), [q][t], of pass "k1"
    int stagger_ns; // version 4: CTA i starts i / gridDim * stagger_ns late so that the CTAs' copy phases do not coincide
    int pf_stride;  // > 0: CTA b asks L2 for the parameters of state b + pf_stride while it computes (co-resident CTA count)
    long long nparams, segsz;
    double law[4];
    ChfLoop loops[CHF_MAX_LOOPS];
    ChfSeg segs[CHF_MAX_SEGS];
    ChfPair pairs[CHF_MAX_SEGS * CHF_MAX_SEGS];  // [src * CHF_MAX_SEGS + tgt]
};
struct ChfArgs {
    const double *vec_in;   // [B][nparams] vector sent through the forward transform (params / dparams), or NULL
    const double *kin_src;  // [B][nparams] added to the adjoint result on k>0 entries (kinetic gradient / Hessian), or NULL
    const double *seg_in;   // [B][segsz]  cached segment positions used as x in the pair phase, or NULL
    double *seg_out;        // [B][segsz]  optional by-product: params_to_segmpos(vec_in)
    double *vec_out;        // [B][nparams] adjoint result (gradient / Hessian-vector), or NULL
    double *scalar_out;     // [B] kinetic - potential, or NULL
    int kin_from;           // 0 none, 1 kinetic energy from vec_in (load phase), 2 from kin_src (store phase)
    long long nstates;      // batch size (bound of the L2 prefetch)
    const unsigned char *skip;  // [B] or NULL: states with a non-zero byte are left out (their outputs are not written)
    const int *live_list;       // or NULL: the states to process (any order, each at most once) ...
    const int *live_count;      // ... and how many of them (device memory: written by the previous kernel of the stream)
};

// XOR swizzle of the low nibble with the next two nibbles: every access pattern of the passes (lanes along any one of the
// three hex digits of a 4096-point loop, the fused load / store passes included) and of the pair phase is conflict-free
CHB_HD int chf_swz(int i, int m) { return i ^ (((i >> 4) ^ (i >> 8)) & m); }
CHB_HD int chf_nbase(int r) { return r == 2 ? 1 : (r <= 4 ? 2 : (r <= 8 ? 3 : 4)); }

// ------------------------------------------------------------------------------------------------
// register butterflies: y_t = sum_q a_q exp(SIGN 2 pi i q t / R), natural order in and out
// ------------------------------------------------------------------------------------------------
template <int SIGN>
CHB_HD cd chf_muli(cd a) { return SIGN > 0 ? mk(-a.y, a.x) : mk(a.y, -a.x); }
// a * (c + i SIGN s)
template <int SIGN>
CHB_HD cd chf_mulcs(cd a, double c, double s) {
    return SIGN > 0 ? mk(a.x * c - a.y * s, a.x * s + a.y * c) : mk(a.x * c + a.y * s, a.y * c - a.x * s);
}
template <int SIGN>
CHB_HD void chf_dft2(cd &a0, cd &a1) {
    cd t = a0;
    a0 = cadd(t, a1);
    a1 = csub(t, a1);
}
template <int SIGN>
CHB_HD void chf_dft4(cd &a0, cd &a1, cd &a2, cd &a3) {
    cd t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), t3 = chf_muli<SIGN>(csub(a1, a3));
    a0 = cadd(t0, t2);
    a1 = cadd(t1, t3);
    a2 = csub(t0, t2);
    a3 = csub(t1, t3);
}
#define CHF_H 0.70710678118654752440
#define CHF_C1 0.92387953251128673848
#define CHF_S1 0.38268343236508978178

template <int R, int SIGN>
struct ChfDft;
template <int SIGN>
struct ChfDft<2, SIGN> {
    static CHB_HD void run(cd *a) { chf_dft2<SIGN>(a[0], a[1]); }
};
template <int SIGN>
struct ChfDft<4, SIGN> {
    static CHB_HD void run(cd *a) { chf_dft4<SIGN>(a[0], a[1], a[2], a[3]); }
};
template <int SIGN>
struct ChfDft<8, SIGN> {
    // q = q0 + 2 q1, t = t1 + 4 t0 : w8^(q t) = w8^(q0 t1) w2^(q0 t0) w4^(q1 t1)
    static CHB_HD void run(cd *a) {
        chf_dft4<SIGN>(a[0], a[2], a[4], a[6]);  // u[0][t1] in a[2 t1]
        chf_dft4<SIGN>(a[1], a[3], a[5], a[7]);  // u[1][t1] in a[1 + 2 t1]
        a[3] = chf_mulcs<SIGN>(a[3], CHF_H, CHF_H);
        a[5] = chf_muli<SIGN>(a[5]);
        a[7] = chf_mulcs<SIGN>(a[7], -CHF_H, CHF_H);
        cd y[8];
#pragma unroll
        for (int t1 = 0; t1 < 4; ++t1) {
            y[t1] = cadd(a[2 * t1], a[2 * t1 + 1]);
            y[t1 + 4] = csub(a[2 * t1], a[2 * t1 + 1]);
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) a[t] = y[t];
    }
};
template <int SIGN>
struct ChfDft<16, SIGN> {
    // q = q0 + 4 q1, t = t1 + 4 t0 : w16^(q t) = w16^(q0 t1) w4^(q0 t0) w4^(q1 t1)
    static CHB_HD void run(cd *a) {
#pragma unroll
        for (int q0 = 0; q0 < 4; ++q0) chf_dft4<SIGN>(a[q0], a[q0 + 4], a[q0 + 8], a[q0 + 12]);  // u[q0][t1] in a[q0 + 4 t1]
        a[1 + 4] = chf_mulcs<SIGN>(a[1 + 4], CHF_C1, CHF_S1);    // w16^1
        a[1 + 8] = chf_mulcs<SIGN>(a[1 + 8], CHF_H, CHF_H);      // w16^2
        a[1 + 12] = chf_mulcs<SIGN>(a[1 + 12], CHF_S1, CHF_C1);  // w16^3
        a[2 + 4] = chf_mulcs<SIGN>(a[2 + 4], CHF_H, CHF_H);      // w16^2
        a[2 + 8] = chf_muli<SIGN>(a[2 + 8]);                     // w16^4
        a[2 + 12] = chf_mulcs<SIGN>(a[2 + 12], -CHF_H, CHF_H);   // w16^6
        a[3 + 4] = chf_mulcs<SIGN>(a[3 + 4], CHF_S1, CHF_C1);    // w16^3
        a[3 + 8] = chf_mulcs<SIGN>(a[3 + 8], -CHF_H, CHF_H);     // w16^6
        a[3 + 12] = chf_mulcs<SIGN>(a[3 + 12], -CHF_C1, -CHF_S1);  // w16^9
#pragma unroll
        for (int t1 = 0; t1 < 4; ++t1) chf_dft4<SIGN>(a[4 * t1], a[4 * t1 + 1], a[4 * t1 + 2], a[4 * t1 + 3]);  // y[t1+4 t0] in a[t0+4 t1]
        cd y[16];
#pragma unroll
        for (int t1 = 0; t1 < 4; ++t1)
#pragma unroll
            for (int t0 = 0; t0 < 4; ++t0) y[t1 + 4 * t0] = a[t0 + 4 * t1];
#pragma unroll
        for (int t = 0; t < 16; ++t) a[t] = y[t];
    }
};
// cos / sin of 2 pi m / R for the odd radices
CHB_HD constexpr double chf_cosR(int R, int m) {
    return R == 3 ? (m == 0 ? 1.0 : -0.5)
         : R == 5 ? (m == 0 ? 1.0 : ((m == 1 || m == 4) ? 0.30901699437494742410 : -0.80901699437494742410))
                  : (m == 0 ? 1.0 : ((m == 1 || m == 6) ? 0.62348980185873353053
                                   : ((m == 2 || m == 5) ? -0.22252093395631440429 : -0.90096886790241912624)));
}
CHB_HD constexpr double chf_sinR(int R, int m) {
    return R == 3 ? (m == 0 ? 0.0 : (m == 1 ? 0.86602540378443864676 : -0.86602540378443864676))
         : R == 5 ? (m == 0 ? 0.0 : (m == 1 ? 0.95105651629515357212 : (m == 2 ? 0.58778525229247312917
                                   : (m == 3 ? -0.58778525229247312917 : -0.95105651629515357212))))
                  : (m == 0 ? 0.0 : (m == 1 ? 0.78183148246802980871 : (m == 2 ? 0.97492791218182360702
                                   : (m == 3 ? 0.43388373911755812048 : (m == 4 ? -0.43388373911755812048
                                   : (m == 5 ? -0.97492791218182360702 : -0.78183148246802980871))))));
}
template <int R, int SIGN>
struct ChfDftOdd {
    static CHB_HD void run(cd *a) {
        constexpr int H = (R - 1) / 2;
        cd t[H + 1], d[H + 1], y[R];
        y[0] = a[0];
#pragma unroll
        for (int p = 1; p <= H; ++p) {
            t[p] = cadd(a[p], a[R - p]);
            d[p] = csub(a[p], a[R - p]);
            y[0] = cadd(y[0], t[p]);
        }
#pragma unroll
        for (int k = 1; k <= H; ++k) {
            cd m = a[0], n = mk(0.0, 0.0);
#pragma unroll
            for (int p = 1; p <= H; ++p) {
                const double c = chf_cosR(R, (k * p) % R), s = chf_sinR(R, (k * p) % R);
                m.x += c * t[p].x;
                m.y += c * t[p].y;
                n.x += s * d[p].x;
                n.y += s * d[p].y;
            }
            n = chf_muli<SIGN>(n);
            y[k] = cadd(m, n);
            y[R - k] = csub(m, n);
        }
#pragma unroll
        for (int k = 0; k < R; ++k) a[k] = y[k];
    }
};
template <int SIGN>
struct ChfDft<3, SIGN> : ChfDftOdd<3, SIGN> {};
template <int SIGN>
struct ChfDft<5, SIGN> : ChfDftOdd<5, SIGN> {};
template <int SIGN>
struct ChfDft<7, SIGN> : ChfDftOdd<7, SIGN> {};

// a[t] *= w^t for t = 1..R-1, w^(2^e) read from the base table (conjugated for SIGN > 0)
template <int R, int SIGN>
CHB_HD void chf_twiddle(cd *a, const cd *base, int stride) {
    cd w[R];
#pragma unroll
    for (int e = 0; (1 << e) < R; ++e) {
        cd v = base[e * stride];  // table layout [e][j]: lanes read consecutive entries
        if (SIGN > 0) v.y = -v.y;
        w[1 << e] = v;
    }
#pragma unroll
    for (int t = 1; t < R; ++t) {
        if ((t & (t - 1)) != 0) {  // not a power of two: product of the top bit and the rest
            int hb = 1;
            while (hb * 2 <= t) hb *= 2;
            w[t] = cmul(w[hb], w[t - hb]);
        }
        a[t] = cmul(a[t], w[t]);
    }
}

// w[t] = w^t for t = 1..R-1 from the base powers w^(2^e) (conjugated for SIGN > 0)
template <int R, int SIGN>
CHB_HD void chf_twiddle_powers(cd *w, const cd *base, int stride) {
#pragma unroll
    for (int e = 0; (1 << e) < R; ++e) {
        cd v = base[e * stride];  // table layout [e][j]: lanes read consecutive entries
        if (SIGN > 0) v.y = -v.y;
        w[1 << e] = v;
    }
#pragma unroll
    for (int t = 3; t < R; ++t) {
        if ((t & (t - 1)) != 0) {
            int hb = 1;
            while (hb * 2 <= t) hb *= 2;
            w[t] = cmul(w[hb], w[t - hb]);
        }
    }
}

// Radix-16 passes of a 16-aligned length with the index arithmetic folded for the two strides every 16^k-type length
// uses (Lp = 256: the swizzle does not depend on q, the 16 slots are compile-time offsets from one address; Lp = 16:
// the swizzle reduces to j ^ q and both butterflies of a thread share their twiddles).  Same arithmetic as chf_pass.
template <int SIGN, bool DIT, class Grp>
CHB_D void chf_pass16_lp256(Grp &grp, cd *buf, int N, const cd *tw) {
    const int nbf = N / 16;
    grp.par([&](int tid, int T) {
        for (int b = tid; b < nbf; b += T) {
            const int j = b & 255;
            cd *p = buf + ((b >> 8) << 12) + (j & 0xF0);
            const int x = (j ^ (j >> 4)) & 15;  // slot nibble = x ^ q: bits 8..11 of the slot index are q
            cd a[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) a[q] = p[q * 256 + (x ^ q)];
            if (DIT) chf_twiddle<16, SIGN>(a, tw + j, 256);
            ChfDft<16, SIGN>::run(a);
            if (!DIT) chf_twiddle<16, SIGN>(a, tw + j, 256);
#pragma unroll
            for (int t = 0; t < 16; ++t) p[t * 256 + (x ^ t)] = a[t];
        }
    });
}
template <int SIGN, bool DIT, class Grp>
CHB_D void chf_pass16_lp16(Grp &grp, cd *buf, int N, const cd *tw) {
    const int nbf = N / 16;
    grp.par([&](int tid, int T) {
        const int j = tid & 15;  // T is a multiple of 32: every butterfly of this thread has the same j
        cd w[16];
        chf_twiddle_powers<16, SIGN>(w, tw + j, 16);
        for (int b = tid; b < nbf; b += T) {
            cd *p = buf + ((b >> 4) << 8);
            const int x = (j ^ (b >> 4)) & 15;  // bits 8..11 of the slot index are the block number
            cd a[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) a[q] = p[q * 16 + (x ^ q)];
            if (DIT) {
#pragma unroll
                for (int t = 1; t < 16; ++t) a[t] = cmul(a[t], w[t]);
            }
            ChfDft<16, SIGN>::run(a);
            if (!DIT) {
#pragma unroll
                for (int t = 1; t < 16; ++t) a[t] = cmul(a[t], w[t]);
            }
#pragma unroll
            for (int t = 0; t < 16; ++t) p[t * 16 + (x ^ t)] = a[t];
        }
    });
}

// one pass over the N/R butterflies of a stage with current length L (DIF: twiddle the outputs, DIT: the inputs)
template <int R, int SIGN, bool DIT, class Grp>
CHB_D void chf_pass(Grp &grp, cd *buf, int swm, int N, int L, const cd *tw) {
    const int Lp = L / R, nbf = N / R;
    grp.par([&](int tid, int T) {
        for (int b = tid; b < nbf; b += T) {
            const int blk = b / Lp, j = b - blk * Lp;
            const int base = blk * L + j;
            cd a[R];
#pragma unroll
            for (int q = 0; q < R; ++q) a[q] = buf[chf_swz(base + q * Lp, swm)];
            if (DIT && Lp > 1) chf_twiddle<R, SIGN>(a, tw + j, Lp);
            ChfDft<R, SIGN>::run(a);
            if (!DIT && Lp > 1) chf_twiddle<R, SIGN>(a, tw + j, Lp);
#pragma unroll
            for (int t = 0; t < R; ++t) buf[chf_swz(base + t * Lp, swm)] = a[t];
        }
    });
}
template <int SIGN, bool DIT, class Grp>
CHB_D void chf_pass_r(Grp &grp, cd *buf, int swm, int N, int L, int r, const cd *tw, int generic = 0) {
    if (r == 16 && swm == 15 && !generic) {
        if (L == 4096) { chf_pass16_lp256<SIGN, DIT>(grp, buf, N, tw); return; }
        if (L == 256) { chf_pass16_lp16<SIGN, DIT>(grp, buf, N, tw); return; }
    }
    switch (r) {
        case 16: chf_pass<16, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        case 8: chf_pass<8, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        case 4: chf_pass<4, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        case 2: chf_pass<2, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        case 3: chf_pass<3, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        case 5: chf_pass<5, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
        default: chf_pass<7, SIGN, DIT>(grp, buf, swm, N, L, tw); break;
    }
}
// backward (SIGN +1) transform, spectrum Z[k] in slot rev[k]  ->  z[n] in slot n  (decimation in time)
template <class Grp>
CHB_D void chf_fft_spec_to_time(Grp &grp, cd *buf, const ChfLoop &L, bool skip_first = false, const cd *twsm = (const cd *)0, int generic = 0) {
    int len = 1;
    for (int s = L.nfac - 1; s >= 0; --s) {
        len *= L.fac[s];
        if (skip_first && s == L.nfac - 1) continue;  // done by chf_load_pass
        const cd *tw = (twsm && L.tw_soff[s] >= 0) ? twsm + L.tw_soff[s] : L.tw[s];
        chf_pass_r<+1, true>(grp, buf, L.swm, L.N, len, L.fac[s], tw, generic);
    }
}
// forward (SIGN -1) transform, g[n] in slot n  ->  Y[k] in slot rev[k]  (decimation in frequency)
template <class Grp>
CHB_D void chf_fft_time_to_spec(Grp &grp, cd *buf, const ChfLoop &L, bool skip_last = false, const cd *twsm = (const cd *)0, int generic = 0) {
    int len = L.N;
    for (int s = 0; s < L.nfac - (skip_last ? 1 : 0); ++s) {  // the last pass is chf_store_pass when skipped
        const cd *tw = (twsm && L.tw_soff[s] >= 0) ? twsm + L.tw_soff[s] : L.tw[s];
        chf_pass_r<-1, false>(grp, buf, L.swm, L.N, len, L.fac[s], tw, generic);
        len /= L.fac[s];
    }
}

// ------------------------------------------------------------------------------------------------
// load phase of one loop: parameters -> spectrum Z in digit-reversed slots (changevar_mom_pos, :4489-4573, fused)
// ------------------------------------------------------------------------------------------------
CHB_HD void chf_load4(const double *p, int aligned16, double *v) {
#if defined(__CUDA_ARCH__)
    if (aligned16) {
        // streamed once: keep the small L1 (228 KB - shared memory) for the twiddle tables
        asm("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v[0]), "=d"(v[1]) : "l"(p));
        asm("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v[2]), "=d"(v[3]) : "l"(p + 2));
        return;
    }
#endif
    (void)aligned16;
    v[0] = p[0]; v[1] = p[1]; v[2] = p[2]; v[3] = p[3];
}
CHB_HD void chf_store4(double *p, int aligned16, const double *v) {
#if defined(__CUDA_ARCH__)
    if (aligned16) {
        *(double2 *)p = make_double2(v[0], v[1]);
        *(double2 *)(p + 2) = make_double2(v[2], v[3]);
        return;
    }
#endif
    (void)aligned16;
    p[0] = v[0]; p[1] = v[1]; p[2] = v[2]; p[3] = v[3];
}

template <class Grp>
CHB_D void chf_load_loop(Grp &grp, const ChfSys &F, const ChfLoop &L, const double *vec_b, cd *buf, double *red,
                          bool want_kin) {
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
                double p[4];
                chf_load4(base + 4LL * k, F.aligned16, p);
                kin += (p[0] * p[0] + p[1] * p[1]) + (p[2] * p[2] + p[3] * p[3]);
                const double sc = L.loopmul / (double)((long long)k * L.ncm);
                // X0 = (p0 + i p1) sc, X1 = (p2 + i p3) sc ; Z[k] = X0 + i X1 ; Z[N-k] = conj(X0) + i conj(X1)
                buf[chf_swz(L.rev[k], L.swm)] = mk((p[0] - p[3]) * sc, (p[1] + p[2]) * sc);
                buf[chf_swz(L.rev[L.N - k], L.swm)] = mk((p[0] + p[3]) * sc, (p[2] - p[1]) * sc);
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}

// store phase of one loop: spectrum Y of the segment gradient -> parameter-space result (changevar_mom_pos_T,
// :4746-4827, and params_to_kin_nrg_grad_daxpy, :4947-4973, fused)
template <class Grp>
CHB_D void chf_store_loop(Grp &grp, const ChfSys &F, const ChfLoop &L, const cd *buf, const double *kin_b, double *out_b,
                           double *red, bool want_kin) {
    grp.par([&](int tid, int T) {
        double kin = 0.0;
        const long long off = L.mom_off - L.nco_in;
        for (int k = tid; k < L.npr; k += T) {
            const cd A = buf[chf_swz(L.rev[k], L.swm)];
            if (k == 0) {
                // f = 1, Y_0[0] = Re, imaginary parts of the DC term get no gradient
                for (int c = 0; c < 4 - L.nco_in; ++c) out_b[L.mom_off + c] = 0.0;
                if (L.co0 >= 0) out_b[L.mom_off + L.co0] = A.x * L.loopmul;
                if (L.co2 >= 0) out_b[L.mom_off + L.co2] = A.y * L.loopmul;
            } else {
                const cd Bm = buf[chf_swz(L.rev[L.N - k], L.swm)];  // conj(Bm) = (Bm.x, -Bm.y)
                const double sc = L.loopmul / (double)((long long)k * L.ncm);
                // Y_0 = (A + conj(Bm))/2, Y_1 = (A - conj(Bm))/(2i); the adjoint carries a factor 2 for k > 0
                double v[4] = {(A.x + Bm.x) * sc, (A.y - Bm.y) * sc, (A.y + Bm.y) * sc, (Bm.x - A.x) * sc};
                if (kin_b) {
                    double p[4];
                    chf_load4(kin_b + off + 4LL * k, F.aligned16, p);
                    kin += (p[0] * p[0] + p[1] * p[1]) + (p[2] * p[2] + p[3] * p[3]);
                    v[0] += p[0]; v[1] += p[1]; v[2] += p[2]; v[3] += p[3];
                }
                chf_store4(out_b + off + 4LL * k, F.aligned16, v);
            }
        }
        if (want_kin) red[grp.g * T + tid] += 0.5 * kin;
    });
}

// ------------------------------------------------------------------------------------------------
// v2: load / store fused into the Lp = 1 pass (the last factor R of the plan, a power of two).
//
// That pass works on butterflies "klo": spectral indices k = klo + q M, q = 0..R-1, M = N/R, which live in the R
// consecutive slots rev[klo] .. rev[klo] + R - 1.  Rows k < npr = (R/2) M of the parameter vector are q < R/2; the upper
// half q >= R/2 is the Hermitian mirror Z[N-k] of the rows (M - klo) + (R-1-q) M, i.e. of the direct rows of butterfly
// M - klo.  Two adjacent lanes own the butterflies klo and M - klo, each reads its R/2 direct rows ONCE from global
// memory (R/2 independent 32-byte loads in flight per thread), computes Z[k] for itself and Z[N-k] for its partner,
// and the mirrors change lanes by warp shuffle.  klo = 0 (DC and Nyquist rows) and klo = M/2 are their own partners.
// The store pass is the transpose: every lane sends the upper half of its butterfly output to its partner and stores
// its R/2 direct rows with the Hermitian split, changevar_mom_pos_T and the kinetic term applied on the way out.
// ------------------------------------------------------------------------------------------------
template <int R>
struct ChfXS {
    cd a[R];
    cd x[(R / 2) > 0 ? (R / 2) : 1];  // exchanged with lane ^ xm
    int xm;
    int klo;  // -1: idle lane
    int base;
};

// 1 / d for a normal positive d: hardware seed (MUFU.RCP64H, about 20 bits) + two Newton steps; no special cases
CHB_HD double chf_rcp(double d) {
#if defined(__CUDA_ARCH__)
    double x0;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x0) : "d"(d));
    double e = fma(-d, x0, 1.0);
    const double x1 = fma(x0, e, x0);
    e = fma(-d, x1, 1.0);
    return fma(x1, e, x1);
#else
    return 1.0 / d;
#endif
}
// d^(-1/2) for a normal positive d: hardware seed (MUFU.RSQ64H) + two Newton steps
CHB_HD double chf_rsqrt(double d) {
#if defined(__CUDA_ARCH__)
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    double e = fma(-d * y, y, 1.0);
    y = fma(y * e, fma(e, 0.375, 0.5), y);
    e = fma(-d * y, y, 1.0);
    return fma(y * e, 0.5, y);
#else
    return 1.0 / sqrt(d);
#endif
}

CHB_HD int chf_pair_klo(int pr, int h, int M, int npairs) {
    if (pr >= npairs) return -1;
    if (pr == 0) return h == 0 ? 0 : ((M & 1) ? -1 : M / 2);
    return h == 0 ? pr : M - pr;
}

template <int R, class Grp>
CHB_D void chf_load_pass(Grp &grp, const ChfSys &F, const ChfLoop &L, const double *vec_b, cd *buf, double *red,
                          bool want_kin) {
    constexpr int H = R / 2;
    const int M = L.N / R, npairs = (M + 1) / 2, half = grp.T / 2;
    const double *rows = vec_b + L.mom_off - L.nco_in;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    for (int p0 = 0; p0 < npairs; p0 += half) {
        grp.template par_xchg<ChfXS<R>, H>(
            [&](int tid, int T, ChfXS<R> &s) {
                (void)T;
                const int pr = p0 + (tid >> 1);
                const int klo = chf_pair_klo(pr, tid & 1, M, npairs);
                s.klo = klo;
                s.xm = pr == 0 ? 0 : 1;
                double p[H][4];
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    p[q][0] = p[q][1] = p[q][2] = p[q][3] = 0.0;
                    if (klo > 0 || (klo == 0 && q > 0)) chf_load4(rows + 4LL * (klo + q * M), F.aligned16, p[q]);
                }
                s.base = klo >= 0 ? L.rev[klo] : 0;
                double kin = 0.0;
                const double dklo = (double)klo;
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    if (klo == 0 && q == 0) {
                        // DC row: the c2r ignores Im of the zero-frequency term; removed parameters are zero
                        const double c0 = L.co0 >= 0 ? vec_b[L.mom_off + L.co0] : 0.0;
                        const double c2 = L.co2 >= 0 ? vec_b[L.mom_off + L.co2] : 0.0;
                        s.a[q] = mk(c0 * L.loopmul, c2 * L.loopmul);
                        s.x[q] = mk(0.0, 0.0);
                    } else {
                        const double sc = klo >= 0 ? lm * chf_rcp(fma((double)q, dM, dklo)) : 0.0;
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
                if (s.klo < 0) return;
                const bool dc = s.klo == 0;
#pragma unroll
                for (int q = H; q < R; ++q) {
                    // mirror of row (R-1-q) of the partner butterfly; butterfly 0 pairs q with R - q and holds the Nyquist zero
                    const cd gen = s.x[R - 1 - q];
                    const cd spc = q == H ? mk(0.0, 0.0) : s.x[(R - q) < H ? (R - q) : 0];
                    s.a[q] = dc ? spc : gen;
                }
                ChfDft<R, +1>::run(s.a);
#pragma unroll
                for (int t = 0; t < R; ++t) buf[chf_swz(s.base + t, L.swm)] = s.a[t];
            });
    }
    grp.sync();
}

template <int R, class Grp>
CHB_D void chf_store_pass(Grp &grp, const ChfSys &F, const ChfLoop &L, const cd *buf, const double *kin_b, double *out_b,
                           double *red, bool want_kin) {
    constexpr int H = R / 2;
    const int M = L.N / R, npairs = (M + 1) / 2, half = grp.T / 2;
    const long long off = L.mom_off - L.nco_in;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    for (int p0 = 0; p0 < npairs; p0 += half) {
        grp.template par_xchg<ChfXS<R>, H>(
            [&](int tid, int T, ChfXS<R> &s) {
                (void)T;
                const int pr = p0 + (tid >> 1);
                const int klo = chf_pair_klo(pr, tid & 1, M, npairs);
                s.klo = klo;
                s.xm = pr == 0 ? 0 : 1;
                s.base = klo >= 0 ? L.rev[klo] : 0;
#pragma unroll
                for (int q = 0; q < R; ++q) s.a[q] = klo >= 0 ? buf[chf_swz(s.base + q, L.swm)] : mk(0.0, 0.0);
                ChfDft<R, -1>::run(s.a);  // a[t] = Y[klo + t M]
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
                if (klo < 0) return;
                double p[H][4];
                if (kin_b) {
#pragma unroll
                    for (int t = 0; t < H; ++t) {
                        p[t][0] = p[t][1] = p[t][2] = p[t][3] = 0.0;
                        if (klo > 0 || t > 0) chf_load4(kin_b + off + 4LL * (klo + t * M), F.aligned16, p[t]);
                    }
                }
                double kin = 0.0;
                const double dklo = (double)klo;
#pragma unroll
                for (int t = 0; t < H; ++t) {
                    const int k = klo + t * M;
                    const cd A = s.a[t], Bm = s.x[t];
                    if (klo == 0 && t == 0) {
                        // f = 1, Y_0[0] = Re, imaginary parts of the DC term get no gradient
                        for (int c = 0; c < 4 - L.nco_in; ++c) out_b[L.mom_off + c] = 0.0;
                        if (L.co0 >= 0) out_b[L.mom_off + L.co0] = A.x * L.loopmul;
                        if (L.co2 >= 0) out_b[L.mom_off + L.co2] = A.y * L.loopmul;
                    } else {
                        const double sc = lm * chf_rcp(fma((double)t, dM, dklo));
                        // Y_0 = (A + conj(Bm))/2, Y_1 = (A - conj(Bm))/(2i); the adjoint carries a factor 2 for k > 0
                        double v[4] = {(A.x + Bm.x) * sc, (A.y - Bm.y) * sc, (A.y + Bm.y) * sc, (Bm.x - A.x) * sc};
                        if (kin_b) {
                            kin += (p[t][0] * p[t][0] + p[t][1] * p[t][1]) + (p[t][2] * p[t][2] + p[t][3] * p[t][3]);
                            v[0] += p[t][0]; v[1] += p[t][1]; v[2] += p[t][2]; v[3] += p[t][3];
                        }
                        chf_store4(out_b + off + 4LL * k, F.aligned16, v);
                    }
                }
                if (want_kin) red[grp.g * grp.T + tid] += 0.5 * kin;
            });
    }
    grp.sync();
}

// ------------------------------------------------------------------------------------------------
// v3: same pass structure as v2, but (measured with tools/memphase_bench.cu: 32-byte rows at a 32-byte lane stride
// move at 2.8 TB/s, contiguous 16-byte lanes at 4.5 TB/s)
//   * every global access is a HALF row per lane: the two lanes of a pair share each row (lane h reads / writes
//     doubles 2h, 2h+1), so a warp-level instruction covers contiguous 512-byte runs; a second shuffle exchange moves
//     the halves to the lane that owns the row;
//   * the parameter rows a thread has loaded are parked in TENSOR MEMORY (tcgen05.st, 64 columns per butterfly) and
//     read back by the same thread in the store pass for the kinetic term (tcgen05.ld): the rows are neither re-read
//     from L2 nor kept in registers across the transforms.  Blackwell's 256 KB of TMEM per SM is otherwise idle here.
// ------------------------------------------------------------------------------------------------
CHB_HD void chf_pair_rows(int pr, int M, int npairs, int *ka, int *kb) {
    *ka = chf_pair_klo(pr, 0, M, npairs);
    *kb = chf_pair_klo(pr, 1, M, npairs);
}
CHB_HD void chf_load2(const double *p, int aligned16, cd *v) {
#if defined(__CUDA_ARCH__)
    if (aligned16) {
        asm("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v->x), "=d"(v->y) : "l"(p));
        return;
    }
#endif
    (void)aligned16;
    *v = mk(p[0], p[1]);
}
CHB_HD void chf_store2(double *p, int aligned16, cd v) {
#if defined(__CUDA_ARCH__)
    if (aligned16) {
        *(double2 *)p = v;
        return;
    }
#endif
    (void)aligned16;
    p[0] = v.x; p[1] = v.y;
}

template <int R, class Grp>
CHB_D void chf_load_pass3(Grp &grp, const ChfSys &F, const ChfLoop &L, const double *vec_b, cd *buf, double *red,
                           bool want_kin, bool stash) {
    constexpr int H = R / 2;
    const int M = L.N / R, npairs = (M + 1) / 2, half = grp.T / 2;
    const double *rows = vec_b + L.mom_off - L.nco_in;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    int it = 0;
    for (int p0 = 0; p0 < npairs; p0 += half, ++it) {
        grp.template par_xchg2<ChfXS<R>, H>(
            [&](int tid, int T, ChfXS<R> &s) {  // half rows of both butterflies of the pair
                (void)T;
                const int pr = p0 + (tid >> 1), h = tid & 1;
                int ka, kb;
                chf_pair_rows(pr, M, npairs, &ka, &kb);
                cd u[H], w[H];
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    u[q] = w[q] = mk(0.0, 0.0);
                    if (ka > 0 || (ka == 0 && q > 0)) chf_load2(rows + 4LL * (ka + q * M) + 2 * h, F.aligned16, &u[q]);
                    if (kb > 0) chf_load2(rows + 4LL * (kb + q * M) + 2 * h, F.aligned16, &w[q]);
                }
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    s.a[q] = h == 0 ? u[q] : w[q];  // half of the own row
                    s.x[q] = h == 0 ? w[q] : u[q];  // half of the partner's row
                }
                s.klo = h == 0 ? ka : kb;
                s.xm = 1;
            },
            [&](int tid, int T, ChfXS<R> &s) {  // whole own rows: stash, spectrum values for this butterfly and its mirror
                (void)T;
                const int pr = p0 + (tid >> 1), h = tid & 1;
                const int klo = s.klo;
                double p[H][4];
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    const cd lo = h == 0 ? s.a[q] : s.x[q], hi = h == 0 ? s.x[q] : s.a[q];
                    p[q][0] = lo.x; p[q][1] = lo.y; p[q][2] = hi.x; p[q][3] = hi.y;
                }
                if (stash) grp.template tmem_st<4 * H>(it * 8 * H, &p[0][0]);
                s.base = klo >= 0 ? L.rev[klo] : 0;
                double kin = 0.0;
                const double dklo = (double)klo;
#pragma unroll
                for (int q = 0; q < H; ++q) {
                    if (klo == 0 && q == 0) {
                        const double c0 = L.co0 >= 0 ? vec_b[L.mom_off + L.co0] : 0.0;
                        const double c2 = L.co2 >= 0 ? vec_b[L.mom_off + L.co2] : 0.0;
                        s.a[q] = mk(c0 * L.loopmul, c2 * L.loopmul);
                        s.x[q] = mk(0.0, 0.0);
                    } else {
                        const double sc = klo >= 0 ? lm * chf_rcp(fma((double)q, dM, dklo)) : 0.0;
                        kin += (p[q][0] * p[q][0] + p[q][1] * p[q][1]) + (p[q][2] * p[q][2] + p[q][3] * p[q][3]);
                        s.a[q] = mk((p[q][0] - p[q][3]) * sc, (p[q][1] + p[q][2]) * sc);
                        s.x[q] = mk((p[q][0] + p[q][3]) * sc, (p[q][2] - p[q][1]) * sc);
                    }
                }
                if (want_kin && klo >= 0) red[grp.g * grp.T + tid] += 0.5 * kin;
                s.xm = pr == 0 ? 0 : 1;
            },
            [&](int tid, int T, ChfXS<R> &s) {
                (void)tid; (void)T;
                if (s.klo < 0) return;
                const bool dc = s.klo == 0;
#pragma unroll
                for (int q = H; q < R; ++q) {
                    const cd gen = s.x[R - 1 - q];
                    const cd spc = q == H ? mk(0.0, 0.0) : s.x[(R - q) < H ? (R - q) : 0];
                    s.a[q] = dc ? spc : gen;
                }
                ChfDft<R, +1>::run(s.a);
#pragma unroll
                for (int t = 0; t < R; ++t) buf[chf_swz(s.base + t, L.swm)] = s.a[t];
            });
    }
    grp.sync();
}

template <int R, class Grp>
CHB_D void chf_store_pass3(Grp &grp, const ChfSys &F, const ChfLoop &L, const cd *buf, const double *kin_b, double *out_b,
                            double *red, bool want_kin, bool stash) {
    constexpr int H = R / 2;
    const int M = L.N / R, npairs = (M + 1) / 2, half = grp.T / 2;
    const long long off = L.mom_off - L.nco_in;
    const double lm = L.loopmul / (double)L.ncm, dM = (double)M;
    int it = 0;
    for (int p0 = 0; p0 < npairs; p0 += half, ++it) {
        grp.template par_xchg2<ChfXS<R>, H>(
            [&](int tid, int T, ChfXS<R> &s) {
                (void)T;
                const int pr = p0 + (tid >> 1);
                const int klo = chf_pair_klo(pr, tid & 1, M, npairs);
                s.klo = klo;
                s.xm = pr == 0 ? 0 : 1;
                s.base = klo >= 0 ? L.rev[klo] : 0;
#pragma unroll
                for (int q = 0; q < R; ++q) s.a[q] = klo >= 0 ? buf[chf_swz(s.base + q, L.swm)] : mk(0.0, 0.0);
                ChfDft<R, -1>::run(s.a);  // a[t] = Y[klo + t M]
                const bool dc = klo == 0;
#pragma unroll
                for (int t = 0; t < H; ++t) {
                    const cd gen = s.a[R - 1 - t];
                    const cd spc = t == 0 ? mk(0.0, 0.0) : s.a[R - t];
                    s.x[t] = dc ? spc : gen;
                }
            },
            [&](int tid, int T, ChfXS<R> &s) {  // result rows of the own butterfly; halves for the partner in x
                (void)T;
                const int h = tid & 1;
                const int klo = s.klo;
                double kin = 0.0;
                const double dklo = (double)klo;
                constexpr int HC = H < 4 ? H : 4;  // rows per tensor-memory read: keeps the live registers below the cap
#pragma unroll
                for (int t0 = 0; t0 < H; t0 += HC) {
                    double p[HC][4];
                    if (stash) {
                        grp.template tmem_ld<4 * HC>(it * 8 * H + 8 * t0, &p[0][0]);
                    } else {
#pragma unroll
                        for (int u = 0; u < HC; ++u) {
                            p[u][0] = p[u][1] = p[u][2] = p[u][3] = 0.0;
                            if (kin_b && (klo > 0 || (klo == 0 && t0 + u > 0)))
                                chf_load4(kin_b + off + 4LL * (klo + (t0 + u) * M), F.aligned16, p[u]);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < HC; ++u) {
                        const int t = t0 + u;
                        const cd A = s.a[t], Bm = s.x[t];
                        double v[4] = {0.0, 0.0, 0.0, 0.0};
                        if (klo == 0 && t == 0) {
                            for (int c = 0; c < 4 - L.nco_in; ++c) out_b[L.mom_off + c] = 0.0;
                            if (L.co0 >= 0) out_b[L.mom_off + L.co0] = A.x * L.loopmul;
                            if (L.co2 >= 0) out_b[L.mom_off + L.co2] = A.y * L.loopmul;
                        } else if (klo >= 0) {
                            const double sc = lm * chf_rcp(fma((double)t, dM, dklo));
                            v[0] = (A.x + Bm.x) * sc; v[1] = (A.y - Bm.y) * sc; v[2] = (A.y + Bm.y) * sc; v[3] = (Bm.x - A.x) * sc;
                            if (kin_b) {
                                kin += (p[u][0] * p[u][0] + p[u][1] * p[u][1]) + (p[u][2] * p[u][2] + p[u][3] * p[u][3]);
                                v[0] += p[u][0]; v[1] += p[u][1]; v[2] += p[u][2]; v[3] += p[u][3];
                            }
                        }
                        // lane 0 of the pair writes doubles 0,1 of both rows, lane 1 doubles 2,3
                        s.a[t] = h == 0 ? mk(v[0], v[1]) : mk(v[2], v[3]);
                        s.x[t] = h == 0 ? mk(v[2], v[3]) : mk(v[0], v[1]);
                    }
                }
                if (want_kin && klo >= 0) red[grp.g * grp.T + tid] += 0.5 * kin;
                s.xm = 1;
            },
            [&](int tid, int T, ChfXS<R> &s) {
                (void)T;
                const int pr = p0 + (tid >> 1), h = tid & 1;
                int ka, kb;
                chf_pair_rows(pr, M, npairs, &ka, &kb);
#pragma unroll
                for (int t = 0; t < H; ++t) {
                    const cd va = h == 0 ? s.a[t] : s.x[t];  // half h of row ka + t M
                    const cd vb = h == 0 ? s.x[t] : s.a[t];  // half h of row kb + t M
                    if (ka > 0 || (ka == 0 && t > 0)) chf_store2(out_b + off + 4LL * (ka + t * M) + 2 * h, F.aligned16, va);
                    if (kb > 0) chf_store2(out_b + off + 4LL * (kb + t * M) + 2 * h, F.aligned16, vb);
                }
            });
    }
    grp.sync();
}

template <class Grp>
CHB_D void chf_load_pass_r(Grp &grp, const ChfSys &F, const ChfLoop &L, const double *vec_b, cd *buf, double *red,
                            bool want_kin, bool stash) {
    if (F.ver >= 3) {
        switch (L.fac[L.nfac - 1]) {
            case 16: chf_load_pass3<16>(grp, F, L, vec_b, buf, red, want_kin, stash); break;
            case 8: chf_load_pass3<8>(grp, F, L, vec_b, buf, red, want_kin, stash); break;
            case 4: chf_load_pass3<4>(grp, F, L, vec_b, buf, red, want_kin, stash); break;
            default: chf_load_pass3<2>(grp, F, L, vec_b, buf, red, want_kin, stash); break;
        }
        return;
    }
#ifdef CHF_KEEP_OLD
    switch (L.fac[L.nfac - 1]) {
        case 16: chf_load_pass<16>(grp, F, L, vec_b, buf, red, want_kin); break;
        case 8: chf_load_pass<8>(grp, F, L, vec_b, buf, red, want_kin); break;
        case 4: chf_load_pass<4>(grp, F, L, vec_b, buf, red, want_kin); break;
        default: chf_load_pass<2>(grp, F, L, vec_b, buf, red, want_kin); break;
    }
#endif
}
template <class Grp>
CHB_D void chf_store_pass_r(Grp &grp, const ChfSys &F, const ChfLoop &L, const cd *buf, const double *kin_b,
                             double *out_b, double *red, bool want_kin, bool stash) {
    if (F.ver >= 3) {
        switch (L.fac[L.nfac - 1]) {
            case 16: chf_store_pass3<16>(grp, F, L, buf, kin_b, out_b, red, want_kin, stash); break;
            case 8: chf_store_pass3<8>(grp, F, L, buf, kin_b, out_b, red, want_kin, stash); break;
            case 4: chf_store_pass3<4>(grp, F, L, buf, kin_b, out_b, red, want_kin, stash); break;
            default: chf_store_pass3<2>(grp, F, L, buf, kin_b, out_b, red, want_kin, stash); break;
        }
        return;
    }
#ifdef CHF_KEEP_OLD
    switch (L.fac[L.nfac - 1]) {
        case 16: chf_store_pass<16>(grp, F, L, buf, kin_b, out_b, red, want_kin); break;
        case 8: chf_store_pass<8>(grp, F, L, buf, kin_b, out_b, red, want_kin); break;
        case 4: chf_store_pass<4>(grp, F, L, buf, kin_b, out_b, red, want_kin); break;
        default: chf_store_pass<2>(grp, F, L, buf, kin_b, out_b, red, want_kin); break;
    }
#endif
}

// ask L2 for the parameter rows of a later state (the CTA that will own it is not resident yet)
CHB_D void chf_prefetch_l2(const double *base, long long first, long long count, const double *arr_end) {
#if defined(__CUDA_ARCH__)
    unsigned long long a0 = ((unsigned long long)(base + first)) & ~15ULL;
    unsigned long long a1 = ((unsigned long long)(base + first + count)) & ~15ULL;
    const unsigned long long lim = ((unsigned long long)arr_end) & ~15ULL;
    if (a1 > lim) a1 = lim;
    while (a0 < a1) {
        const unsigned n = (unsigned)((a1 - a0) < 16384ULL ? (a1 - a0) : 16384ULL);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(n) : "memory");
        a0 += n;
    }
#else
    (void)base; (void)first; (void)count; (void)arr_end;
#endif
}

// ------------------------------------------------------------------------------------------------
// pair phase, in place: PMODE 0 energy, 1 energy + gradient, 2 Hessian-vector, 3 none (segment positions only)
// ------------------------------------------------------------------------------------------------
CHB_HD void chf_inter_law(int kind, const double *law, double xsq, double *V, double *V1, double *V2) {
    double a, b;
    if (kind == 0) {
#if defined(__CUDA_ARCH__)
        double r = rsqrt(xsq);
#else
        double r = 1.0 / sqrt(xsq);
#endif
        double r2 = r * r;
        a = r2 * r2 * r;  // xsq^-2.5 (inline_gravity_pot, :111-118)
        b = xsq * a;
        *V = -xsq * b;
        *V1 = 0.5 * b;
        *V2 = -0.75 * a;
    } else {
        a = law[3] * pow(xsq, law[2]);  // power_law_pot, :125-133
        b = xsq * a;
        *V = -xsq * b;
        *V1 = law[0] * b;
        *V2 = law[1] * a;
    }
}

template <int NSEG, int PMODE, class Ctx>
CHB_D void chf_pair(Ctx &ctx, const ChfSys &F, cd *sm, bool have_fwd, const double *seg_in_b, double *seg_out_b,
                     double *red) {
    ctx.par([&](int tid, int nthr) {
        double e = 0.0;
        const int size = F.segm_size;
        const double inv_size = 1.0 / (double)size;
        for (int i = tid; i < size; i += nthr) {
            cd x[NSEG], dx[NSEG], g[NSEG];
            int slot[NSEG];
#pragma unroll
            for (int s = 0; s < NSEG; ++s) {
                const ChfSeg &sg = F.segs[s];
                const ChfLoop &L = F.loops[sg.loop];
                const int row = sg.row0 + (sg.timerev > 0 ? i : size - 1 - i);
                slot[s] = L.smem_off + chf_swz(row, L.swm);
                cd v = mk(0.0, 0.0);
                if (have_fwd) {
                    const cd u = sm[slot[s]];
                    v = sg.isid ? u : mk(u.x * sg.rot[0] + u.y * sg.rot[2], u.x * sg.rot[1] + u.y * sg.rot[3]);  // x @ R
                }
                if (PMODE == 2) {
                    dx[s] = v;
                    x[s] = ((const cd *)seg_in_b)[(long long)s * size + i];
                } else {
                    x[s] = seg_in_b ? ((const cd *)seg_in_b)[(long long)s * size + i] : v;
                    dx[s] = mk(0.0, 0.0);
                }
                if (seg_out_b) ((cd *)seg_out_b)[(long long)s * size + i] = v;
                g[s] = mk(0.0, 0.0);
            }
            if (PMODE != 3) {
#pragma unroll
                for (int s = 0; s < NSEG; ++s) {
#pragma unroll
                    for (int t = 0; t < NSEG; ++t) {
                        const ChfPair &P = F.pairs[s * CHF_MAX_SEGS + t];
                        if (!P.on) continue;
                        cd xs = x[s], ds = dx[s];
                        if (!P.isid) {
                            xs = mk(P.rot[0] * x[s].x + P.rot[1] * x[s].y, P.rot[2] * x[s].x + P.rot[3] * x[s].y);
                            if (PMODE == 2)
                                ds = mk(P.rot[0] * dx[s].x + P.rot[1] * dx[s].y, P.rot[2] * dx[s].x + P.rot[3] * dx[s].y);
                        }
                        const cd d = csub(xs, x[t]);
                        const double xsq = d.x * d.x + d.y * d.y;
                        double V, V1, V2;
                        chf_inter_law(F.law_kind, F.law, xsq, &V, &V1, &V2);
                        const double binfac = P.fac * inv_size;
                        if (PMODE <= 1) e += V * binfac;
                        if (PMODE >= 1) {
                            const double f = -2.0 * binfac;  // globalmul = -1 (:2843)
                            cd gv;
                            if (PMODE == 1) {
                                gv = mk(f * V1 * d.x, f * V1 * d.y);
                            } else {
                                const cd dd = csub(ds, dx[t]);
                                const double bb = 2.0 * V2 * (d.x * dd.x + d.y * dd.y);
                                gv = mk(f * (bb * d.x + V1 * dd.x), f * (bb * d.y + V1 * dd.y));
                            }
                            if (P.isid) {
                                g[s] = cadd(g[s], gv);
                            } else {  // R^T g
                                g[s].x += P.rot[0] * gv.x + P.rot[2] * gv.y;
                                g[s].y += P.rot[1] * gv.x + P.rot[3] * gv.y;
                            }
                            g[t] = csub(g[t], gv);
                        }
                    }
                }
            }
            if (PMODE == 1 || PMODE == 2) {
#pragma unroll
                for (int s = 0; s < NSEG; ++s) {
                    const ChfSeg &sg = F.segs[s];
                    // segmpos_to_pos_slice_T (:5896-5976): pos_slice row += R . g
                    sm[slot[s]] = sg.isid ? g[s]
                                          : mk(sg.rot[0] * g[s].x + sg.rot[1] * g[s].y, sg.rot[2] * g[s].x + sg.rot[3] * g[s].y);
                }
            }
        }
        if (PMODE <= 1) red[tid] -= e;
    });
}

// Pair phase of the common layout (F.pair_fast): gravity, one interaction per unordered segment pair, no rotations.
// V = -r, V' = r^3 / 2, V'' = -3 r^5 / 4 with r = (x^2)^(-1/2) (inline_gravity_pot, :111-118, in closed form), so a pair
// sample costs one rsqrt and ~17 fp64 instructions; the interaction weights sit in registers for the whole phase.
template <int NSEG, int PMODE, class Ctx>
CHB_D void chf_pair_fast(Ctx &ctx, const ChfSys &F, cd *sm, bool have_fwd, const double *seg_in_b, double *seg_out_b,
                          double *red) {
    ctx.par([&](int tid, int nthr) {
        const int size = F.segm_size;
        const double inv_size = 1.0 / (double)size;
        double w[NSEG][NSEG];
        int off[NSEG], swm[NSEG];
#pragma unroll
        for (int s = 0; s < NSEG; ++s) {
            const ChfLoop &L = F.loops[F.segs[s].loop];
            off[s] = L.smem_off;
            swm[s] = L.swm;
#pragma unroll
            for (int t = 0; t < NSEG; ++t) w[s][t] = t > s ? F.pairs[s * CHF_MAX_SEGS + t].fac * inv_size : 0.0;
        }
        double e = 0.0;
        for (int i = tid; i < size; i += nthr) {
            cd x[NSEG], dx[NSEG], g[NSEG];
            int slot[NSEG];
#pragma unroll
            for (int s = 0; s < NSEG; ++s) {
                slot[s] = off[s] + chf_swz(F.segs[s].row0 + i, swm[s]);
                cd v = mk(0.0, 0.0);
                if (have_fwd) v = sm[slot[s]];
                if (PMODE == 2) {
                    dx[s] = v;
                    x[s] = ((const cd *)seg_in_b)[(long long)s * size + i];
                } else {
                    x[s] = seg_in_b ? ((const cd *)seg_in_b)[(long long)s * size + i] : v;
                    dx[s] = mk(0.0, 0.0);
                }
                if (seg_out_b) ((cd *)seg_out_b)[(long long)s * size + i] = v;
                g[s] = mk(0.0, 0.0);
            }
            if (PMODE != 3) {
#pragma unroll
                for (int s = 0; s < NSEG; ++s) {
#pragma unroll
                    for (int t = s + 1; t < NSEG; ++t) {
                        const cd d = csub(x[s], x[t]);
                        const double r = chf_rsqrt(fma(d.x, d.x, d.y * d.y));
                        if (PMODE <= 1) e = fma(w[s][t], r, e);
                        if (PMODE >= 1) {
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
                }
            }
            if (PMODE == 1 || PMODE == 2) {
#pragma unroll
                for (int s = 0; s < NSEG; ++s) sm[slot[s]] = g[s];
            }
        }
        if (PMODE <= 1) red[tid] += e;  // - potential
    });
}

// once per (persistent) CTA: copy the twiddle tables of the passes into the shared-memory region behind the loop buffers
template <class Ctx>
CHB_D void cta_fused_prologue(Ctx &ctx, const ChfSys &F, cd *sm) {
    if (F.tw_smem_len <= 0) return;
    const ChfLoop &L = F.loops[F.tw_src_loop];
    ctx.par([&](int tid, int nthr) {
        for (int s = 0; s < L.nfac; ++s) {
            if (L.tw_soff[s] < 0) continue;
            cd *dst = sm + F.tw_smem_off + L.tw_soff[s];
            for (int i = tid; i < L.tw_len[s]; i += nthr) dst[i] = L.tw[s][i];
        }
    });
}

// ------------------------------------------------------------------------------------------------
// the CTA program
// ------------------------------------------------------------------------------------------------
template <int NSEG, int PMODE, class Ctx>
CHB_D void cta_fused(Ctx &ctx, const ChfSys &F, const ChfArgs &A, long long b, long long b_pf, double *red, cd *sm) {
    const double *vec_b = A.vec_in ? A.vec_in + b * F.nparams : (const double *)0;
    const double *kin_b = A.kin_src ? A.kin_src + b * F.nparams : (const double *)0;
    const double *seg_in_b = A.seg_in ? A.seg_in + b * F.segsz : (const double *)0;
    double *seg_out_b = A.seg_out ? A.seg_out + b * F.segsz : (double *)0;
    double *out_b = A.vec_out ? A.vec_out + b * F.nparams : (double *)0;
    const bool have_fwd = vec_b != (const double *)0;
    const cd *twsm = (F.tw_smem_len > 0 && !(F.dbg & 128)) ? sm + F.tw_smem_off : (const cd *)0;
    // the rows loaded by the forward pass are the rows of the kinetic term: keep them in tensor memory
    const bool stash = F.ver >= 3 && F.tmem_cols > 0 && have_fwd && kin_b == vec_b && out_b != (double *)0 && !(F.dbg & 64);
    ctx.groups(F.nloop, F.T, [&](auto &grp) {
        grp.par([&](int tid, int T) { red[grp.g * T + tid] = 0.0; });
        if (have_fwd) {
            const ChfLoop &L = F.loops[grp.g];
            cd *buf = sm + L.smem_off;
            if (F.ver >= 2) {
                if (!(F.dbg & 1)) chf_load_pass_r(grp, F, L, vec_b, buf, red, A.kin_from == 1, stash);
                if (F.pf_stride > 0 && b_pf < A.nstates && !(F.dbg & 512))
                    grp.one([&]() {
                        chf_prefetch_l2(A.vec_in + b_pf * F.nparams, L.mom_off - L.nco_in, 4LL * L.npr,
                                        A.vec_in + A.nstates * F.nparams);
                    });
                if (!(F.dbg & 2)) chf_fft_spec_to_time(grp, buf, L, true, twsm, F.dbg & 256);
            } else {
#ifdef CHF_KEEP_OLD
                chf_load_loop(grp, F, L, vec_b, buf, red, A.kin_from == 1);
                chf_fft_spec_to_time(grp, buf, L);
#endif
            }
        }
    });
    if (F.dbg & 4) {
    } else if (F.pair_fast)
        chf_pair_fast<NSEG, PMODE>(ctx, F, sm, have_fwd, seg_in_b, seg_out_b, red);
    else
        chf_pair<NSEG, PMODE>(ctx, F, sm, have_fwd, seg_in_b, seg_out_b, red);
    if (out_b) {
        ctx.groups(F.nloop, F.T, [&](auto &grp) {
            const ChfLoop &L = F.loops[grp.g];
            cd *buf = sm + L.smem_off;
            if (F.ver >= 2) {
                if (!(F.dbg & 8)) chf_fft_time_to_spec(grp, buf, L, true, twsm, F.dbg & 256);
                if (!(F.dbg & 16)) chf_store_pass_r(grp, F, L, buf, kin_b, out_b, red, A.kin_from == 2, stash);
            } else {
#ifdef CHF_KEEP_OLD
                chf_fft_time_to_spec(grp, buf, L);
                chf_store_loop(grp, F, L, buf, kin_b, out_b, red, A.kin_from == 2);
#endif
            }
        });
    }
    if (A.scalar_out && !(F.dbg & 32)) {
        cta_reduce(ctx, red);
        ctx.par([&](int tid, int nthr) {
            if (tid == 0) A.scalar_out[b] = red[0];
        });
    }
}
