// One Arnoldi step of the batched lock-step GMRES of the Newton-Krylov driver (choreographies2_b200/newton.py), for every
// state of the batch in ONE launch: w' = w - v_j (Arnoldi on the shifted operator A - I), Gram-Schmidt against v_0..v_j,
// re-orthogonalisation of the rows that lost more than a factor 10, norm, v_(j+1) = w''/|w''|.
//
// It replaces what scipy's gmres does per state inside the reference's solver (choreo/scipy_plus/nonlin.py:55-134 ->
// scipy.optimize.nonlin.KrylovJacobian -> scipy.sparse.linalg.gmres, restart 20): there the Krylov basis of ONE state lives
// in cache; here the bases of thousands of states are [restart+1, batch, nparams] in HBM and the step is HBM-bound, so the
// kernel reads every basis vector exactly once per step (one CTA owns one state; the candidate vector stays in registers /
// shared memory; modified Gram-Schmidt: the dot product with v_i uses the vector already purged of v_0..v_(i-1)).
// Algorithmic bytes per state and step: 8 n (j + 3)  [v_0..v_j and w read, v_(j+1) written].
#pragma once
#ifndef CHB_HOSTSIM
#include <cooperative_groups.h>
#endif
#include <cmath>
#include <cstdint>

#define CHK_THREADS 512
#define CHK_MAX_NE 16

#if defined(__CUDACC__) && !defined(CHB_HOSTSIM)
#define CHK_HD __host__ __device__ __forceinline__
#else
#define CHK_HD inline
#endif

// Least-squares side of GMRES, kept per state (all optional: R == nullptr leaves only the Hessenberg column in h):
//   R [batch][m+1][m] triangular factor, g [batch][m+1] rotated right-hand side, cs [batch][m][2] the Givens rotations,
//   target [batch] stopping threshold on |g[k+1]|, kdone [batch] inner iterations done, done [batch] stop flags (also the
//   skip mask of the step).
struct ChkGivens {
    double *R, *g, *cs;
    const double *target;
    long long *kdone;
    unsigned char *done;
    int m;
    int *live_list, *live_count;  // optional: the states still running after this step are appended (any order) for the next one
    int split;  // 1: the step kernel only stores the raw Hessenberg column into R[:, j]; k_givens_tail (one THREAD per state)
                // does the update -- the serial per-state tail then runs for all states at once instead of once per CTA wave
};

// one thread per state: previous rotations on the new column, the new rotation, R / g update, stopping test
// (the tail of one inner iteration of scipy.sparse.linalg.gmres)
CHK_HD void chk_givens_update(const ChkGivens &G, long long a, int j, double *hcol /* [j + 2] */) {
    const int m = G.m;
    const double *cs = G.cs + (size_t)a * m * 2;
    for (int k = 0; k < j; ++k) {
        const double c = cs[2 * k], s = cs[2 * k + 1];
        const double t = c * hcol[k] + s * hcol[k + 1];
        hcol[k + 1] = -s * hcol[k] + c * hcol[k + 1];
        hcol[k] = t;
    }
    double den = sqrt(hcol[j] * hcol[j] + hcol[j + 1] * hcol[j + 1]);
    if (!(den > 0.0)) den += 1.0;
    const double c = hcol[j] / den, s = hcol[j + 1] / den;
    hcol[j] = c * hcol[j] + s * hcol[j + 1];
    hcol[j + 1] = 0.0;
    G.cs[(size_t)a * m * 2 + 2 * j] = c;
    G.cs[(size_t)a * m * 2 + 2 * j + 1] = s;
    double *R = G.R + (size_t)a * (m + 1) * m;
    for (int i = 0; i <= j + 1; ++i) R[(size_t)i * m + j] = hcol[i];
    double *g = G.g + (size_t)a * (m + 1);
    const double gj = g[j];
    g[j + 1] = -s * gj;
    g[j] = c * gj;
    G.kdone[a] = j + 1;
    if (fabs(g[j + 1]) <= G.target[a]) {
        G.done[a] = 1;
    } else if (G.live_list) {
#ifdef __CUDA_ARCH__
        G.live_list[atomicAdd(G.live_count, 1)] = (int)a;
#else
        G.live_list[(*G.live_count)++] = (int)a;
#endif
    }
}

// y = R[:k,:k]^-1 g[:k] for one state (a zero pivot is replaced by 1, like the array-level code)
CHK_HD void chk_back_subst(const double *R, const double *g, int m, int k, double *y) {
    for (int j = k - 1; j >= 0; --j) {
        double acc = g[j];
        for (int i = j + 1; i < k; ++i) acc -= R[(size_t)j * m + i] * y[i];
        const double d = R[(size_t)j * m + j];
        y[j] = acc / (d != 0.0 ? d : 1.0);
    }
}

#ifndef CHB_HOSTSIM
__device__ __forceinline__ double chk_block_sum(double v, double (*sred)[CHK_THREADS / 32], int &ph) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sred[ph][threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < CHK_THREADS / 32; ++k) s += sred[ph][k];
    ph ^= 1;  // the buffer used two sums ago is free again: every thread passed the barrier in between after reading it
    return s;
}

// two sums at once (independent dependency chains: the FP64 add / shuffle latencies overlap)
__device__ __forceinline__ void chk_block_sum2(double &u, double &v, double (*sred2)[2 * (CHK_THREADS / 32)], int &ph) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        u += __shfl_xor_sync(0xffffffffu, u, o);
        v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    if ((threadIdx.x & 31) == 0) {
        sred2[ph][2 * (threadIdx.x >> 5)] = u;
        sred2[ph][2 * (threadIdx.x >> 5) + 1] = v;
    }
    __syncthreads();
    // lane l takes the partial sums of warp l % 16, four xor steps leave the total in every lane (same order everywhere)
    const double2 uv = *reinterpret_cast<const double2 *>(&sred2[ph][2 * (threadIdx.x & (CHK_THREADS / 32 - 1))]);
    u = uv.x;
    v = uv.y;
#pragma unroll
    for (int o = CHK_THREADS / 64; o > 0; o >>= 1) {
        u += __shfl_xor_sync(0xffffffffu, u, o);
        v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    ph ^= 1;
}

// n <= CHK_THREADS * NE: the candidate vector and two basis rows (current + prefetched next) live in registers
template <int NE>
__global__ void __launch_bounds__(CHK_THREADS, 1)
k_arnoldi_reg(long long rows, long long n, int j, int ldh, double *V, const double *w, const unsigned char *skip, double *h,
              ChkGivens G) {
    __shared__ double sred[2][CHK_THREADS / 32];
    __shared__ double sh[64];
    const long long a = blockIdx.x;
    const int tid = threadIdx.x;
    const size_t vs = (size_t)rows * (size_t)n;
    double *vout = V + (size_t)(j + 1) * vs + (size_t)a * n;
    if (skip && skip[a]) {  // (inside the library's GMRES nobody reads the vectors of a stopped state: no zero fill)
        if (!G.R)
            for (long long e = tid; e < n; e += CHK_THREADS) vout[e] = 0.0;
        if (h && tid <= j + 1) h[(size_t)a * ldh + tid] = 0.0;
        return;
    }
    if (tid < 64) sh[tid] = 0.0;
    int ph = 0;
    double acc[NE], vc[NE], vn[NE];
    const double *vj = V + (size_t)j * vs + (size_t)a * n;
    const double *wr = w + (size_t)a * n;
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const long long e = tid + (long long)CHK_THREADS * k;
        vc[k] = e < n ? vj[e] : 0.0;
        acc[k] = e < n ? wr[e] - vc[k] : 0.0;
    }
    double part = 0.0;
#pragma unroll
    for (int k = 0; k < NE; ++k) part = fma(acc[k], acc[k], part);
    const double n0 = sqrt(chk_block_sum(part, sred, ph));
    double hn = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        // basis order j, 0, 1, .., j-1 (v_j is in registers already)
        for (int t = 0; t <= j; ++t) {
            const int i = t == 0 ? j : t - 1;
            if (t < j) {
                const double *vr = V + (size_t)t * vs + (size_t)a * n;  // next one: index t
#pragma unroll
                for (int k = 0; k < NE; ++k) {
                    const long long e = tid + (long long)CHK_THREADS * k;
                    vn[k] = e < n ? vr[e] : 0.0;
                }
            }
            part = 0.0;
#pragma unroll
            for (int k = 0; k < NE; ++k) part = fma(vc[k], acc[k], part);
            const double hi = chk_block_sum(part, sred, ph);
#pragma unroll
            for (int k = 0; k < NE; ++k) acc[k] = fma(-hi, vc[k], acc[k]);
            if (tid == 0) sh[i] += hi;
#pragma unroll
            for (int k = 0; k < NE; ++k) vc[k] = vn[k];
        }
        part = 0.0;
#pragma unroll
        for (int k = 0; k < NE; ++k) part = fma(acc[k], acc[k], part);
        hn = sqrt(chk_block_sum(part, sred, ph));
        if (pass == 0 && hn < 0.1 * n0) {
#pragma unroll
            for (int k = 0; k < NE; ++k) {
                const long long e = tid + (long long)CHK_THREADS * k;
                vc[k] = e < n ? vj[e] : 0.0;
            }
            continue;
        }
        break;
    }
    const double inv = hn > 0.0 ? 1.0 / hn : 1.0;
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const long long e = tid + (long long)CHK_THREADS * k;
        if (e < n) vout[e] = acc[k] * inv;
    }
    if (tid == 0) {
        sh[j] += 1.0;
        sh[j + 1] = hn;
        if (h)
            for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = sh[i];
        if (G.R) {
            if (G.split) {
                double *Rc = G.R + (size_t)a * (G.m + 1) * G.m + j;
                for (int i = 0; i <= j + 1; ++i) Rc[(size_t)i * G.m] = sh[i];
            } else {
                chk_givens_update(G, a, j, sh);
            }
        }
    }
}

// the same step with the basis rows streamed through a shared-memory ring by bulk asynchronous copies (one cp.async.bulk per
// row, completion on an mbarrier per stage): up to CHK_RING_MAX rows in flight per SM instead of the one prefetched row of
// k_arnoldi_reg, which is what the HBM latency x bandwidth product asks for with a single resident CTA.  Needs even n (16-byte
// rows) and at least two rows of shared memory.
#define CHK_RING_MAX 6
__device__ __forceinline__ void chk_ring_issue(double *dst, const double *src, unsigned bytes, unsigned long long *bar) {
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"(b) : "memory");
}
// every lane polls (measured: one polling lane per warp + __syncwarp is 15 % slower at batch 8192)
__device__ __forceinline__ void chk_ring_wait(unsigned long long *bar, unsigned parity) {
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "CHK_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra CHK_DONE;\n"
        "bra CHK_WAIT;\n"
        "CHK_DONE:\n"
        "}\n" ::"r"(b), "r"(parity) : "memory");
}

template <int NE>
__global__ void __launch_bounds__(CHK_THREADS, 1)
k_arnoldi_ring(long long rows, long long n, int j, int ldh, double *V, const double *w, const unsigned char *skip, double *h,
               ChkGivens G, int nstage, unsigned rowbytes) {
    extern __shared__ __align__(128) unsigned char chk_ring[];
    __shared__ double sred[2][CHK_THREADS / 32];
    __shared__ __align__(16) double sred2[2][2 * (CHK_THREADS / 32)];
    __shared__ double sh[64];
    __shared__ __align__(8) unsigned long long bars[CHK_RING_MAX];
    const long long a = blockIdx.x;
    const int tid = threadIdx.x;
    const size_t vs = (size_t)rows * (size_t)n;
    double *vout = V + (size_t)(j + 1) * vs + (size_t)a * n;
    if (skip && skip[a]) {  // (inside the library's GMRES nobody reads the vectors of a stopped state: no zero fill)
        if (!G.R)
            for (long long e = tid; e < n; e += CHK_THREADS) vout[e] = 0.0;
        if (h && tid <= j + 1) h[(size_t)a * ldh + tid] = 0.0;
        return;
    }
    if (tid < 64) sh[tid] = 0.0;
    for (int s = 0; s < nstage; ++s)
        for (long long e = n + tid; e < (long long)CHK_THREADS * NE; e += CHK_THREADS)
            ((double *)chk_ring)[(size_t)s * (CHK_THREADS * NE) + e] = 0.0;  // the pad of every stage, once
    if (tid == 0) {
        for (int s = 0; s < nstage; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&bars[s])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const double *va = V + (size_t)a * n;  // row i of this state: va + i * vs
    auto row_of = [&](int t) { return t == 0 ? j : t - 1; };  // basis order j, 0, 1, .., j-1
    // a stage holds CHK_THREADS * NE doubles: the copies fill the first n, the rest is zeroed once (rows are read unpredicated)
    auto stage = [&](int s) { return (double *)(chk_ring + (size_t)s * (sizeof(double) * CHK_THREADS * NE)); };
    // consumer position in the ring: stage su, phase parity pu (kept incrementally: no integer division in the row loop);
    // a consumed stage is refilled in place, and at the start of a pass the ring is empty with the producer at (su, pu) too
    int su = 0;
    unsigned pu = 0;
    auto fill_ring = [&]() {  // thread 0: the first rows of a pass
        int si = su;
        for (int t = 0; t <= j && t < nstage; ++t) {
            chk_ring_issue(stage(si), va + (size_t)row_of(t) * vs, rowbytes, &bars[si]);
            if (++si == nstage) si = 0;
        }
    };
    if (tid == 0) fill_ring();
    int ph = 0, ph2 = 0;
    double acc[NE], vc0[NE], vc1[NE];
    const double *wr = w + (size_t)a * n;
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const long long e = tid + (long long)CHK_THREADS * k;
        acc[k] = e < n ? wr[e] : 0.0;
    }
    double n0 = 0.0, hn = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        // rows are taken two at a time (one reduction barrier per pair, two independent dependency chains): inside a pair the
        // second coefficient is computed before the first row is subtracted -- the rows are orthonormal, the difference is
        // rounding, and the second pass below covers the rest
        for (int t = 0; t <= j;) {
            const bool first = pass == 0 && t == 0;           // v_j: the shift w - v_j and its norm come first, alone
            const int nr = (first || t == j) ? 1 : 2;
            const int s0 = su;
            const unsigned par0 = pu;
            if (++su == nstage) { su = 0; pu ^= 1u; }
            const int s1 = su;
            const unsigned par1 = pu;
            if (nr == 2 && ++su == nstage) { su = 0; pu ^= 1u; }
            chk_ring_wait(&bars[s0], par0);
            const double *r0 = stage(s0);
#pragma unroll
            for (int k = 0; k < NE; ++k) vc0[k] = r0[tid + CHK_THREADS * k];
            if (nr == 2) {
                chk_ring_wait(&bars[s1], par1);
                const double *r1 = stage(s1);
#pragma unroll
                for (int k = 0; k < NE; ++k) vc1[k] = r1[tid + CHK_THREADS * k];
            } else {
#pragma unroll
                for (int k = 0; k < NE; ++k) vc1[k] = 0.0;
            }
            if (first) {
                double pa = 0.0, pb = 0.0;
#pragma unroll
                for (int k = 0; k < NE; k += 2) {
                    acc[k] -= vc0[k];
                    pa = fma(acc[k], acc[k], pa);
                    if (k + 1 < NE) {
                        acc[k + 1] -= vc0[k + 1];
                        pb = fma(acc[k + 1], acc[k + 1], pb);
                    }
                }
                n0 = sqrt(chk_block_sum(pa + pb, sred, ph));
            }
            double p0a = 0.0, p0b = 0.0, p1a = 0.0, p1b = 0.0;
#pragma unroll
            for (int k = 0; k < NE; k += 2) {
                p0a = fma(vc0[k], acc[k], p0a);
                p1a = fma(vc1[k], acc[k], p1a);
                if (k + 1 < NE) {
                    p0b = fma(vc0[k + 1], acc[k + 1], p0b);
                    p1b = fma(vc1[k + 1], acc[k + 1], p1b);
                }
            }
            double h0 = p0a + p0b, h1 = p1a + p1b;
            chk_block_sum2(h0, h1, sred2, ph2);  // its barrier: every thread has read the stage(s)
            if (tid == 0) {
                sh[row_of(t)] += h0;
                if (t + nstage <= j) chk_ring_issue(stage(s0), va + (size_t)row_of(t + nstage) * vs, rowbytes, &bars[s0]);
                if (nr == 2) {
                    sh[row_of(t + 1)] += h1;
                    if (t + 1 + nstage <= j)
                        chk_ring_issue(stage(s1), va + (size_t)row_of(t + 1 + nstage) * vs, rowbytes, &bars[s1]);
                }
            }
#pragma unroll
            for (int k = 0; k < NE; ++k) acc[k] = fma(-h1, vc1[k], fma(-h0, vc0[k], acc[k]));
            t += nr;
        }
        double pa = 0.0, pb = 0.0;
#pragma unroll
        for (int k = 0; k < NE; k += 2) {
            pa = fma(acc[k], acc[k], pa);
            if (k + 1 < NE) pb = fma(acc[k + 1], acc[k + 1], pb);
        }
        hn = sqrt(chk_block_sum(pa + pb, sred, ph));
        if (pass == 0 && hn < 0.1 * n0) {  // second pass: stream the rows again (every stage is free)
            if (tid == 0) fill_ring();
            continue;
        }
        break;
    }
    const double inv = hn > 0.0 ? 1.0 / hn : 1.0;
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const long long e = tid + (long long)CHK_THREADS * k;
        if (e < n) vout[e] = acc[k] * inv;
    }
    if (tid == 0) {
        sh[j] += 1.0;
        sh[j + 1] = hn;
        if (h)
            for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = sh[i];
        if (G.R) {
            if (G.split) {
                double *Rc = G.R + (size_t)a * (G.m + 1) * G.m + j;
                for (int i = 0; i <= j + 1; ++i) Rc[(size_t)i * G.m] = sh[i];
            } else {
                chk_givens_update(G, a, j, sh);
            }
        }
    }
}

// per-row sum of squares and max |.| in one pass (the norms of the Newton driver: 2-norm of the gradient for the forcing terms
// and the line search, max norm for the termination test; a non-finite entry makes the sum non-finite).  One CTA per row,
// fixed reduction order: a row's result does not depend on the batch it is in.
__global__ void __launch_bounds__(256)
k_row_stats(long long n, const double *a, double *sumsq, double *maxabs) {
    __shared__ double s1[8], s2[8];
    const double *r = a + (size_t)blockIdx.x * n;
    double p0 = 0.0, p1 = 0.0, m = 0.0;
    long long e = threadIdx.x;
    for (; e + 256 < n; e += 512) {
        const double x = r[e], y = r[e + 256];
        p0 = fma(x, x, p0);
        p1 = fma(y, y, p1);
        m = fmax(m, fmax(fabs(x), fabs(y)));
    }
    if (e < n) {
        const double x = r[e];
        p0 = fma(x, x, p0);
        m = fmax(m, fabs(x));
    }
    double p = p0 + p1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        p += __shfl_xor_sync(0xffffffffu, p, o);
        m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s1[threadIdx.x >> 5] = p;
        s2[threadIdx.x >> 5] = m;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ps = 0.0, ms = 0.0;
        for (int k = 0; k < 8; ++k) {
            ps += s1[k];
            ms = fmax(ms, s2[k]);
        }
        sumsq[blockIdx.x] = ps;
        if (maxabs) maxabs[blockIdx.x] = ms;
    }
}

// the per-state tail of the step for all states at once (see ChkGivens::split): one thread per state
__global__ void __launch_bounds__(128)
k_givens_tail(long long rows, int j, ChkGivens G) {
    const long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= rows || G.done[a]) return;
    double hcol[64];
    const double *Rc = G.R + (size_t)a * (G.m + 1) * G.m + j;
    for (int i = 0; i <= j + 1; ++i) hcol[i] = Rc[(size_t)i * G.m];
    chk_givens_update(G, a, j, hcol);
}

// larger rows: the candidate vector lives in shared memory (8 n bytes), every basis row is streamed once from HBM for the dot
// product and once more for the update (an L2 hit: the same CTA just read it).  GMEM: rows longer than shared memory (the deep
// refinement levels of a search, n > 25600): the candidate lives in its destination row V[j+1] (every thread only ever touches
// its own elements, so plain loads / stores suffice).
template <bool GMEM>
__global__ void __launch_bounds__(CHK_THREADS, 1)
k_arnoldi_smem(long long rows, long long n, int j, int ldh, double *V, const double *w, const unsigned char *skip, double *h,
               ChkGivens G) {
    extern __shared__ double chk_acc_s[];
    __shared__ double sred[2][CHK_THREADS / 32];
    __shared__ double sh[64];
    const long long a = blockIdx.x;
    const int tid = threadIdx.x;
    const size_t vs = (size_t)rows * (size_t)n;
    double *vout = V + (size_t)(j + 1) * vs + (size_t)a * n;
    double *chk_acc = GMEM ? vout : chk_acc_s;
    if (skip && skip[a]) {  // (inside the library's GMRES nobody reads the vectors of a stopped state: no zero fill)
        if (!G.R)
            for (long long e = tid; e < n; e += CHK_THREADS) vout[e] = 0.0;
        if (h && tid <= j + 1) h[(size_t)a * ldh + tid] = 0.0;
        return;
    }
    if (tid < 64) sh[tid] = 0.0;
    int ph = 0;
    const double *vj = V + (size_t)j * vs + (size_t)a * n;
    const double *wr = w + (size_t)a * n;
    double part = 0.0;
#pragma unroll 4
    for (long long e = tid; e < n; e += CHK_THREADS) {
        const double d = wr[e] - vj[e];
        chk_acc[e] = d;
        part = fma(d, d, part);
    }
    const double n0 = sqrt(chk_block_sum(part, sred, ph));
    double hn = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        for (int t = 0; t <= j; ++t) {
            const int i = t == 0 ? j : t - 1;
            const double *vr = V + (size_t)i * vs + (size_t)a * n;
            part = 0.0;
#pragma unroll 8
            for (long long e = tid; e < n; e += CHK_THREADS) part = fma(vr[e], chk_acc[e], part);
            const double hi = chk_block_sum(part, sred, ph);
#pragma unroll 8
            for (long long e = tid; e < n; e += CHK_THREADS) chk_acc[e] = fma(-hi, vr[e], chk_acc[e]);
            if (tid == 0) sh[i] += hi;
        }
        part = 0.0;
#pragma unroll 4
        for (long long e = tid; e < n; e += CHK_THREADS) part = fma(chk_acc[e], chk_acc[e], part);
        hn = sqrt(chk_block_sum(part, sred, ph));
        if (pass == 0 && hn < 0.1 * n0) continue;
        break;
    }
    const double inv = hn > 0.0 ? 1.0 / hn : 1.0;
    for (long long e = tid; e < n; e += CHK_THREADS) vout[e] = chk_acc[e] * inv;
    if (tid == 0) {
        sh[j] += 1.0;
        sh[j + 1] = hn;
        if (h)
            for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = sh[i];
        if (G.R) {
            if (G.split) {
                double *Rc = G.R + (size_t)a * (G.m + 1) * G.m + j;
                for (int i = 0; i <= j + 1; ++i) Rc[(size_t)i * G.m] = sh[i];
            } else {
                chk_givens_update(G, a, j, sh);
            }
        }
    }
}

// Rows longer than one CTA's shared memory (the deep refinement levels of a search: 3C at nint 49152 has n = 98302) and few states:
// a thread-block CLUSTER of CHK_CL CTAs per state.  The candidate vector is distributed over the cluster's shared memory (CTA r
// holds the elements [r len, (r + 1) len)), every CTA streams its slice of the basis rows, and the inner products are completed
// through distributed shared memory: each CTA writes its partial sum into slot r of every CTA of the cluster, one cluster barrier,
// then all add the CHK_CL partial sums in the same order (deterministic, identical on every CTA).  Same modified Gram-Schmidt
// recurrence as k_arnoldi_smem.
#define CHK_CL 8
__device__ __forceinline__ double chk_cluster_sum(double v, double (*sred)[CHK_THREADS / 32], int &ph, double (*cpart)[CHK_CL], int &cph,
                                                  cooperative_groups::cluster_group &cl, unsigned rank) {
    const double s = chk_block_sum(v, sred, ph);
    if (threadIdx.x < CHK_CL) *cl.map_shared_rank(&cpart[cph][rank], threadIdx.x) = s;
    cl.sync();
    double t = 0.0;
#pragma unroll
    for (int r = 0; r < CHK_CL; ++r) t += cpart[cph][r];
    cph ^= 1;  // the buffer of two sums ago is free: every CTA of the cluster passed a cluster barrier after reading it
    return t;
}
__global__ void __cluster_dims__(CHK_CL, 1, 1) __launch_bounds__(CHK_THREADS, 1)
k_arnoldi_cluster(long long rows, long long n, long long len, int j, int ldh, double *V, const double *w, const unsigned char *skip,
                  double *h, ChkGivens G) {
    extern __shared__ double chk_acc[];
    __shared__ double sred[2][CHK_THREADS / 32];
    __shared__ double cpart[2][CHK_CL];
    __shared__ double sh[64];
    cooperative_groups::cluster_group cl = cooperative_groups::this_cluster();
    const unsigned rank = cl.block_rank();
    const long long a = blockIdx.x / CHK_CL;
    const int tid = threadIdx.x;
    const size_t vs = (size_t)rows * (size_t)n;
    const long long lo = (long long)rank * len;
    const long long cnt = lo < n ? (n - lo < len ? n - lo : len) : 0;  // elements of this CTA
    double *vout = V + (size_t)(j + 1) * vs + (size_t)a * n + lo;
    if (skip && skip[a]) {  // the whole cluster leaves together
        if (!G.R)
            for (long long e = tid; e < cnt; e += CHK_THREADS) vout[e] = 0.0;
        if (rank == 0 && h && tid <= j + 1) h[(size_t)a * ldh + tid] = 0.0;
        return;
    }
    if (tid < 64) sh[tid] = 0.0;
    int ph = 0, cph = 0;
    const double *vj = V + (size_t)j * vs + (size_t)a * n + lo;
    const double *wr = w + (size_t)a * n + lo;
    double part = 0.0;
#pragma unroll 4
    for (long long e = tid; e < cnt; e += CHK_THREADS) {
        const double d = wr[e] - vj[e];
        chk_acc[e] = d;
        part = fma(d, d, part);
    }
    const double n0 = sqrt(chk_cluster_sum(part, sred, ph, cpart, cph, cl, rank));
    double hn = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        for (int t = 0; t <= j; ++t) {
            const int i = t == 0 ? j : t - 1;
            const double *vr = V + (size_t)i * vs + (size_t)a * n + lo;
            part = 0.0;
#pragma unroll 8
            for (long long e = tid; e < cnt; e += CHK_THREADS) part = fma(vr[e], chk_acc[e], part);
            const double hi = chk_cluster_sum(part, sred, ph, cpart, cph, cl, rank);
#pragma unroll 8
            for (long long e = tid; e < cnt; e += CHK_THREADS) chk_acc[e] = fma(-hi, vr[e], chk_acc[e]);
            if (tid == 0) sh[i] += hi;
        }
        part = 0.0;
#pragma unroll 4
        for (long long e = tid; e < cnt; e += CHK_THREADS) part = fma(chk_acc[e], chk_acc[e], part);
        hn = sqrt(chk_cluster_sum(part, sred, ph, cpart, cph, cl, rank));
        if (pass == 0 && hn < 0.1 * n0) continue;
        break;
    }
    const double inv = hn > 0.0 ? 1.0 / hn : 1.0;
    for (long long e = tid; e < cnt; e += CHK_THREADS) vout[e] = chk_acc[e] * inv;
    if (rank == 0 && tid == 0) {
        sh[j] += 1.0;
        sh[j + 1] = hn;
        if (h)
            for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = sh[i];
        if (G.R) {
            if (G.split) {
                double *Rc = G.R + (size_t)a * (G.m + 1) * G.m + j;
                for (int i = 0; i <= j + 1; ++i) Rc[(size_t)i * G.m] = sh[i];
            } else {
                chk_givens_update(G, a, j, sh);
            }
        }
    }
}

// solution of the least-squares problem: grid = rows x nchunk CTAs, every CTA repeats the (tiny) back substitution of its row
__global__ void __launch_bounds__(256)
k_gmres_solution(long long rows, long long n, int m, const double *V, const double *R, const double *g, const long long *kdone,
                 double scale, double *sol, int nchunk) {
    __shared__ double y[64];
    const long long a = blockIdx.x / nchunk;
    const int c = (int)(blockIdx.x % nchunk);
    int k = (int)kdone[a];
    k = k < 0 ? 0 : (k > m ? m : k);
    if (threadIdx.x == 0) chk_back_subst(R + (size_t)a * (m + 1) * m, g + (size_t)a * (m + 1), m, k, y);
    __syncthreads();
    const size_t vs = (size_t)rows * (size_t)n;
    const double *va = V + (size_t)a * n;
    const long long per = (n + nchunk - 1) / nchunk, e0 = c * per, e1 = e0 + per < n ? e0 + per : n;
    for (long long e = e0 + threadIdx.x; e < e1; e += 256) {
        double s = 0.0;
#pragma unroll 4
        for (int j = 0; j < k; ++j) s = fma(y[j], va[(size_t)j * vs + e], s);
        sol[(size_t)a * n + e] = scale * s;
    }
}
#else
static void chk_solution_host(long long rows, long long n, int m, const double *V, const double *R, const double *g,
                              const long long *kdone, double scale, double *sol) {
    const size_t vs = (size_t)rows * (size_t)n;
    for (long long a = 0; a < rows; ++a) {
        double y[64];
        int k = (int)kdone[a];
        k = k < 0 ? 0 : (k > m ? m : k);
        chk_back_subst(R + (size_t)a * (m + 1) * m, g + (size_t)a * (m + 1), m, k, y);
        for (long long e = 0; e < n; ++e) {
            double s = 0.0;
            for (int j = 0; j < k; ++j) s += y[j] * V[(size_t)j * vs + (size_t)a * n + e];
            sol[(size_t)a * n + e] = scale * s;
        }
    }
}
static void chk_row_stats_host(long long rows, long long n, const double *a, double *sumsq, double *maxabs) {
    for (long long b = 0; b < rows; ++b) {
        double p = 0.0, m = 0.0;
        for (long long e = 0; e < n; ++e) {
            const double x = a[(size_t)b * n + e];
            p += x * x;
            m = std::fmax(m, std::fabs(x));
        }
        sumsq[b] = p;
        if (maxabs) maxabs[b] = m;
    }
}
// host emulation (tests/hostsim): the same arithmetic, one state after the other
static void chk_arnoldi_host(long long rows, long long n, int j, int ldh, double *V, const double *w,
                             const unsigned char *skip, double *h, ChkGivens G) {
    const size_t vs = (size_t)rows * (size_t)n;
    std::vector<double> acc((size_t)n);
    for (long long a = 0; a < rows; ++a) {
        double *vout = V + (size_t)(j + 1) * vs + (size_t)a * n;
        double hloc[64];
        double *hr = hloc;
        if (skip && skip[a]) {
            for (long long e = 0; e < n; ++e) vout[e] = 0.0;
            if (h)
                for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = 0.0;
            continue;
        }
        const double *vj = V + (size_t)j * vs + (size_t)a * n;
        double s = 0.0;
        for (long long e = 0; e < n; ++e) {
            acc[e] = w[(size_t)a * n + e] - vj[e];
            s += acc[e] * acc[e];
        }
        const double n0 = std::sqrt(s);
        for (int i = 0; i <= j; ++i) hr[i] = 0.0;
        double hn = 0.0;
        for (int pass = 0; pass < 2; ++pass) {
            for (int t = 0; t <= j; ++t) {
                const int i = t == 0 ? j : t - 1;
                const double *vr = V + (size_t)i * vs + (size_t)a * n;
                double hi = 0.0;
                for (long long e = 0; e < n; ++e) hi += vr[e] * acc[e];
                for (long long e = 0; e < n; ++e) acc[e] -= hi * vr[e];
                hr[i] += hi;
            }
            s = 0.0;
            for (long long e = 0; e < n; ++e) s += acc[e] * acc[e];
            hn = std::sqrt(s);
            if (pass == 0 && hn < 0.1 * n0) continue;
            break;
        }
        const double inv = hn > 0.0 ? 1.0 / hn : 1.0;
        for (long long e = 0; e < n; ++e) vout[e] = acc[e] * inv;
        hr[j] += 1.0;
        hr[j + 1] = hn;
        if (h)
            for (int i = 0; i <= j + 1; ++i) h[(size_t)a * ldh + i] = hr[i];
        if (G.R) chk_givens_update(G, a, j, hr);  // (the emulation always runs the tail in place)
    }
}
#endif
