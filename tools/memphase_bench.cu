// tools/memphase_bench.cu -- which global<->shared access pattern feeds one 200 KB CTA per SM fastest?
//
// Stand-alone micro-benchmark behind the fused kernel's load / store pass design:
// 4096 "states" of 3 x 8190 doubles (the 3B nint=4096 parameter vector), one CTA of 384 threads and 200 KB of dynamic
// shared memory per state, so exactly one CTA is resident per SM like k_fused.  Each variant moves one state between
// global and shared memory and does nothing else.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/memphase_bench tools/memphase_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr int NPR = 2048, NLOOP = 3, T = 128, NTHR = 384;
constexpr long long NPARAMS = 24570;  // 3 * (4 * 2048 - 2)
constexpr int M = 256, H = 8;

__device__ __forceinline__ void ld16(const double *p, double &a, double &b) {
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p));
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// rows of loop g start at element g*8190 - 2 (the two removed DC parameters), row k = 4 doubles at +4k
__device__ __forceinline__ const double *rows_of(const double *state, int g) { return state + (long long)g * 8190 - 2; }

// V0: the pattern of chf_load_pass<16>: lane pair (klo, M-klo), 8 whole rows per thread, two 16-byte loads per row
__global__ void __launch_bounds__(NTHR, 1) k_load_rows(const double *in, double *sink, int iters) {
    extern __shared__ __align__(1024) double sm[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        const double *rows = rows_of(in + b * NPARAMS, g);
        double *buf = sm + g * 8192;
        for (int it = 0; it < 2; ++it) {
            const int pr = it * 64 + (tid >> 1), h = tid & 1;
            const int klo = pr == 0 ? (h ? M / 2 : 0) : (h ? M - pr : pr);
            double p[H][4];
#pragma unroll
            for (int q = 0; q < H; ++q) {
                const int k = klo + q * M;
                if (k > 0) { ld16(rows + 4LL * k, p[q][0], p[q][1]); ld16(rows + 4LL * k + 2, p[q][2], p[q][3]); }
                else { p[q][0] = p[q][1] = p[q][2] = p[q][3] = 0.0; }
            }
#pragma unroll
            for (int q = 0; q < H; ++q) {
                double2 *d = (double2 *)(buf + 4 * ((klo * H + q) ^ (klo & 7)));
                d[0] = make_double2(p[q][0], p[q][1]);
                d[1] = make_double2(p[q][2], p[q][3]);
            }
        }
        __syncthreads();
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = sm[threadIdx.x];
}

// V1: half rows: lane pair shares a row, every warp-level load covers contiguous 512-byte runs
__global__ void __launch_bounds__(NTHR, 1) k_load_half(const double *in, double *sink, int iters) {
    extern __shared__ __align__(1024) double sm[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        const double *rows = rows_of(in + b * NPARAMS, g);
        double *buf = sm + g * 8192;
        for (int it = 0; it < 2; ++it) {
            const int pr = it * 64 + (tid >> 1), h = tid & 1;
            const int ka = pr == 0 ? 0 : pr, kb = pr == 0 ? M / 2 : M - pr;
            double p[2 * H][2];
#pragma unroll
            for (int q = 0; q < H; ++q) {
                const int k0 = ka + q * M, k1 = kb + q * M;
                if (k0 > 0) ld16(rows + 4LL * k0 + 2 * h, p[q][0], p[q][1]); else p[q][0] = p[q][1] = 0.0;
                ld16(rows + 4LL * k1 + 2 * h, p[H + q][0], p[H + q][1]);
            }
#pragma unroll
            for (int q = 0; q < 2 * H; ++q)
                *(double2 *)(buf + 2 * (((pr * 2 * H + q) * 2 + h) ^ (pr & 7))) = make_double2(p[q][0], p[q][1]);
        }
        __syncthreads();
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = sm[threadIdx.x];
}

// V2: plain coalesced copy, 16 bytes per lane, UNR loads in flight per thread
template <int UNR>
__global__ void __launch_bounds__(NTHR, 1) k_load_plain(const double *in, double *sink, int iters) {
    extern __shared__ __align__(1024) double sm[];
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        const double *src = in + b * NPARAMS;  // 24570 doubles = 12285 double2
        for (int i0 = 0; i0 < 12285; i0 += NTHR * UNR) {
            double p[UNR][2];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int i = i0 + u * NTHR + threadIdx.x;
                if (i < 12285) ld16(src + 2LL * i, p[u][0], p[u][1]);
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int i = i0 + u * NTHR + threadIdx.x;
                if (i < 12285) *(double2 *)(sm + 2 * i) = make_double2(p[u][0], p[u][1]);
            }
        }
        __syncthreads();
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = sm[threadIdx.x];
}

// V3: TMA bulk copy global -> shared, one 65520-byte copy per loop, completion on an mbarrier
__global__ void __launch_bounds__(NTHR, 1) k_load_bulk(const double *in, double *sink, int iters) {
    extern __shared__ __align__(1024) double sm[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    unsigned phase = 0;
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        if (threadIdx.x == 0) {
            const unsigned bytes = 3u * 65520u;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes));
            for (int g = 0; g < NLOOP; ++g) {
                const double *src = in + b * NPARAMS + (long long)g * 8190;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(sm + g * 8192)), "l"(src), "r"(65520u), "r"(smem_u32(&bar)) : "memory");
            }
        }
        unsigned done = 0;
        while (!done) {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
        }
        phase ^= 1;
        __syncthreads();
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = sm[threadIdx.x];
}


// V4: TMA tensor load with a 5-D view [state][k0][k1][q][4 doubles] of one loop's rows: one box = the whole loop, landing
// in shared memory as [k0][k1][q][4], i.e. the 8 direct rows of butterfly klo = k0 + 16 k1 sit in the 16 slots
// (256 bytes) where that butterfly's first-pass output goes.  check != 0: verify the layout against global memory.
struct Maps3 { CUtensorMap m[NLOOP]; };
__global__ void __launch_bounds__(NTHR, 1) k_load_tma5d(const __grid_constant__ Maps3 maps, const double *in, double *sink,
                                                         int iters, int check, unsigned long long *errs, int lds, int swz128) {
    extern __shared__ __align__(1024) double sm[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    unsigned phase = 0;
    double acc = 0.0;
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(3u * 65536u));
            for (int g = 0; g < NLOOP; ++g) {
                asm volatile(
                    "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                    ::"r"(smem_u32(sm + g * 8192)), "l"(&maps.m[g]), "r"(0), "r"(0), "r"(0), "r"(0), "r"((int)b),
                      "r"(smem_u32(&bar)) : "memory");
            }
        }
        unsigned done = 0;
        while (!done) {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
        }
        phase ^= 1;
        // every thread reads its two 256-byte blocks like the in-place load pass would
        const int g = threadIdx.x / T, tid = threadIdx.x % T;
        for (int it = 0; it < 2 && lds; ++it) {
            const int pr = it * 64 + (tid >> 1), h = tid & 1;
            const int klo = pr == 0 ? (h ? M / 2 : 0) : (h ? M - pr : pr);
            const int blk = klo;  // layout [k1][k0][q][4 doubles]: block index = klo
            const double2 *src = (const double2 *)(sm + g * 8192 + blk * 32);
            const int xr = swz128 ? ((2 * blk) & 7) : 0;  // SWIZZLE_128B: 16-byte chunk index bits 0..2 ^= address bits 7..9
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const double2 v = src[(c & 8) | ((c & 7) ^ ((xr + (c >> 3)) & 7))];
                acc += v.x + v.y;
                if (check) {
                    const int q = c >> 1;
                    const long long e = b * NPARAMS + (long long)g * 8190 - 2 + 4LL * (klo + q * M) + 2 * (c & 1);
                    if (e >= 0 && (v.x != in[e] || v.y != in[e + 1])) atomicAdd(errs, 1ULL);
                }
            }
        }
        __syncthreads();
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = acc;
}

// TMEM as a per-thread stash: every thread parks 64 doubles (128 columns of its lane) and reads them back
__global__ void __launch_bounds__(NTHR, 1) k_tmem_stash(double *sink, int iters, unsigned long long *errs) {
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = tmem_base_s;
    // lane quarter of this warp in bits 16.., column block by warp / 4
    const uint32_t addr = base + (((uint32_t)(warp & 3) * 32u) << 16) + (uint32_t)(warp >> 2) * 128u;
    double acc = 0.0;
    for (int itn = blockIdx.x; itn < iters; itn += gridDim.x) {
        uint32_t v[64], w[64];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = (uint32_t)(threadIdx.x * 1315423911u + i * 2654435761u + half * 97u + itn);
            asm volatile(
                "tcgen05.st.sync.aligned.32x32b.x64.b32 [%0], "
                "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
                "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, "
                "%33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, "
                "%49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63, %64};"
                ::"r"(addr + half * 64),
                  "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                  "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                  "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                  "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]),
                  "r"(v[32]), "r"(v[33]), "r"(v[34]), "r"(v[35]), "r"(v[36]), "r"(v[37]), "r"(v[38]), "r"(v[39]),
                  "r"(v[40]), "r"(v[41]), "r"(v[42]), "r"(v[43]), "r"(v[44]), "r"(v[45]), "r"(v[46]), "r"(v[47]),
                  "r"(v[48]), "r"(v[49]), "r"(v[50]), "r"(v[51]), "r"(v[52]), "r"(v[53]), "r"(v[54]), "r"(v[55]),
                  "r"(v[56]), "r"(v[57]), "r"(v[58]), "r"(v[59]), "r"(v[60]), "r"(v[61]), "r"(v[62]), "r"(v[63])
                : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
                "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
                "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
                : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                  "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15]),
                  "=r"(w[16]), "=r"(w[17]), "=r"(w[18]), "=r"(w[19]), "=r"(w[20]), "=r"(w[21]), "=r"(w[22]), "=r"(w[23]),
                  "=r"(w[24]), "=r"(w[25]), "=r"(w[26]), "=r"(w[27]), "=r"(w[28]), "=r"(w[29]), "=r"(w[30]), "=r"(w[31]),
                  "=r"(w[32]), "=r"(w[33]), "=r"(w[34]), "=r"(w[35]), "=r"(w[36]), "=r"(w[37]), "=r"(w[38]), "=r"(w[39]),
                  "=r"(w[40]), "=r"(w[41]), "=r"(w[42]), "=r"(w[43]), "=r"(w[44]), "=r"(w[45]), "=r"(w[46]), "=r"(w[47]),
                  "=r"(w[48]), "=r"(w[49]), "=r"(w[50]), "=r"(w[51]), "=r"(w[52]), "=r"(w[53]), "=r"(w[54]), "=r"(w[55]),
                  "=r"(w[56]), "=r"(w[57]), "=r"(w[58]), "=r"(w[59]), "=r"(w[60]), "=r"(w[61]), "=r"(w[62]), "=r"(w[63])
                : "r"(addr + half * 64) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 64; ++i) {
                const uint32_t expect = (uint32_t)(threadIdx.x * 1315423911u + i * 2654435761u + half * 97u + itn);
                if (w[i] != expect) atomicAdd(errs, 1ULL);
                acc += (double)(w[i] & 1u);
            }
        }
    }
    if (sink) sink[blockIdx.x * NTHR + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "n"(512));
}

// ---- stores ---------------------------------------------------------------------------------------------
// S0: the pattern of chf_store_pass<16>: 8 whole rows per thread, two 16-byte stores per row
__global__ void __launch_bounds__(NTHR, 1) k_store_rows(double *out, int iters) {
    extern __shared__ __align__(1024) double sm[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    for (int i = threadIdx.x; i < 3 * 8192; i += NTHR) sm[i] = (double)i;
    __syncthreads();
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        double *rows = out + b * NPARAMS + (long long)g * 8190 - 2;
        const double *buf = sm + g * 8192;
        for (int it = 0; it < 2; ++it) {
            const int pr = it * 64 + (tid >> 1), h = tid & 1;
            const int klo = pr == 0 ? (h ? M / 2 : 0) : (h ? M - pr : pr);
#pragma unroll
            for (int q = 0; q < H; ++q) {
                const int k = klo + q * M;
                const double2 *s = (const double2 *)(buf + 4 * ((klo * H + q) ^ (klo & 7)));
                if (k > 0) {
                    *(double2 *)(rows + 4LL * k) = s[0];
                    *(double2 *)(rows + 4LL * k + 2) = s[1];
                }
            }
        }
        __syncthreads();
    }
}
// S1: half rows per lane
__global__ void __launch_bounds__(NTHR, 1) k_store_half(double *out, int iters) {
    extern __shared__ __align__(1024) double sm[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    for (int i = threadIdx.x; i < 3 * 8192; i += NTHR) sm[i] = (double)i;
    __syncthreads();
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        double *rows = out + b * NPARAMS + (long long)g * 8190 - 2;
        const double *buf = sm + g * 8192;
        for (int it = 0; it < 2; ++it) {
            const int pr = it * 64 + (tid >> 1), h = tid & 1;
            const int ka = pr == 0 ? 0 : pr, kb = pr == 0 ? M / 2 : M - pr;
#pragma unroll
            for (int q = 0; q < H; ++q) {
                const int k0 = ka + q * M, k1 = kb + q * M;
                const double2 v0 = *(const double2 *)(buf + 2 * (((pr * 2 * H + q) * 2 + h) ^ (pr & 7)));
                const double2 v1 = *(const double2 *)(buf + 2 * (((pr * 2 * H + H + q) * 2 + h) ^ (pr & 7)));
                if (k0 > 0) *(double2 *)(rows + 4LL * k0 + 2 * h) = v0;
                *(double2 *)(rows + 4LL * k1 + 2 * h) = v1;
            }
        }
        __syncthreads();
    }
}
// S2: plain coalesced
__global__ void __launch_bounds__(NTHR, 1) k_store_plain(double *out, int iters) {
    extern __shared__ __align__(1024) double sm[];
    for (int i = threadIdx.x; i < 3 * 8192; i += NTHR) sm[i] = (double)i;
    __syncthreads();
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        double *dst = out + b * NPARAMS;
#pragma unroll 4
        for (int i = threadIdx.x; i < 12285; i += NTHR) *(double2 *)(dst + 2LL * i) = *(const double2 *)(sm + 2 * i);
        __syncthreads();
    }
}
// S3: TMA bulk store shared -> global
__global__ void __launch_bounds__(NTHR, 1) k_store_bulk(double *out, int iters) {
    extern __shared__ __align__(1024) double sm[];
    for (int i = threadIdx.x; i < 3 * 8192; i += NTHR) sm[i] = (double)i;
    __syncthreads();
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        if (threadIdx.x == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            for (int g = 0; g < NLOOP; ++g) {
                double *dst = out + b * NPARAMS + (long long)g * 8190;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(dst), "r"(smem_u32(sm + g * 8192)), "r"(65520u) : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}


// S4: TMA tensor store shared -> global through the same 5-D view (scatter of 32-byte rows)
__global__ void __launch_bounds__(NTHR, 1) k_store_tma5d(const __grid_constant__ Maps3 maps, int iters) {
    extern __shared__ __align__(1024) double sm[];
    for (int i = threadIdx.x; i < 3 * 8192; i += NTHR) sm[i] = (double)i;
    __syncthreads();
    for (long long b = blockIdx.x; b < iters; b += gridDim.x) {
        if (threadIdx.x == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            for (int g = 0; g < NLOOP; ++g)
                asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3, %4, %5}], [%6];"
                             ::"l"(&maps.m[g]), "r"(0), "r"(0), "r"(0), "r"(0), "r"((int)b), "r"(smem_u32(sm + g * 8192)) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class F>
static float time_it(F launch, int reps = 5) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r) launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms / reps;
}

int main() {
    const int B = 4096;
    const size_t smem = 200 * 1024;
    double *in, *out, *sink;
    double *in_alloc;
    CK(cudaMalloc(&in_alloc, (size_t)B * NPARAMS * 8 + 1024));
    in = in_alloc + 32;  // the rows of loop 0 start 16 bytes before the vector
    CK(cudaMalloc(&out, (size_t)B * NPARAMS * 8 + 1024));
    CK(cudaMalloc(&sink, (size_t)B * NTHR * 8));
    CK(cudaMemset(in, 0, (size_t)B * NPARAMS * 8));
    const double gb = (double)B * NPARAMS * 8 / 1e9;
#define RUN(name, kern, grid, ...)                                                                      \
    do {                                                                                                \
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));         \
        float ms = time_it([&]() { kern<<<grid, NTHR, smem>>>(__VA_ARGS__); });                         \
        printf("%-34s grid %5d : %.3f ms  %.0f GB/s\n", name, grid, ms, gb / (ms * 1e-3));              \
    } while (0)
    for (int grid : {4096, 148}) {
        RUN("load rows (current v2 pattern)", k_load_rows, grid, in, sink, B);
        RUN("load half rows", k_load_half, grid, in, sink, B);
        RUN("load plain unroll 4", k_load_plain<4>, grid, in, sink, B);
        RUN("load plain unroll 8", k_load_plain<8>, grid, in, sink, B);
        RUN("load plain unroll 16", k_load_plain<16>, grid, in, sink, B);
        RUN("load TMA bulk", k_load_bulk, grid, in, sink, B);
        RUN("store rows (current v2 pattern)", k_store_rows, grid, out, B);
        RUN("store half rows", k_store_half, grid, out, B);
        RUN("store plain", k_store_plain, grid, out, B);
        RUN("store TMA bulk", k_store_bulk, grid, out, B);
    }

    // ---- TMA tensor maps (driver entry point through the runtime: no -lcuda needed) ----
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    EncodeFn encode = nullptr;
    {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        encode = (EncodeFn)fn;
    }
    unsigned long long *errs;
    CK(cudaMalloc(&errs, 8));
    CK(cudaMemset(errs, 0, 8));
    {   // recognisable input for the layout check
        double *h = (double *)malloc((size_t)B * NPARAMS * 8);
        for (long long i = 0; i < (long long)B * NPARAMS; ++i) h[i] = (double)(i % 1000003) + 0.25;
        CK(cudaMemcpy(in, h, (size_t)B * NPARAMS * 8, cudaMemcpyHostToDevice));
        free(h);
    }
    for (int swz = 0; swz < 2; ++swz) {
        Maps3 maps, omaps;
        bool ok = true;
        for (int g = 0; g < NLOOP; ++g) {
            // dims inner -> outer: 4 doubles | q (256 rows apart) | k0 (1 row apart) | k1 (16 rows apart) | state
            cuuint64_t dims[5] = {4, 8, 16, 16, (cuuint64_t)B};
            cuuint64_t strides[4] = {8192, 32, 512, (cuuint64_t)NPARAMS * 8};  // bytes, dims 1..4
            cuuint32_t box[5] = {4, 8, 16, 16, 1};
            cuuint32_t estr[5] = {1, 1, 1, 1, 1};
            CUresult r = encode(&maps.m[g], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 5, (void *)(in + (long long)g * 8190 - 2), dims,
                                strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d (loop %d swz %d)\n", (int)r, g, swz); ok = false; }
            r = encode(&omaps.m[g], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 5, (void *)(out + 32 + (long long)g * 8190 - 2), dims,
                       strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled (store) failed: %d\n", (int)r); ok = false; }
        }
        if (!ok) continue;
        CK(cudaFuncSetAttribute(k_load_tma5d, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CK(cudaFuncSetAttribute(k_store_tma5d, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        {
            CK(cudaMemset(errs, 0, 8));
            k_load_tma5d<<<64, NTHR, smem>>>(maps, in, sink, 64, 1, errs, 1, swz);
            CK(cudaDeviceSynchronize());
            unsigned long long he = 0;
            CK(cudaMemcpy(&he, errs, 8, cudaMemcpyDeviceToHost));
            printf("TMA 5-D layout check (swizzle128 %d): %llu mismatches\n", swz, he);
        }
        for (int lds = 0; lds < 2; ++lds) {
            float ms = time_it([&]() { k_load_tma5d<<<4096, NTHR, smem>>>(maps, in, sink, B, 0, errs, lds, swz); });
            printf("load TMA 5-D tensor swz128=%d lds=%d     grid  4096 : %.3f ms  %.0f GB/s\n", swz, lds, ms, gb / (ms * 1e-3));
        }
        float ms = time_it([&]() { k_store_tma5d<<<4096, NTHR, smem>>>(omaps, B); });
        printf("store TMA 5-D tensor swz128=%d           grid  4096 : %.3f ms  %.0f GB/s\n", swz, ms, gb / (ms * 1e-3));
    }
    {
        CK(cudaMemset(errs, 0, 8));
        k_tmem_stash<<<148, NTHR>>>(sink, 148, errs);
        CK(cudaDeviceSynchronize());
        unsigned long long he = 0;
        CK(cudaMemcpy(&he, errs, 8, cudaMemcpyDeviceToHost));
        printf("TMEM stash check: %llu mismatches\n", he);
        float ms = time_it([&]() { k_tmem_stash<<<148, NTHR>>>(sink, B, errs); });
        printf("TMEM stash 196 KB st + ld per state, %d states on 148 CTAs: %.3f ms\n", B, ms);
    }
    return 0;
}
