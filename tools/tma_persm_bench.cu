// tools/tma_persm_bench.cu -- how fast can ONE SM move 196 KB with cp.async.bulk, as a function of how many SMs do it at once?
// The version-4 fused kernel (chb_fused4.cuh) moves 196 KB in and 196 KB out per state per SM; this probe separates the per-SM
// port limit from the chip-wide limits.  Source region is small (L2 resident) or large (HBM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/tma_persm_bench tools/tma_persm_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)
constexpr long long NPARAMS = 24570;
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// mode 0: loads only, 1: stores only, 2: store then load (dependent, like the kernel).  nrun copies per loop.
__global__ void __launch_bounds__(384, 1) k_tma(const double *in, double *out, int reps, long long nstates_wrap, int mode, int nrun) {
    extern __shared__ __align__(1024) double sm[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    unsigned phase = 0;
    const unsigned per = 65520u / nrun / 16u * 16u;
    for (int r = 0; r < reps; ++r) {
        const long long b = ((long long)blockIdx.x + (long long)r * gridDim.x) % nstates_wrap;
        if (threadIdx.x == 0) {
            if (mode >= 1) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                for (int g = 0; g < 3; ++g)
                    for (int c = 0; c < nrun; ++c)
                        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                     ::"l"(out + b * NPARAMS + g * 8190 + c * (per / 8)), "r"(smem_u32(sm + g * 8192 + c * (per / 8))), "r"(per) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            if (mode != 1) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(3u * per * nrun));
                for (int g = 0; g < 3; ++g)
                    for (int c = 0; c < nrun; ++c)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(smem_u32(sm + g * 8192 + c * (per / 8))), "l"(in + b * NPARAMS + g * 8190 + c * (per / 8)), "r"(per), "r"(smem_u32(&bar)) : "memory");
            }
        }
        if (mode != 1) {
            unsigned done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
            phase ^= 1;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
    const int B = 4096;
    const size_t smem = 200 * 1024;
    double *in, *out;
    CK(cudaMalloc(&in, (size_t)B * NPARAMS * 8 + 1024));
    CK(cudaMalloc(&out, (size_t)B * NPARAMS * 8 + 1024));
    CK(cudaMemset(in, 0, (size_t)B * NPARAMS * 8));
    CK(cudaFuncSetAttribute(k_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char *mname[3] = {"load", "store", "store->load"};
    for (int nrun : {1, 8, 32})
    for (long long wrap : {64LL, 4096LL})
        for (int mode = 0; mode < 3; ++mode)
            for (int grid : {1, 8, 37, 74, 148}) {
                const int reps = 200;
                k_tma<<<grid, 384, smem>>>(in, out, 20, wrap, mode, nrun);
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0));
                k_tma<<<grid, 384, smem>>>(in, out, reps, wrap, mode, nrun);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaGetLastError());
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double us_per = ms * 1e3 / reps;
                const double bytes = 3.0 * (65520 / nrun / 16 * 16) * nrun * (mode == 2 ? 2 : 1);
                printf("%-12s runs/loop %2d region %4lld states (%s) grid %3d : %6.2f us per state per SM, %6.1f GB/s per SM, %7.0f GB/s chip\n",
                       mname[mode], nrun, wrap, wrap == 64 ? "L2" : "HBM", grid, us_per, bytes / us_per * 1e-3, bytes / us_per * 1e-3 * grid);
            }
    return 0;
}
