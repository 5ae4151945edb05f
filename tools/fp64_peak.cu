// tools/fp64_peak.cu -- measures the denominators MEASURED_PEAKS.json does not hold:
//   * FP64 FMA peak (non-tensor DFMA pipe), 8 independent chains per thread
//   * pinned host<->device copy bandwidth, each direction alone and both together (PCIe ceiling of the e2e leg)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; } } while (0)

__global__ void dfma_kernel(double *out, double a, double b, int iters) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

int main() {
    cudaDeviceProp pr;
    CK(cudaGetDeviceProperties(&pr, 0));
    const int threads = 512, blocks = pr.multiProcessorCount * 4, iters = 4096;
    double *out;
    CK(cudaMalloc(&out, sizeof(double) * threads * blocks));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(e0));
        dfma_kernel<<<blocks, threads>>>(out, 0.999999, 1e-9, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double flops = 2.0 * 64.0 * iters * (double)threads * blocks;
    const double tf = flops / (best * 1e-3) / 1e12;
    // sustained: back to back for ~2 s
    CK(cudaEventRecord(e0));
    int nrep = (int)(2000.0f / best) + 1;
    for (int rep = 0; rep < nrep; ++rep) dfma_kernel<<<blocks, threads>>>(out, 0.999999, 1e-9, iters);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms_s; CK(cudaEventElapsedTime(&ms_s, e0, e1));
    const double tf_s = flops * nrep / (ms_s * 1e-3) / 1e12;

    // PCIe
    const size_t nbytes = (size_t)512 << 20;
    void *h0, *h1, *d0, *d1;
    CK(cudaMallocHost(&h0, nbytes)); CK(cudaMallocHost(&h1, nbytes));
    CK(cudaMalloc(&d0, nbytes)); CK(cudaMalloc(&d1, nbytes));
    cudaStream_t s0, s1;
    CK(cudaStreamCreate(&s0)); CK(cudaStreamCreate(&s1));
    double h2d = 0, d2h = 0, both = 0;
    for (int rep = 0; rep < 3; ++rep) {
        float ms;
        CK(cudaEventRecord(e0, s0)); CK(cudaMemcpyAsync(d0, h0, nbytes, cudaMemcpyHostToDevice, s0)); CK(cudaEventRecord(e1, s0));
        CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        if (nbytes / (ms * 1e-3) / 1e9 > h2d) h2d = nbytes / (ms * 1e-3) / 1e9;
        CK(cudaEventRecord(e0, s0)); CK(cudaMemcpyAsync(h1, d1, nbytes, cudaMemcpyDeviceToHost, s0)); CK(cudaEventRecord(e1, s0));
        CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        if (nbytes / (ms * 1e-3) / 1e9 > d2h) d2h = nbytes / (ms * 1e-3) / 1e9;
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, 0));
        CK(cudaMemcpyAsync(d0, h0, nbytes, cudaMemcpyHostToDevice, s0));
        CK(cudaMemcpyAsync(h1, d1, nbytes, cudaMemcpyDeviceToHost, s1));
        CK(cudaStreamSynchronize(s0)); CK(cudaStreamSynchronize(s1));
        CK(cudaEventRecord(e1, 0)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        if (nbytes / (ms * 1e-3) / 1e9 > both) both = nbytes / (ms * 1e-3) / 1e9;
    }
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"fp64_fma_tflops_burst\": %.2f, \"fp64_fma_tflops_sustained\": %.2f, "
           "\"pinned_h2d_gbs\": %.1f, \"pinned_d2h_gbs\": %.1f, \"pinned_bidir_gbs_per_direction\": %.1f}\n",
           pr.name, pr.multiProcessorCount, tf, tf_s, h2d, d2h, both);
    return 0;
}
