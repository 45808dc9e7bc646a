// fp64_peak.cu -- measured FP64 (DFMA) issue peak of the GPU this runs on: the secondary roofline of the push kernels
// ("FP64 pipe is the secondary bound and must be reported").  Every thread runs ILP independent DFMA
// chains; 8 warps x 4 blocks per SM keep the FP64 pipe saturated.  Prints one JSON line:
//   burst:     best of 10 launches of ~20 ms each (clock at its boost value),
//   sustained: mean over back-to-back launches for --seconds S (default 4; the part then sits on its power cap).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

template <int ILP>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
    double v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; k++) v[k] = (double)threadIdx.x * 1e-3 + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < ILP; k++) v[k] = fma(v[k], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < ILP; k++) s += v[k];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the chains alive
}

int main(int argc, char** argv) {
    double seconds = 4.0;
    for (int i = 1; i + 1 < argc; i++) if (!strcmp(argv[i], "--seconds")) seconds = atof(argv[i + 1]);
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256, ILP = 8;
    double* out;
    cudaMalloc(&out, (size_t)blocks * threads * sizeof(double));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 1 << 14;
    k_dfma<ILP><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
    cudaDeviceSynchronize();
    const double fma_per_launch = (double)blocks * threads * ILP * iters;
    double best = 0.0;
    for (int r = 0; r < 10; r++) {
        cudaEventRecord(e0);
        k_dfma<ILP><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double rate = fma_per_launch / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    // sustained: back to back for `seconds`
    float ms1; cudaEventRecord(e0); k_dfma<ILP><<<blocks, threads>>>(out, iters, 0.999999, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms1, e0, e1);
    const int n = (int)(seconds * 1e3 / ms1) + 1;
    // the second half of the run is the sustained figure (the power cap has bitten by then)
    cudaEvent_t h0; cudaEventCreate(&h0);
    for (int i = 0; i < n / 2; i++) k_dfma<ILP><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
    cudaEventRecord(h0);
    for (int i = n / 2; i < n; i++) k_dfma<ILP><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float msh; cudaEventElapsedTime(&msh, h0, e1);
    const double sustained = fma_per_launch * (n - n / 2) / (msh * 1e-3);
    if (cudaGetLastError() != cudaSuccess) { fprintf(stderr, "cuda error\n"); return 1; }
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_per_s_burst\": %.4e, \"dfma_per_s_sustained\": %.4e, \"fp64_tflops_burst\": %.3f, "
           "\"fp64_tflops_sustained\": %.3f, \"dfma_per_clk_per_sm_at_max_clock\": %.2f, \"sm_max_mhz\": %d, \"sustained_seconds\": %.1f}\n",
           p.name, sms, best, sustained, 2e-12 * best, 2e-12 * sustained, best / ((double)p.clockRate * 1e3) / sms, p.clockRate / 1000, seconds / 2);
    return 0;
}
