// rt_peak.cu -- FP32 FMA peak of this GPU, measured: register-resident FFMA chains on every SM.
// This is the denominator of the roofline fraction of the brute-force intersection kernels
// (MEASURED_PEAKS.json carries HBM and BF16 tensor peaks only).
#include "rt_kernels.cuh"

namespace rtb {

#define PEAK_CHAINS 16
#define PEAK_ITERS  2048
#define PEAK_THREADS 512

__global__ void __launch_bounds__(PEAK_THREADS) k_ffma_chain(float* sink, float a, float b)
{
    float x[PEAK_CHAINS];
    #pragma unroll
    for (int i = 0; i < PEAK_CHAINS; ++i) x[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < PEAK_ITERS; ++it) {
        #pragma unroll
        for (int r = 0; r < 4; ++r) {
            #pragma unroll
            for (int i = 0; i < PEAK_CHAINS; ++i) x[i] = __fmaf_rn(x[i], a, b);
        }
    }
    float s = 0.f;
    #pragma unroll
    for (int i = 0; i < PEAK_CHAINS; ++i) s += x[i];
    if (s == 123.456f) sink[0] = s;            // never true; keeps the chains alive
}

cudaError_t run_fp32_peak(int sm_count, double* tflops, cudaStream_t st)
{
    float* sink = nullptr;
    cudaError_t e = cudaMalloc(&sink, 64);
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sm_count * 4;           // 4 x 512 threads = 2048 resident threads per SM
    const double flop = (double)blocks * PEAK_THREADS * (double)PEAK_ITERS * 4.0 * PEAK_CHAINS * 2.0;
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0, st);
        k_ffma_chain<<<blocks, PEAK_THREADS, 0, st>>>(sink, 0.999f, 1e-4f);
        cudaEventRecord(e1, st);
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 1 && ms > 0.f) best = fmax(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(sink);
    if (e == cudaSuccess) e = cudaGetLastError();
    *tflops = best;
    return e;
}

} // namespace rtb
