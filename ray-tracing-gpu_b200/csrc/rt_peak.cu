// rt_peak.cu -- FP32 FMA peak of this GPU, measured: register-resident FFMA chains on every SM.
// This is the denominator of the roofline fraction of the brute-force intersection kernels
// (MEASURED_PEAKS.json carries HBM and BF16 tensor peaks only).
#include <algorithm>
#include <vector>

#include "rt_kernels.cuh"

namespace rtb {

#define PEAK_CHAINS 16
#define PEAK_ITERS  2048
#define PEAK_THREADS 512

// `clk` (may be null): per block, SM cycles (clock64) and nanoseconds (globaltimer) thread 0 spent in the chain --
// the clock the chain really ran at and the FFMA issue rate per SM can be recomputed from them
__global__ void __launch_bounds__(PEAK_THREADS, 4) k_ffma_chain(float* sink, float a, float b, unsigned long long* clk)
{
    __shared__ unsigned long long s_start[2];       // (kept out of the registers: 4 blocks x 512 threads need <= 32 per thread)
    if (threadIdx.x == 0) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        s_start[0] = (unsigned long long)clock64(); s_start[1] = t0;
    }
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
    if (threadIdx.x == 0 && clk) {
        unsigned long long t1;
        const unsigned long long c1 = (unsigned long long)clock64();
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        clk[2 * blockIdx.x] = c1 - s_start[0] + (s == 123.456f ? 1ull : 0ull);     // (depends on the chains: not hoisted above them)
        clk[2 * blockIdx.x + 1] = t1 - s_start[1];
    }
}

// detail (may be null, 8 doubles): FFMA instructions per thread, threads per block, blocks, best elapsed ms (CUDA
// events), median per-block SM cycles and nanoseconds of that launch, SM clock = cycles / ns (MHz), FFMA per clock
// and SM = all FFMA / (elapsed x that clock x SMs) (the hardware limit is 128)
cudaError_t run_fp32_peak(int sm_count, double* tflops, cudaStream_t st, double* detail)
{
    float* sink = nullptr;
    cudaError_t e = cudaMalloc(&sink, 64);
    if (e != cudaSuccess) return e;
    const int blocks = sm_count * 4;           // 4 x 512 threads = 2048 resident threads per SM
    unsigned long long* clk = nullptr;
    std::vector<unsigned long long> h_clk((size_t)2 * blocks), best_clk;
    if (detail && (e = cudaMalloc(&clk, sizeof(unsigned long long) * 2 * blocks)) != cudaSuccess) { cudaFree(sink); return e; }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double flop = (double)blocks * PEAK_THREADS * (double)PEAK_ITERS * 4.0 * PEAK_CHAINS * 2.0;
    double best = 0.0;
    float best_ms = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0, st);
        k_ffma_chain<<<blocks, PEAK_THREADS, 0, st>>>(sink, 0.999f, 1e-4f, clk);
        cudaEventRecord(e1, st);
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 1 && ms > 0.f && flop / (ms * 1e-3) / 1e12 > best) {
            best = flop / (ms * 1e-3) / 1e12;
            best_ms = ms;
            if (clk && cudaMemcpy(h_clk.data(), clk, sizeof(unsigned long long) * 2 * blocks, cudaMemcpyDeviceToHost) == cudaSuccess) best_clk = h_clk;
        }
    }
    if (detail) {
        std::vector<double> cyc, ns;
        for (size_t i = 0; i + 1 < best_clk.size(); i += 2) { cyc.push_back((double)best_clk[i]); ns.push_back((double)best_clk[i + 1]); }
        std::sort(cyc.begin(), cyc.end()); std::sort(ns.begin(), ns.end());
        const double mc = cyc.empty() ? 0.0 : cyc[cyc.size() / 2], mn = ns.empty() ? 0.0 : ns[ns.size() / 2];
        const double ffma_per_thread = (double)PEAK_ITERS * 4.0 * PEAK_CHAINS;
        detail[0] = ffma_per_thread; detail[1] = PEAK_THREADS; detail[2] = blocks; detail[3] = best_ms;
        detail[4] = mc; detail[5] = mn; detail[6] = mn > 0 ? mc / mn * 1e3 : 0.0;
        // FFMA issued per clock and SM over the WHOLE launch: all FFMA / (elapsed time x measured clock x SMs).  (Per block,
        // 4 x 512 threads x FFMA / block cycles over-states it: the four blocks of an SM do not overlap for their whole life.)
        const double clk_hz = detail[6] * 1.0e6;
        detail[7] = (best_ms > 0.f && clk_hz > 0.0) ? ffma_per_thread * PEAK_THREADS * (double)blocks / (best_ms * 1.0e-3 * clk_hz * sm_count) : 0.0;
    }
    if (clk) cudaFree(clk);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(sink);
    if (e == cudaSuccess) e = cudaGetLastError();
    *tflops = best;
    return e;
}

} // namespace rtb
