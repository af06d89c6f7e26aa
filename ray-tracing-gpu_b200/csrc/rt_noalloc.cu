// rt_noalloc.cu -- rtb200_trace_nearest_static: the nearest-hit ray query of rtb200_trace_nearest with its scratch in
// STATIC device memory, for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it).  No counterpart in the reference.
//
// rtb200_trace_nearest stages the caller's host arrays in growable device buffers (rtb200_ctx::q_in0 .. q_out2,
// cudaMalloc on first use / growth).  Here the same five buffers are slices of ONE file-scope
//     static __device__ __align__(256) unsigned char g_query_scratch[RTB200_STATIC_QUERY_MAX * 44]     (2.75 MB)
// which every GPU that loads the library holds until the process exits.  The entry point calls no cudaMalloc, cudaFree
// or allocating library routine: two host-to-device copies, the query kernel of the resident path (rt_resident.cu,
// unchanged), three device-to-host copies, one stream synchronisation.
//
// In host code `g_query_scratch` is a host-side stand-in for the symbol, NOT a device address; the address is taken
// once per device, with that device current, by cudaGetSymbolAddress (scratch_base) and handed to the kernel.
//
// Every call on a device shares the array, and it keeps the last call's contents.  Nothing in it has to start at zero
// (inputs [0, n) are overwritten by the copies, outputs [0, n) by the kernel, nothing beyond n is read), so nothing is
// cleared.  Calls that overlap on the same GPU -- two contexts on one device used from two threads -- RACE on it:
// serialise them.
#include <mutex>

#include "rt_context.h"

using namespace rtb;

#define STATIC_MAX_DEVICES 64
#define STATIC_BYTES_PER_RAY (12 + 12 + 4 + 4 + 12)        // origin, direction | order, t, normal

static __device__ __align__(256) unsigned char g_query_scratch[(size_t)RTB200_STATIC_QUERY_MAX * STATIC_BYTES_PER_RAY];

// device address of the scratch array on `device` (which must be current); cached per device
static cudaError_t scratch_base(int device, unsigned char** out)
{
    static std::mutex mu;
    static unsigned char* base[STATIC_MAX_DEVICES] = {nullptr};
    std::lock_guard<std::mutex> lock(mu);
    if (device < 0 || device >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    if (!base[device]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_query_scratch);
        if (e != cudaSuccess) return e;
        base[device] = p;
    }
    *out = base[device];
    return cudaSuccess;
}

#define NCU(call)                                                        \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) return set_cuda_error(ctx, e__, #call);  \
    } while (0)

extern "C" int rtb200_trace_nearest_static(rtb200_ctx* ctx, int n, const float* o3, const float* d3, uint32_t flags,
                                           int32_t* out_order, float* out_t, float* out_n3)
{
    if (!ctx || n < 0 || (n && (!o3 || !d3 || !out_order || !out_t || !out_n3))) return RTB200_E_ARG;
    if (!ctx->have_scene) { ctx->last_error = "no scene"; return RTB200_E_STATE; }
    if (n > RTB200_STATIC_QUERY_MAX) {
        ctx->last_error = "rtb200_trace_nearest_static: n exceeds RTB200_STATIC_QUERY_MAX (65536 rays per call); split the batch or use rtb200_trace_nearest";
        return RTB200_E_ARG;
    }
    // only the scene-resident query kernel works on caller-sized scratch alone: the tiled / culled / LBVH paths keep
    // per-scene structures and per-query lists of their own
    // (the path selection of rtb200_trace_nearest, rt_api.cu)
    const bool q_stream = ((size_t)ctx->ds.rl.bytes > 64 * 1024 || (flags & (RTB200_F_FORCE_STREAM | RTB200_F_USE_BVH))) && !(flags & RTB200_F_FORCE_RESIDENT);
    if (q_stream) {
        ctx->last_error = "rtb200_trace_nearest_static: only scenes of the scene-resident path (packed scene <= 64 KB or RTB200_F_FORCE_RESIDENT; no FORCE_STREAM / USE_BVH)";
        return RTB200_E_ARG;
    }
    if (n == 0) return RTB200_OK;
    NCU(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    NCU(scratch_base(ctx->device, &base));
    const size_t N = (size_t)RTB200_STATIC_QUERY_MAX, b3 = sizeof(float) * 3 * (size_t)n;
    float* in_o   = reinterpret_cast<float*>(base);
    float* in_d   = reinterpret_cast<float*>(base + 12 * N);
    int*   o_ord  = reinterpret_cast<int*>(base + 24 * N);
    float* o_t    = reinterpret_cast<float*>(base + 28 * N);
    float* o_n    = reinterpret_cast<float*>(base + 32 * N);
    NCU(cudaMemcpyAsync(in_o, o3, b3, cudaMemcpyHostToDevice, ctx->stream));
    NCU(cudaMemcpyAsync(in_d, d3, b3, cudaMemcpyHostToDevice, ctx->stream));
    NCU(launch_query_nearest(ctx->ds, n, in_o, in_d, o_ord, o_t, o_n, (flags & RTB200_F_FUSED_FP) == 0,
                             (flags & RTB200_F_EXACT_FP) == 0, ctx->stream));
    NCU(cudaMemcpyAsync(out_order, o_ord, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    NCU(cudaMemcpyAsync(out_t, o_t, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    NCU(cudaMemcpyAsync(out_n3, o_n, b3, cudaMemcpyDeviceToHost, ctx->stream));
    NCU(cudaStreamSynchronize(ctx->stream));
    return RTB200_OK;
}
