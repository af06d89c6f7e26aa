// rt_comm.cu -- multi-GPU: one context per GPU (one process per GPU), NVLink / NVSwitch.
// The image is sharded in interleaved row bands (rank = band mod n_ranks, balances sky /
// geometry / mirror cost), the scene is replicated, and the only exchange is the framebuffer
// reaching rank 0.  Two ways:
//   * peer frame (default where the GPUs have peer access): rank 0 exports its full-frame buffers
//     with CUDA IPC, every other rank maps them, and the kernels that PRODUCE the pixels (the
//     fused resident level kernel / k_resolve) store their rows straight into rank 0's memory over
//     NVLink -- compute and "gather" are one kernel, no staging buffer, no de-interleave pass.  One
//     tiny stream-ordered ncclAllReduce tells rank 0 that every rank's stores have landed.
//   * fallback: grouped ncclSend/ncclRecv of each rank's compact band buffer + a de-interleave kernel.
// NCCL is loaded lazily with dlopen so that single-GPU use has no NCCL dependency.
#include <dlfcn.h>
#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>

#include "rt_context.h"

namespace rtb {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi* nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return nullptr;
        api.handle = h;
        bool ok = true;
#define LOAD(field, sym) do { *(void**)(&api.field) = dlsym(h, sym); ok = ok && api.field; } while (0)
        LOAD(GetUniqueId, "ncclGetUniqueId");
        LOAD(CommInitRank, "ncclCommInitRank");
        LOAD(CommDestroy, "ncclCommDestroy");
        LOAD(Send, "ncclSend");
        LOAD(Recv, "ncclRecv");
        LOAD(Broadcast, "ncclBroadcast");
        LOAD(AllReduce, "ncclAllReduce");
        LOAD(GroupStart, "ncclGroupStart");
        LOAD(GroupEnd, "ncclGroupEnd");
        LOAD(GetErrorString, "ncclGetErrorString");
#undef LOAD
        if (!ok) { api.handle = nullptr; }
    }
    return api.handle ? &api : nullptr;
}

struct Comm {
    ncclComm_t comm = nullptr;
    DevBuf recv_f32, recv_rgba8;     // rank 0: compact buffers of ranks 1..n-1, concatenated
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    // peer frame: rank 0's full-frame buffers mapped into every rank (CUDA IPC over NVLink)
    bool   p2p_tried = false, p2p = false;
    void*  map_rgba8 = nullptr;      // ranks > 0: rank 0's full RGBA8 frame in this process
    void*  map_f32 = nullptr;        //            rank 0's full float frame (only when requested)
    size_t cap_rgba8 = 0, cap_f32 = 0;
    DevBuf sync;                     // 2 ints: all-reduce in / out
    DevBuf handles;                  // 2 IPC handles travelling through ncclBroadcast
};

static int nccl_fail(rtb200_ctx* c, ncclResult_t r, const char* where)
{
    NcclApi* a = nccl_api();
    char buf[256];
    snprintf(buf, sizeof buf, "%s: %s", where, a && a->GetErrorString ? a->GetErrorString(r) : "nccl error");
    if (c) c->last_error = buf;
    return RTB200_E_NCCL;
}

int comm_unique_id(uint8_t* id)
{
    if (!id) return RTB200_E_ARG;
    NcclApi* a = nccl_api();
    if (!a) return RTB200_E_NCCL;
    static_assert(sizeof(ncclUniqueId) <= RTB200_UNIQUE_ID_BYTES, "unique id size");
    ncclUniqueId u;
    if (a->GetUniqueId(&u) != ncclSuccess) return RTB200_E_NCCL;
    memset(id, 0, RTB200_UNIQUE_ID_BYTES);
    memcpy(id, &u, sizeof u);
    return RTB200_OK;
}

int comm_init(rtb200_ctx* c, int rank, int n_ranks, const uint8_t* id)
{
    comm_destroy(c);
    c->rank = 0; c->n_ranks = 1;
    if (n_ranks == 1) return RTB200_OK;
    NcclApi* a = nccl_api();
    if (!a) { c->last_error = "libnccl.so.2 not loadable"; return RTB200_E_NCCL; }
    if (cudaSetDevice(c->device) != cudaSuccess) return set_cuda_error(c, cudaGetLastError(), "cudaSetDevice");
    Comm* cm = new Comm();
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    ncclResult_t r = a->CommInitRank(&cm->comm, n_ranks, u, rank);
    if (r != ncclSuccess) { delete cm; return nccl_fail(c, r, "ncclCommInitRank"); }
    cudaEventCreate(&cm->e0); cudaEventCreate(&cm->e1);
    c->comm = cm; c->rank = rank; c->n_ranks = n_ranks;
    return RTB200_OK;
}

static void peer_unmap(Comm* cm)
{
    if (cm->map_rgba8) { cudaIpcCloseMemHandle(cm->map_rgba8); cm->map_rgba8 = nullptr; }
    if (cm->map_f32) { cudaIpcCloseMemHandle(cm->map_f32); cm->map_f32 = nullptr; }
    (void)cudaGetLastError();
}

void comm_destroy(rtb200_ctx* c)
{
    if (!c || !c->comm) return;
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    if (c->stream) cudaStreamSynchronize(c->stream);
    peer_unmap(cm);
    cm->sync.release(); cm->handles.release();
    if (a && cm->comm) a->CommDestroy(cm->comm);
    cm->recv_f32.release(); cm->recv_rgba8.release();
    if (cm->e0) cudaEventDestroy(cm->e0);
    if (cm->e1) cudaEventDestroy(cm->e1);
    delete cm;
    c->comm = nullptr; c->rank = 0; c->n_ranks = 1;
}

static int rows_of_rank(int H, int band, int n, int r)
{
    int rows = 0;
    for (int y0 = r * band; y0 < H; y0 += band * n) rows += std::min(band, H - y0);
    return rows;
}

// full[y] <- compact buffer of the rank that owns row y
__global__ void __launch_bounds__(256) k_deinterleave(int W, int H, int band, int n_ranks,
                                                      const float* own_f32, const uchar4* own_u8,
                                                      const float* recv_f32, const uchar4* recv_u8,
                                                      const long long* rank_row_off,   // first row of rank r inside recv (r >= 1)
                                                      float* full_f32, uchar4* full_u8)
{
    const int y = blockIdx.y;
    const int b = y / band, r = b % n_ranks;
    const int lrow = (b / n_ranks) * band + (y - b * band);
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < W; x += gridDim.x * blockDim.x) {
        const size_t dst = (size_t)y * W + x;
        if (r == 0) {
            const size_t src = (size_t)lrow * W + x;
            if (full_f32) { full_f32[3 * dst] = own_f32[3 * src]; full_f32[3 * dst + 1] = own_f32[3 * src + 1]; full_f32[3 * dst + 2] = own_f32[3 * src + 2]; }
            if (full_u8) full_u8[dst] = own_u8[src];
        } else {
            const size_t src = (size_t)(rank_row_off[r] + lrow) * W + x;
            if (full_f32) { full_f32[3 * dst] = recv_f32[3 * src]; full_f32[3 * dst + 1] = recv_f32[3 * src + 1]; full_f32[3 * dst + 2] = recv_f32[3 * src + 2]; }
            if (full_u8) full_u8[dst] = recv_u8[src];
        }
    }
}

int comm_gather_frame(rtb200_ctx* c, int W, int H, int band, bool want_f32, bool want_u8, double* ms)
{
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    if (!a || !cm) return RTB200_E_STATE;
    const int n = c->n_ranks, me = c->rank;
    cudaError_t e;
    std::vector<long long> off((size_t)n + 1, 0);
    long long acc = 0;
    for (int r = 1; r < n; ++r) { off[r] = acc; acc += rows_of_rank(H, band, n, r); }
    const int my_rows = rows_of_rank(H, band, n, me);
    cudaEventRecord(cm->e0, c->stream);
    ncclResult_t r;
    if (me == 0) {
        if (want_f32 && (e = cm->recv_f32.ensure(sizeof(float) * 3 * (size_t)std::max<long long>(acc, 1) * W)) != cudaSuccess) return set_cuda_error(c, e, "recv_f32");
        if (want_u8 && (e = cm->recv_rgba8.ensure(4 * (size_t)std::max<long long>(acc, 1) * W)) != cudaSuccess) return set_cuda_error(c, e, "recv_rgba8");
        if (want_f32 && (e = c->full_f32.ensure(sizeof(float) * 3 * (size_t)W * H)) != cudaSuccess) return set_cuda_error(c, e, "full_f32");
        if (want_u8 && (e = c->full_rgba8.ensure(4 * (size_t)W * H)) != cudaSuccess) return set_cuda_error(c, e, "full_rgba8");
        if ((r = a->GroupStart()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupStart");
        for (int s = 1; s < n; ++s) {
            const size_t rows = (size_t)rows_of_rank(H, band, n, s);
            if (!rows) continue;
            if (want_f32 && (r = a->Recv(cm->recv_f32.as<float>() + 3 * (size_t)off[s] * W, rows * W * 3, ncclFloat, s, cm->comm, c->stream)) != ncclSuccess) return nccl_fail(c, r, "ncclRecv");
            if (want_u8 && (r = a->Recv(cm->recv_rgba8.as<unsigned char>() + 4 * (size_t)off[s] * W, rows * W * 4, ncclUint8, s, cm->comm, c->stream)) != ncclSuccess) return nccl_fail(c, r, "ncclRecv");
        }
        if ((r = a->GroupEnd()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupEnd");
        // rank_row_off lives in the tail of the stats buffer's pinned mirror -> use a small device copy
        DevBuf& cnt = c->counters;
        if ((e = cnt.ensure(sizeof(long long) * ((size_t)n + 1) + 64)) != cudaSuccess) return set_cuda_error(c, e, "offsets");
        long long* h_off = reinterpret_cast<long long*>(c->h_small.as<unsigned char>() + 6144);
        memcpy(h_off, off.data(), sizeof(long long) * ((size_t)n + 1));
        if ((e = cudaMemcpyAsync(cnt.p, h_off, sizeof(long long) * ((size_t)n + 1), cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "offsets copy");
        dim3 grid((unsigned)std::min((W + 255) / 256, 64), (unsigned)H);
        k_deinterleave<<<grid, 256, 0, c->stream>>>(W, H, band, n, c->fb_f32.as<float>(), c->fb_rgba8.as<uchar4>(),
                                                    cm->recv_f32.as<float>(), cm->recv_rgba8.as<uchar4>(), cnt.as<long long>(),
                                                    want_f32 ? c->full_f32.as<float>() : nullptr, want_u8 ? c->full_rgba8.as<uchar4>() : nullptr);
        if ((e = cudaGetLastError()) != cudaSuccess) return set_cuda_error(c, e, "k_deinterleave");
    } else {
        if ((r = a->GroupStart()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupStart");
        if (my_rows) {
            if (want_f32 && (r = a->Send(c->fb_f32.p, (size_t)my_rows * W * 3, ncclFloat, 0, cm->comm, c->stream)) != ncclSuccess) return nccl_fail(c, r, "ncclSend");
            if (want_u8 && (r = a->Send(c->fb_rgba8.p, (size_t)my_rows * W * 4, ncclUint8, 0, cm->comm, c->stream)) != ncclSuccess) return nccl_fail(c, r, "ncclSend");
        }
        if ((r = a->GroupEnd()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupEnd");
    }
    cudaEventRecord(cm->e1, c->stream);
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return set_cuda_error(c, e, "gather sync");
    float t = 0.f;
    cudaEventElapsedTime(&t, cm->e0, cm->e1);
    if (ms) *ms = t;
    return RTB200_OK;
}

// ---------------------------------------------------------------------------------------
// Peer frame
// ---------------------------------------------------------------------------------------
// min over ranks of `v` (stream-ordered, then waited for): agreement on success / failure
static int all_min(rtb200_ctx* c, int v, int* out)
{
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    cudaError_t e;
    if ((e = cm->sync.ensure(2 * sizeof(int))) != cudaSuccess) return set_cuda_error(c, e, "sync buffer");
    if ((e = cudaMemcpyAsync(cm->sync.p, &v, sizeof(int), cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "sync h2d");
    ncclResult_t r = a->AllReduce(cm->sync.p, cm->sync.as<int>() + 1, 1, ncclInt, ncclMin, cm->comm, c->stream);
    if (r != ncclSuccess) return nccl_fail(c, r, "ncclAllReduce");
    int res = 0;
    if ((e = cudaMemcpyAsync(&res, cm->sync.as<int>() + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "sync d2h");
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return set_cuda_error(c, e, "sync");
    *out = res;
    return RTB200_OK;
}

// Collective.  Makes rank 0's full-frame buffers (W x H RGBA8, plus float RGB when asked) visible
// in every rank.  On return *rgba8 / *f32 are the pointers THIS rank's kernels must store to
// (rank 0: its own buffers), or p2p is off for good and the caller uses the NCCL gather.
int comm_peer_frame(rtb200_ctx* c, int W, int H, bool want_f32, void** rgba8, void** f32)
{
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    *rgba8 = nullptr; *f32 = nullptr;
    if (!a || !cm) return RTB200_E_STATE;
    if (cm->p2p_tried && !cm->p2p) return RTB200_OK;
    if (getenv("RTB200_NO_P2P")) { cm->p2p_tried = true; cm->p2p = false; return RTB200_OK; }
    const size_t need8 = 4 * (size_t)W * H, need32 = want_f32 ? sizeof(float) * 3 * (size_t)W * H : 0;
    const int me = c->rank;
    int rc, agreed = 0;
    if (!(cm->p2p && need8 <= cm->cap_rgba8 && need32 <= cm->cap_f32)) {
        // (re)establish the mapping: peers let go of the old buffers before rank 0 reallocates them
        cm->p2p_tried = true;
        if (me != 0) peer_unmap(cm);
        if ((rc = all_min(c, 1, &agreed)) != RTB200_OK) return rc;
        cudaError_t e = cudaSuccess;
        cudaIpcMemHandle_t h[2];
        memset(h, 0, sizeof h);
        int ok = 1;
        if (me == 0) {
            if ((e = c->full_rgba8.ensure(need8)) != cudaSuccess) ok = 0;
            if (ok && want_f32 && (e = c->full_f32.ensure(need32)) != cudaSuccess) ok = 0;
            if (ok && cudaIpcGetMemHandle(&h[0], c->full_rgba8.p) != cudaSuccess) ok = 0;
            if (ok && want_f32 && cudaIpcGetMemHandle(&h[1], c->full_f32.p) != cudaSuccess) ok = 0;
            (void)cudaGetLastError();
        }
        if ((e = cm->handles.ensure(sizeof h)) != cudaSuccess) return set_cuda_error(c, e, "handle buffer");
        if (me == 0 && (e = cudaMemcpyAsync(cm->handles.p, h, sizeof h, cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle h2d");
        ncclResult_t r = a->Broadcast(cm->handles.p, cm->handles.p, sizeof h, ncclUint8, 0, cm->comm, c->stream);
        if (r != ncclSuccess) return nccl_fail(c, r, "ncclBroadcast");
        if (me != 0) {
            if ((e = cudaMemcpyAsync(h, cm->handles.p, sizeof h, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle d2h");
            if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle sync");
            if (cudaIpcOpenMemHandle(&cm->map_rgba8, h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cm->map_rgba8 = nullptr; }
            if (ok && want_f32 && cudaIpcOpenMemHandle(&cm->map_f32, h[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cm->map_f32 = nullptr; }
            (void)cudaGetLastError();
        }
        if ((rc = all_min(c, ok, &agreed)) != RTB200_OK) return rc;
        if (!agreed) {                       // some rank cannot reach rank 0's memory: NCCL gather from now on
            if (me != 0) peer_unmap(cm);
            cm->p2p = false; cm->cap_rgba8 = cm->cap_f32 = 0;
            return RTB200_OK;
        }
        cm->p2p = true;
        cm->cap_rgba8 = need8;
        cm->cap_f32 = need32;
    }
    *rgba8 = me == 0 ? c->full_rgba8.p : cm->map_rgba8;
    *f32 = want_f32 ? (me == 0 ? c->full_f32.p : cm->map_f32) : nullptr;
    return RTB200_OK;
}

// After the kernels that stored into rank 0's frame: one tiny all-reduce, ordered after them on
// every rank's stream.  When it completes on rank 0, every rank's stores have landed.
int comm_peer_frame_done(rtb200_ctx* c, double* ms, bool wait, cudaStream_t on)
{
    cudaStream_t s = on ? on : c->stream;
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    if (!a || !cm) return RTB200_E_STATE;
    cudaError_t e;
    if ((e = cm->sync.ensure(2 * sizeof(int))) != cudaSuccess) return set_cuda_error(c, e, "sync buffer");
    if (wait) cudaEventRecord(cm->e0, s);
    ncclResult_t r = a->AllReduce(cm->sync.p, cm->sync.as<int>() + 1, 1, ncclInt, ncclMin, cm->comm, s);
    if (r != ncclSuccess) return nccl_fail(c, r, "ncclAllReduce");
    if (!wait) return RTB200_OK;
    cudaEventRecord(cm->e1, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return set_cuda_error(c, e, "peer frame sync");
    float t = 0.f;
    cudaEventElapsedTime(&t, cm->e0, cm->e1);
    if (ms) *ms = t;
    return RTB200_OK;
}

} // namespace rtb
