// rt_comm.cu -- multi-GPU: one context per GPU (one process per GPU), NVLink / NVSwitch.
// The image is sharded in interleaved row bands (rank = band mod n_ranks, balances sky /
// geometry / mirror cost), the scene is replicated, and the only exchange is the framebuffer
// reaching rank 0.  Two ways:
//   * peer frame (default where the GPUs have peer access): rank 0 exports its full-frame buffers
//     with CUDA IPC, every other rank maps them, and the kernels that PRODUCE the pixels (the
//     fused resident level kernel / k_resolve) store their rows straight into rank 0's memory over
//     NVLink -- compute and "gather" are one kernel, no staging buffer, no de-interleave pass.  The
//     last CTA of that kernel writes a sequence number into rank 0's flag word (release.sys); rank
//     0's stream waits on its own memory (cuStreamWaitValue32).  No collective per frame.
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
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
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
        LOAD(AllGather, "ncclAllGather");
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
    // peer frame (see the section below): rank 0's double-buffered full frame + flag words mapped
    // into every rank, every rank's ack word mapped into rank 0 (CUDA IPC over NVLink)
    bool   p2p_tried = false, p2p = false;
    void*  map_rgba8 = nullptr;      // ranks > 0: rank 0's RGBA8 arena in this process
    void*  map_f32 = nullptr;        //            rank 0's float arena (only when requested)
    unsigned int* peer_ack[16] = {nullptr};   // rank 0: the ack word of rank r
    size_t stride8 = 0, stride32 = 0;         // bytes of one frame half
    unsigned int seq = 0;            // sequence number of the last peer frame (identical on every rank)
    DevBuf ack, done;                // ranks > 0: ack word written by rank 0; CTA counter of the in-kernel signal
    DevBuf sync;                     // 2 ints: all-reduce in / out
    DevBuf handles;                  // PeerSlot per rank, all-gathered
};

static void peer_unmap(Comm* cm);

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

static int all_min(rtb200_ctx* c, int v, int* out);

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
    // RTB200_NO_P2P is read ONCE, here, and agreed on by all ranks: every later branch of the frame path must depend
    // on replicated state only (a rank that alone skipped the peer frame would desynchronise the frame counter)
    int agreed = 1;
    const int rc = all_min(c, getenv("RTB200_NO_P2P") ? 0 : 1, &agreed);
    if (rc != RTB200_OK) return rc;
    if (!agreed) { cm->p2p_tried = true; cm->p2p = false; }
    return RTB200_OK;
}

void comm_destroy(rtb200_ctx* c)
{
    if (!c || !c->comm) return;
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    peer_unmap(cm);
    cm->sync.release(); cm->handles.release(); cm->ack.release(); cm->done.release();
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
    c->full_off8 = c->full_off32 = 0;
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
        long long* h_off = reinterpret_cast<long long*>(c->h_small.as<unsigned char>() + 12288);
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
// Memory, all plain cudaMalloc exported with CUDA IPC (handles travel once through ncclAllGather):
//   rank 0   full_rgba8 = [ frame half 0 | frame half 1 | flag words, one per rank ]   (+ full_f32, two halves)
//   rank r   ack word (256 B), mapped by rank 0
// Frame `seq` (a counter every rank advances identically) lives in half seq & 1.
//   rank r > 0:  stream waits until ack >= seq - 2        (rank 0 is done with the half about to be overwritten)
//                kernels store their rows into rank 0's half over NVLink
//                the last CTA of the producing kernel writes flag[r] = seq (release.sys)   -- peer_frame_signal()
//   rank 0:      writes ack = seq - 1 into every peer (k_peer_ack; it entered frame seq, so frame seq - 1 is consumed),
//                renders its own rows, then its stream waits for flag[r] >= seq, r = 1..n-1 (cuStreamWaitValue32
//                on its OWN memory), and only then reads the frame.
// No collective and no NCCL kernel on the per-frame path; peers can run at most one frame ahead of
// rank 0's consumer.  Words start at PEER_SEQ0 (no clearing, no second barrier); comparisons are cyclic.
static const unsigned int PEER_SEQ0 = 0x5EED0000u;
static const size_t PEER_FLAG_BYTES = 256;
static const int PEER_MAX_RANKS = 16;

typedef int (*StreamWaitValue32Fn)(cudaStream_t, unsigned long long, unsigned int, unsigned int);   // CUresult cuStreamWaitValue32(CUstream, CUdeviceptr, cuuint32_t, unsigned)

static StreamWaitValue32Fn stream_wait_value32()
{
    static StreamWaitValue32Fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<StreamWaitValue32Fn>(p);
        (void)cudaGetLastError();
    }
    return fn;
}

struct AckTargets { unsigned int* p[PEER_MAX_RANKS]; int n; };

__global__ void k_peer_ack(AckTargets t, unsigned int seq)
{
    const int i = threadIdx.x;
    if (i < t.n && t.p[i]) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(t.p[i]), "r"(seq) : "memory");
    }
}

__global__ void k_peer_signal(unsigned int* flag, unsigned int seq)
{
    // earlier kernels of this stream stored the rows (stream order); make them visible system-wide first
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(flag), "r"(seq) : "memory");
}

__global__ void k_fill_words(unsigned int* w, int n, unsigned int v)
{
    if ((int)threadIdx.x < n) w[threadIdx.x] = v;
}

static size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

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

static void peer_unmap(Comm* cm)
{
    if (cm->map_rgba8) { cudaIpcCloseMemHandle(cm->map_rgba8); cm->map_rgba8 = nullptr; }
    if (cm->map_f32) { cudaIpcCloseMemHandle(cm->map_f32); cm->map_f32 = nullptr; }
    for (int r = 0; r < PEER_MAX_RANKS; ++r)
        if (cm->peer_ack[r]) { cudaIpcCloseMemHandle(cm->peer_ack[r]); cm->peer_ack[r] = nullptr; }
    (void)cudaGetLastError();
}

struct PeerSlot {                     // what one rank contributes to the handle all-gather
    cudaIpcMemHandle_t h[3];          // rank 0: RGBA8 arena, float arena; ranks > 0: [2] = ack word
    int ok;
    int pad[15];
};
static_assert(sizeof(PeerSlot) == 256, "PeerSlot layout");

// (Re)establish the mapping for frames of at least need8 / need32 bytes.  Collective.
static int peer_establish(rtb200_ctx* c, size_t need8, size_t need32)
{
    NcclApi* a = nccl_api();
    Comm* cm = c->comm;
    const int me = c->rank, n = c->n_ranks;
    int rc, agreed = 0;
    cudaError_t e = cudaSuccess;
    cm->p2p_tried = true;
    cm->p2p = false;
    // nobody may still be using the old mapping; peers let go before rank 0 reallocates
    cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->copy_stream);
    peer_unmap(cm);
    if ((rc = all_min(c, 1, &agreed)) != RTB200_OK) return rc;

    const size_t stride8 = align256(need8 + need8 / 8), stride32 = need32 ? align256(need32 + need32 / 8) : 0;
    int ok = (n <= PEER_MAX_RANKS && stream_wait_value32() != nullptr) ? 1 : 0;
    PeerSlot mine;
    memset(&mine, 0, sizeof mine);
    if (me == 0) {
        if (ok && c->full_rgba8.ensure(2 * stride8 + PEER_FLAG_BYTES) != cudaSuccess) ok = 0;
        if (ok && stride32 && c->full_f32.ensure(2 * stride32) != cudaSuccess) ok = 0;
        if (ok) k_fill_words<<<1, 64, 0, c->stream>>>(reinterpret_cast<unsigned int*>(c->full_rgba8.as<unsigned char>() + 2 * stride8), (int)(PEER_FLAG_BYTES / 4), PEER_SEQ0);
        if (ok && cudaIpcGetMemHandle(&mine.h[0], c->full_rgba8.p) != cudaSuccess) ok = 0;
        if (ok && stride32 && cudaIpcGetMemHandle(&mine.h[1], c->full_f32.p) != cudaSuccess) ok = 0;
    } else {
        if (ok && (cm->ack.ensure(256) != cudaSuccess || cm->done.ensure(256) != cudaSuccess)) ok = 0;
        if (ok) {
            k_fill_words<<<1, 64, 0, c->stream>>>(cm->ack.as<unsigned int>(), 64, PEER_SEQ0);
            cudaMemsetAsync(cm->done.p, 0, 256, c->stream);
        }
        if (ok && cudaIpcGetMemHandle(&mine.h[2], cm->ack.p) != cudaSuccess) ok = 0;
    }
    (void)cudaGetLastError();
    mine.ok = ok;

    // all-gather of the slots (in place), then back to the host
    std::vector<PeerSlot> all((size_t)n);
    if ((e = cm->handles.ensure(sizeof(PeerSlot) * (size_t)n)) != cudaSuccess) return set_cuda_error(c, e, "handle buffer");
    PeerSlot* d_all = cm->handles.as<PeerSlot>();
    if ((e = cudaMemcpyAsync(d_all + me, &mine, sizeof mine, cudaMemcpyHostToDevice, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle h2d");
    ncclResult_t r = a->AllGather(d_all + me, d_all, sizeof(PeerSlot), ncclUint8, cm->comm, c->stream);
    if (r != ncclSuccess) return nccl_fail(c, r, "ncclAllGather");
    if ((e = cudaMemcpyAsync(all.data(), d_all, sizeof(PeerSlot) * (size_t)n, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle d2h");
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return set_cuda_error(c, e, "handle sync");
    for (int q = 0; q < n; ++q) ok = ok && all[(size_t)q].ok;

    // map, and check every mapping by reading the word the owner initialised through it
    unsigned int probe = 0;
    if (ok && me != 0) {
        if (cudaIpcOpenMemHandle(&cm->map_rgba8, all[0].h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cm->map_rgba8 = nullptr; }
        if (ok && stride32 && cudaIpcOpenMemHandle(&cm->map_f32, all[0].h[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cm->map_f32 = nullptr; }
        if (ok && (cudaMemcpy(&probe, static_cast<unsigned char*>(cm->map_rgba8) + 2 * stride8 + 4 * (size_t)me, 4, cudaMemcpyDeviceToHost) != cudaSuccess || probe != PEER_SEQ0)) ok = 0;
    }
    if (ok && me == 0) {
        for (int q = 1; q < n && ok; ++q) {
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[(size_t)q].h[2], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
            cm->peer_ack[q] = static_cast<unsigned int*>(p);
            if (cudaMemcpy(&probe, p, 4, cudaMemcpyDeviceToHost) != cudaSuccess || probe != PEER_SEQ0) ok = 0;
        }
    }
    (void)cudaGetLastError();
    if ((rc = all_min(c, ok, &agreed)) != RTB200_OK) return rc;
    if (!agreed) {                           // some rank cannot reach another's memory: NCCL gather from now on
        peer_unmap(cm);
        cm->stride8 = cm->stride32 = 0;
        return RTB200_OK;
    }
    cm->p2p = true;
    cm->stride8 = stride8; cm->stride32 = stride32;
    cm->seq = PEER_SEQ0;
    return RTB200_OK;
}

// Start of a frame on every rank.  Collective only when the mapping has to be (re)established (first
// frame, a larger frame, the float frame asked for the first time) -- every rank takes that branch
// together because the decision depends on replicated state only.  pf->rgba8 == null on return:
// p2p is off for good, the caller uses the NCCL gather.
int comm_peer_begin(rtb200_ctx* c, int W, int H, bool want_f32, bool ack_previous, bool wait_in_kernel, PeerFrame* pf, cudaStream_t s)
{
    Comm* cm = c->comm;
    *pf = PeerFrame();
    if (!nccl_api() || !cm) return RTB200_E_STATE;
    if (cm->p2p_tried && !cm->p2p) return RTB200_OK;
    const size_t need8 = 4 * (size_t)W * H, need32 = want_f32 ? sizeof(float) * 3 * (size_t)W * H : 0;
    if (!(cm->p2p && need8 <= cm->stride8 && need32 <= cm->stride32)) {
        const int rc = peer_establish(c, std::max(need8, cm->stride8), want_f32 ? std::max(need32, cm->stride32) : cm->stride32);
        if (rc != RTB200_OK) return rc;
        if (!cm->p2p) return RTB200_OK;
    }
    const unsigned int seq = ++cm->seq;
    const size_t off8 = (seq & 1u) * cm->stride8, off32 = (seq & 1u) * cm->stride32;
    pf->seq = seq;
    if (c->rank == 0) {
        pf->rgba8 = c->full_rgba8.as<unsigned char>() + off8;
        pf->f32 = want_f32 ? c->full_f32.as<unsigned char>() + off32 : nullptr;
        c->full_off8 = off8; c->full_off32 = off32;
        if (ack_previous) return comm_peer_ack(c, seq - 1u, s);
        return RTB200_OK;
    }
    unsigned char* base = static_cast<unsigned char*>(cm->map_rgba8);
    pf->rgba8 = base + off8;
    pf->f32 = want_f32 ? static_cast<unsigned char*>(cm->map_f32) + off32 : nullptr;
    pf->flag = reinterpret_cast<unsigned int*>(base + 2 * cm->stride8) + c->rank;
    pf->done = cm->done.as<unsigned int>();
    // the half this frame overwrites held frame seq - 2: wait until rank 0 has let go of it
    if (wait_in_kernel) { pf->ack = cm->ack.as<unsigned int>(); pf->ack_min = seq - 2u; return RTB200_OK; }
    if (stream_wait_value32()(s, (unsigned long long)(uintptr_t)cm->ack.p, seq - 2u, 0u /* GEQ, cyclic */) != 0) {
        c->last_error = "cuStreamWaitValue32 (ack) failed";
        return RTB200_E_CUDA;
    }
    return RTB200_OK;
}

// ranks > 0 that asked comm_peer_begin not to wait (wait_in_kernel) and then do not use a kernel for it
int comm_peer_wait_ack(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s)
{
    if (!pf.ack) return RTB200_OK;
    if (stream_wait_value32()(s, (unsigned long long)(uintptr_t)pf.ack, pf.ack_min, 0u /* GEQ, cyclic */) != 0) {
        c->last_error = "cuStreamWaitValue32 (ack) failed";
        return RTB200_E_CUDA;
    }
    return RTB200_OK;
}

int comm_peer_signal(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s)
{
    if (!pf.flag) return RTB200_OK;
    k_peer_signal<<<1, 1, 0, s>>>(pf.flag, pf.seq);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? RTB200_OK : set_cuda_error(c, e, "k_peer_signal");
}

int comm_peer_wait(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s)
{
    Comm* cm = c->comm;
    if (!cm || c->rank != 0) return RTB200_OK;
    const unsigned int* flags = reinterpret_cast<const unsigned int*>(c->full_rgba8.as<unsigned char>() + 2 * cm->stride8);
    for (int r = 1; r < c->n_ranks; ++r)
        if (stream_wait_value32()(s, (unsigned long long)(uintptr_t)(flags + r), pf.seq, 0u) != 0) {
            c->last_error = "cuStreamWaitValue32 (flag) failed";
            return RTB200_E_CUDA;
        }
    return RTB200_OK;
}

int comm_peer_ack(rtb200_ctx* c, unsigned int seq, cudaStream_t s)
{
    Comm* cm = c->comm;
    if (!cm || c->rank != 0) return RTB200_OK;
    AckTargets t;
    memset(&t, 0, sizeof t);
    t.n = c->n_ranks;
    for (int r = 1; r < c->n_ranks; ++r) t.p[r] = cm->peer_ack[r];
    k_peer_ack<<<1, 32, 0, s>>>(t, seq);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? RTB200_OK : set_cuda_error(c, e, "k_peer_ack");
}

} // namespace rtb
