// rt_context.h -- the context behind the C ABI (host side of librtb200.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/rtb200.h"
#include "rt_kernels.cuh"

namespace rtb {

// growable device buffer (never shrinks; steady-state frames allocate nothing)
struct DevBuf {
    void*  p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinnedBuf {
    void*  p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct LevelBuf {                 // one wavefront level (rays of one bounce count)
    DevBuf od0, od1, colour, links;
};

struct Comm;                      // rt_bvh.cu
int  bvh_build(rtb200_ctx* c);
void bvh_release(rtb200_ctx* c);
void bvh_invalidate(rtb200_ctx* c);                                         // the scene changed
bool bvh_ready(const rtb200_ctx* c);
int  bvh_prepare(rtb200_ctx* c, const float* origins3, int n);              // pad the boxes for rays starting this far out
// rt_comm.cu (NCCL, loaded lazily with dlopen)

} // namespace rtb

struct rtb200_ctx {
    int device = 0;
    int sm_count = 0, l2_bytes = 0;
    size_t smem_optin = 0;
    char name[128] = {0};
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // 6/7: wait for the peers
    cudaEvent_t timer[4] = {nullptr, nullptr, nullptr, nullptr};   // rtb200_timer_mark
    std::string last_error;

    // scene
    bool have_scene = false;
    rtb::DeviceScene ds{};
    rtb::DevBuf scene_buf;            // resident blob + streamed arrays, one allocation
    rtb::PinnedBuf stage;             // pinned staging for uploads
    rtb::DevBuf raw_buf;              // large scenes: the caller's raw arrays on the device, input of the pack kernel
    std::vector<rtb::QuadRec> h_quad; // host copy of the quadrics and their centred forms (rt_bvh.cu decides which
    std::vector<float4> h_quad_r3;    // of them are bounded and go into the tree)
    int max_children = 0;             // 0, 1 or 2 secondary rays per hit at most
    bool any_reflect = false, any_refract = false;
    double last_h2d_ms = 0.0;
    bool upload_pending = false;      // an asynchronous scene upload may still be reading `stage`

    // frame buffers (local rows of this rank)
    rtb::DevBuf c0, fb_f32, fb_rgba8, stats, counters;
    bool stats_clean = false;         // the device counters were zeroed at the end of the previous frame
    rtb::DevBuf full_f32, full_rgba8; // rank 0 with a communicator: gathered frame (peer frame: two halves + flag words)
    size_t full_off8 = 0, full_off32 = 0;   // byte offset of the LAST frame inside them
    rtb::PinnedBuf h_small;           // counters / stats read-back
    std::vector<rtb::LevelBuf> levels;
    float*  out_f32_dev = nullptr;    // where THIS frame's kernels store pixels: the local band buffers, or
    uchar4* out_rgba8_dev = nullptr;  // rank 0's full frame (own or peer-mapped over NVLink, FrameParams::fb_global)
    // overlapped read-back (one GPU, host output pointers given): finished row chunks travel to the
    // host on a second stream while the next chunk renders
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t  copy_ev[4] = {nullptr, nullptr, nullptr, nullptr};   // 0/1: chunk done (alternating), 2/3: copy begin / end
    float*   host_f32 = nullptr;      // non-null during a render that overlaps its read-back
    uint8_t* host_rgba8 = nullptr;
    int      copy_chunks = 0;         // chunks already handed to the copy stream in this frame
    bool     host_bands = false;      // RTB200_F_HOST_BANDS frame: the host pointers are the FULL frame, this rank writes its bands
    int      host_band_rows = 16;
    // level loop without host round trips (resident path with secondary rays): what the last frame
    // of this shape needed per level, the guess the next one is sized by
    struct LevelGuess {
        bool valid = false;
        unsigned long long scene_serial = 0;
        int W = 0, H = 0, k = 0, max_bounces = 0, children = 0, rows_per_chunk = 0, local_rows = 0;
        std::vector<unsigned> rays;   // [l] = most rays any chunk had at level l (l >= 1)
    } level_guess;
    unsigned long long scene_serial = 0;   // bumped by rtb200_set_scene
    unsigned long long level_fallbacks = 0;   // frames redone through the synchronous level loop (a queue guess was too small)
    rtb::PinnedBuf h_levels;               // per-chunk copies of the level counters
    rtb::DevBuf tail_dev;                  // table of the deep levels one tail-mode launch walks
    rtb::PinnedBuf tail_host;
    rtb::DevBuf seq_fb[2], seq_stats;  // rtb200_render_sequence: double-buffered RGBA8 frames, per-frame counters
    rtb::PinnedBuf seq_host;
    int last_W = 0, last_H = 0, last_local_rows = 0;
    bool last_gathered = false;

    // query scratch
    rtb::DevBuf q_in0, q_in1, q_out0, q_out1, q_out2;

    // multi-GPU
    rtb::Comm* comm = nullptr;
    int rank = 0, n_ranks = 1;
    const void* peer_cur = nullptr;   // rtb::PeerFrame of the frame being rendered (null: none)
    bool peer_signalled = false;      // a kernel of this frame carries the completion signal itself

    // streamed path scratch (rt_stream.cu) and host-side test counts of the last frame
    void* stream_scratch = nullptr;
    void* bvh_state = nullptr;        // rt_bvh.cu
    void* cull_state = nullptr;       // rt_cullscan.cu: Morton-sorted triangles + sphere hierarchy of the current scene
    unsigned long long host_tests[3] = {0, 0, 0};

    // peak cache
    double peak_tflops = 0.0, peak_mhz = 0.0;
};

namespace rtb {
int  set_cuda_error(rtb200_ctx* c, cudaError_t e, const char* where);
int  local_row_count(int H, int band_rows, int n_ranks, int rank);          // rt_api.cu
void fill_frame_params(const rtb200_frame* fr, int n_ranks, int rank, FrameParams& f);
int  chunk_copy_out(rtb200_ctx* c, int W, int py0, int prows);              // rt_api.cu: overlapped read-back of finished rows
int  render_streamed(rtb200_ctx* c, const rtb200_frame* fr, const FrameParams& fp, int local_rows,
                     bool exact, rtb200_stats* st, uint32_t* launches);     // rt_stream.cu
void stream_release(rtb200_ctx* c);
struct CullScene;
int  cull_prepare(rtb200_ctx* c, CullScene* out);                            // rt_cullscan.cu
void cull_release(rtb200_ctx* c);
int  stream_query_nearest(rtb200_ctx* c, int n, uint32_t flags);            // inputs q_in0/q_in1, outputs q_out0..2 (device)
int  stream_query_shadow(rtb200_ctx* c, int n, uint32_t flags);
// rt_comm.cu
int  comm_unique_id(uint8_t* id);
int  comm_init(rtb200_ctx* c, int rank, int n_ranks, const uint8_t* id);
void comm_destroy(rtb200_ctx* c);
int  comm_gather_frame(rtb200_ctx* c, int W, int H, int band_rows, bool want_f32, bool want_rgba8, double* ms);
// peer frame: what one frame's kernels on THIS rank need (see rt_comm.cu)
struct PeerFrame {
    void* rgba8 = nullptr;            // this frame's half of rank 0's double-buffered full frame (null: use the NCCL gather)
    void* f32 = nullptr;              // same for the float frame (null unless asked for)
    unsigned int* flag = nullptr;     // ranks > 0: this rank's completion word in rank 0's memory
    unsigned int* done = nullptr;     // ranks > 0: local CTA counter for the in-kernel signal
    unsigned int  seq = 0;            // frame sequence number (same on every rank)
    const unsigned int* ack = nullptr;   // ranks > 0, wait_in_kernel: local word that must reach ack_min before the first store
    unsigned int  ack_min = 0;
};
// ack_previous (rank 0): tell the peers that the frame before this one is consumed.  wait_in_kernel (ranks > 0): the
// caller's first kernel waits for the acknowledgement itself (LevelParams::peer_ack) instead of the stream.
int  comm_peer_begin(rtb200_ctx* c, int W, int H, bool want_f32, bool ack_previous, bool wait_in_kernel, PeerFrame* pf, cudaStream_t s);
int  comm_peer_wait_ack(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s);  // ranks > 0: the stream waits for pf.ack >= pf.ack_min
int  comm_peer_signal(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s);   // ranks > 0, when no kernel signalled by itself
int  comm_peer_wait(rtb200_ctx* c, const PeerFrame& pf, cudaStream_t s);     // rank 0: s waits for every peer's word
int  comm_peer_ack(rtb200_ctx* c, unsigned int seq, cudaStream_t s);         // rank 0: frames <= seq are consumed
}
