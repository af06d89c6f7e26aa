// rt_sequence.cu -- rtb200_render_sequence: a camera path rendered as ONE pipelined submission.
//
// The reference renders once per process (Main.cpp:181) and its Dessiner() only re-blits the
// texture (Main.cpp:229-250); "ms/frame" on a GPU implies repeated frames with the scene resident
// (SURVEY.md 8f, rank 2).  rtb200_render is synchronous -- it returns a finished frame and its
// statistics -- so a loop over it pays one host round trip per frame (~35 us next to an 83 us
// kernel on BASELINE config 2).  A sequence removes the round trips:
//   * frames whose level-0 kernel is the whole frame (scene-resident path, no secondary rays, one
//     sample per pixel) are enqueued back to back: kernel i+1 starts when kernel i ends;
//   * with a host destination the RGBA8 frames alternate between two device buffers and travel to
//     the host on the copy stream while the next frame renders (PCIe and the SMs overlap);
//   * several GPUs (peer frame, rt_comm.cu): every rank enqueues its kernels without waiting on the
//     host; a peer's side stream pushes each finished frame into rank 0's frame half with the copy
//     engine (one 2-D copy of its bands) and writes the completion word, rank 0's side stream waits
//     for the peers' words, copies the frame out and acknowledges, so neither the transfer nor the
//     signalling nor the read-back is on any rank's kernel-to-kernel path (peers run at most two
//     frames ahead);
//   * per-frame counters live in per-frame slots and come back in one copy at the end.
// Anything else (secondary rays, supersampling, streamed / LBVH paths, NCCL-gather fallback) is
// rendered frame by frame through rtb200_render -- same results, no pipelining.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "rt_context.h"

using namespace rtb;

#define QCU(call)                                                        \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) return set_cuda_error(ctx, e__, #call);  \
    } while (0)

static int seq_fail(rtb200_ctx* ctx, int code, const char* msg)
{
    if (ctx) ctx->last_error = msg;
    return code;
}

extern "C" int rtb200_render_sequence(rtb200_ctx* ctx, const rtb200_frame* frames, int n_frames,
                                      uint8_t* out_rgba8, rtb200_stats* stats)
{
    if (!ctx || !frames || n_frames < 1) return RTB200_E_ARG;
    if (!ctx->have_scene) return seq_fail(ctx, RTB200_E_STATE, "rtb200_render_sequence before rtb200_set_scene");
    const rtb200_frame& f0 = frames[0];
    for (int i = 1; i < n_frames; ++i)
        if (frames[i].width != f0.width || frames[i].height != f0.height || frames[i].spp_grid != f0.spp_grid ||
            frames[i].flags != f0.flags || frames[i].band_rows != f0.band_rows)
            return seq_fail(ctx, RTB200_E_ARG, "all frames of a sequence must share size, sampling, flags and banding");
    if (f0.width <= 0 || f0.height <= 0) return seq_fail(ctx, RTB200_E_ARG, "bad frame size");
    QCU(cudaSetDevice(ctx->device));
    const size_t frame_bytes = 4 * (size_t)f0.width * f0.height;
    const bool multi = ctx->comm && ctx->n_ranks > 1;

    // ---- can the whole sequence be pipelined? ------------------------------------------------------
    const int k = f0.spp_grid > 1 ? f0.spp_grid : 1;
    bool fast = k == 1 && !(f0.flags & (RTB200_F_FORCE_STREAM | RTB200_F_USE_BVH | RTB200_F_GATHER_F32)) &&
                ((size_t)ctx->ds.rl.bytes <= 64 * 1024 || (f0.flags & RTB200_F_FORCE_RESIDENT)) &&
                (size_t)ctx->ds.rl.bytes + 2048 <= ctx->smem_optin;
    for (int i = 0; i < n_frames && fast; ++i) {
        const rtb200_frame& f = frames[i];
        int children = ctx->max_children;
        if (f.min_energy < 0.f) children = 2;
        if ((f.flags & RTB200_F_NO_SECONDARY) || f.max_bounces <= 0) children = 0;
        if (children != 0) fast = false;                      // secondary rays need the host between levels
    }
    FrameParams fp0;
    fill_frame_params(&f0, ctx->n_ranks, ctx->rank, fp0);
    const int band = fp0.band_lrows / fp0.k;
    const int local_rows = local_row_count(fp0.H, band, ctx->n_ranks, ctx->rank);
    if ((size_t)std::max(local_rows, 1) * fp0.LW > (16u << 20)) fast = false;       // frames that need chunking
    // NOTE every condition above depends on replicated state only (out_rgba8 may differ between ranks
    // and must not enter the decision): all ranks pipeline, or none does.
    PeerFrame pf;
    if (fast && multi) {
        // first frame of the path; rank 0 has consumed whatever it rendered before (ack).  Collective
        // only when the mapping must be (re)established -- the same decision on every rank.
        const int prc = comm_peer_begin(ctx, fp0.W, fp0.H, false, true, local_rows > 0, &pf, ctx->stream);
        if (prc != RTB200_OK) return prc;
        if (!pf.rgba8) fast = false;
    }
    if (!fast) {
        for (int i = 0; i < n_frames; ++i) {
            const int rc = rtb200_render(ctx, &frames[i], nullptr, out_rgba8 ? out_rgba8 + (size_t)i * frame_bytes : nullptr,
                                         stats ? &stats[i] : nullptr);
            if (rc != RTB200_OK) return rc;
        }
        return RTB200_OK;
    }

    // ---- buffers ----------------------------------------------------------------------------------------
    cudaError_t e;
    const size_t stat_bytes = sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT;
    const size_t lpx = (size_t)std::max(local_rows, 1) * fp0.W;
    if ((e = ctx->seq_stats.ensure(stat_bytes * (size_t)n_frames)) != cudaSuccess) return set_cuda_error(ctx, e, "sequence counters");
    if ((e = ctx->seq_host.ensure(stat_bytes * (size_t)n_frames)) != cudaSuccess) return set_cuda_error(ctx, e, "sequence counters (host)");
    // Several GPUs, how a peer's rows reach rank 0 in a pipelined path.  "copy" (default): the kernel writes a local
    // compact frame and the copy engine pushes it on the side stream while the next frame renders -- ONE 2-D copy, a
    // band of rows is one contiguous piece on both sides -- followed by the completion word.  "store"
    // (RTB200_PEER_PUSH=store, and what synchronous rtb200_render calls do): the kernel writes its rows over NVLink
    // itself, 32 bytes at a time; rank 0 then takes in at most ~400 GB/s, which 8 GPUs on config 2 exceed
    // (0.145 ms per frame against 0.093 with the copy engine's large packets; 4 GPUs 0.090 / 0.083; 2 GPUs 0.090 / 0.086).
    bool push_copy = multi && ctx->rank != 0;
    if (const char* pp = getenv("RTB200_PEER_PUSH")) push_copy = push_copy && pp[0] != 's';
    uchar4* target[2] = {nullptr, nullptr};
    if (multi && !push_copy) { /* the frame halves of rank 0, chosen per frame by comm_peer_begin */ }
    else if (out_rgba8 || push_copy) {
        for (int b = 0; b < 2; ++b) {
            if ((e = ctx->seq_fb[b].ensure(lpx * 4)) != cudaSuccess) return set_cuda_error(ctx, e, "sequence frame buffer");
            target[b] = ctx->seq_fb[b].as<uchar4>();
        }
    } else {
        if ((e = ctx->fb_rgba8.ensure(lpx * 4)) != cudaSuccess) return set_cuda_error(ctx, e, "fb_rgba8");
        target[0] = target[1] = ctx->fb_rgba8.as<uchar4>();
    }
    QCU(cudaMemsetAsync(ctx->seq_stats.p, 0, stat_bytes * (size_t)n_frames, ctx->stream));
    ctx->stats_clean = false;

    // ---- enqueue -------------------------------------------------------------------------------------------
    const bool exact = (f0.flags & RTB200_F_FUSED_FP) == 0;
    cudaEvent_t done[2] = {ctx->copy_ev[0], ctx->copy_ev[1]};
    cudaEvent_t copied[2] = {ctx->ev[2], ctx->ev[3]};
    QCU(cudaEventRecord(ctx->ev[0], ctx->stream));
    for (int i = 0; i < n_frames; ++i) {
        if (multi && i > 0) {
            const int prc = comm_peer_begin(ctx, fp0.W, fp0.H, false, false, local_rows > 0, &pf, ctx->stream);
            if (prc != RTB200_OK || !pf.rgba8) return prc != RTB200_OK ? prc : seq_fail(ctx, RTB200_E_STATE, "peer frame lost inside a sequence");
        }
        const int b = multi ? (int)(pf.seq & 1u) : (i & 1);
        const bool push = push_copy && pf.flag;
        const bool consumer = out_rgba8 && (!multi || ctx->rank == 0);
        FrameParams fp;
        fill_frame_params(&frames[i], ctx->n_ranks, ctx->rank, fp);
        fp.secondary = 0;
        fp.fb_global = (multi && !push) ? 1 : 0;
        LevelParams p;
        memset(&p, 0, sizeof p);
        p.f = fp;
        p.blob = ctx->ds.blob; p.rl = ctx->ds.rl;
        p.level = 0;
        p.chunk_ly0 = 0; p.chunk_rows = local_rows; p.local_LH = local_rows;
        p.stats = ctx->seq_stats.as<unsigned long long>() + (size_t)i * RT_STAT_SLOTS * ST_COUNT;
        p.direct = 1;
        p.cull = (frames[i].flags & RTB200_F_EXACT_FP) ? 0 : 1;
        p.fb_f32 = nullptr;                                   // a sequence delivers RGBA8 only
        p.fb_rgba8 = (multi && !push) ? static_cast<uchar4*>(pf.rgba8) : target[b];
        if (multi && pf.flag && local_rows > 0 && !push) {    // wait for the half, store, signal: all in the one kernel
            p.peer_flag = pf.flag; p.peer_done = pf.done; p.peer_seq = pf.seq;
            p.peer_ack = pf.ack; p.peer_ack_min = pf.ack_min;
        }
        if ((consumer || push) && i >= 2) QCU(cudaStreamWaitEvent(ctx->stream, copied[b], 0));   // buffer b is free again
        if (local_rows > 0) QCU(launch_resident_level(p, exact, ctx->sm_count, ctx->stream));
        else if (multi && pf.flag && !push) { const int rc = comm_peer_signal(ctx, pf, ctx->stream); if (rc != RTB200_OK) return rc; }
        if (push) {
            // side stream: wait for the kernel and for rank 0 to have let go of the half, push the bands, signal
            QCU(cudaEventRecord(done[b], ctx->stream));
            QCU(cudaStreamWaitEvent(ctx->copy_stream, done[b], 0));
            int rc = comm_peer_wait_ack(ctx, pf, ctx->copy_stream);
            if (rc != RTB200_OK) return rc;
            const size_t row_bytes = 4 * (size_t)fp0.W, band_bytes = (size_t)band * row_bytes;
            const int n_full = local_rows / band, tail = local_rows - n_full * band;
            unsigned char* dst = static_cast<unsigned char*>(pf.rgba8);
            const unsigned char* src = reinterpret_cast<const unsigned char*>(target[b]);
            if (n_full > 0)
                QCU(cudaMemcpy2DAsync(dst + (size_t)ctx->rank * band_bytes, (size_t)ctx->n_ranks * band_bytes, src, band_bytes,
                                      band_bytes, (size_t)n_full, cudaMemcpyDeviceToDevice, ctx->copy_stream));
            if (tail > 0)
                QCU(cudaMemcpyAsync(dst + ((size_t)n_full * ctx->n_ranks + ctx->rank) * band_bytes, src + (size_t)n_full * band_bytes,
                                    (size_t)tail * row_bytes, cudaMemcpyDeviceToDevice, ctx->copy_stream));
            rc = comm_peer_signal(ctx, pf, ctx->copy_stream);
            if (rc != RTB200_OK) return rc;
            QCU(cudaEventRecord(copied[b], ctx->copy_stream));
        }
        if (consumer || (multi && ctx->rank == 0)) {
            // side stream: (several GPUs: wait for the peers' words,) copy the frame out, acknowledge
            QCU(cudaEventRecord(done[b], ctx->stream));
            QCU(cudaStreamWaitEvent(ctx->copy_stream, done[b], 0));
            if (multi) { const int rc = comm_peer_wait(ctx, pf, ctx->copy_stream); if (rc != RTB200_OK) return rc; }
            if (consumer) {
                QCU(cudaMemcpyAsync(out_rgba8 + (size_t)i * frame_bytes, p.fb_rgba8, frame_bytes, cudaMemcpyDeviceToHost, ctx->copy_stream));
                QCU(cudaEventRecord(copied[b], ctx->copy_stream));
            }
            // acknowledge = "the peers may overwrite this half"; a last frame that stays on the device for
            // rtb200_read_frame is acknowledged by rank 0's next render call instead
            if (multi && (consumer || i + 1 < n_frames)) { const int rc = comm_peer_ack(ctx, pf.seq, ctx->copy_stream); if (rc != RTB200_OK) return rc; }
        }
    }
    QCU(cudaEventRecord(ctx->ev[1], ctx->stream));
    QCU(cudaMemcpyAsync(ctx->seq_host.p, ctx->seq_stats.p, stat_bytes * (size_t)n_frames, cudaMemcpyDeviceToHost, ctx->stream));
    QCU(cudaStreamSynchronize(ctx->stream));
    QCU(cudaStreamSynchronize(ctx->copy_stream));
    float total = 0.f;
    QCU(cudaEventElapsedTime(&total, ctx->ev[0], ctx->ev[1]));

    // ---- statistics -----------------------------------------------------------------------------------------
    ctx->last_W = fp0.W; ctx->last_H = fp0.H; ctx->last_local_rows = local_rows;
    ctx->last_gathered = multi && ctx->rank == 0;
    if (out_rgba8) ctx->last_W = 0;                           // the frames went to the host; nothing to read back later
    {   // a CTA that gave up waiting for rank 0's acknowledgement: the sequence may hold torn frames
        const unsigned long long* h = ctx->seq_host.as<unsigned long long>();
        unsigned long long timeouts = 0;
        for (size_t i = 0; i < (size_t)n_frames * RT_STAT_SLOTS; ++i) timeouts += h[i * ST_COUNT + ST_PEER_TIMEOUT];
        if (timeouts) { ctx->last_error = "peer frame: rank 0 did not acknowledge within 2 s; frames of this sequence may be torn"; return RTB200_E_NCCL; }
    }
    if (stats) {
        const unsigned long long* h = ctx->seq_host.as<unsigned long long>();
        const double n_surf = std::max(1, ctx->ds.rl.n_surf);
        for (int i = 0; i < n_frames; ++i) {
            unsigned long long sum[ST_COUNT] = {0};
            for (int s = 0; s < RT_STAT_SLOTS; ++s)
                for (int c = 0; c < ST_COUNT; ++c) sum[c] += h[((size_t)i * RT_STAT_SLOTS + s) * ST_COUNT + c];
            rtb200_stats st;
            memset(&st, 0, sizeof st);
            st.ms_render = total / n_frames;                  // pipelined: only the whole sequence has a duration
            st.ms_primary_kernel = st.ms_trace_nearest = total / n_frames;
            st.rays_primary = (uint64_t)local_rows * (uint64_t)fp0.LW;
            st.rays_shadow = sum[ST_SHADOW];
            st.nan_pixels = sum[ST_NAN];
            st.tests_triangle = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_tri / n_surf);
            st.tests_quadric  = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_quad / n_surf);
            st.tests_plane    = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_plane / n_surf);
            st.kernel_launches = local_rows > 0 ? 1 : 0;
            st.depth_levels = 1; st.chunks = 1; st.path = 0;
            stats[i] = st;
        }
    }
    return RTB200_OK;
}
