// rt_stream.cu -- host side of the streamed (tiled brute-force) wavefront path.
//
// Per image chunk and per recursion level k:
//   trace_nearest  every ray x every surface           -> (t, order)
//   shade_a        hit point, normal, shadow queries, children (level k+1 queue), links
//   trace_shadow   every shadow query x every surface  -> filter RGB
//   shade_b        ambient + per-light diffuse/Phong * filter, in light order -> ray colour
// then the same bottom-up fold (k_combine) and resolve (k_resolve) as the resident path.
// The host reads three counters per level (children, links, shadow queries) to size the next
// launches exactly; nothing else crosses the bus during a frame.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "rt_context.h"
#include "rt_stream.cuh"
#include "rt_bvh.cuh"
#include "rt_cullscan.cuh"

using namespace rtb;

#define SCU(call)                                                       \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return set_cuda_error(c, e__, #call);   \
    } while (0)

// Samples traced per pass over the levels (a frame larger than this is rendered in row chunks; config 5: 32 per GPU).
// Every chunk costs three host synchronisations and the tails of ~18 launches: config 5 took 500 / 483 / 474 ms per
// frame with 8 / 16 / 32 M samples per chunk (same frame).  32 M samples hold ~8 GB of queues.  RTB200_CHUNK_SAMPLES
// overrides it (tests render small frames in many chunks).
#ifndef STREAM_CHUNK_SAMPLES
#define STREAM_CHUNK_SAMPLES (32u << 20)
#endif

namespace rtb {

struct StreamBufs {                // per-context scratch of the streamed path (lives in the ctx levels + these)
    DevBuf hit, hitP, hitN, q_base, sq0, sq1, sq_src, sq_dst, filter;
    DevBuf tasks, fix, far, lane;  // per-level counters (12 words per level), fix-up list, LBVH far-ray list, lane-walk task list
    DevBuf ckey, cidx, ckey2, cperm, qkey, qidx, qkey2, qperm, sort_tmp;   // culled scan: (key, index) pairs of the children / queries of a level and their sorted forms
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // per-stage stopwatch
};

// RTB200_NO_BUNDLE_CULL=1: no culling of whole triangles per warp -- the plain tiled scan (the roofline measurement)
static int bundle_cull_enabled()
{
    const char* e = getenv("RTB200_NO_BUNDLE_CULL");
    return (e && atoi(e) != 0) ? 0 : 1;
}

// The culled scan (rt_cullscan.cu) is the default; RTB200_CULLSCAN=0 (like RTB200_NO_BUNDLE_CULL=1) runs the plain tiled scan.
static bool cullscan_enabled()
{
    if (bundle_cull_enabled() == 0) return false;
    const char* e = getenv("RTB200_CULLSCAN");
    return !(e && atoi(e) == 0);
}

// queue order of the tiled path (see render_streamed); RTB200_QUERY_GROUPING=0 restores the ray order for A/B runs
static bool query_grouping_enabled()
{
    const char* e = getenv("RTB200_QUERY_GROUPING");
    return !(e && atoi(e) == 0);
}

static StreamBufs* stream_bufs(rtb200_ctx* c)
{
    if (!c->stream_scratch) c->stream_scratch = new StreamBufs();
    return static_cast<StreamBufs*>(c->stream_scratch);
}

void stream_release(rtb200_ctx* c)
{
    if (!c->stream_scratch) return;
    StreamBufs* b = static_cast<StreamBufs*>(c->stream_scratch);
    b->hit.release(); b->hitP.release(); b->hitN.release(); b->q_base.release(); b->sq0.release(); b->sq1.release(); b->sq_src.release(); b->sq_dst.release(); b->filter.release(); b->tasks.release(); b->fix.release(); b->far.release(); b->lane.release();
    b->ckey.release(); b->cidx.release(); b->ckey2.release(); b->cperm.release(); b->qkey.release(); b->qidx.release(); b->qkey2.release(); b->qperm.release(); b->sort_tmp.release();
    for (auto& e : b->ev) if (e) cudaEventDestroy(e);
    delete b;
    c->stream_scratch = nullptr;
}

int render_streamed(rtb200_ctx* c, const rtb200_frame* fr, const FrameParams& fp, int local_rows,
                    bool exact, rtb200_stats* st, uint32_t* launches)
{
    StreamBufs* sb = stream_bufs(c);
    const int mode = (fr->flags & RTB200_F_FUSED_FP) ? 2 : ((fr->flags & RTB200_F_EXACT_FP) ? 1 : 0);
    const bool use_bvh = (fr->flags & RTB200_F_USE_BVH) != 0;
    // warp <-> shadow query mapping of the LBVH kernel: by light when >= 64 lattice samples start on the average
    // primitive and the light count lets a stride of 8 separate the lights (RTB200_SHADOW_BY_LIGHT=0/1 forces it)
    bool shadow_by_light = false;
    if (use_bvh) {
        const int L = c->ds.rl.n_light;
        const double prims = (double)std::max(c->ds.bvh_n_prims, 1);
        const double samples = (double)fr->width * fr->height * std::max(fr->spp_grid, 1) * std::max(fr->spp_grid, 1);
        shadow_by_light = L > 1 && (8 % L == 0 || L % 8 == 0) && samples / prims >= 64.0;
        if (const char* e = getenv("RTB200_SHADOW_BY_LIGHT")) shadow_by_light = atoi(e) != 0;
    }
    // tiled kernels: level-0 rays by 16x16 patches and each CTA's shadow queries light by light, so that a warp of the
    // trace kernels holds a compact bundle (changes the ORDER of the queues, no value: frames stay bit-identical;
    // config 3: shadow stage 112 -> 60 ms, profiles/r1_c3_ab_query_grouping.log)
    // (LBVH: only the patches -- its shadow kernel has a warp mapping of its own that relies on the queries of a ray
    // being contiguous; measured in round 2: config 5 507 -> 480 ms, config 4 unchanged)
    bool tile_rays = query_grouping_enabled();
    if (use_bvh) if (const char* e = getenv("RTB200_BVH_TILE_RAYS")) tile_rays = tile_rays && atoi(e) != 0;   // (LBVH: the patches only; =0 for A/B runs)
    const bool group_queries = tile_rays && !use_bvh && c->ds.rl.n_light >= 1 && c->ds.rl.n_light <= 8;
    if (use_bvh) {                               // boxes must be padded for rays that start at the camera
        const int rc = bvh_prepare(c, fr->cam_pos, 1);
        if (rc != RTB200_OK) return rc;
    }
    // culled scan: Morton-sorted triangles + sphere hierarchy of this scene (built once per rtb200_set_scene)
    const bool cullscan = !use_bvh && mode != 1 && cullscan_enabled();
    CullScene cull;
    memset(&cull, 0, sizeof cull);
    if (cullscan) {
        const int rc = cull_prepare(c, &cull);
        if (rc != RTB200_OK) return rc;
    }
    // levels >= 1 of the culled scan: rays sorted by the surface they leave, queries by (light, surface of the hit point)
    // -- such groups are compact bundles, queue neighbours are not (RTB200_SORT_SECONDARY=0: queue order, for A/B runs)
    bool sort_secondary = cullscan;
    if (const char* e = getenv("RTB200_SORT_SECONDARY")) sort_secondary = sort_secondary && atoi(e) != 0;
    if (cull.key_inv <= 0.f) sort_secondary = false;
    int light_bits = 0;
    while ((1 << light_bits) < std::max(c->ds.rl.n_light, 1)) ++light_bits;
    // a level with fewer rays than this is traced one ray per lane (tasks of 32): see CullTraceParams::task_shift
    int small_level = c->sm_count * 16 * 128 * 3;
    if (const char* e = getenv("RTB200_SMALL_LEVEL")) small_level = atoi(e);
    // cones wider than this are not worth a warp walk (CullTraceParams::lane_sin).  Measured on config 3 (two sweeps inside
    // one call each): flat between 0.05 and 0.15 for nearest-hit rays and between 0.2 and 0.5 for shadow queries
    // (21.6 - 22.2 ms per frame), worse outside (shadow 0.12: 22.9 ms; everything to the warps: 25.3 ms)
    float lane_sin = 0.1f, lane_sin_shadow = 0.25f;
    if (const char* e = getenv("RTB200_LANE_SIN")) lane_sin = (float)atof(e);
    if (const char* e = getenv("RTB200_LANE_SIN_SHADOW")) lane_sin_shadow = (float)atof(e);
    bool rays_sorted = false;                   // the rays of the level about to be traced have a permutation in sb->cperm
    const int k = fp.k;
    const int local_LH = local_rows * k;
    const int L = c->ds.rl.n_light;
    int children = c->max_children;
    if (fr->min_energy < 0.f) children = 2;
    if (!fp.secondary || fr->max_bounces <= 0) children = 0;

    unsigned chunk_samples = STREAM_CHUNK_SAMPLES;
    if (const char* e = getenv("RTB200_CHUNK_SAMPLES")) { const long v = atol(e); if (v > 0 && v <= (1l << 30)) chunk_samples = (unsigned)v; }
    int rows_per_chunk = (int)((chunk_samples / (unsigned)fp.LW) / (32u * k)) * 32 * k;
    if (rows_per_chunk < 32 * k) rows_per_chunk = 32 * k;
    if (rows_per_chunk > local_LH) rows_per_chunk = ((local_LH + 32 * k - 1) / (32 * k)) * 32 * k;

    cudaError_t e;
    if ((e = c->c0.ensure(sizeof(float) * 3 * (size_t)rows_per_chunk * fp.LW)) != cudaSuccess) return set_cuda_error(c, e, "c0");
    const size_t n_counters = 4 * ((size_t)std::max(fr->max_bounces, 0) + 2);
    if ((e = c->counters.ensure(sizeof(unsigned) * n_counters)) != cudaSuccess) return set_cuda_error(c, e, "counters");
    unsigned* h_cnt = c->h_small.as<unsigned>();
    unsigned* d_cnt = c->counters.as<unsigned>();
    // per level: nearest tasks, shadow tasks, fix-ups, far rays, far queries, -, lane-walk tasks (nearest: count, counter; shadow: count, counter)
    const size_t n_tasks_words = 12 * ((size_t)std::max(fr->max_bounces, 0) + 2);
    if ((e = sb->tasks.ensure(sizeof(unsigned) * n_tasks_words)) != cudaSuccess) return set_cuda_error(c, e, "task counters");
    unsigned* d_tasks = sb->tasks.as<unsigned>();
    unsigned max_level_seen = 0, chunks = 0;
    unsigned long long rays_nearest = 0;
    cudaEvent_t ev_a = c->ev[2], ev_b = c->ev[3];
    bool timed_first = false;
    // per-stage device time: E0 nearest E1 shade-A E2 | host reads the counters | E3 shadow E4 shade-B E5;
    // the second triple is read after the NEXT synchronisation point
    for (auto& ev : sb->ev) if (!ev) SCU(cudaEventCreate(&ev));
    cudaEvent_t* E = sb->ev;
    double ms_nearest = 0.0, ms_shadow = 0.0, ms_shade = 0.0;
    bool pending_b = false;
    auto flush_b = [&]() {
        if (!pending_b) return;
        float a = 0.f, b = 0.f;
        if (cudaEventElapsedTime(&a, E[3], E[4]) == cudaSuccess) ms_shadow += a;
        if (cudaEventElapsedTime(&b, E[4], E[5]) == cudaSuccess) ms_shade += b;
        pending_b = false;
    };

    for (int ly0 = 0; ly0 < local_LH; ly0 += rows_per_chunk, ++chunks) {
        const int rows = std::min(rows_per_chunk, local_LH - ly0);
        SCU(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned) * n_counters, c->stream));
        SCU(cudaMemsetAsync(d_tasks, 0, sizeof(unsigned) * n_tasks_words, c->stream));
        std::vector<int> n_rays, n_links;
        n_rays.push_back(rows * fp.LW);
        int level = 0;
        for (;; ++level) {
            const int n_in = n_rays[level];
            if (n_in == 0) break;
            const bool spawn = children > 0 && level < fr->max_bounces;
            if ((int)c->levels.size() < level + 2) c->levels.resize((size_t)level + 2);
            LevelBuf& cur = c->levels[level];
            LevelBuf& nxt = c->levels[level + 1];
            const size_t n = (size_t)n_in;
            if ((e = sb->hit.ensure(sizeof(float2) * n)) != cudaSuccess || (e = sb->hitP.ensure(sizeof(float4) * n)) != cudaSuccess ||
                (e = sb->hitN.ensure(sizeof(float4) * n)) != cudaSuccess || (e = sb->q_base.ensure(sizeof(int) * n)) != cudaSuccess ||
                (e = sb->sq_src.ensure(sizeof(int2) * n * std::max(L, 1))) != cudaSuccess ||
                (group_queries && (e = sb->sq_dst.ensure(sizeof(int) * n * std::max(L, 1))) != cudaSuccess) ||
                (e = sb->filter.ensure(sizeof(float4) * n * std::max(L, 1))) != cudaSuccess)
                return set_cuda_error(c, e, "stream scratch");
            if (level > 0 && (e = cur.colour.ensure(sizeof(float4) * n)) != cudaSuccess) return set_cuda_error(c, e, "colour");
            if (spawn) {
                const size_t cap = n * children;
                if ((e = nxt.od0.ensure(sizeof(float4) * cap)) != cudaSuccess || (e = nxt.od1.ensure(sizeof(float4) * cap)) != cudaSuccess ||
                    (e = cur.links.ensure(sizeof(int4) * n)) != cudaSuccess)
                    return set_cuda_error(c, e, "queue");
            }
            // ---- nearest hit -------------------------------------------------------------------
            TraceParams tp;
            memset(&tp, 0, sizeof tp);
            tp.f = fp; tp.s = c->ds; tp.primary = (level == 0);
            tp.chunk_ly0 = ly0; tp.chunk_rows = rows; tp.local_LH = local_LH;
            tp.in_od0 = cur.od0.as<float4>(); tp.in_od1 = cur.od1.as<float4>();
            tp.n_rays = n_in; tp.hit = sb->hit.as<float2>(); tp.mode = mode; tp.cull_bundles = bundle_cull_enabled();
            const bool timed = !timed_first && level == 0;
            if (timed) SCU(cudaEventRecord(ev_a, c->stream));
            SCU(cudaEventRecord(E[0], c->stream));
            tp.stats = c->stats.as<unsigned long long>();
            if (use_bvh) {
                if (level > 0) {                   // secondary rays may start beyond the radius the boxes are padded for
                    if ((e = sb->far.ensure(sizeof(int) * n)) != cudaSuccess) return set_cuda_error(c, e, "far list");
                    tp.far_list = sb->far.as<int>(); tp.far_count = d_tasks + 12 * level + 3; tp.far_cap = n_in;
                }
                SCU(launch_bvh_nearest(tp, c->stream));
                if (level > 0) { SCU(launch_bvh_far_nearest(tp, c->sm_count, c->stream)); ++*launches; }
            }
            else if (cullscan) {
                CullTraceParams ctp;
                ctp.t = tp; ctp.cs = cull; ctp.task_counter = d_tasks + 12 * level;
                ctp.perm = (level > 0 && rays_sorted) ? sb->cperm.as<int>() : nullptr;
                ctp.task_shift = (level > 0 && n_in < small_level) ? 5 : 7;
                {
                    const int n_tasks = tp.primary ? ((fp.LW + 15) / 16) * ((rows + 7) / 8) : (n_in + (1 << ctp.task_shift) - 1) >> ctp.task_shift;
                    if ((e = sb->lane.ensure(sizeof(int) * (size_t)std::max(n_tasks, 1))) != cudaSuccess) return set_cuda_error(c, e, "lane list");
                    ctp.lane_list = sb->lane.as<int>(); ctp.lane_cap = n_tasks;
                    ctp.lane_count = d_tasks + 12 * level + 6; ctp.lane_task_counter = d_tasks + 12 * level + 7; ctp.lane_sin = lane_sin;
                    ++*launches;
                }
                SCU(launch_cull_nearest(ctp, c->sm_count, c->stream));
            }
            else         SCU(launch_trace_nearest(tp, c->sm_count, c->stream));
            SCU(cudaEventRecord(E[1], c->stream));
            if (timed) { SCU(cudaEventRecord(ev_b, c->stream)); timed_first = true; }
            ++*launches;
            rays_nearest += (unsigned long long)n_in;
            // ---- shade A ------------------------------------------------------------------------
            ShadeAParams sa;
            memset(&sa, 0, sizeof sa);
            sa.f = fp; if (!spawn) sa.f.secondary = 0;
            sa.s = c->ds; sa.level = level;
            sa.chunk_ly0 = ly0; sa.chunk_rows = rows; sa.local_LH = local_LH;
            sa.in_od0 = tp.in_od0; sa.in_od1 = tp.in_od1; sa.n_rays = n_in; sa.hit = tp.hit;
            sa.hitP = sb->hitP.as<float4>(); sa.hitN = sb->hitN.as<float4>(); sa.q_base = sb->q_base.as<int>();
            sa.sq_src = sb->sq_src.as<int2>();
            sa.tile_rays = tile_rays ? 1 : 0;
            sa.sq_dst = group_queries ? sb->sq_dst.as<int>() : nullptr;
            sa.out_od0 = nxt.od0.as<float4>(); sa.out_od1 = nxt.od1.as<float4>();
            sa.out_count = d_cnt + 4 * level; sa.link_count = d_cnt + 4 * level + 1; sa.sq_count = d_cnt + 4 * level + 2;
            sa.links = cur.links.as<int4>();
            sa.stats = c->stats.as<unsigned long long>();
            sa.exact = exact ? 1 : 0;
            sa.mark_far = use_bvh ? 1 : 0;
            const bool sort_children = sort_secondary && spawn;
            const bool sort_queries = sort_secondary && level > 0 && group_queries;
            if (sort_children) {
                const size_t cap = n * children;
                if ((e = sb->ckey.ensure(4 * cap)) != cudaSuccess || (e = sb->cidx.ensure(4 * cap)) != cudaSuccess ||
                    (e = sb->ckey2.ensure(4 * cap)) != cudaSuccess || (e = sb->cperm.ensure(4 * cap)) != cudaSuccess) return set_cuda_error(c, e, "sort buffers");
                sa.child_key = sb->ckey.as<unsigned int>(); sa.child_idx = sb->cidx.as<int>();
            }
            if (sort_queries) {
                const size_t cap = n * std::max(L, 1);
                if ((e = sb->qkey.ensure(4 * cap)) != cudaSuccess || (e = sb->qidx.ensure(4 * cap)) != cudaSuccess ||
                    (e = sb->qkey2.ensure(4 * cap)) != cudaSuccess || (e = sb->qperm.ensure(4 * cap)) != cudaSuccess) return set_cuda_error(c, e, "sort buffers");
                sa.sq_key = sb->qkey.as<unsigned int>(); sa.sq_idx = sb->qidx.as<int>();
            }
            for (int a = 0; a < 3; ++a) sa.key_lo[a] = cull.key_lo[a];
            sa.key_inv = cull.key_inv;
            SCU(launch_shade_a(sa, c->stream));
            SCU(cudaEventRecord(E[2], c->stream));
            ++*launches;
            SCU(cudaMemcpyAsync(h_cnt, d_cnt + 4 * level, 4 * sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
            SCU(cudaStreamSynchronize(c->stream));
            flush_b();
            {
                float a = 0.f, b = 0.f;
                SCU(cudaEventElapsedTime(&a, E[0], E[1]));
                SCU(cudaEventElapsedTime(&b, E[1], E[2]));
                ms_nearest += a; ms_shade += b;
            }
            const int n_children = (int)h_cnt[0], n_lk = (int)h_cnt[1], n_sq = (int)h_cnt[2];
            st->rays_shadow += (uint64_t)n_sq;
            rays_sorted = false;
            if (sort_children && n_children > 256) {           // the next level's rays, by the surface they leave
                SCU(cull_sort_pairs(&sb->sort_tmp, sb->ckey.as<unsigned int>(), sb->ckey2.as<unsigned int>(), sb->cidx.as<int>(), sb->cperm.as<int>(),
                                    n_children, 30, c->stream));
                rays_sorted = true;
                *launches += 3;
            }
            const bool queries_sorted = sort_queries && n_sq > 256;
            if (queries_sorted) {
                SCU(cull_sort_pairs(&sb->sort_tmp, sb->qkey.as<unsigned int>(), sb->qkey2.as<unsigned int>(), sb->qidx.as<int>(), sb->qperm.as<int>(),
                                    n_sq, 27 + light_bits, c->stream));
                *launches += 3;
            }
            // ---- shadow filters --------------------------------------------------------------
            SCU(cudaEventRecord(E[3], c->stream));
            if (n_sq > 0) {
                ShadowParams sp;
                memset(&sp, 0, sizeof sp);
                sp.s = c->ds; sp.sq_src = sa.sq_src; sp.hitP = sa.hitP; sp.n_queries = n_sq;
                sp.sq_dst = sa.sq_dst;
                sp.by_light = shadow_by_light ? 1 : 0;
                sp.filter = sb->filter.as<float4>(); sp.stats = sa.stats; sp.mode = mode; sp.cull_bundles = bundle_cull_enabled();
                {
                    // queries whose product order matters go on a list and are redone in insertion order (every path)
                    if ((e = sb->fix.ensure(sizeof(int) * (size_t)n_sq)) != cudaSuccess) return set_cuda_error(c, e, "fix-up list");
                    unsigned* tk = sb->tasks.as<unsigned>() + 12 * level;
                    sp.fix_count = tk + 2; sp.fix_list = sb->fix.as<int>(); sp.fix_cap = n_sq;
                    CullShadowParams csp;
                    csp.s = sp; csp.cs = cull; csp.task_counter = tk + 1;
                    csp.perm = queries_sorted ? sb->qperm.as<int>() : nullptr;
                    csp.task_shift = (level > 0 && n_sq < small_level) ? 5 : 7;
                    csp.lane_list = nullptr; csp.lane_count = nullptr; csp.lane_task_counter = nullptr; csp.lane_cap = 0; csp.lane_sin = lane_sin_shadow;
                    if (cullscan && group_queries) {
                        const int n_tasks = (n_sq + (1 << csp.task_shift) - 1) >> csp.task_shift;
                        if ((e = sb->lane.ensure(sizeof(int) * (size_t)std::max(n_tasks, 1))) != cudaSuccess) return set_cuda_error(c, e, "lane list");
                        csp.lane_list = sb->lane.as<int>(); csp.lane_cap = n_tasks;
                        csp.lane_count = tk + 8; csp.lane_task_counter = tk + 9;
                        ++*launches;
                    }
                    if (use_bvh) {
                        if ((e = sb->far.ensure(sizeof(int) * (size_t)n_sq)) != cudaSuccess) return set_cuda_error(c, e, "far list");
                        sp.far_list = sb->far.as<int>(); sp.far_count = tk + 4; sp.far_cap = n_sq;
                        csp.s = sp;
                        SCU(launch_bvh_shadow(sp, c->stream));
                        SCU(launch_bvh_far_shadow(sp, c->sm_count, c->stream));
                        ++*launches;
                    }
                    else if (cullscan && group_queries) SCU(launch_cull_shadow(csp, c->sm_count, c->stream));
                    else                                SCU(launch_trace_shadow(sp, c->sm_count, c->stream));
                    SCU(launch_shadow_fixup(csp, c->sm_count, c->stream));     // (reads the count itself; nearly always 0)
                }
                *launches += 2;
            }
            SCU(cudaEventRecord(E[4], c->stream));
            // ---- shade B -------------------------------------------------------------------------
            ShadeBParams sbp;
            memset(&sbp, 0, sizeof sbp);
            sbp.f = fp; sbp.s = c->ds; sbp.level = level;
            sbp.chunk_ly0 = ly0; sbp.chunk_rows = rows; sbp.local_LH = local_LH;
            sbp.in_od1 = tp.in_od1; sbp.n_rays = n_in;
            sbp.hitP = sa.hitP; sbp.hitN = sa.hitN; sbp.q_base = sa.q_base; sbp.filter = sb->filter.as<float4>();
            sbp.c0 = c->c0.as<float>(); sbp.colour = cur.colour.as<float4>(); sbp.exact = exact ? 1 : 0;
            SCU(launch_shade_b(sbp, c->stream));
            SCU(cudaEventRecord(E[5], c->stream));
            pending_b = true;
            ++*launches;
            n_links.push_back(spawn ? n_lk : 0);
            if (!spawn) { ++level; break; }
            n_rays.push_back(n_children);
        }
        const int levels_used = (int)n_links.size();
        max_level_seen = std::max<unsigned>(max_level_seen, (unsigned)levels_used);
        for (int l = levels_used - 1; l >= 0; --l) {
            if (n_links[l] <= 0) continue;
            CombineParams cp;
            memset(&cp, 0, sizeof cp);
            cp.links = c->levels[l].links.as<int4>(); cp.n_links = n_links[l];
            cp.parent_is_level0 = (l == 0);
            cp.parent_c0 = c->c0.as<float>(); cp.parent_colour = c->levels[l].colour.as<float4>();
            cp.child_colour = c->levels[l + 1].colour.as<float4>(); cp.mat = c->ds.mat;
            SCU(launch_combine(cp, exact, c->stream));
            ++*launches;
        }
        ResolveParams r;
        memset(&r, 0, sizeof r);
        r.f = fp; r.c0 = c->c0.as<float>();
        r.chunk_py0 = ly0 / k; r.chunk_prows = std::min(rows / k, local_rows - r.chunk_py0);
        r.fb_f32 = c->out_f32_dev; r.fb_rgba8 = c->out_rgba8_dev;
        r.stats = c->stats.as<unsigned long long>();
        SCU(launch_resolve(r, c->stream));
        ++*launches;
        {
            const int rc2 = chunk_copy_out(c, fp.W, r.chunk_py0, r.chunk_prows);
            if (rc2 != RTB200_OK) return rc2;
        }
    }
    SCU(cudaStreamSynchronize(c->stream));
    flush_b();
    st->ms_trace_nearest = ms_nearest; st->ms_trace_shadow = ms_shadow; st->ms_shade = ms_shade;
    st->depth_levels = max_level_seen;
    st->chunks = chunks;
    st->path = use_bvh ? 2 : 1;
    st->ms_primary_kernel = -1.0;             // first level-0 nearest-hit launch, read after the final sync
    // nearest-hit queries test every surface by construction (no early-out): exact host-side count;
    // shadow queries are counted on the device (warps leave early once their filters are zero)
    c->host_tests[0] = use_bvh ? 0ull : rays_nearest * (unsigned long long)c->ds.n_tri;   // LBVH: counted on the device
    c->host_tests[1] = use_bvh ? 0ull : rays_nearest * (unsigned long long)c->ds.n_quad;   // LBVH: counted on the device
    c->host_tests[2] = rays_nearest * (unsigned long long)c->ds.n_plane;
    return RTB200_OK;
}

// rays in c->q_in0 (origins) / c->q_in1 (directions) -> c->q_out0 (order) / q_out1 (t) / q_out2 (normal)
int stream_query_nearest(rtb200_ctx* c, int n, uint32_t flags)
{
    StreamBufs* sb = stream_bufs(c);
    const int mode = (flags & RTB200_F_FUSED_FP) ? 2 : ((flags & RTB200_F_EXACT_FP) ? 1 : 0);
    cudaError_t e;
    if ((e = sb->sq0.ensure(sizeof(float4) * (size_t)n)) != cudaSuccess || (e = sb->sq1.ensure(sizeof(float4) * (size_t)n)) != cudaSuccess ||
        (e = sb->hit.ensure(sizeof(float2) * (size_t)n)) != cudaSuccess) return set_cuda_error(c, e, "query scratch");
    const size_t stat_bytes = sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT;
    if ((e = c->stats.ensure(stat_bytes)) != cudaSuccess) return set_cuda_error(c, e, "stats");
    if (!c->stats_clean) SCU(cudaMemsetAsync(c->stats.p, 0, stat_bytes, c->stream));
    c->stats_clean = false;                      // the traversal reports a stack overflow through the frame counters
    SCU(launch_query_pack_rays(n, c->q_in0.as<float>(), c->q_in1.as<float>(), sb->sq0.as<float4>(), sb->sq1.as<float4>(), c->stream));
    TraceParams tp;
    memset(&tp, 0, sizeof tp);
    tp.s = c->ds; tp.primary = 0; tp.in_od0 = sb->sq0.as<float4>(); tp.in_od1 = sb->sq1.as<float4>();
    tp.n_rays = n; tp.hit = sb->hit.as<float2>(); tp.mode = mode; tp.cull_bundles = bundle_cull_enabled();
    tp.stats = c->stats.as<unsigned long long>();
    if (flags & RTB200_F_USE_BVH) SCU(launch_bvh_nearest(tp, c->stream));
    else if (mode != 1 && cullscan_enabled()) {
        CullTraceParams ctp;
        ctp.t = tp;
        const int rc = cull_prepare(c, &ctp.cs);
        if (rc != RTB200_OK) return rc;
        if ((e = sb->tasks.ensure(64)) != cudaSuccess) return set_cuda_error(c, e, "task counters");
        SCU(cudaMemsetAsync(sb->tasks.p, 0, 4, c->stream));
        ctp.task_counter = sb->tasks.as<unsigned>();
        ctp.perm = nullptr; ctp.task_shift = 7;
        const int n_tasks = (n + 127) / 128;
        if ((e = sb->lane.ensure(sizeof(int) * (size_t)n_tasks)) != cudaSuccess) return set_cuda_error(c, e, "lane list");
        SCU(cudaMemsetAsync(sb->tasks.p, 0, 16, c->stream));
        ctp.lane_list = sb->lane.as<int>(); ctp.lane_cap = n_tasks;
        ctp.lane_count = sb->tasks.as<unsigned>() + 1; ctp.lane_task_counter = sb->tasks.as<unsigned>() + 2; ctp.lane_sin = 0.1f;
        SCU(launch_cull_nearest(ctp, c->sm_count, c->stream));
    }
    else                          SCU(launch_trace_nearest(tp, c->sm_count, c->stream));
    SCU(launch_query_finish_nearest(c->ds, n, tp.in_od0, tp.in_od1, tp.hit, c->q_out0.as<int>(), c->q_out1.as<float>(),
                                    c->q_out2.as<float>(), mode != 2, c->stream));
    return RTB200_OK;
}

// points in c->q_in0 / light positions in c->q_in1 -> filters in c->q_out2
int stream_query_shadow(rtb200_ctx* c, int n, uint32_t flags)
{
    StreamBufs* sb = stream_bufs(c);
    const int mode = (flags & RTB200_F_FUSED_FP) ? 2 : ((flags & RTB200_F_EXACT_FP) ? 1 : 0);
    cudaError_t e;
    if ((e = sb->sq0.ensure(sizeof(float4) * (size_t)n)) != cudaSuccess || (e = sb->sq1.ensure(sizeof(float4) * (size_t)n)) != cudaSuccess ||
        (e = sb->filter.ensure(sizeof(float4) * (size_t)n)) != cudaSuccess) return set_cuda_error(c, e, "query scratch");
    const size_t stat_bytes = sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT;
    if ((e = c->stats.ensure(stat_bytes)) != cudaSuccess) return set_cuda_error(c, e, "stats");
    if (!c->stats_clean) SCU(cudaMemsetAsync(c->stats.p, 0, stat_bytes, c->stream));
    c->stats_clean = false;                      // the query kernels add to the frame counters
    SCU(launch_query_pack_shadow(n, c->q_in0.as<float>(), c->q_in1.as<float>(), sb->sq0.as<float4>(), sb->sq1.as<float4>(), mode != 2, c->stream));
    ShadowParams sp;
    memset(&sp, 0, sizeof sp);
    sp.s = c->ds; sp.sq0 = sb->sq0.as<float4>(); sp.sq1 = sb->sq1.as<float4>(); sp.n_queries = n;
    sp.filter = sb->filter.as<float4>(); sp.stats = c->stats.as<unsigned long long>(); sp.mode = mode;
    {
        if ((e = sb->fix.ensure(sizeof(int) * (size_t)n)) != cudaSuccess || (e = sb->tasks.ensure(64)) != cudaSuccess) return set_cuda_error(c, e, "fix-up list");
        SCU(cudaMemsetAsync(sb->tasks.p, 0, 12, c->stream));
        sp.fix_count = sb->tasks.as<unsigned>() + 2; sp.fix_list = sb->fix.as<int>(); sp.fix_cap = n;
        if (flags & RTB200_F_USE_BVH) SCU(launch_bvh_shadow(sp, c->stream));
        else                          SCU(launch_trace_shadow(sp, c->sm_count, c->stream));
        CullShadowParams csp;
        memset(&csp, 0, sizeof csp);
        csp.s = sp;
        SCU(launch_shadow_fixup(csp, c->sm_count, c->stream));
    }
    SCU(launch_query_unpack_filter(n, sp.filter, c->q_out2.as<float>(), c->stream));
    return RTB200_OK;
}

} // namespace rtb
