// rt_trace.cu -- the PLAIN brute-force intersection kernels of the streamed path (large scenes) and the shading
// stages of its wavefront.
//
// Every ray of a wavefront level is tested against EVERY surface (this is the reference's
// algorithm: Scene.cpp:1710-1715 and :1850-1858 are linear scans).  What makes it fast:
//   * triangles stream through shared memory in 12 KB tiles (256 x 48 B: p0|order, e1|material,
//     e2) fetched by 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx), double
//     buffered; every warp reads a triangle as three broadcast LDS.128;
//   * a persistent CTA holds 1024 rays (256 threads x 4 rays in registers) so one staged
//     triangle serves 4 rays per thread (ILP) and 1024 rays per CTA (arithmetic intensity
//     ~ 1000 x 46 flop per 48 B: HBM/L2 are irrelevant, the FMA pipe is the roof);
//   * per (ray, triangle) a two-phase CULLING FILTER in fused arithmetic (rt_cull.cuh): phase 1 computes
//     det and S.P only -- no division, compared in numerator space -- and a warp vote skips the rest of the
//     test when no lane of the warp can be accepted; phase 2 finishes v and t for the survivors of phase 1.
//     The filter is conservative by construction (margins that bound the fused / reference difference of every
//     numerator); only rays that pass it run the reference's own rounding sequence (triangle_t<true>), so the
//     accepted distances are bit-identical to the CPU renderer.
//   * shadow queries leave at warp granularity once every ray of the warp holds a zero filter.
// These kernels are the roofline kernels (bench.py times them with RTB200_NO_BUNDLE_CULL=1), the RTB200_F_EXACT_FP
// path, and what explicit shadow queries and scenes with more than 8 lights run through.  The DEFAULT path of a frame
// is the culled scan of rt_cullscan.cu, which skips whole triangles per warp (round 1 did that inside these tile
// loops, meeting at two CTA barriers per tile: 91 ms for config 3 against 47 ms now) and returns the same bits.
#include "rt_stream.cuh"
#include "rt_cull.cuh"
#include "rt_bundle.cuh"
#include "rt_tma.cuh"

namespace rtb {

#define FULL_MASK 0xffffffffu

struct TraceSmem {
    TriRec tile[ST_STAGES][ST_TILE];                 // 24 KB
    unsigned char side[ST_SIDE_MAX];                 // quadric + plane records (when they fit)
    uint64_t full[ST_STAGES];
};

// ---------------------------------------------------------------------------------------
// Nearest hit.  One CTA = 1024 rays.  PRIMARY: the rays are a 32x32 tile of the sample lattice
// (warp = 16x8 pixels, thread = a 2x2 quad) generated on the fly; otherwise 1024 consecutive
// queue entries.
// ---------------------------------------------------------------------------------------
template <int MODE, bool PRIMARY>
__global__ void __launch_bounds__(ST_THREADS, 2)
k_trace_nearest(const __grid_constant__ TraceParams p)
{
    __shared__ __align__(128) TraceSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const DeviceScene& s = p.s;
    const int n_tiles = s.n_tri_padded / ST_TILE;
    const int side_bytes = (int)(sizeof(QuadRec) * s.n_quad + sizeof(PlaneRec) * s.n_plane);
    const bool side_in_smem = side_bytes <= ST_SIDE_MAX;

    if (tid == 0) {
        for (int i = 0; i < ST_STAGES; ++i) mbar_init(&sm.full[i], 1);
        mbar_fence_init();
    }
    if (side_in_smem) {                              // quadrics then planes, contiguous in the scene buffer
        const float4* src = reinterpret_cast<const float4*>(s.quad);
        float4* dst = reinterpret_cast<float4*>(sm.side);
        for (int i = tid; i < side_bytes / 16; i += ST_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const QuadRec*  quads  = side_in_smem ? reinterpret_cast<const QuadRec*>(sm.side) : s.quad;
    const PlaneRec* planes = side_in_smem ? reinterpret_cast<const PlaneRec*>(sm.side + sizeof(QuadRec) * s.n_quad) : s.plane;

    const int tiles_x = (p.f.LW + 31) >> 5;
    const int n_blocks = PRIMARY ? tiles_x * ((p.chunk_rows + 31) >> 5) : (p.n_rays + ST_BLOCK_RAYS - 1) / ST_BLOCK_RAYS;
    uint32_t phase_bits = 0;                         // parity of each stage's barrier
    unsigned long long exec_rays = 0;                // (ray, triangle) tests entered
    for (int blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        // ---- this thread's 4 rays ------------------------------------------------------------
        V3 o[ST_RPT], d[ST_RPT];
        float best_t[ST_RPT], tlim[ST_RPT];
        int best_id[ST_RPT], slot[ST_RPT];
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) {
            bool valid;
            if (PRIMARY) {
                const int tx = blk % tiles_x, ty = blk / tiles_x;
                const int lx = (tx << 5) + ((warp & 1) << 4) + ((lane & 7) << 1) + (r & 1);
                const int cy = (ty << 5) + ((warp >> 1) << 3) + ((lane >> 3) << 1) + (r >> 1);
                const int ly = p.chunk_ly0 + cy;
                valid = lx < p.f.LW && cy < p.chunk_rows && ly < p.local_LH;
                slot[r] = valid ? cy * p.f.LW + lx : -1;
                primary_ray<(MODE != 2)>(p.f, valid ? lx : 0, valid ? global_lattice_row(p.f, ly) : 0, o[r], d[r]);
            } else {
                const int i = blk * ST_BLOCK_RAYS + warp * (32 * ST_RPT) + r * 32 + lane;   // a warp = 128 consecutive rays
                valid = i < p.n_rays;
                slot[r] = valid ? i : -1;
                const float4 a = valid ? p.in_od0[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 b = valid ? p.in_od1[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                o[r] = f4_xyz(a); d[r] = f4_xyz(b);
            }
            if (!valid) d[r] = mk3(0.f, 0.f, 0.f);     // zero direction: det = 0, nothing can hit
            best_t[r] = -1.0f; best_id[r] = -1; tlim[r] = 3.0e38f;
        }
        float dmax = 0.f;                             // largest |direction component| of this thread's rays (filter margins)
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) dmax = fmaxf(dmax, max_abs3(d[r]));
        // ---- triangles: double-buffered TMA tiles ---------------------------------------------
        if (tid == 0 && n_tiles > 0) {
            mbar_arrive_expect_tx(&sm.full[0], (uint32_t)sizeof(TriRec) * ST_TILE);
            tma_load_1d(sm.tile[0], s.tri, (uint32_t)sizeof(TriRec) * ST_TILE, &sm.full[0]);
        }
        for (int t = 0; t < n_tiles; ++t) {
            const int st = t & (ST_STAGES - 1);
            if (tid == 0 && t + 1 < n_tiles) {       // the other stage was released by the barrier below
                const int nx = (t + 1) & (ST_STAGES - 1);
                mbar_arrive_expect_tx(&sm.full[nx], (uint32_t)sizeof(TriRec) * ST_TILE);
                tma_load_1d(sm.tile[nx], s.tri + (size_t)(t + 1) * ST_TILE, (uint32_t)sizeof(TriRec) * ST_TILE, &sm.full[nx]);
            }
            mbar_wait(&sm.full[st], (phase_bits >> st) & 1u);
            phase_bits ^= (1u << st);
            const TriRec* __restrict__ tile = sm.tile[st];
            if (MODE == 1) {                         // no filter: the reference sequence for every test
                #pragma unroll 2
                for (int j = 0; j < ST_TILE; ++j) {
                    const float4 ra = tile[j].a, rb = tile[j].b, rc = tile[j].c;
                    const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
                    #pragma unroll
                    for (int r = 0; r < ST_RPT; ++r) {
                        const float tt = triangle_t<true>(o[r], d[r], p0, e1, e2);
                        if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f)) { best_t[r] = tt; best_id[r] = __float_as_int(ra.w); }
                    }
                }
            } else {
                // filter + reference-rounded survivors for triangle j of the tile, all 4 rays of the thread
                auto test_triangle = [&](int j) {
                    const float4 ra = tile[j].a, rb = tile[j].b, rc = tile[j].c;
                    const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
                    const TriMargins mg = tri_margins(e1, e2, dmax);
                    bool cand[ST_RPT];
                    bool any = false;
                    #pragma unroll
                    for (int r = 0; r < ST_RPT; ++r) {       // phase 1: det and u, fused, no division
                        cand[r] = triangle_phase1(o[r], d[r], p0, e1, e2, mg);    // branch-free: the 4 rays interleave
                        any |= cand[r];
                    }
                    if (__any_sync(FULL_MASK, any)) {
                        #pragma unroll
                        for (int r = 0; r < ST_RPT; ++r) {
                            if (cand[r] && slot[r] >= 0) {
                                const float tt = triangle_survivor<MODE>(o[r], d[r], p0, e1, e2, tlim[r], mg);
                                if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f)) {
                                    best_t[r] = tt; best_id[r] = __float_as_int(ra.w);
                                    tlim[r] = fmaf(tt, CULL_TREL, CULL_TABS);
                                }
                            }
                        }
                    }
                };
                #pragma unroll 2
                for (int j = 0; j < ST_TILE; ++j) test_triangle(j);
            }
            __syncthreads();                         // everyone is done with this stage
        }

        // ---- quadrics and planes (reference sequence; ties go to the lower insertion order) ----
        for (int q = 0; q < s.n_quad; ++q) {
            const float4 q0 = quads[q].q0, q1 = quads[q].q1, q2 = quads[q].q2;
            const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
            const int ord = __float_as_int(q0.w);
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) {
                bool lin;
                const float tt = MODE == 2 ? quadric_t<false>(o[r], d[r], Q, lin) : quadric_t<true>(o[r], d[r], Q, lin);
                if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f || (tt == best_t[r] && ord < best_id[r]))) { best_t[r] = tt; best_id[r] = ord; }
            }
        }
        for (int q = 0; q < s.n_plane; ++q) {
            const float4 nc = planes[q].nc;
            const int ord = planes[q].order;
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) {
                const float tt = MODE == 2 ? plane_t<false>(o[r], d[r], f4_xyz(nc), nc.w) : plane_t<true>(o[r], d[r], f4_xyz(nc), nc.w);
                if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f || (tt == best_t[r] && ord < best_id[r]))) { best_t[r] = tt; best_id[r] = ord; }
            }
        }
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r)
            if (slot[r] >= 0) { p.hit[slot[r]] = make_float2(best_t[r], __int_as_float(best_id[r])); exec_rays += (unsigned long long)s.n_tri; }
    }
    if (MODE != 1 && p.stats) {                      // per-ray triangle tests entered: all of them
        #pragma unroll
        for (int off = 16; off > 0; off >>= 1) exec_rays += __shfl_xor_sync(FULL_MASK, exec_rays, off);
        if (lane == 0 && exec_rays) atomicAdd(&p.stats[(size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT + ST_TEST_TRI_EXEC], exec_rays);
    }
}

// ---------------------------------------------------------------------------------------
// Shadow filter (Scene.cpp:1842-1861): product of colour*kt over every surface hit strictly
// between EPSILON and the light.  One CTA = 1024 consecutive shadow queries.
// The scan meets the occluders type by type (triangles, quadrics, planes), the reference in insertion order
// (Scene.cpp:1856).  Up to two factors commute exactly; a query that met three or more, not in increasing insertion
// order, with a non-zero product goes on the fix-up list and is redone in order by k_shadow_fixup (rt_cullscan.cu).
// ---------------------------------------------------------------------------------------
#define OCC_ORDER_LIMIT (1 << 29)      // per-ray state word: (largest order + 1) << 3 | met out of order << 2 | occluders (saturating at 3)

template <int MODE>
__global__ void __launch_bounds__(ST_THREADS, 2)
k_trace_shadow(const __grid_constant__ ShadowParams p)
{
    __shared__ __align__(128) TraceSmem sm;
    __shared__ unsigned int s_occ[ST_RPT][ST_THREADS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const DeviceScene& s = p.s;
    const int n_tiles = s.n_tri_padded / ST_TILE;
    const int side_bytes = (int)(sizeof(QuadRec) * s.n_quad + sizeof(PlaneRec) * s.n_plane);
    const bool side_in_smem = side_bytes <= ST_SIDE_MAX;
    const bool track = s.rl.n_surf < OCC_ORDER_LIMIT;      // (beyond: every query with >= 3 occluders is redone)

    if (tid == 0) {
        for (int i = 0; i < ST_STAGES; ++i) mbar_init(&sm.full[i], 1);
        mbar_fence_init();
    }
    if (side_in_smem) {
        const float4* src = reinterpret_cast<const float4*>(s.quad);
        float4* dst = reinterpret_cast<float4*>(sm.side);
        for (int i = tid; i < side_bytes / 16; i += ST_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const QuadRec*  quads  = side_in_smem ? reinterpret_cast<const QuadRec*>(sm.side) : s.quad;
    const PlaneRec* planes = side_in_smem ? reinterpret_cast<const PlaneRec*>(sm.side + sizeof(QuadRec) * s.n_quad) : s.plane;

    const int n_blocks = (p.n_queries + ST_BLOCK_RAYS - 1) / ST_BLOCK_RAYS;
    uint32_t phase_bits = 0;
    unsigned long long tests_tri = 0, tests_side_q = 0, tests_side_p = 0, fixups = 0;

    for (int blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        V3 o[ST_RPT], d[ST_RPT];
        float dist[ST_RPT], tlim[ST_RPT], fr[ST_RPT], fg[ST_RPT], fb[ST_RPT];
        int slot[ST_RPT];
        int n_valid = 0;
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) {
            const int i = blk * ST_BLOCK_RAYS + warp * (32 * ST_RPT) + r * 32 + lane;   // a warp = 128 consecutive queries
            const bool valid = i < p.n_queries;
            slot[r] = valid ? i : -1;
            n_valid += valid ? 1 : 0;
            o[r] = d[r] = mk3(0.f, 0.f, 0.f); dist[r] = 0.f;
            if (valid) shadow_query<(MODE != 2)>(p, i, o[r], d[r], dist[r]);
            tlim[r] = fmaf(dist[r], CULL_TREL, CULL_TABS);
            fr[r] = fg[r] = fb[r] = 1.0f;
            if (!valid) { fr[r] = fg[r] = fb[r] = 0.0f; }        // counts as finished
            s_occ[r][tid] = 0u;
        }
        bool warp_done = false;
        float dmax = 0.f;                             // largest |direction component| of this thread's queries (filter margins)
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) dmax = fmaxf(dmax, max_abs3(d[r]));
        auto occluder = [&](int r, int ord, int mat) {
            const float4 m0 = s.mat[mat].m0;
            const float kt = s.mat[mat].m2.x;
            fr[r] = __fmul_rn(fr[r], __fmul_rn(m0.x, kt));
            fg[r] = __fmul_rn(fg[r], __fmul_rn(m0.y, kt));
            fb[r] = __fmul_rn(fb[r], __fmul_rn(m0.z, kt));
            unsigned int st = s_occ[r][tid];
            const unsigned int key = (unsigned int)(ord + 1);
            unsigned int flags = st & 7u;
            if (!track || key < (st >> 3)) flags |= 4u;
            if ((flags & 3u) < 3u) ++flags;
            s_occ[r][tid] = (max(st >> 3, track ? key : 0u) << 3) | flags;
        };

        if (tid == 0 && n_tiles > 0) {
            mbar_arrive_expect_tx(&sm.full[0], (uint32_t)sizeof(TriRec) * ST_TILE);
            tma_load_1d(sm.tile[0], s.tri, (uint32_t)sizeof(TriRec) * ST_TILE, &sm.full[0]);
        }
        for (int t = 0; t < n_tiles; ++t) {
            const int st = t & (ST_STAGES - 1);
            if (tid == 0 && t + 1 < n_tiles) {
                const int nx = (t + 1) & (ST_STAGES - 1);
                mbar_arrive_expect_tx(&sm.full[nx], (uint32_t)sizeof(TriRec) * ST_TILE);
                tma_load_1d(sm.tile[nx], s.tri + (size_t)(t + 1) * ST_TILE, (uint32_t)sizeof(TriRec) * ST_TILE, &sm.full[nx]);
            }
            mbar_wait(&sm.full[st], (phase_bits >> st) & 1u);
            phase_bits ^= (1u << st);
            const TriRec* __restrict__ tile = sm.tile[st];

            if (!warp_done) {
                const int real = min(ST_TILE, s.n_tri - t * ST_TILE);
                tests_tri += (unsigned long long)n_valid * (unsigned long long)real;
                if (MODE == 1) {
                    #pragma unroll 2
                    for (int j = 0; j < ST_TILE; ++j) {
                        const float4 ra = tile[j].a, rb = tile[j].b, rc = tile[j].c;
                        const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
                        #pragma unroll
                        for (int r = 0; r < ST_RPT; ++r) {
                            const float tt = triangle_t<true>(o[r], d[r], p0, e1, e2);
                            if (slot[r] >= 0 && tt > RT_EPS && tt < dist[r]) occluder(r, __float_as_int(ra.w), __float_as_int(rb.w));
                        }
                    }
                } else {
                    #pragma unroll 2
                    for (int j = 0; j < ST_TILE; ++j) {
                        const float4 ra = tile[j].a, rb = tile[j].b, rc = tile[j].c;
                        const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
                        const TriMargins mg = tri_margins(e1, e2, dmax);
                        bool cand[ST_RPT];
                        bool any = false;
                        #pragma unroll
                        for (int r = 0; r < ST_RPT; ++r) {
                            cand[r] = triangle_phase1(o[r], d[r], p0, e1, e2, mg);    // branch-free: the 4 rays interleave
                            any |= cand[r];
                        }
                        if (__any_sync(FULL_MASK, any)) {
                            #pragma unroll
                            for (int r = 0; r < ST_RPT; ++r) {
                                if (cand[r] && slot[r] >= 0) {
                                    const float tt = triangle_survivor<MODE>(o[r], d[r], p0, e1, e2, tlim[r], mg);
                                    if (tt > RT_EPS && tt < dist[r]) occluder(r, __float_as_int(ra.w), __float_as_int(rb.w));
                                }
                            }
                        }
                    }
                }
                bool done = true;
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) done = done && (fr[r] == 0.f && fg[r] == 0.f && fb[r] == 0.f);
                warp_done = __all_sync(FULL_MASK, done);     // 0 * finite stays 0: the warp may stop testing
            }
            __syncthreads();
        }

        if (!warp_done) {
            tests_side_q += (unsigned long long)n_valid * (unsigned long long)s.n_quad;
            tests_side_p += (unsigned long long)n_valid * (unsigned long long)s.n_plane;
            for (int q = 0; q < s.n_quad; ++q) {
                const float4 q0 = quads[q].q0, q1 = quads[q].q1, q2 = quads[q].q2;
                const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    bool lin;
                    const float tt = MODE == 2 ? quadric_t<false>(o[r], d[r], Q, lin) : quadric_t<true>(o[r], d[r], Q, lin);
                    if (slot[r] >= 0 && tt > RT_EPS && tt < dist[r]) occluder(r, __float_as_int(q0.w), __float_as_int(q1.w));
                }
            }
            for (int q = 0; q < s.n_plane; ++q) {
                const float4 nc = planes[q].nc;
                const int m = planes[q].material, ord = planes[q].order;
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    const float tt = MODE == 2 ? plane_t<false>(o[r], d[r], f4_xyz(nc), nc.w) : plane_t<true>(o[r], d[r], f4_xyz(nc), nc.w);
                    if (slot[r] >= 0 && tt > RT_EPS && tt < dist[r]) occluder(r, ord, m);
                }
            }
        }
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r)
            if (slot[r] >= 0) {
                p.filter[p.sq_dst ? p.sq_dst[slot[r]] : slot[r]] = make_float4(fr[r], fg[r], fb[r], 0.f);
                const unsigned int st = s_occ[r][tid];
                if ((st & 4u) && (st & 3u) >= 3u && (fr[r] != 0.f || fg[r] != 0.f || fb[r] != 0.f) && p.fix_count) {
                    const unsigned int k = atomicAdd(p.fix_count, 1u);
                    if ((int)k < p.fix_cap) p.fix_list[k] = slot[r];
                    ++fixups;
                }
            }
    }

    // ---- statistics: tests entered (per thread -> warp -> one atomic per warp) ----------------
    for (int off = 16; off > 0; off >>= 1) {
        tests_tri += __shfl_xor_sync(FULL_MASK, tests_tri, off);
        tests_side_q += __shfl_xor_sync(FULL_MASK, tests_side_q, off);
        tests_side_p += __shfl_xor_sync(FULL_MASK, tests_side_p, off);
        fixups += __shfl_xor_sync(FULL_MASK, fixups, off);
    }
    if ((tid & 31) == 0) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (tests_tri) { atomicAdd(&st[ST_TEST_TRI_SHADOW], tests_tri); atomicAdd(&st[ST_TEST_TRI_EXEC], tests_tri); }
        if (tests_side_q) atomicAdd(&st[ST_TEST_QUAD], tests_side_q);
        if (tests_side_p) atomicAdd(&st[ST_TEST_PLANE], tests_side_p);
        if (fixups) atomicAdd(&st[ST_SHADOW_FIXUP], fixups);
    }
}

// ---------------------------------------------------------------------------------------
// shade-A: per ray -- hit point and normal (as CIntersection would carry them), one shadow
// query per light facing the stored normal (Scene.cpp:1751-1759), reflected / refracted
// children (Scene.cpp:1779-1823).  Filter slots of one ray are contiguous, in light order (q_base); the queries
// themselves are queued in that order too, or -- sq_dst given -- light by light within the CTA.
// ---------------------------------------------------------------------------------------
template <bool E>
RT_DEV void level_ray(const FrameParams& f, int level, int chunk_ly0, int i, const float4* od0, const float4* od1,
                      V3& o, V3& d, float& energy, float& ior)
{
    if (level == 0) {
        const int x = i % f.LW, cy = i / f.LW;
        primary_ray<E>(f, x, global_lattice_row(f, chunk_ly0 + cy), o, d);
        energy = 1.0f; ior = 1.0f;
    } else {
        const float4 a = od0[i], b = od1[i];
        o = f4_xyz(a); d = f4_xyz(b); energy = a.w; ior = b.w;
    }
}

template <bool E>
__global__ void __launch_bounds__(256) k_shade_a(const __grid_constant__ ShadeAParams p)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < p.n_rays;
    if (p.tile_rays && p.level == 0) {
        // 16x16 samples per CTA: the hit points of a warp's queries and the origins of its children lie in a compact
        // patch, so the cones the tiled kernels cull with are narrow in both directions (a scanline is narrow in one)
        const int tiles_x = (p.f.LW + 15) >> 4;
        const int x = (blockIdx.x % tiles_x) * 16 + (threadIdx.x & 15), y = (blockIdx.x / tiles_x) * 16 + (threadIdx.x >> 4);
        valid = x < p.f.LW && y < p.chunk_rows;
        i = y * p.f.LW + x;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool grouped = p.sq_dst != nullptr;
    const DeviceScene& s = p.s;
    V3 o = mk3(0.f, 0.f, 0.f), d = o, P = o, N = o;
    float energy = 0.f, ior = 0.f, t = -1.f;
    int id = -1, mat = 0;
    if (valid) {
        level_ray<E>(p.f, p.level, p.chunk_ly0, i, p.in_od0, p.in_od1, o, d, energy, ior);
        const float2 h = p.hit[i];
        t = h.x; id = __float_as_int(h.y);
    }
    const bool hit = valid && id >= 0;
    float kr = 0.f, kt = 0.f, mior = 0.f;
    if (hit) {
        const float4 r0 = s.surf[id].r0, r1 = s.surf[id].r1;
        const int type = __float_as_int(r0.w);
        mat = __float_as_int(r1.w);
        if (type == SURF_TRIANGLE) N = f4_xyz(s.surf[id].r3);
        else if (type == SURF_PLANE) N = f4_xyz(r0);
        else {
            const float4 r2 = s.surf[id].r2;
            const Quadric Q = { r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z, r2.w };
            bool lin;
            (void)quadric_t<E>(o, d, Q, lin);              // which branch produced t
            N = quadric_normal<E>(o, d, t, Q, lin);
        }
        P = axpy3<E>(o, d, t);
        const float4 m1 = s.mat[mat].m1, m2 = s.mat[mat].m2;
        kr = m1.w; kt = m2.x; mior = m2.y;
    }
    // ---- shadow queries ---------------------------------------------------------------------
    const int far_bit = (p.mark_far && hit && s.bvh_n_prims > 0 && bvh_is_far(s, P)) ? SQ_FAR_BIT : 0;
    int n_face = 0;
    unsigned int face_bits = 0;                              // bit l: light l faces the stored normal
    if (hit)
        for (int l = 0; l < s.rl.n_light; ++l) {
            const V3 Ld = sub3<E>(f4_xyz(s.light[l].l0), P);
            const bool facing = dot3<E>(Ld, N) > 0.f;
            n_face += facing ? 1 : 0;
            face_bits |= facing ? (1u << (l & 31)) : 0u;
        }
    // ---- queue slots: ONE atomic per counter and CTA ------------------------------------------------
    // (one per warp made a quarter of a million atomics on a single address per 4K frame: the kernel spent
    // a third of its time waiting for them, ncu source page of config 4)
    bool want_refl = false, want_refr = false;
    float e_refl = 0.f, e_refr = 0.f;
    if (hit && p.f.secondary && p.level < p.f.max_bounces) {
        e_refl = FP<E>::mul(kr, energy); e_refr = FP<E>::mul(kt, energy);
        want_refl = e_refl > p.f.min_energy; want_refr = e_refr > p.f.min_energy;
    }
    const unsigned int m_refl = __ballot_sync(FULL_MASK, want_refl), m_refr = __ballot_sync(FULL_MASK, want_refr);
    const unsigned int n_refl = __popc(m_refl), n_refr = __popc(m_refr), m_any = m_refl | m_refr;
    int incl = n_face;
    for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(FULL_MASK, incl, off); if (lane >= off) incl += v; }
    const int total = __shfl_sync(FULL_MASK, incl, 31);
    __shared__ uint4 s_cnt[8];                               // per warp: queries, reflected, refracted, links
    __shared__ uint4 s_base[8];
    __shared__ unsigned short s_lcnt[BUNDLE_MAX_LIGHTS * 8]; // grouped queues: facing rays per (light, warp) ...
    __shared__ unsigned short s_loff[BUNDLE_MAX_LIGHTS * 8]; // ... and their exclusive scan in that order
    if (lane == 0) s_cnt[warp] = make_uint4((unsigned int)total, n_refl, n_refr, (unsigned int)__popc(m_any));
    if (grouped)
        for (int l = 0; l < BUNDLE_MAX_LIGHTS; ++l) {
            const unsigned int m = __ballot_sync(FULL_MASK, (face_bits >> l) & 1u);
            if (lane == 0) s_lcnt[l * 8 + warp] = (unsigned short)__popc(m);
        }
    __syncthreads();
    if (warp == 0) {
        const uint4 c = lane < 8 ? s_cnt[lane] : make_uint4(0u, 0u, 0u, 0u);
        const unsigned int kids = c.y + c.z;
        unsigned int iq = c.x, ik = kids, il = c.w;           // inclusive scans over the 8 warps
        for (int off = 1; off < 8; off <<= 1) {
            const unsigned int vq = __shfl_up_sync(FULL_MASK, iq, off), vk = __shfl_up_sync(FULL_MASK, ik, off), vl = __shfl_up_sync(FULL_MASK, il, off);
            if (lane >= off) { iq += vq; ik += vk; il += vl; }
        }
        const unsigned int tq = __shfl_sync(FULL_MASK, iq, 7), tk = __shfl_sync(FULL_MASK, ik, 7), tl = __shfl_sync(FULL_MASK, il, 7);
        const unsigned int t_refl = __reduce_add_sync(FULL_MASK, c.y), t_refr = __reduce_add_sync(FULL_MASK, c.z);
        unsigned int bq = 0, bk = 0, bl = 0;
        if (lane == 0) {
            if (tq) bq = atomicAdd(p.sq_count, tq);
            if (tk) {
                bk = atomicAdd(p.out_count, tk);
                bl = atomicAdd(p.link_count, tl);
                unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
                if (t_refl) atomicAdd(&st[ST_REFLECT], (unsigned long long)t_refl);
                if (t_refr) atomicAdd(&st[ST_REFRACT], (unsigned long long)t_refr);
            }
        }
        bq = __shfl_sync(FULL_MASK, bq, 0); bk = __shfl_sync(FULL_MASK, bk, 0); bl = __shfl_sync(FULL_MASK, bl, 0);
        if (lane < 8) s_base[lane] = make_uint4(bq + iq - c.x, bk + ik - kids, bl + il - c.w, bq);
        if (grouped) {                                        // 64 (light, warp) counts, two per lane
            const unsigned int v0 = s_lcnt[2 * lane], v1 = s_lcnt[2 * lane + 1];
            unsigned int inc = v0 + v1;
            for (int off = 1; off < 32; off <<= 1) { const unsigned int v = __shfl_up_sync(FULL_MASK, inc, off); if (lane >= off) inc += v; }
            const unsigned int ex = inc - (v0 + v1);
            s_loff[2 * lane] = (unsigned short)ex; s_loff[2 * lane + 1] = (unsigned short)(ex + v0);
        }
    }
    __syncthreads();
    const uint4 wb = s_base[warp];
    const unsigned int base = wb.x;
    int q = (int)base + incl - n_face;
    if (valid) {
        p.q_base[i] = q;
        p.hitP[i] = make_float4(P.x, P.y, P.z, t);
        p.hitN[i] = make_float4(N.x, N.y, N.z, __int_as_float(id));
    }
    if (grouped) {
        // the CTA's queries light by light: a warp of the shadow kernel then holds (up to) 128 hit points of ONE
        // light -- one narrow cone to cull with, and every kept triangle is one its 128 queries may really need
        const unsigned int below = (1u << lane) - 1u;
        for (int l = 0; l < s.rl.n_light; ++l) {
            const bool facing = (face_bits >> l) & 1u;
            const unsigned int m = __ballot_sync(FULL_MASK, facing);
            if (facing) {
                const int k = (int)(wb.w + s_loff[l * 8 + warp] + __popc(m & below));
                RT_CHECK(k >= 0 && k < p.n_rays * s.rl.n_light && q >= 0 && q < p.n_rays * s.rl.n_light);
                p.sq_src[k] = make_int2(i, l | far_bit);
                p.sq_dst[k] = q++;
                if (p.sq_key) { p.sq_key[k] = ((unsigned int)l << 27) | (morton30(P.x, P.y, P.z, p.key_lo[0], p.key_lo[1], p.key_lo[2], p.key_inv) >> 3); p.sq_idx[k] = k; }
            }
        }
    } else if (hit)
        for (int l = 0; l < s.rl.n_light; ++l) {
            const V3 Ld = sub3<E>(f4_xyz(s.light[l].l0), P);
            if (dot3<E>(Ld, N) > 0.f) {                                   // direction and distance: shadow_query()
                if (p.sq_key) { p.sq_key[q] = ((unsigned int)l << 27) | (morton30(P.x, P.y, P.z, p.key_lo[0], p.key_lo[1], p.key_lo[2], p.key_inv) >> 3); p.sq_idx[q] = q; }
                p.sq_src[q++] = make_int2(i, l | far_bit);
            }
        }
    // ---- children --------------------------------------------------------------------------------
    if (m_any) {
        const unsigned int cbase = wb.y, lbase = wb.z;
        const unsigned int below = (1u << lane) - 1u;
        int s_refl = -1, s_refr = -1;
        if (want_refl) {
            s_refl = (int)(cbase + __popc(m_refl & below));
            RT_CHECK(s_refl >= 0 && s_refl < 2 * p.n_rays);
            const V3 rd = reflect3<E>(d, N);
            p.out_od0[s_refl] = make_float4(P.x, P.y, P.z, e_refl);
            p.out_od1[s_refl] = make_float4(rd.x, rd.y, rd.z, 0.0f);
            if (p.child_key) { p.child_key[s_refl] = morton30(P.x, P.y, P.z, p.key_lo[0], p.key_lo[1], p.key_lo[2], p.key_inv); p.child_idx[s_refl] = s_refl; }
        }
        if (want_refr) {
            s_refr = (int)(cbase + n_refl + __popc(m_refr & below));
            RT_CHECK(s_refr >= 0 && s_refr < 2 * p.n_rays);
            V3 n2 = N; float ratio, nior;
            if (ior == mior) { nior = p.f.scene_ior; ratio = FP<E>::div(mior, p.f.scene_ior); n2 = neg3(N); }
            else             { nior = mior;          ratio = FP<E>::div(p.f.scene_ior, mior); }
            const V3 rd = refract3<E>(d, n2, ratio);
            p.out_od0[s_refr] = make_float4(P.x, P.y, P.z, e_refr);
            p.out_od1[s_refr] = make_float4(rd.x, rd.y, rd.z, nior);
            if (p.child_key) { p.child_key[s_refr] = morton30(P.x, P.y, P.z, p.key_lo[0], p.key_lo[1], p.key_lo[2], p.key_inv); p.child_idx[s_refr] = s_refr; }
        }
        if (want_refl || want_refr) p.links[lbase + __popc(m_any & below)] = make_int4(i, s_refl, s_refr, mat);
    }
}

// ---------------------------------------------------------------------------------------
// shade-B: the ray's own colour (Scene.cpp:1742-1776) with the shadow filters now known.
// ---------------------------------------------------------------------------------------
template <bool E>
__global__ void __launch_bounds__(256) k_shade_b(const __grid_constant__ ShadeBParams p)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_rays) return;
    const DeviceScene& s = p.s;
    const float4 hn = p.hitN[i];
    const int id = __float_as_int(hn.w);
    float cr = p.f.background[0], cg = p.f.background[1], cb = p.f.background[2];
    if (id >= 0) {
        const V3 N = f4_xyz(hn), P = f4_xyz(p.hitP[i]);
        V3 d;
        if (p.level == 0) { V3 o; primary_ray<E>(p.f, i % p.f.LW, global_lattice_row(p.f, p.chunk_ly0 + i / p.f.LW), o, d); }
        else d = f4_xyz(p.in_od1[i]);
        const int mat = __float_as_int(s.surf[id].r1.w);
        const float4 m0 = s.mat[mat].m0, m1 = s.mat[mat].m1;
        cr = FP<E>::mul(m0.x, m0.w); cg = FP<E>::mul(m0.y, m0.w); cb = FP<E>::mul(m0.z, m0.w);
        int q = p.q_base[i];
        for (int l = 0; l < s.rl.n_light; ++l) {
            const float4 l0 = s.light[l].l0, l1 = s.light[l].l1;
            V3 Ld = sub3<E>(f4_xyz(l0), P);
            if (!(dot3<E>(Ld, N) > 0.f)) continue;
            const float4 fl = p.filter[q++];
            const float dist = norm3<E>(Ld);
            Ld = scale3<E>(Ld, FP<E>::rcp(dist));
            const float lcr = FP<E>::mul(l1.x, fl.x), lcg = FP<E>::mul(l1.y, fl.y), lcb = FP<E>::mul(l1.z, fl.z);
            const float g = FP<E>::mul(FP<E>::mul(l0.w, m1.x), dot3<E>(N, Ld));
            cr = FP<E>::madd(FP<E>::mul(m0.x, g), lcr, cr);
            cg = FP<E>::madd(FP<E>::mul(m0.y, g), lcg, cg);
            cb = FP<E>::madd(FP<E>::mul(m0.z, g), lcb, cb);
            const V3 R = reflect3<E>(Ld, N);
            const float ps = dot3<E>(R, d);
            if (ps > 0.f) {
                const float ph = FP<E>::mul(FP<E>::mul(l0.w, m1.y), powf(ps, m1.z));
                cr = FP<E>::madd(lcr, ph, cr); cg = FP<E>::madd(lcg, ph, cg); cb = FP<E>::madd(lcb, ph, cb);
            }
        }
    }
    if (p.level == 0) { float* c = p.c0 + 3 * (size_t)i; c[0] = cr; c[1] = cg; c[2] = cb; }
    else p.colour[i] = make_float4(cr, cg, cb, 0.f);
}

// ---------------------------------------------------------------------------------------
// Ray-level query plumbing (rtb200_trace_nearest / rtb200_trace_shadow through the tiled kernels)
// ---------------------------------------------------------------------------------------
__global__ void k_query_pack_rays(int n, const float* o3, const float* d3, float4* od0, float4* od1)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    od0[i] = make_float4(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2], 1.0f);
    od1[i] = make_float4(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2], 1.0f);
}

template <bool E>
__global__ void k_query_finish_nearest(DeviceScene s, int n, const float4* od0, const float4* od1, const float2* hit,
                                       int* out_order, float* out_t, float* out_n3)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 h = hit[i];
    const int id = __float_as_int(h.y);
    V3 N = mk3(0.f, 0.f, 0.f);
    if (id >= 0) {
        const V3 o = f4_xyz(od0[i]), d = f4_xyz(od1[i]);
        const float4 r0 = s.surf[id].r0, r1 = s.surf[id].r1;
        const int type = __float_as_int(r0.w);
        if (type == SURF_TRIANGLE) N = f4_xyz(s.surf[id].r3);
        else if (type == SURF_PLANE) N = f4_xyz(r0);
        else {
            const float4 r2 = s.surf[id].r2;
            const Quadric Q = { r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z, r2.w };
            bool lin;
            (void)quadric_t<E>(o, d, Q, lin);
            N = quadric_normal<E>(o, d, h.x, Q, lin);
        }
    }
    out_order[i] = id; out_t[i] = h.x;
    out_n3[3 * i] = N.x; out_n3[3 * i + 1] = N.y; out_n3[3 * i + 2] = N.z;
}

template <bool E>
__global__ void k_query_pack_shadow(int n, const float* p3, const float* l3, float4* sq0, float4* sq1)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 P = ld3(p3 + 3 * (size_t)i);
    V3 Ld = sub3<E>(ld3(l3 + 3 * (size_t)i), P);
    const float dist = norm3<E>(Ld);
    Ld = scale3<E>(Ld, FP<E>::rcp(dist));
    sq0[i] = make_float4(P.x, P.y, P.z, dist);
    sq1[i] = make_float4(Ld.x, Ld.y, Ld.z, 0.f);
}

__global__ void k_query_unpack_filter(int n, const float4* filter, float* out3)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 f = filter[i];
    out3[3 * i] = f.x; out3[3 * i + 1] = f.y; out3[3 * i + 2] = f.z;
}

cudaError_t launch_query_pack_rays(int n, const float* o3, const float* d3, float4* od0, float4* od1, cudaStream_t st)
{
    k_query_pack_rays<<<(n + 255) / 256, 256, 0, st>>>(n, o3, d3, od0, od1);
    return cudaGetLastError();
}
cudaError_t launch_query_finish_nearest(const DeviceScene& s, int n, const float4* od0, const float4* od1, const float2* hit,
                                        int* out_order, float* out_t, float* out_n3, bool exact, cudaStream_t st)
{
    if (exact) k_query_finish_nearest<true><<<(n + 255) / 256, 256, 0, st>>>(s, n, od0, od1, hit, out_order, out_t, out_n3);
    else       k_query_finish_nearest<false><<<(n + 255) / 256, 256, 0, st>>>(s, n, od0, od1, hit, out_order, out_t, out_n3);
    return cudaGetLastError();
}
cudaError_t launch_query_pack_shadow(int n, const float* p3, const float* l3, float4* sq0, float4* sq1, bool exact, cudaStream_t st)
{
    if (exact) k_query_pack_shadow<true><<<(n + 255) / 256, 256, 0, st>>>(n, p3, l3, sq0, sq1);
    else       k_query_pack_shadow<false><<<(n + 255) / 256, 256, 0, st>>>(n, p3, l3, sq0, sq1);
    return cudaGetLastError();
}
cudaError_t launch_query_unpack_filter(int n, const float4* filter, float* out3, cudaStream_t st)
{
    k_query_unpack_filter<<<(n + 255) / 256, 256, 0, st>>>(n, filter, out3);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------
template <int MODE, bool PRIMARY>
static cudaError_t launch_nearest_t(const TraceParams& p, int blocks, cudaStream_t st)
{
    k_trace_nearest<MODE, PRIMARY><<<blocks, ST_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_trace_nearest(const TraceParams& p, int sm_count, cudaStream_t st)
{
    const int tiles_x = (p.f.LW + 31) / 32;
    const int n_blocks = p.primary ? tiles_x * ((p.chunk_rows + 31) / 32) : (p.n_rays + ST_BLOCK_RAYS - 1) / ST_BLOCK_RAYS;
    if (n_blocks <= 0) return cudaSuccess;
    const int blocks = n_blocks < 2 * sm_count ? n_blocks : 2 * sm_count;      // persistent: 2 CTAs per SM
    if (p.primary) {
        if (p.mode == 1) return launch_nearest_t<1, true>(p, blocks, st);
        if (p.mode == 2) return launch_nearest_t<2, true>(p, blocks, st);
        return launch_nearest_t<0, true>(p, blocks, st);
    }
    if (p.mode == 1) return launch_nearest_t<1, false>(p, blocks, st);
    if (p.mode == 2) return launch_nearest_t<2, false>(p, blocks, st);
    return launch_nearest_t<0, false>(p, blocks, st);
}

cudaError_t launch_trace_shadow(const ShadowParams& p, int sm_count, cudaStream_t st)
{
    const int n_blocks = (p.n_queries + ST_BLOCK_RAYS - 1) / ST_BLOCK_RAYS;
    if (n_blocks <= 0) return cudaSuccess;
    const int blocks = n_blocks < 2 * sm_count ? n_blocks : 2 * sm_count;
    if (p.mode == 1) k_trace_shadow<1><<<blocks, ST_THREADS, 0, st>>>(p);
    else if (p.mode == 2) k_trace_shadow<2><<<blocks, ST_THREADS, 0, st>>>(p);
    else k_trace_shadow<0><<<blocks, ST_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_shade_a(const ShadeAParams& p, cudaStream_t st)
{
    if (p.n_rays <= 0) return cudaSuccess;
    int blocks = (p.n_rays + 255) / 256;
    if (p.tile_rays && p.level == 0) blocks = ((p.f.LW + 15) / 16) * ((p.chunk_rows + 15) / 16);
    if (p.exact) k_shade_a<true><<<blocks, 256, 0, st>>>(p);
    else         k_shade_a<false><<<blocks, 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_shade_b(const ShadeBParams& p, cudaStream_t st)
{
    if (p.n_rays <= 0) return cudaSuccess;
    const int blocks = (p.n_rays + 255) / 256;
    if (p.exact) k_shade_b<true><<<blocks, 256, 0, st>>>(p);
    else         k_shade_b<false><<<blocks, 256, 0, st>>>(p);
    return cudaGetLastError();
}

} // namespace rtb
