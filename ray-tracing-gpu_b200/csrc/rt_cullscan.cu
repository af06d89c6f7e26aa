// rt_cullscan.cu -- the culled scan: the default kernels of the streamed path (large scenes, brute-force semantics).
//
// The reference tests every ray against every surface (Scene.cpp:1710-1715, :1850-1858).  The plain tiled kernels of
// rt_trace.cu do exactly that at 2/3 of the FP32 peak; this file gets the SAME answers -- every kept value is still
// produced by the reference's rounding sequence -- without entering most of those tests:
//   * once per scene (cull_prepare, all on the device): the triangles are sorted by the Morton code of their centroid
//     and a 32-ary hierarchy of bounding spheres is laid over the sorted array -- level 0 one sphere per triangle,
//     level l one sphere per block of 32 level-(l-1) entries (no pointers: entry i of level l covers entries
//     32 i .. 32 i + 31 of level l - 1);
//   * a WARP holds 128 rays (4 per lane) and the cone(s) that contain them (rt_bundle.cuh: primary rays one cone from
//     the camera, secondary rays up to four, shadow queries up to two per light with the light as apex).  It walks
//     the hierarchy top-down: in a round lane j tests entry j of a block against the cones, the ballot of the
//     survivors is the list of blocks to descend into -- one coalesced 512-byte load and ~50 instructions per lane
//     decide 32 spheres;
//   * the triangles that survive level 0 are queued per warp (shared memory) and then tested ray by ray with the
//     two-phase filter + reference sequence of rt_cull.cuh, four rays per lane interleaved, the next triangle's
//     record prefetched while the current one is tested;
//   * warps never meet: no CTA barrier anywhere (the tile loop of rt_trace.cu met at two barriers per 256 triangles
//     with survivor counts that differed by orders of magnitude: ncu, profiles/r1_c3_shadow_culled_ncu_full.txt);
//     they take tasks of 128 rays from an atomic counter, so a slow bundle delays nobody;
//   * quadrics of centred form (spheres, ellipsoids) are culled per warp the same way, lane j looking at quadric j.
// A sphere that is culled contains no point at which any ray of the bundle can be REPORTED to hit the triangle by
// CTriangle::Intersection -- the contract and the pad of rt_bundle.cuh; a block's sphere contains the inflated spheres
// of everything below it (cull_prepare), so culling a block culls only what its members' own tests would cull.
// Shadow filters: Scene.cpp:1856 multiplies in insertion order; the walk meets the occluders in Morton order.  Up to
// two factors commute exactly; a query with three or more occluders met out of order and a non-zero product is redone
// by k_shadow_fixup in insertion order (counted in rtb200_stats.shadow_fixups).
#include <algorithm>
#include <cfloat>
#include <climits>

#include <cub/cub.cuh>

#include "rt_context.h"
#include "rt_cullscan.cuh"
#include "rt_bundle.cuh"

namespace rtb {

#define CS_THREADS    128
#define CS_WARPS      (CS_THREADS / 32)
#ifndef CS_PAIR
#define CS_PAIR       1            // nearest-hit walk, level 0: two blocks of 32 spheres per round (see cs_keep2)
#endif
#define CS_QCAP       192          // per-warp queue of surviving triangles; drained when a round's worth of slots (32 or 64) is not free
#define CS_MAX_CONES  16           // shadow: 8 lights x 2; nearest: 4
#define CS_NEAREST_CONES 4
#ifndef CS_MIN_CTAS
#define CS_MIN_CTAS   4
#endif

// offsets of the levels inside the one array that holds them all: [level 3][level 2][level 1][level 0]
RT_DEV int cs_level_offset(int n3, int level)
{
    return level == 0 ? 1057 * n3 : (level == 1 ? 33 * n3 : (level == 2 ? n3 : 0));
}

struct CsWarpSmem {
    Bundle cones[CS_MAX_CONES];
    int    queue[CS_QCAP];
    TriRec stage[32];              // the next 32 queued triangles, fetched by 32 lanes at once (one L2 latency per 32
                                   // triangles instead of one per triangle: a lone warp scanning everything -- a bundle
                                   // no cone fits -- was bound by exactly that latency)
};

template <bool LIGHT>
RT_DEV bool cs_keep(const Bundle* cones, int n_c, bool keep_all, float4 sph, float w)
{
    bool keep = keep_all;
    for (int k = 0; k < n_c; ++k) {
        const Bundle& b = cones[k];
        const bool culled = LIGHT ? light_cone_culls(b, sph, w) : bundle_culls(b, mk3(sph.x, sph.y, sph.z), sph.w, w);
        keep = keep || !culled;
    }
    return keep && w > 0.f;
}
// the same for two spheres at once: two independent dependency chains per lane (the cone test is one chain of ~35
// operations, and with four warps per scheduler the walk waits on exactly those latencies: ncu, stall_wait ~2 per issue).
// Used by the nearest-hit walk (config 3: 5.30 -> 4.98 ms, a 20 k-triangle height field 2.55 -> 2.36 ms); the shadow
// walk, whose light cones are cheaper to test, lost 4 % with it on the height field and keeps one block per round.
template <bool LIGHT>
RT_DEV void cs_keep2(const Bundle* cones, int n_c, bool keep_all, float4 sph_a, float w_a, float4 sph_b, float w_b, bool& keep_a, bool& keep_b)
{
    bool ka = keep_all, kb = keep_all;
    for (int k = 0; k < n_c; ++k) {
        const Bundle& b = cones[k];
        const bool ca = LIGHT ? light_cone_culls(b, sph_a, w_a) : bundle_culls(b, mk3(sph_a.x, sph_a.y, sph_a.z), sph_a.w, w_a);
        const bool cb = LIGHT ? light_cone_culls(b, sph_b, w_b) : bundle_culls(b, mk3(sph_b.x, sph_b.y, sph_b.z), sph_b.w, w_b);
        ka = ka || !ca; kb = kb || !cb;
    }
    keep_a = ka && w_a > 0.f; keep_b = kb && w_b > 0.f;
}

// Top-down walk of the sphere hierarchy with the warp's cones, RESUMABLE: cs_fill appends the triangles that survive
// level 0 to the warp's queue until fewer than 32 slots are free (returns false: drain the queue and call again) or
// the hierarchy is exhausted (returns true).  The walk's state -- per level the mask of blocks still to descend into
// and the index of their first entry -- is warp-uniform and lives in registers.  One copy of the cone test, one copy
// of the drain in the caller: the first version inlined both per level and per call site, 105 KB of SASS, and the
// warps stalled on instruction fetch more than on anything else (ncu: no_instruction 3.4 warps per issue).
struct CsWalk {
    int lvl, top_b, qn;              // lvl: level whose mask is being consumed (CS_LEVELS = fetch the next top block)
    unsigned int m1, m2, m3;
    int b1, b2, b3;
};
RT_DEV void cs_walk_init(CsWalk& w) { w.lvl = CS_LEVELS; w.top_b = 0; w.qn = 0; w.m1 = w.m2 = w.m3 = 0u; w.b1 = w.b2 = w.b3 = 0; }

template <bool LIGHT>
RT_DEV bool cs_fill(const CullScene& cs, CsWarpSmem& ws, int n_c, bool keep_all, int lane, CsWalk& w)
{
    constexpr bool PAIR = CS_PAIR && !LIGHT;
    [[maybe_unused]] constexpr int ROUND = PAIR ? 64 : 32;  // queue slots a round can fill; the queue is drained above 128 entries
    for (;;) {
        int level, block, block2 = -1;                      // the block(s) of 32 entries to test now
        if (w.lvl == CS_LEVELS) {
            if (w.top_b >= cs.top_blocks) return true;
            level = CS_LEVELS - 1; block = w.top_b++;
        } else {
            const unsigned int m = w.lvl == 1 ? w.m1 : (w.lvl == 2 ? w.m2 : w.m3);
            if (m == 0u) { ++w.lvl; continue; }
            const int base = w.lvl == 1 ? w.b1 : (w.lvl == 2 ? w.b2 : w.b3);
            block = base + __ffs((int)m) - 1;
            unsigned int rest = m & (m - 1u);
            if (PAIR && w.lvl == 1 && rest) { block2 = base + __ffs((int)rest) - 1; rest &= rest - 1u; }
            if (w.lvl == 1) w.m1 = rest; else if (w.lvl == 2) w.m2 = rest; else w.m3 = rest;
            level = w.lvl - 1;
        }
        const int i = cs_level_offset(cs.n3, level) + block * 32 + lane;
        RT_CHECK(level >= 0 && level < CS_LEVELS && block >= 0 && i >= 0 && i < cs.n3 * (1057 + 32768));
        bool keep = false, keep2 = false;
        const bool have = level < CS_LEVELS - 1 || block * 32 + lane < cs.n3;           // (the top level holds n3 entries, not a multiple of 32)
        if (PAIR) {
            // ONE call site: a lone block is paired with a sphere of weight 0, which is never kept
            const int i2 = block2 >= 0 ? cs_level_offset(cs.n3, 0) + block2 * 32 + lane : i;
            const float4 sa = __ldg(cs.sph + (have ? i : 0)), sb = __ldg(cs.sph + (have ? i2 : 0));
            const float wa = have ? __ldg(cs.w + i) : 0.f, wb = (have && block2 >= 0) ? __ldg(cs.w + i2) : 0.f;
            cs_keep2<LIGHT>(ws.cones, n_c, keep_all, sa, wa, sb, wb, keep, keep2);
        } else if (have)
            keep = cs_keep<LIGHT>(ws.cones, n_c, keep_all, __ldg(cs.sph + i), __ldg(cs.w + i));
        const unsigned int mm = __ballot_sync(FULL_MASK, keep);
        if (level == 0) {
            RT_CHECK(w.qn >= 0 && w.qn + ROUND <= CS_QCAP);
            const unsigned int mm2 = PAIR ? __ballot_sync(FULL_MASK, keep2) : 0u;
            const unsigned int below = (1u << lane) - 1u;
            if (keep) ws.queue[w.qn + __popc(mm & below)] = block * 32 + lane;
            if (PAIR && keep2) ws.queue[w.qn + __popc(mm) + __popc(mm2 & below)] = block2 * 32 + lane;
            w.qn += __popc(mm) + __popc(mm2);
            if (w.qn > 128) return false;
        } else {
            if (level == 1) { w.m1 = mm; w.b1 = block * 32; } else if (level == 2) { w.m2 = mm; w.b2 = block * 32; } else { w.m3 = mm; w.b3 = block * 32; }
            w.lvl = level;
        }
    }
}

// ---------------------------------------------------------------------------------------
// Per-RAY walk -- for the bundles no cone fits (rays of a deep wavefront level: neighbours in space that fly apart in
// every direction).  Walking the hierarchy with such a bundle keeps everything, i.e. a scan of all triangles by one
// warp; instead every ray descends the same hierarchy ALONE (cs_ray_misses: the cone tests of rt_bundle.cuh for a cone
// of zero opening, same pad), carried by a group of lanes (cs_lane_walk below).  The groups of a warp step in lock
// step -- each over the 32 entries of its own current block.  ~50 blocks instead of 10 000 triangles per ray.
// ---------------------------------------------------------------------------------------
// One ray as the lanes see it: the half line from `o` along the unit vector `u`, hits within [0, far] of o.
// LIGHT: o = the light, u = from the light towards the hit point, far = their distance (widened); the shadow ray's own
// origin -- what the pad of a triangle scales with -- lies within dadd = far of o.
struct LaneRay {
    V3 o, u;
    float far, dadd, padk, slack;
    bool ok;                          // false: non-finite or degenerate ray -> the lane keeps everything
};
RT_DEV LaneRay cs_lane_ray(V3 o, V3 d)
{
    LaneRay r;
    const float l2 = dot3<false>(d, d);
    r.ok = finite3(o) && finite3(d) && l2 > 1.0e-20f && l2 < 1.0e20f;
    const float il = r.ok ? rsqrtf(l2) : 0.f;
    r.o = o; r.u = mk3(d.x * il, d.y * il, d.z * il);
    r.far = 3.0e38f;
    r.slack = 4.0e-7f * (fabsf(o.x) + fabsf(o.y) + fabsf(o.z)) + 1.0e-6f;
    r.dadd = r.slack;
    r.padk = BUNDLE_PADK * fmaxf(1.0f, sqrtf(l2) * 1.00001f);
    return r;
}
RT_DEV LaneRay cs_lane_query(V3 L, V3 d, float dist)          // d = unit direction TOWARDS the light L
{
    LaneRay r;
    r.ok = finite3(L) && finite3(d) && dist >= 0.f && dist <= 3.0e38f;
    r.o = L; r.u = neg3(d);
    r.far = fmaf(dist, 1.0001f, 1.0e-4f);
    r.slack = 4.0e-7f * (fabsf(L.x) + fabsf(L.y) + fabsf(L.z)) + 1.0e-6f;
    r.dadd = r.far + r.slack;
    r.padk = BUNDLE_PADK * 1.0001f;
    return r;
}
// true: no point of the INFLATED sphere (pad of rt_bundle.cuh for origins within |c - o| + dadd + radius) lies on the ray.
// The tests of bundle_culls / light_cone_culls for a cone of zero opening: distance of the centre to the line against
// the inflated radius (the cancellation of |v|^2 - s^2 and |u| = 1 +- 3e-7 are covered by 4e-6 |v|^2), sphere entirely
// behind the origin, sphere entirely beyond the far end.
RT_DEV bool cs_ray_misses(const LaneRay& r, float4 sph, float w)
{
    const float vx = sph.x - r.o.x, vy = sph.y - r.o.y, vz = sph.z - r.o.z;
    const float s = fmaf(vx, r.u.x, fmaf(vy, r.u.y, vz * r.u.z));
    const float vv = fmaf(vx, vx, fmaf(vy, vy, vz * vz));
    const float dist = vv * rsqrtf(fmaxf(vv, 1.0e-30f)) * 1.00001f;
    const float D = dist + r.dadd + sph.w;
    const float pad = fmaxf(r.padk * D * w, 3.8e-6f * D);
    const float rp = fmaf(sph.w + pad, 1.0001f, 1.0e-5f + r.slack + 2.0e-6f * dist);
    const float perp2 = fmaf(-s, s, vv) - 4.0e-6f * vv;
    return perp2 > rp * rp || s < -rp || s > r.far + rp;
}

// A ray is walked by a GROUP of CS_LANE_G adjacent lanes (a warp holds 32 / CS_LANE_G rays): the lanes of a group
// share the ray and the walk's state (every lane keeps a copy), each tests its share of a block's 32 spheres -- entry
// k G + g for lane g, so that neighbours read neighbours -- and the union of their masks (log2 G shuffles) is the
// block's mask.  The triangles that survive level 0 are dealt out by rank, every lane testing its own.  Why: with one
// ray per lane a warp's walk is ~100 lock-step rounds of 32 dependent sphere tests + the leaf tests, 0.8 ms for ONE
// slice of 32 rays whatever the level's size -- and a kernel cannot end before its slowest warp: the launch list of
// config 3 showed the lane kernels of levels 1, 2 and 3 at 0.84 / 0.81 / 0.83 ms for 8 : 2.4 : 1 rays.
#ifndef CS_LANE_G
#define CS_LANE_G 8
#endif
static_assert(CS_LANE_G == 1 || CS_LANE_G == 2 || CS_LANE_G == 4 || CS_LANE_G == 8 || CS_LANE_G == 16, "group size");

// leaf(first, mask): the group's ray may touch triangles first + j for the bits j of mask (called by every lane of the
// group with the same arguments; the lanes diverge inside).  stop(): evaluated by every lane at the top of a round, the
// group leaves the walk when any of its lanes says so.
template <class Leaf, class Stop>
RT_DEV void cs_lane_walk(const CullScene& cs, const LaneRay& b, bool active, int lane, Leaf&& leaf, Stop&& stop)
{
    constexpr int G = CS_LANE_G;
    const int g = lane & (G - 1);
    const unsigned int group_bits = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << (lane & ~(G - 1));
    int lvl = CS_LEVELS, top_b = 0;
    unsigned int m1 = 0u, m2 = 0u, m3 = 0u;
    int b1 = 0, b2 = 0, b3 = 0;
    bool busy = active;
    for (;;) {
        const unsigned int stops = __ballot_sync(FULL_MASK, busy && stop());
        if (stops & group_bits) busy = false;
        if (!__any_sync(FULL_MASK, busy)) break;
        // ---- which block does this group test now? -----------------------------------------------
        int level = -1, block = 0;
        while (busy && level < 0) {
            if (lvl == CS_LEVELS) {
                if (top_b >= cs.top_blocks) { busy = false; break; }
                level = CS_LEVELS - 1; block = top_b++;
            } else {
                const unsigned int m = lvl == 1 ? m1 : (lvl == 2 ? m2 : m3);
                if (m == 0u) { ++lvl; continue; }
                const int base = lvl == 1 ? b1 : (lvl == 2 ? b2 : b3);
                block = base + __ffs((int)m) - 1;
                const unsigned int rest = m & (m - 1u);
                if (lvl == 1) m1 = rest; else if (lvl == 2) m2 = rest; else m3 = rest;
                level = lvl - 1;
            }
        }
        // ---- the block's 32 entries against the group's ray, 32 / G per lane (lock step over the warp) ---
        unsigned int mm = 0u;
        if (level >= 0) {
            const int i0 = cs_level_offset(cs.n3, level) + block * 32;
            const int n_in = level == CS_LEVELS - 1 ? min(32, cs.n3 - block * 32) : 32;
            #pragma unroll 4
            for (int k = 0; k < 32 / G; ++k) {
                const int j = k * G + g;
                if (j < n_in) {
                    const float4 sph = __ldg(cs.sph + i0 + j);
                    const float w = __ldg(cs.w + i0 + j);
                    const bool keep = w > 0.f && !(b.ok && cs_ray_misses(b, sph, w));
                    mm |= keep ? (1u << j) : 0u;
                }
            }
        }
        #pragma unroll
        for (int off = 1; off < G; off <<= 1) mm |= __shfl_xor_sync(FULL_MASK, mm, off);     // (every lane: groups without a block pass 0)
        if (level >= 0) {
            if (level == 0) { if (mm) leaf(block * 32, mm); }
            else {
                if (level == 1) { m1 = mm; b1 = block * 32; } else if (level == 2) { m2 = mm; b2 = block * 32; } else { m3 = mm; b3 = block * 32; }
                lvl = level;
            }
        }
    }
}
// the bits of `mask` this lane of the group takes: every G-th one by rank
RT_DEV unsigned int cs_lane_share(unsigned int mask, int g)
{
    constexpr int G = CS_LANE_G;
    if (G == 1) return mask;
    unsigned int mine = 0u;
    for (int rank = 0; mask; ++rank) {
        const unsigned int low = mask & (0u - mask);
        if ((rank & (G - 1)) == g) mine |= low;
        mask ^= low;
    }
    return mine;
}

// nearest hit of ONE ray per group over the triangles: every lane returns the best of ITS share (the caller combines).
// Kept out of line: rare path, and the kernels are large enough.
template <int MODE>
__device__ __noinline__ int cs_lane_nearest(const CullScene& cs, V3 o, V3 d, bool active, int lane, float& best_t, int& best_id)
{
    const LaneRay b = cs_lane_ray(o, d);
    const int g = lane & (CS_LANE_G - 1);
    float bt = best_t;
    int bi = best_id, tests = 0;
    const float dmax = max_abs3(d);
    cs_lane_walk(cs, b, active, lane, [&](int first, unsigned int mask) {
        mask = cs_lane_share(mask, g);
        while (mask) {
            const TriRec* t = cs.tri + first + __ffs((int)mask) - 1;
            mask &= mask - 1u;
            const float4 ra = __ldg(&t->a), rb = __ldg(&t->b), rc = __ldg(&t->c);
            const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
            const TriMargins mg = tri_margins(e1, e2, dmax);
            ++tests;
            if (!triangle_phase1(o, d, p0, e1, e2, mg)) continue;
            const float tt = triangle_survivor<MODE>(o, d, p0, e1, e2, bt < 0.f ? 3.0e38f : fmaf(bt, CULL_TREL, CULL_TABS), mg);
            const int ord = __float_as_int(ra.w);
            if (tt > RT_EPS && (tt < bt || bt < 0.f || (tt == bt && ord < bi))) { bt = tt; bi = ord; }
        }
    }, []() { return false; });
    best_t = bt; best_id = bi;
    return tests;
}

// shadow filter of ONE query per group over the triangles: every lane returns the product over the occluders of ITS
// share and their number (saturating at 3); a lane that holds a zero product stops the group
template <int MODE>
__device__ __noinline__ int cs_lane_shadow(const CullScene& cs, const MatRec* __restrict__ mat, V3 L, V3 o, V3 d, float dist, bool active, int lane,
                                           float& fr, float& fg, float& fb, unsigned int& occ)
{
    const LaneRay b = cs_lane_query(L, d, dist);
    const int g = lane & (CS_LANE_G - 1);
    float r_ = fr, g_ = fg, b_ = fb;
    unsigned int n_occ = occ;
    int tests = 0;
    const float dmax = max_abs3(d), tlim = fmaf(dist, CULL_TREL, CULL_TABS);
    cs_lane_walk(cs, b, active, lane, [&](int first, unsigned int mask) {
        mask = cs_lane_share(mask, g);
        while (mask) {
            const TriRec* t = cs.tri + first + __ffs((int)mask) - 1;
            mask &= mask - 1u;
            if (r_ == 0.f && g_ == 0.f && b_ == 0.f) return;             // 0 * finite stays 0
            const float4 ra = __ldg(&t->a), rb = __ldg(&t->b), rc = __ldg(&t->c);
            const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
            const TriMargins mg = tri_margins(e1, e2, dmax);
            ++tests;
            if (!triangle_phase1(o, d, p0, e1, e2, mg)) continue;
            const float tt = triangle_survivor<MODE>(o, d, p0, e1, e2, tlim, mg);
            if (tt > RT_EPS && tt < dist) {
                const int m = __float_as_int(rb.w);
                const float4 m0 = mat[m].m0;
                const float kt = mat[m].m2.x;
                r_ = __fmul_rn(r_, __fmul_rn(m0.x, kt)); g_ = __fmul_rn(g_, __fmul_rn(m0.y, kt)); b_ = __fmul_rn(b_, __fmul_rn(m0.z, kt));
                if (n_occ < 3u) ++n_occ;
            }
        }
    }, [&]() { return r_ == 0.f && g_ == 0.f && b_ == 0.f; });
    fr = r_; fg = g_; fb = b_; occ = n_occ;
    return tests;
}

// quadrics of centred form  sum q_i (x_i - c_i)^2 = rhs, q_i > 0 (r3 = c | rhs, rhs > 0): bounding sphere of the region
// outside of which the reference's discriminant cannot come out positive for any origin within D of the centre
// (the margin of quadric_miss, rt_resident.cu / DESIGN.md 5.3: rhs * 1.05 + 2e-5 * M, M = sum q_i (o_i^2 + c_i^2) with
// |o_i| <= |c_i| + D).  Returns false when the quadric has no such sphere (cylinders, cones, rotated ones: always kept).
RT_DEV bool cs_quadric_sphere(float4 q0, float4 r3, float D, float& radius)
{
    if (!(r3.w > 0.f) || !(q0.x > 0.f) || !(q0.y > 0.f) || !(q0.z > 0.f)) return false;
    const float fx = fabsf(r3.x) + D, fy = fabsf(r3.y) + D, fz = fabsf(r3.z) + D;
    const float M = q0.x * (fx * fx + r3.x * r3.x) + q0.y * (fy * fy + r3.y * r3.y) + q0.z * (fz * fz + r3.z * r3.z);
    const float lim = fmaf(r3.w, 1.05f, 2.0e-5f * M * 1.001f);
    const float qmin = fminf(q0.x, fminf(q0.y, q0.z));
    radius = sqrtf(lim / qmin) * 1.0001f;
    return radius <= 3.0e37f;
}

template <bool LIGHT>
RT_DEV bool cs_keep_quadric(const Bundle* cones, int n_c, bool keep_all, float4 q0, float4 r3)
{
    if (keep_all) return true;
    bool keep = false;
    for (int k = 0; k < n_c; ++k) {
        const Bundle& b = cones[k];
        const float vx = r3.x - b.ax, vy = r3.y - b.ay, vz = r3.z - b.az;
        const float D = sqrtf(fmaf(vx, vx, fmaf(vy, vy, vz * vz))) * 1.0001f + b.dadd;      // every origin lies within D of the centre
        float radius;
        if (!cs_quadric_sphere(q0, r3, D, radius)) return true;
        const float4 sph = make_float4(r3.x, r3.y, r3.z, radius);
        const bool culled = LIGHT ? light_cone_culls(b, sph, 0.f) : bundle_culls(b, mk3(sph.x, sph.y, sph.z), sph.w, 0.f);
        keep = keep || !culled;
    }
    return keep;
}

// the cones of a warp's 128 nearest-hit rays: one, or -- when that one is wide -- up to four after two splits around the
// most distant directions.  Returns the number of cones written (0 with keep_all = true: nothing can be culled).
// (Loops that are NOT unrolled on purpose: every copy of bundle_of_rays / split_members is ~300 instructions, and the
// first version of this function was two thirds of the kernel's 128 KB of SASS.)
RT_DEV int cs_nearest_cones(const V3* o, const V3* d, const bool* use, Bundle* cones, int lane, bool& keep_all)
{
    keep_all = false;
    int grp[ST_RPT];                                            // group of each ray after the splits (2 bits)
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) grp[r] = 0;
    Bundle one;
    int n = 0;
    bool usable = true, split = false;
    #pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        // pass 0: the cone of all rays; pass 1 (only when that one is wide): the cones of the four groups
        const int n_groups = pass == 0 ? 1 : CS_NEAREST_CONES;
        #pragma unroll 1
        for (int k = 0; k < n_groups; ++k) {
            bool in[ST_RPT];
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) in[r] = use[r] && (pass == 0 || grp[r] == k);
            const Bundle b = bundle_of_rays(o, d, in);
            if (pass == 0) one = b;
            else {
                usable = usable && b.state != 2;
                if (b.state == 1) { if (lane == 0) cones[n] = b; ++n; }
            }
        }
        if (pass == 0) {
            if (one.state == 0) return 0;                       // no rays at all
            if (one.state == 1 && !(one.sin_t > BUNDLE_SPLIT_SIN)) break;
            split = true;
            V3 dn[ST_RPT];
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) {
                const float l2 = dot3<false>(d[r], d[r]);
                const float il = (l2 > 1.0e-20f && l2 < 1.0e20f) ? rsqrtf(l2) : 0.f;
                dn[r] = mk3(d[r].x * il, d[r].y * il, d[r].z * il);
            }
            #pragma unroll 1
            for (int it = 0; it < 3; ++it) {                    // all rays in two, then each half in two
                bool in[ST_RPT], in_a[ST_RPT], in_b[ST_RPT];
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) in[r] = use[r] && (it == 0 || (grp[r] & 1) == it - 1);
                split_members(dn, in, in_a, in_b);
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) if (in_b[r]) grp[r] |= (it == 0 ? 1 : 2);
            }
        }
    }
    if (!split || !usable) {
        if (one.state == 1) { if (lane == 0) cones[0] = one; __syncwarp(); return 1; }
        keep_all = true;
        return 0;
    }
    __syncwarp();
    return n;
}

// ---------------------------------------------------------------------------------------
// Nearest hit (Scene.cpp:1705-1720).  One warp task = 128 rays: PRIMARY a 16 x 8 block of the sample lattice
// (thread = a 2 x 2 quad, rays generated on the fly), otherwise 128 consecutive queue entries.
// ---------------------------------------------------------------------------------------
// A task no (narrow) cone fits is put on cp.lane_list instead of being traced: k_lane_nearest traces those ray by ray
// -- its own small kernel, so that the rare path costs this one neither registers nor instruction cache.
template <int MODE, bool PRIMARY>
__global__ void __launch_bounds__(CS_THREADS, CS_MIN_CTAS)
k_cull_nearest(const __grid_constant__ CullTraceParams cp)
{
    __shared__ CsWarpSmem sm[CS_WARPS];
    const TraceParams& p = cp.t;
    const CullScene& cs = cp.cs;
    const DeviceScene& s = p.s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    CsWarpSmem& ws = sm[warp];
    const int tiles_x = (p.f.LW + 15) >> 4;
    const int shift = PRIMARY ? 7 : cp.task_shift, rpt = 1 << (shift - 5);       // rays per lane in this launch: 4 or 1
    const int n_tasks = PRIMARY ? tiles_x * ((p.chunk_rows + 7) >> 3) : (p.n_rays + (1 << shift) - 1) >> shift;
    unsigned long long exec_rays = 0;

    for (;;) {
        int task = 0;
        if (lane == 0) task = (int)atomicAdd(cp.task_counter, 1u);
        task = __shfl_sync(FULL_MASK, task, 0);
        if (task >= n_tasks) break;

        V3 o[ST_RPT], d[ST_RPT];
        float best_t[ST_RPT], tlim[ST_RPT];
        int best_id[ST_RPT], slot[ST_RPT];
        bool use[ST_RPT];
        int n_valid = 0;
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) {
            bool valid;
            if (PRIMARY) {
                const int tx = task % tiles_x, ty = task / tiles_x;
                const int lx = (tx << 4) + ((lane & 7) << 1) + (r & 1);
                const int cy = (ty << 3) + ((lane >> 3) << 1) + (r >> 1);
                const int ly = p.chunk_ly0 + cy;
                valid = lx < p.f.LW && cy < p.chunk_rows && ly < p.local_LH;
                slot[r] = valid ? cy * p.f.LW + lx : -1;
                primary_ray<(MODE != 2)>(p.f, valid ? lx : 0, valid ? global_lattice_row(p.f, ly) : 0, o[r], d[r]);
            } else {
                const int i0 = (task << shift) + r * 32 + lane;
                valid = r < rpt && i0 < p.n_rays;
                const int i = (valid && cp.perm) ? cp.perm[i0] : i0;
                RT_CHECK(!valid || (i >= 0 && i < p.n_rays));
                slot[r] = valid ? i : -1;
                const float4 a = valid ? p.in_od0[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 b = valid ? p.in_od1[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                o[r] = f4_xyz(a); d[r] = f4_xyz(b);
            }
            if (!valid) d[r] = mk3(0.f, 0.f, 0.f);     // zero direction: det = 0, nothing can hit
            best_t[r] = -1.0f; best_id[r] = -1; tlim[r] = 3.0e38f;
            use[r] = valid;
            n_valid += valid ? 1 : 0;
        }
        float dmax = 0.f;                             // largest |direction component| of this thread's rays (filter margins)
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) dmax = fmaxf(dmax, max_abs3(d[r]));
        bool keep_all = true;
        const int n_c = cs_nearest_cones(o, d, use, ws.cones, lane, keep_all);
        n_valid = __reduce_add_sync(FULL_MASK, n_valid);
        bool to_lanes = keep_all;                                          // no cone fits, or only wide ones: the lane kernel takes this task
        for (int k = 0; k < n_c; ++k) to_lanes = to_lanes || ws.cones[k].sin_t > cp.lane_sin;
        if (to_lanes && n_valid > 0 && cp.lane_list) {
            unsigned int k = 0;
            if (lane == 0) k = atomicAdd(cp.lane_count, 1u);
            k = __shfl_sync(FULL_MASK, k, 0);
            if ((int)k < cp.lane_cap) { if (lane == 0) cp.lane_list[k] = task; continue; }
        }

        // ---- triangles ------------------------------------------------------------------------
        auto test_triangle = [&](const float4 ra, const float4 rb, const float4 rc) {
            const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
            const TriMargins mg = tri_margins(e1, e2, dmax);
            bool cand[ST_RPT];
            cand[0] = triangle_phase1(o[0], d[0], p0, e1, e2, mg);
            bool any = cand[0];
            #pragma unroll
            for (int r = 1; r < ST_RPT; ++r) cand[r] = false;
            if (rpt > 1) {                                  // (tasks of 32 rays -- the small levels -- fill slot 0 only)
                #pragma unroll
                for (int r = 1; r < ST_RPT; ++r) {
                    cand[r] = triangle_phase1(o[r], d[r], p0, e1, e2, mg);
                    any |= cand[r];
                }
            }
            if (__any_sync(FULL_MASK, any)) {
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    if (cand[r] && slot[r] >= 0) {
                        const float tt = triangle_survivor<MODE>(o[r], d[r], p0, e1, e2, tlim[r], mg);
                        const int ord = __float_as_int(ra.w);
                        // Morton order is not insertion order: equal distances go to the lower order (Scene.cpp:1713)
                        if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f || (tt == best_t[r] && ord < best_id[r]))) {
                            best_t[r] = tt; best_id[r] = ord;
                            tlim[r] = fmaf(tt, CULL_TREL, CULL_TABS);
                        }
                    }
                }
            }
        };
        if (n_c > 0 || keep_all) {
            CsWalk wk;
            cs_walk_init(wk);
            bool finished;
            do {
                finished = cs_fill<false>(cs, ws, n_c, keep_all, lane, wk);
                __syncwarp();
                const int qn = wk.qn;
                #pragma unroll 1
                for (int e0 = 0; e0 < qn; e0 += 32) {
                    const int cnt = min(32, qn - e0);
                    if (lane < cnt) {
                        RT_CHECK(ws.queue[e0 + lane] >= 0 && ws.queue[e0 + lane] < 32768 * cs.n3);
                        const TriRec* t = cs.tri + ws.queue[e0 + lane];
                        const float4 a = __ldg(&t->a), b = __ldg(&t->b), c = __ldg(&t->c);
                        ws.stage[lane].a = a; ws.stage[lane].b = b; ws.stage[lane].c = c;
                    }
                    __syncwarp();
                    #pragma unroll 1
                    for (int e = 0; e < cnt; ++e) test_triangle(ws.stage[e].a, ws.stage[e].b, ws.stage[e].c);
                    __syncwarp();
                }
                exec_rays += (unsigned long long)qn * (unsigned long long)n_valid;
                wk.qn = 0;
                __syncwarp();
            } while (!finished);
        }

        // ---- quadrics (lane j looks at quadric j for the warp) and planes; ties go to the lower insertion order ----
        for (int q0i = 0; q0i < s.n_quad; q0i += 32) {
            const int qi = q0i + lane;
            bool keep = false;
            if (qi < s.n_quad && n_valid > 0) {
                const float4 q0 = s.quad[qi].q0;
                keep = cs_keep_quadric<false>(ws.cones, n_c, keep_all || n_c == 0, q0, s.surf[__float_as_int(q0.w)].r3);
            }
            unsigned int mask = __ballot_sync(FULL_MASK, keep);
            while (mask) {
                const int q = q0i + __ffs((int)mask) - 1;
                mask &= mask - 1u;
                const float4 q0 = s.quad[q].q0, q1 = s.quad[q].q1, q2 = s.quad[q].q2;
                const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                const int ord = __float_as_int(q0.w);
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    if (r >= rpt) break;                    // (small levels: slot 0 only)
                    bool lin;
                    const float tt = MODE == 2 ? quadric_t<false>(o[r], d[r], Q, lin) : quadric_t<true>(o[r], d[r], Q, lin);
                    if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f || (tt == best_t[r] && ord < best_id[r]))) { best_t[r] = tt; best_id[r] = ord; }
                }
            }
        }
        for (int q = 0; q < s.n_plane; ++q) {
            const float4 nc = s.plane[q].nc;
            const int ord = s.plane[q].order;
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) {
                if (r >= rpt) break;
                const float tt = MODE == 2 ? plane_t<false>(o[r], d[r], f4_xyz(nc), nc.w) : plane_t<true>(o[r], d[r], f4_xyz(nc), nc.w);
                if (tt > RT_EPS && (tt < best_t[r] || best_t[r] < 0.f || (tt == best_t[r] && ord < best_id[r]))) { best_t[r] = tt; best_id[r] = ord; }
            }
        }
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r)
            if (slot[r] >= 0) p.hit[slot[r]] = make_float2(best_t[r], __int_as_float(best_id[r]));
    }
    if (p.stats && lane == 0 && exec_rays)
        atomicAdd(&p.stats[(size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT + ST_TEST_TRI_EXEC], exec_rays);
}

// ---------------------------------------------------------------------------------------
// Shadow filter (Scene.cpp:1842-1861): product of colour * kt over every surface hit strictly between EPSILON
// and the light.  One warp task = 128 consecutive (ray, light) queries of the queue shade-A wrote light by light.
// ---------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(CS_THREADS, CS_MIN_CTAS)
k_cull_shadow(const __grid_constant__ CullShadowParams cp)
{
    __shared__ CsWarpSmem sm[CS_WARPS];
    // per ray, rarely touched: largest insertion order among its occluders so far | occluders (saturating) and "met out of order"
    __shared__ int s_ord[ST_RPT][CS_THREADS];
    __shared__ unsigned char s_occ[ST_RPT][CS_THREADS];
    const ShadowParams& p = cp.s;
    const CullScene& cs = cp.cs;
    const DeviceScene& s = p.s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    CsWarpSmem& ws = sm[warp];
    const int n_lights = s.rl.n_light;
    const int shift = cp.task_shift, rpt = 1 << (shift - 5);
    const int n_tasks = (p.n_queries + (1 << shift) - 1) >> shift;
    unsigned long long tests_tri = 0, tests_side_q = 0, tests_side_p = 0, exec_rays = 0, fixups = 0;

    for (;;) {
        int task = 0;
        if (lane == 0) task = (int)atomicAdd(cp.task_counter, 1u);
        task = __shfl_sync(FULL_MASK, task, 0);
        if (task >= n_tasks) break;

        V3 o[ST_RPT], d[ST_RPT];
        float dist[ST_RPT], fr[ST_RPT], fg[ST_RPT], fb[ST_RPT];
        int slot[ST_RPT], lid[ST_RPT];
        int n_valid = 0;
        unsigned int light_bits = 0;
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) {
            const int i0 = (task << shift) + r * 32 + lane;
            const bool valid = r < rpt && i0 < p.n_queries;
            const int i = (valid && cp.perm) ? cp.perm[i0] : i0;
            RT_CHECK(!valid || (i >= 0 && i < p.n_queries));
            slot[r] = valid ? i : -1;
            lid[r] = valid ? p.sq_src[i].y : -1;
            n_valid += valid ? 1 : 0;
            o[r] = d[r] = mk3(0.f, 0.f, 0.f); dist[r] = 0.f;
            if (valid) { shadow_query<(MODE != 2)>(p, i, o[r], d[r], dist[r]); light_bits |= 1u << (lid[r] & 31); }
            fr[r] = fg[r] = fb[r] = valid ? 1.0f : 0.0f;             // an invalid slot counts as finished
            s_ord[r][tid] = -1; s_occ[r][tid] = 0;
        }
        n_valid = __reduce_add_sync(FULL_MASK, n_valid);
        light_bits = __reduce_or_sync(FULL_MASK, light_bits);
        float dmax = 0.f;
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r) dmax = fmaxf(dmax, max_abs3(d[r]));

        // ---- the warp's cone(s) per light present: one, or two around the two most distant directions ----------
        int n_c = 0;
        bool keep_all = n_lights > 32;
        for (int l = 0; l < n_lights && !keep_all; ++l) {
            if (!((light_bits >> l) & 1u)) continue;
            const V3 L = f4_xyz(s.light[l].l0);
            bool in[ST_RPT];
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) in[r] = lid[r] == l;
            Bundle b = bundle_of_light(L, d, dist, in), b2;
            b2 = b; b2.state = 0;
            if (b.state == 2 || (b.state == 1 && b.sin_t > BUNDLE_SPLIT_SIN)) {
                V3 nd[ST_RPT];
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) nd[r] = neg3(d[r]);
                bool in_a[ST_RPT], in_b[ST_RPT];
                split_members(nd, in, in_a, in_b);
                Bundle half[2];
                #pragma unroll 1
                for (int h = 0; h < 2; ++h) {
                    bool ih[ST_RPT];
                    #pragma unroll
                    for (int r = 0; r < ST_RPT; ++r) ih[r] = h == 0 ? in_a[r] : in_b[r];
                    const Bundle t = bundle_of_light(L, d, dist, ih);
                    if (h == 0) half[0] = t; else half[1] = t;
                }
                if (half[0].state != 2 && half[1].state != 2) { b = half[0]; b2 = half[1]; }
            }
            if (b.state == 2 || n_c + 2 > CS_MAX_CONES) { keep_all = true; break; }     // no usable cone: every triangle is tested
            RT_CHECK(n_c + 2 <= CS_MAX_CONES);
            if (b.state == 1) { if (lane == 0) ws.cones[n_c] = b; ++n_c; }
            if (b2.state == 1) { if (lane == 0) ws.cones[n_c] = b2; ++n_c; }
        }
        if (keep_all) n_c = 0;
        __syncwarp();
        bool to_lanes = keep_all;                                          // no cone fits, or only wide ones: the lane kernel takes this task
        for (int k = 0; k < n_c; ++k) to_lanes = to_lanes || ws.cones[k].sin_t > cp.lane_sin;
        if (to_lanes && n_valid > 0 && cp.lane_list) {
            unsigned int k = 0;
            if (lane == 0) k = atomicAdd(cp.lane_count, 1u);
            k = __shfl_sync(FULL_MASK, k, 0);
            if ((int)k < cp.lane_cap) { if (lane == 0) cp.lane_list[k] = task; continue; }
        }

        auto occluder = [&](int r, int ord, int mat) {
            const float4 m0 = s.mat[mat].m0;
            const float kt = s.mat[mat].m2.x;
            fr[r] = __fmul_rn(fr[r], __fmul_rn(m0.x, kt));
            fg[r] = __fmul_rn(fg[r], __fmul_rn(m0.y, kt));
            fb[r] = __fmul_rn(fb[r], __fmul_rn(m0.z, kt));
            const int prev = s_ord[r][tid];
            unsigned int st = s_occ[r][tid];
            if (ord < prev) st |= 4u;
            if ((st & 3u) < 3u) ++st;
            s_occ[r][tid] = (unsigned char)st;
            s_ord[r][tid] = max(prev, ord);
        };
        auto test_triangle = [&](const float4 ra, const float4 rb, const float4 rc) {
            const V3 p0 = f4_xyz(ra), e1 = f4_xyz(rb), e2 = f4_xyz(rc);
            const TriMargins mg = tri_margins(e1, e2, dmax);
            bool cand[ST_RPT];
            cand[0] = triangle_phase1(o[0], d[0], p0, e1, e2, mg);
            bool any = cand[0];
            #pragma unroll
            for (int r = 1; r < ST_RPT; ++r) cand[r] = false;
            if (rpt > 1) {                                  // (tasks of 32 queries -- the small levels -- fill slot 0 only)
                #pragma unroll
                for (int r = 1; r < ST_RPT; ++r) {
                    cand[r] = triangle_phase1(o[r], d[r], p0, e1, e2, mg);
                    any |= cand[r];
                }
            }
            if (__any_sync(FULL_MASK, any)) {
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    if (cand[r] && slot[r] >= 0) {
                        const float tt = triangle_survivor<MODE>(o[r], d[r], p0, e1, e2, fmaf(dist[r], CULL_TREL, CULL_TABS), mg);
                        if (tt > RT_EPS && tt < dist[r]) occluder(r, __float_as_int(ra.w), __float_as_int(rb.w));
                    }
                }
            }
        };
        auto all_zero = [&]() {
            bool done = true;
            #pragma unroll
            for (int r = 0; r < ST_RPT; ++r) done = done && (fr[r] == 0.f && fg[r] == 0.f && fb[r] == 0.f);
            return __all_sync(FULL_MASK, done);                       // 0 * finite stays 0: the warp may stop testing
        };
        bool warp_done = false;
        tests_tri += (unsigned long long)n_valid * (unsigned long long)cs.n_tri;
        if (n_valid > 0) {
            CsWalk wk;
            cs_walk_init(wk);
            bool finished;
            do {
                finished = cs_fill<true>(cs, ws, n_c, keep_all, lane, wk);
                __syncwarp();
                const int qn = wk.qn;
                #pragma unroll 1
                for (int e0 = 0; e0 < qn; e0 += 32) {
                    const int cnt = min(32, qn - e0);
                    if (lane < cnt) {
                        RT_CHECK(ws.queue[e0 + lane] >= 0 && ws.queue[e0 + lane] < 32768 * cs.n3);
                        const TriRec* t = cs.tri + ws.queue[e0 + lane];
                        const float4 a = __ldg(&t->a), b = __ldg(&t->b), c = __ldg(&t->c);
                        ws.stage[lane].a = a; ws.stage[lane].b = b; ws.stage[lane].c = c;
                    }
                    __syncwarp();
                    #pragma unroll 1
                    for (int e = 0; e < cnt; ++e) test_triangle(ws.stage[e].a, ws.stage[e].b, ws.stage[e].c);
                    __syncwarp();
                }
                exec_rays += (unsigned long long)qn * (unsigned long long)n_valid;
                wk.qn = 0;
                __syncwarp();
                warp_done = all_zero();
            } while (!finished && !warp_done);
        }

        if (!warp_done) {
            tests_side_q += (unsigned long long)n_valid * (unsigned long long)s.n_quad;
            tests_side_p += (unsigned long long)n_valid * (unsigned long long)s.n_plane;
            for (int q0i = 0; q0i < s.n_quad; q0i += 32) {
                const int qi = q0i + lane;
                bool keep = false;
                if (qi < s.n_quad && n_valid > 0) {
                    const float4 q0 = s.quad[qi].q0;
                    keep = cs_keep_quadric<true>(ws.cones, n_c, keep_all || n_c == 0, q0, s.surf[__float_as_int(q0.w)].r3);
                }
                unsigned int mask = __ballot_sync(FULL_MASK, keep);
                while (mask) {
                    const int q = q0i + __ffs((int)mask) - 1;
                    mask &= mask - 1u;
                    const float4 q0 = s.quad[q].q0, q1 = s.quad[q].q1, q2 = s.quad[q].q2;
                    const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                    #pragma unroll
                    for (int r = 0; r < ST_RPT; ++r) {
                        if (r >= rpt) break;                // (small levels: slot 0 only)
                        bool lin;
                        const float tt = MODE == 2 ? quadric_t<false>(o[r], d[r], Q, lin) : quadric_t<true>(o[r], d[r], Q, lin);
                        if (slot[r] >= 0 && tt > RT_EPS && tt < dist[r]) occluder(r, __float_as_int(q0.w), __float_as_int(q1.w));
                    }
                }
            }
            for (int q = 0; q < s.n_plane; ++q) {
                const float4 nc = s.plane[q].nc;
                const int m = s.plane[q].material, ord = s.plane[q].order;
                #pragma unroll
                for (int r = 0; r < ST_RPT; ++r) {
                    if (r >= rpt) break;
                    const float tt = MODE == 2 ? plane_t<false>(o[r], d[r], f4_xyz(nc), nc.w) : plane_t<true>(o[r], d[r], f4_xyz(nc), nc.w);
                    if (slot[r] >= 0 && tt > RT_EPS && tt < dist[r]) occluder(r, ord, m);
                }
            }
        }
        #pragma unroll
        for (int r = 0; r < ST_RPT; ++r)
            if (slot[r] >= 0) {
                p.filter[p.sq_dst ? p.sq_dst[slot[r]] : slot[r]] = make_float4(fr[r], fg[r], fb[r], 0.f);
                // three or more factors multiplied in another order than Scene.cpp:1856's: redo this query in order
                const unsigned int st = s_occ[r][tid];
                if ((st & 4u) && (st & 3u) >= 3u && (fr[r] != 0.f || fg[r] != 0.f || fb[r] != 0.f) && p.fix_count) {
                    const unsigned int k = atomicAdd(p.fix_count, 1u);
                    if ((int)k < p.fix_cap) p.fix_list[k] = slot[r];
                    ++fixups;
                }
            }
    }

    for (int off = 16; off > 0; off >>= 1) {
        tests_tri += __shfl_xor_sync(FULL_MASK, tests_tri, off);
        tests_side_q += __shfl_xor_sync(FULL_MASK, tests_side_q, off);
        tests_side_p += __shfl_xor_sync(FULL_MASK, tests_side_p, off);
        fixups += __shfl_xor_sync(FULL_MASK, fixups, off);
    }
    if (lane == 0) {
        // (tests_* were accumulated per warp, not per lane: undo the 32-fold sum of the uniform values)
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (tests_tri) atomicAdd(&st[ST_TEST_TRI_SHADOW], tests_tri >> 5);
        if (exec_rays) atomicAdd(&st[ST_TEST_TRI_EXEC], exec_rays);
        if (tests_side_q) atomicAdd(&st[ST_TEST_QUAD], tests_side_q >> 5);
        if (tests_side_p) atomicAdd(&st[ST_TEST_PLANE], tests_side_p >> 5);
        if (fixups) atomicAdd(&st[ST_SHADOW_FIXUP], fixups);
    }
}

// ---------------------------------------------------------------------------------------
// The lane kernels: the tasks the kernels above listed (no narrow cone fits their rays), one ray / query per GROUP of
// CS_LANE_G lanes (cs_lane_walk).  A listed task of 2^shift rays is split into pieces of 32 / G rays, one per warp
// (dynamic).  Each group walks the sphere hierarchy with its ray, then its lanes share out the quadrics (quadric_t
// rejects a miss without a division) and the planes, and combine: the nearest hit is the best of the lanes' bests
// (ties to the lower insertion order, Scene.cpp:1713); a shadow filter is the product of the lanes' products -- up to two
// factors commute exactly, a query with three or more occluders and a non-zero product is redone in insertion order by
// k_shadow_fixup.  Small kernels on purpose: 64 registers, twice the resident warps of the kernels above.
// ---------------------------------------------------------------------------------------
#define CS_LANE_MIN_CTAS 8
#define CS_LANE_RAYS (32 / CS_LANE_G)          // rays per warp

template <int MODE>
__global__ void __launch_bounds__(CS_THREADS, CS_LANE_MIN_CTAS)
k_lane_nearest(const __grid_constant__ CullTraceParams cp)
{
    constexpr int G = CS_LANE_G;
    const TraceParams& p = cp.t;
    const CullScene& cs = cp.cs;
    const DeviceScene& s = p.s;
    const int lane = threadIdx.x & 31, g = lane & (G - 1);
    const int shift = cp.task_shift, pieces = (1 << shift) / CS_LANE_RAYS;       // pieces per listed task
    const int n_listed = (int)min(*cp.lane_count, (unsigned int)cp.lane_cap);
    unsigned long long exec_rays = 0;
    for (;;) {
        int w = 0;
        if (lane == 0) w = (int)atomicAdd(cp.lane_task_counter, 1u);
        w = __shfl_sync(FULL_MASK, w, 0);
        if (w >= n_listed * pieces) break;
        const int i0 = (cp.lane_list[w / pieces] << shift) + (w % pieces) * CS_LANE_RAYS + lane / G;
        const bool valid = i0 < p.n_rays;
        const int i = (valid && cp.perm) ? cp.perm[i0] : i0;
        RT_CHECK(!valid || (i >= 0 && i < p.n_rays));
        V3 o = mk3(0.f, 0.f, 0.f), d = o;
        if (valid) { o = f4_xyz(p.in_od0[i]); d = f4_xyz(p.in_od1[i]); }
        float best_t = -1.0f;
        int best_id = -1;
        const int tests = cs_lane_nearest<MODE>(cs, o, d, valid, lane, best_t, best_id);
        exec_rays += (unsigned long long)__reduce_add_sync(FULL_MASK, tests);
        if (valid) {
            for (int q = g; q < s.n_quad; q += G) {
                const float4 q0 = __ldg(&s.quad[q].q0), q1 = __ldg(&s.quad[q].q1), q2 = __ldg(&s.quad[q].q2);
                const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                const int ord = __float_as_int(q0.w);
                bool lin;
                const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
            }
            for (int q = g; q < s.n_plane; q += G) {
                const float4 nc = __ldg(&s.plane[q].nc);
                const int ord = s.plane[q].order;
                const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
                if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
            }
        }
        #pragma unroll
        for (int off = 1; off < G; off <<= 1) {             // the group's best (every lane takes part: no divergence here)
            const float ot = __shfl_xor_sync(FULL_MASK, best_t, off);
            const int oi = __shfl_xor_sync(FULL_MASK, best_id, off);
            if (ot > RT_EPS && (ot < best_t || best_t < 0.f || (ot == best_t && oi < best_id))) { best_t = ot; best_id = oi; }
        }
        if (valid && g == 0) p.hit[i] = make_float2(best_t, __int_as_float(best_id));
    }
    if (p.stats && lane == 0 && exec_rays)
        atomicAdd(&p.stats[(size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT + ST_TEST_TRI_EXEC], exec_rays);
}

template <int MODE>
__global__ void __launch_bounds__(CS_THREADS, CS_LANE_MIN_CTAS)
k_lane_shadow(const __grid_constant__ CullShadowParams cp)
{
    constexpr int G = CS_LANE_G;
    const ShadowParams& p = cp.s;
    const CullScene& cs = cp.cs;
    const DeviceScene& s = p.s;
    const int lane = threadIdx.x & 31, g = lane & (G - 1);
    const int shift = cp.task_shift, pieces = (1 << shift) / CS_LANE_RAYS;
    const int n_listed = (int)min(*cp.lane_count, (unsigned int)cp.lane_cap);
    unsigned long long exec_rays = 0;
    unsigned int n_queries = 0, tests_q = 0, tests_p = 0, fixups = 0;              // per lane; summed over the warp at the end
    for (;;) {
        int w = 0;
        if (lane == 0) w = (int)atomicAdd(cp.lane_task_counter, 1u);
        w = __shfl_sync(FULL_MASK, w, 0);
        if (w >= n_listed * pieces) break;
        const int i0 = (cp.lane_list[w / pieces] << shift) + (w % pieces) * CS_LANE_RAYS + lane / G;
        const bool valid = i0 < p.n_queries;
        const int i = (valid && cp.perm) ? cp.perm[i0] : i0;
        RT_CHECK(!valid || (i >= 0 && i < p.n_queries));
        V3 o = mk3(0.f, 0.f, 0.f), d = o;
        float dist = 0.f;
        int lid = 0;
        if (valid) { shadow_query<(MODE != 2)>(p, i, o, d, dist); lid = p.sq_src[i].y & ~SQ_FAR_BIT; }
        float fr = valid ? 1.0f : 0.0f, fg = fr, fb = fr;
        unsigned int occ = 0;                                                     // occluders in this lane's product, saturating at 3
        const int tests = cs_lane_shadow<MODE>(cs, s.mat, f4_xyz(s.light[lid].l0), o, d, dist, valid, lane, fr, fg, fb, occ);
        exec_rays += (unsigned long long)__reduce_add_sync(FULL_MASK, tests);
        // is the group's product zero already?  (then nobody looks at the side lists)
        bool zero = fr == 0.f && fg == 0.f && fb == 0.f;
        #pragma unroll
        for (int off = 1; off < G; off <<= 1) {
            const int other = __shfl_xor_sync(FULL_MASK, (int)zero, off);          // (not inside the ||: every lane must shuffle)
            zero = zero || other != 0;
        }
        if (valid && g == 0) ++n_queries;
        if (valid && !zero) {
            auto occluder = [&](int mat) {
                const float4 m0 = s.mat[mat].m0;
                const float kt = s.mat[mat].m2.x;
                fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
                if (occ < 3u) ++occ;
            };
            if (g == 0) { tests_q += (unsigned int)s.n_quad; tests_p += (unsigned int)s.n_plane; }
            for (int q = g; q < s.n_quad; q += G) {
                const float4 q0 = __ldg(&s.quad[q].q0), q1 = __ldg(&s.quad[q].q1), q2 = __ldg(&s.quad[q].q2);
                const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                bool lin;
                const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                if (tt > RT_EPS && tt < dist) occluder(__float_as_int(q1.w));
            }
            for (int q = g; q < s.n_plane; q += G) {
                const float4 nc = __ldg(&s.plane[q].nc);
                const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
                if (tt > RT_EPS && tt < dist) occluder(s.plane[q].material);
            }
        }
        // the group's product and number of occluders (1 * x = x exactly: the lanes without occluders do not disturb it)
        #pragma unroll
        for (int off = 1; off < G; off <<= 1) {
            fr = __fmul_rn(fr, __shfl_xor_sync(FULL_MASK, fr, off));
            fg = __fmul_rn(fg, __shfl_xor_sync(FULL_MASK, fg, off));
            fb = __fmul_rn(fb, __shfl_xor_sync(FULL_MASK, fb, off));
            occ = min(occ + __shfl_xor_sync(FULL_MASK, occ, off), 3u);
        }
        if (valid && g == 0) {
            p.filter[p.sq_dst ? p.sq_dst[i] : i] = make_float4(fr, fg, fb, 0.f);
            if (occ >= 3u && (fr != 0.f || fg != 0.f || fb != 0.f) && p.fix_count) {   // order of the factors unknown: redone in order
                const unsigned int k = atomicAdd(p.fix_count, 1u);
                if ((int)k < p.fix_cap) p.fix_list[k] = i;
                ++fixups;
            }
        }
    }
    n_queries = __reduce_add_sync(FULL_MASK, n_queries);
    tests_q = __reduce_add_sync(FULL_MASK, tests_q);
    tests_p = __reduce_add_sync(FULL_MASK, tests_p);
    fixups = __reduce_add_sync(FULL_MASK, fixups);
    if (lane == 0 && p.stats) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (n_queries) atomicAdd(&st[ST_TEST_TRI_SHADOW], (unsigned long long)n_queries * (unsigned long long)cs.n_tri);
        if (exec_rays) atomicAdd(&st[ST_TEST_TRI_EXEC], exec_rays);
        if (tests_q) atomicAdd(&st[ST_TEST_QUAD], (unsigned long long)tests_q);
        if (tests_p) atomicAdd(&st[ST_TEST_PLANE], (unsigned long long)tests_p);
        if (fixups) atomicAdd(&st[ST_SHADOW_FIXUP], (unsigned long long)fixups);
    }
}

// ---------------------------------------------------------------------------------------
// The rare query whose product order matters: one warp per query walks the unified list (slot = insertion order,
// Scene.cpp:107) 32 surfaces at a time; the occluders of a round are multiplied in slot order.
// ---------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) k_shadow_fixup(const __grid_constant__ CullShadowParams cp)
{
    // one CTA per listed query: the 8 warps test surfaces 256 at a time, the materials of the occluders of a round are
    // gathered in slot order (a prefix sum over the 8 ballots) and thread 0 multiplies them in that order
    __shared__ int s_mat[256];
    __shared__ int s_cnt[8];
    __shared__ float s_f[3];
    const ShadowParams& p = cp.s;
    const DeviceScene& s = p.s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min((int)*p.fix_count, p.fix_cap);
    const int n_surf = s.rl.n_surf;
    for (int w = blockIdx.x; w < n; w += gridDim.x) {
        const int qi = p.fix_list[w];
        RT_CHECK(qi >= 0 && qi < p.n_queries);
        V3 o, d;
        float dist;
        shadow_query<(MODE != 2)>(p, qi, o, d, dist);
        if (tid == 0) { s_f[0] = 1.0f; s_f[1] = 1.0f; s_f[2] = 1.0f; }
        for (int k0 = 0; k0 < n_surf; k0 += 256) {
            const int k = k0 + tid;
            bool hit = false;
            int mat = 0;
            if (k < n_surf) {
                const float4 r0 = s.surf[k].r0, r1 = s.surf[k].r1, r2 = s.surf[k].r2;
                const int type = __float_as_int(r0.w);
                mat = __float_as_int(r1.w);
                float tt;
                if (type == SURF_TRIANGLE) tt = MODE == 2 ? triangle_t<false>(o, d, f4_xyz(r0), f4_xyz(r1), f4_xyz(r2)) : triangle_t<true>(o, d, f4_xyz(r0), f4_xyz(r1), f4_xyz(r2));
                else if (type == SURF_PLANE) tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(r0), r1.x) : plane_t<true>(o, d, f4_xyz(r0), r1.x);
                else {
                    const Quadric Q = { r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z, r2.w };
                    bool lin;
                    tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                }
                hit = tt > RT_EPS && tt < dist;
            }
            const unsigned int mask = __ballot_sync(FULL_MASK, hit);
            if (lane == 0) s_cnt[warp] = __popc(mask);
            __syncthreads();
            int base = 0, total = 0;
            #pragma unroll
            for (int q = 0; q < 8; ++q) { if (q < warp) base += s_cnt[q]; total += s_cnt[q]; }
            if (hit) s_mat[base + __popc(mask & ((1u << lane) - 1u))] = mat;
            __syncthreads();
            if (tid == 0 && total > 0) {
                float fr = s_f[0], fg = s_f[1], fb = s_f[2];
                for (int e = 0; e < total; ++e) {
                    const float4 m0 = s.mat[s_mat[e]].m0;
                    const float kt = s.mat[s_mat[e]].m2.x;
                    fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
                }
                s_f[0] = fr; s_f[1] = fg; s_f[2] = fb;
            }
            __syncthreads();
        }
        if (tid == 0) p.filter[p.sq_dst ? p.sq_dst[qi] : qi] = make_float4(s_f[0], s_f[1], s_f[2], 0.f);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------
static int cs_grid(int n_tasks, int sm_count)
{
    const int ctas = (n_tasks + CS_WARPS - 1) / CS_WARPS;
    return ctas < CS_MIN_CTAS * sm_count ? ctas : CS_MIN_CTAS * sm_count;
}

static int lane_grid(int n_tasks, int rpt, int sm_count)
{
    const long long warps = (long long)n_tasks * rpt;
    const long long ctas = (warps + CS_WARPS - 1) / CS_WARPS;
    return (int)std::min<long long>(ctas, (long long)CS_LANE_MIN_CTAS * sm_count);
}

cudaError_t launch_cull_nearest(const CullTraceParams& p_in, int sm_count, cudaStream_t st)
{
    CullTraceParams p = p_in;
    if (p.t.primary) p.lane_list = nullptr;              // (primary rays always have the camera's cone; a degenerate camera scans)
    const int tiles_x = (p.t.f.LW + 15) / 16;
    const int per = 1 << p.task_shift;
    const int n_tasks = p.t.primary ? tiles_x * ((p.t.chunk_rows + 7) / 8) : (p.t.n_rays + per - 1) / per;
    if (n_tasks <= 0) return cudaSuccess;
    const int blocks = cs_grid(n_tasks, sm_count);
    if (p.t.primary) {
        if (p.t.mode == 2) k_cull_nearest<2, true><<<blocks, CS_THREADS, 0, st>>>(p);
        else               k_cull_nearest<0, true><<<blocks, CS_THREADS, 0, st>>>(p);
    } else {
        if (p.t.mode == 2) k_cull_nearest<2, false><<<blocks, CS_THREADS, 0, st>>>(p);
        else               k_cull_nearest<0, false><<<blocks, CS_THREADS, 0, st>>>(p);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || !p.lane_list) return e;
    // the tasks no cone fitted (the count is read on the device; usually none, or the rays of the deep levels)
    const int lb = lane_grid(n_tasks, per / CS_LANE_RAYS, sm_count);
    if (p.t.mode == 2) k_lane_nearest<2><<<lb, CS_THREADS, 0, st>>>(p);
    else               k_lane_nearest<0><<<lb, CS_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_cull_shadow(const CullShadowParams& p, int sm_count, cudaStream_t st)
{
    const int per = 1 << p.task_shift;
    const int n_tasks = (p.s.n_queries + per - 1) / per;
    if (n_tasks <= 0) return cudaSuccess;
    const int blocks = cs_grid(n_tasks, sm_count);
    if (p.s.mode == 2) k_cull_shadow<2><<<blocks, CS_THREADS, 0, st>>>(p);
    else               k_cull_shadow<0><<<blocks, CS_THREADS, 0, st>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || !p.lane_list) return e;
    const int lb = lane_grid(n_tasks, per / CS_LANE_RAYS, sm_count);
    if (p.s.mode == 2) k_lane_shadow<2><<<lb, CS_THREADS, 0, st>>>(p);
    else               k_lane_shadow<0><<<lb, CS_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_shadow_fixup(const CullShadowParams& p, int sm_count, cudaStream_t st)
{
    if (p.s.mode == 2) k_shadow_fixup<2><<<sm_count * 4, 256, 0, st>>>(p);
    else               k_shadow_fixup<0><<<sm_count * 4, 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t cull_sort_pairs(DevBuf* tmp, const unsigned int* keys_in, unsigned int* keys_out, const int* idx_in, int* idx_out,
                            int n, int bits, cudaStream_t st)
{
    size_t bytes = 0;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, idx_in, idx_out, n, 0, bits, st);
    if (e != cudaSuccess) return e;
    if ((e = tmp->ensure(bytes)) != cudaSuccess) return e;
    return cub::DeviceRadixSort::SortPairs(tmp->p, bytes, keys_in, keys_out, idx_in, idx_out, n, 0, bits, st);
}

// ---------------------------------------------------------------------------------------
// cull_prepare: Morton sort + sphere hierarchy, on the device, once per scene (cached until rtb200_set_scene)
// ---------------------------------------------------------------------------------------
RT_DEV int cs_float_order(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
static inline float cs_order_float(int i) { const int b = i >= 0 ? i : i ^ 0x7fffffff; float f; memcpy(&f, &b, 4); return f; }

__global__ void k_cs_bounds(const TriRec* __restrict__ tri, int n, int* __restrict__ bounds)
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = tri[i].a, b = tri[i].b, c = tri[i].c;
        const float ctr[3] = {a.x + (b.x + c.x) * (1.0f / 3.0f), a.y + (b.y + c.y) * (1.0f / 3.0f), a.z + (b.z + c.z) * (1.0f / 3.0f)};
        #pragma unroll
        for (int k = 0; k < 3; ++k) if (fabsf(ctr[k]) <= 3.0e38f) { lo[k] = fminf(lo[k], ctr[k]); hi[k] = fmaxf(hi[k], ctr[k]); }
    }
    #pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int off = 16; off > 0; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(FULL_MASK, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(FULL_MASK, hi[k], off));
        }
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(&bounds[k], cs_float_order(lo[k])); atomicMax(&bounds[3 + k], cs_float_order(hi[k])); }
    }
}

// 30-bit Morton code of the centroid in CUBIC cells (one scale for the three axes: a flat scene is not cut into sheets)
__global__ void k_cs_keys(const TriRec* __restrict__ tri, int n, float lx, float ly, float lz, float inv, unsigned int* __restrict__ keys, int* __restrict__ idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = tri[i].a, b = tri[i].b, c = tri[i].c;
    const float cx = a.x + (b.x + c.x) * (1.0f / 3.0f), cy = a.y + (b.y + c.y) * (1.0f / 3.0f), cz = a.z + (b.z + c.z) * (1.0f / 3.0f);
    keys[i] = morton30(cx, cy, cz, lx, ly, lz, inv);
    idx[i] = i;
}

// sorted triangle array + level 0 (one thread per slot of the padded array)
__global__ void k_cs_gather(const TriRec* __restrict__ tri, const int* __restrict__ idx, int n, int n0, TriRec* __restrict__ out,
                            float4* __restrict__ sph0, float* __restrict__ w0)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n0) return;
    if (i >= n) {
        TriRec z;
        z.a = make_float4(0.f, 0.f, 0.f, __int_as_float(0x7fffffff)); z.b = make_float4(0.f, 0.f, 0.f, 0.f); z.c = z.b;
        out[i] = z; sph0[i] = make_float4(0.f, 0.f, 0.f, 0.f); w0[i] = 0.f;
        return;
    }
    const TriRec r = tri[idx[i]];
    out[i] = r;
    V3 c; float rad, w;
    triangle_sphere(f4_xyz(r.a), f4_xyz(r.b), f4_xyz(r.c), c, rad, w);
    const bool ok = fabsf(c.x) <= 3.0e38f && fabsf(c.y) <= 3.0e38f && fabsf(c.z) <= 3.0e38f && rad <= 3.0e38f && w <= 3.0e38f;
    // a triangle with non-finite numbers keeps a sphere nothing can cull; w must stay > 0 for every real triangle
    sph0[i] = ok ? make_float4(c.x, c.y, c.z, rad) : make_float4(0.f, 0.f, 0.f, 3.0e38f);
    w0[i] = ok ? fmaxf(w, 1.0e-30f) : 3.0e38f;
}

// level l from level l - 1: sphere around the mean centre of the non-empty children that contains their spheres
__global__ void k_cs_group(const float4* __restrict__ sph_in, const float* __restrict__ w_in, int n_out, float4* __restrict__ sph_out, float* __restrict__ w_out)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_out) return;
    float cx = 0.f, cy = 0.f, cz = 0.f, wmax = 0.f;
    int cnt = 0;
    bool huge = false;
    for (int j = 0; j < 32; ++j) {
        const float w = w_in[g * 32 + j];
        if (!(w > 0.f)) continue;
        const float4 s = sph_in[g * 32 + j];
        if (s.w > 1.0e37f) { huge = true; }
        cx += s.x; cy += s.y; cz += s.z; ++cnt; wmax = fmaxf(wmax, w);
    }
    if (cnt == 0) { sph_out[g] = make_float4(0.f, 0.f, 0.f, 0.f); w_out[g] = 0.f; return; }
    const float inv = 1.0f / (float)cnt;
    cx *= inv; cy *= inv; cz *= inv;
    float R = 0.f;
    for (int j = 0; j < 32; ++j) {
        const float w = w_in[g * 32 + j];
        if (!(w > 0.f)) continue;
        const float4 s = sph_in[g * 32 + j];
        const float dx = s.x - cx, dy = s.y - cy, dz = s.z - cz;
        R = fmaxf(R, sqrtf(dx * dx + dy * dy + dz * dz) * 1.00001f + s.w);
    }
    const float m1 = fabsf(cx) + fabsf(cy) + fabsf(cz);
    R = R * 1.00001f + 4.0e-7f * m1;
    if (huge || !(R <= 3.0e38f) || !(m1 <= 3.0e38f)) { cx = cy = cz = 0.f; R = 3.0e38f; }
    sph_out[g] = make_float4(cx, cy, cz, R);
    w_out[g] = wmax;
}

struct CullState {
    DevBuf tri, sph, w, keys, keys2, idx, idx2, tmp, bounds;
    PinnedBuf h_bounds;
    unsigned long long serial = ~0ull;       // scene the hierarchy was built for
    CullScene cs{};
    double build_ms = 0.0;
    float lo[3] = {0.f, 0.f, 0.f}, ext = 0.f;  // centroid bounds of the triangles (origin and size of the Morton grid)
};

static CullState* cull_state(rtb200_ctx* c)
{
    if (!c->cull_state) c->cull_state = new CullState();
    return static_cast<CullState*>(c->cull_state);
}

void cull_release(rtb200_ctx* c)
{
    if (!c->cull_state) return;
    CullState* b = static_cast<CullState*>(c->cull_state);
    b->tri.release(); b->sph.release(); b->w.release(); b->keys.release(); b->keys2.release();
    b->idx.release(); b->idx2.release(); b->tmp.release(); b->bounds.release(); b->h_bounds.release();
    delete b;
    c->cull_state = nullptr;
}

#define KCU(call)                                                       \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return set_cuda_error(c, e__, #call);   \
    } while (0)

int cull_prepare(rtb200_ctx* c, CullScene* out)
{
    CullState* b = cull_state(c);
    if (b->serial == c->scene_serial) { *out = b->cs; return RTB200_OK; }
    const int T = c->ds.n_tri;
    const int n3 = std::max(1, (T + 32767) / 32768);
    const int n[CS_LEVELS] = {32768 * n3, 1024 * n3, 32 * n3, n3};
    const size_t total = (size_t)n3 * (1057 + 32768);
    cudaStream_t st = c->stream;
    cudaEvent_t e0 = c->ev[6], e1 = c->ev[7];
    KCU(b->tri.ensure(sizeof(TriRec) * (size_t)n[0]));
    KCU(b->sph.ensure(sizeof(float4) * total));
    KCU(b->w.ensure(sizeof(float) * total));
    KCU(b->bounds.ensure(64));
    KCU(b->h_bounds.ensure(64));
    KCU(cudaEventRecord(e0, st));
    int valid[CS_LEVELS];
    valid[0] = T;
    for (int l = 1; l < CS_LEVELS; ++l) valid[l] = (valid[l - 1] + 31) / 32;
    float4* sph = b->sph.as<float4>();
    float* w = b->w.as<float>();
    const int off[CS_LEVELS] = {1057 * n3, 33 * n3, n3, 0};
    KCU(cudaMemsetAsync(w, 0, sizeof(float) * total, st));
    KCU(cudaMemsetAsync(sph, 0, sizeof(float4) * (size_t)off[0], st));
    if (T > 0) {
        KCU(b->keys.ensure(sizeof(unsigned int) * (size_t)T)); KCU(b->keys2.ensure(sizeof(unsigned int) * (size_t)T));
        KCU(b->idx.ensure(sizeof(int) * (size_t)T)); KCU(b->idx2.ensure(sizeof(int) * (size_t)T));
        int* h = b->h_bounds.as<int>();
        for (int k = 0; k < 3; ++k) { h[k] = INT_MAX; h[3 + k] = INT_MIN; }
        KCU(cudaMemcpyAsync(b->bounds.p, h, 24, cudaMemcpyHostToDevice, st));
        k_cs_bounds<<<std::min((T + 255) / 256, 1184), 256, 0, st>>>(c->ds.tri, T, b->bounds.as<int>());
        KCU(cudaGetLastError());
        KCU(cudaMemcpyAsync(h, b->bounds.p, 24, cudaMemcpyDeviceToHost, st));
        KCU(cudaStreamSynchronize(st));
        float lo[3], ext = 0.f;
        b->ext = 0.f;
        for (int k = 0; k < 3; ++k) {
            lo[k] = cs_order_float(h[k]);
            const float hi = cs_order_float(h[3 + k]);
            if (!(lo[k] <= hi)) { lo[k] = 0.f; continue; }          // no finite centroid on this axis
            ext = std::max(ext, hi - lo[k]);
        }
        const float inv = ext > 0.f ? 1023.999f / ext : 0.f;
        b->ext = ext;
        for (int k = 0; k < 3; ++k) b->lo[k] = lo[k];
        k_cs_keys<<<(T + 255) / 256, 256, 0, st>>>(c->ds.tri, T, lo[0], lo[1], lo[2], inv, b->keys.as<unsigned int>(), b->idx.as<int>());
        KCU(cudaGetLastError());
        size_t tmp_bytes = 0;
        KCU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, b->keys.as<unsigned int>(), b->keys2.as<unsigned int>(), b->idx.as<int>(), b->idx2.as<int>(), T, 0, 30, st));
        KCU(b->tmp.ensure(tmp_bytes));
        KCU(cub::DeviceRadixSort::SortPairs(b->tmp.p, tmp_bytes, b->keys.as<unsigned int>(), b->keys2.as<unsigned int>(), b->idx.as<int>(), b->idx2.as<int>(), T, 0, 30, st));
    }
    k_cs_gather<<<(n[0] + 255) / 256, 256, 0, st>>>(c->ds.tri, b->idx2.as<int>(), T, n[0], b->tri.as<TriRec>(), sph + off[0], w + off[0]);
    KCU(cudaGetLastError());
    for (int l = 1; l < CS_LEVELS; ++l) {
        if (valid[l] <= 0) break;
        k_cs_group<<<(valid[l] + 127) / 128, 128, 0, st>>>(sph + off[l - 1], w + off[l - 1], valid[l], sph + off[l], w + off[l]);
        KCU(cudaGetLastError());
    }
    KCU(cudaEventRecord(e1, st));
    KCU(cudaStreamSynchronize(st));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) b->build_ms = ms;
    CullScene& cs = b->cs;
    cs.tri = b->tri.as<TriRec>(); cs.sph = sph; cs.w = w;
    cs.n3 = n3; cs.n_tri = T;
    // sort keys of secondary rays / queries: the cells of the triangles' Morton grid, the box widened 4 x around its
    // centre (reflections reach the ground plane far outside the mesh; what lies farther is clamped to the border cells)
    for (int k = 0; k < 3; ++k) cs.key_lo[k] = 0.f;
    cs.key_inv = 0.f;
    if (T > 0 && b->ext > 0.f) {
        for (int k = 0; k < 3; ++k) cs.key_lo[k] = b->lo[k] - 1.5f * b->ext;
        cs.key_inv = 1023.999f / (4.0f * b->ext);
    }
    cs.top_blocks = (n3 + 31) / 32;
    b->serial = c->scene_serial;
    *out = cs;
    return RTB200_OK;
}

} // namespace rtb
