// rt_resident.cu -- scene-resident fused wavefront kernels (small scenes: the whole scene
// sits in shared memory, staged with ONE bulk TMA copy per CTA).
//
// One launch per recursion level.  A level-k thread owns one ray of bounce count k and does
// what CScene::ObtenirCouleur + ObtenirCouleurSurIntersection + ObtenirFiltreDeSurface do for
// it (Scene.cpp:1705-1861), except that the two recursive calls of Scene.cpp:1779-1823 become
// two rays appended (warp-aggregated) to the level-(k+1) queue plus one "link" record, so the
// colours are folded parent <- children bottom-up, reflection before refraction, exactly in
// the reference's accumulation order and deterministically.  No per-thread stack, no
// recursion; every level runs on a compacted queue so warps never diverge on depth.
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <utility>

#include "rt_kernels.cuh"
#include "rt_cull.cuh"
#include "rt_tma.cuh"

namespace rtb {

#ifndef RES_THREADS
#define RES_THREADS 256            // 256: 16x16 lattice tile per CTA (8 warps of 8x4), 128: 16x8
#endif
#define RES_TILE_H (RES_THREADS / 16)
#ifndef RES_MIN_CTAS
#define RES_MIN_CTAS 3            // 4 (<= 64 registers) was measured slower: spills, 86.0 vs 83.4 us on config 2
#endif

struct SceneView {                 // resident blob seen through shared (or global) memory
    const SurfRec*  surf;
    const MatRec*   mat;
    const LightRec* light;
    int n_surf, n_light;
    bool cull;                     // warp-vote culling allowed (off with RTB200_F_EXACT_FP)
    // primary rays all start at the camera: the origin-dependent half of quadric_miss, once per CTA and surface
    const float4* pre_qo;          // q * (c - cam) | e = (c - cam).q.(c - cam)      (null: not available)
    const float*  pre_lim;         // rhs * 1.05 + 2e-5 * M(cam)
};

// Conservative "this ray cannot hit the quadric" test for quadrics rtb200_set_scene could write in
// centred form  sum_i q_i (x_i - c_i)^2 = rhs  with q_i >= 0, rhs > 0 (spheres, ellipsoids, axis-aligned
// cylinders): r3 = (c, rhs).  In that metric the ray passes the centre at squared distance
// dist2 = e - b^2 / a  (a = q.d.d, b = q.oc.d, e = q.oc.oc) and misses iff dist2 > rhs.  The reference
// decides on the sign of delta = ka^2 - kb (Quadrique.cpp:199) evaluated in UN-centred coordinates,
// whose rounding error in these units is below ~12 * 2^-24 * M, M = sum q_i (o_i^2 + c_i^2); the warp
// only culls beyond rhs * 1.05 + 2e-5 * M (> 25 x that bound), or when the inflated quadric lies
// entirely behind the origin.  a > 1e-20 keeps clear of the reference's A == 0 branch
// (Quadrique.cpp:241), which is an unconditional "hit".
RT_DEV bool quadric_miss(V3 o, V3 d, float4 r0, float K, float4 r3)
{
    const V3 q = mk3(r0.x, r0.y, r0.z);
    const V3 oc = mk3(r3.x - o.x, r3.y - o.y, r3.z - o.z);
    const V3 qd = mk3(q.x * d.x, q.y * d.y, q.z * d.z);
    const V3 qo = mk3(q.x * oc.x, q.y * oc.y, q.z * oc.z);
    const float a = dot3<false>(qd, d);
    const float b = dot3<false>(qo, d);
    const float e = dot3<false>(qo, oc);
    const float M = fmaf(q.x * o.x, o.x, fmaf(q.y * o.y, o.y, fmaf(q.z * o.z, o.z, r3.w + K)));   // sum q c^2 = rhs + K
    const float lim = fmaf(r3.w, 1.05f, 2.0e-5f * M);
    return a > 1.0e-20f && (fmaf(a, e, -b * b) > a * lim || (b < 0.f && e > lim));
}

// the same test with the origin-dependent terms taken from the per-CTA table (primary rays)
RT_DEV void quadric_miss_origin_terms(V3 o, float4 r0, float K, float4 r3, float4& qo_e, float& lim)
{
    const V3 q = mk3(r0.x, r0.y, r0.z);
    const V3 oc = mk3(r3.x - o.x, r3.y - o.y, r3.z - o.z);
    const V3 qo = mk3(q.x * oc.x, q.y * oc.y, q.z * oc.z);
    const float M = fmaf(q.x * o.x, o.x, fmaf(q.y * o.y, o.y, fmaf(q.z * o.z, o.z, r3.w + K)));
    qo_e = make_float4(qo.x, qo.y, qo.z, dot3<false>(qo, oc));
    lim = fmaf(r3.w, 1.05f, 2.0e-5f * M);
}
RT_DEV bool quadric_miss_pre(V3 d, float4 r0, float4 qo_e, float lim)
{
    const V3 qd = mk3(r0.x * d.x, r0.y * d.y, r0.z * d.z);
    const float a = dot3<false>(qd, d);
    const float b = fmaf(qo_e.z, d.z, fmaf(qo_e.y, d.y, qo_e.x * d.x));
    return a > 1.0e-20f && (fmaf(a, qo_e.w, -b * b) > a * lim || (b < 0.f && qo_e.w > lim));
}

// distance of surface record `k` along (o,d); quadrics also report the A==0 branch
//
// Culling with warp votes (k, hence the type, is uniform over the warp): a centred-form quadric
// no ACTIVE lane's ray can touch (quadric_miss), or a triangle no active lane passes the fused
// det/u filter of rt_cull.cuh for, is skipped by the whole warp -- the reference sequence below
// only runs where its result can be a hit, so every value that is kept is still the reference's.
template <bool E> RT_DEV float surface_t(const SurfRec* __restrict__ s, int k, V3 o, V3 d, bool active, bool cull,
                                         int& type, bool& linear, const float4* pre_qo = nullptr, const float* pre_lim = nullptr)
{
    const float4 r0 = s[k].r0, r1 = s[k].r1, r2 = s[k].r2;
    type = __float_as_int(r0.w);
    linear = false;
    if (type == SURF_TRIANGLE) {
        if (cull && !__any_sync(0xffffffffu, active && triangle_candidate(o, d, f4_xyz(r0), f4_xyz(r1), f4_xyz(r2)))) return -1.0f;
        return triangle_t<E>(o, d, f4_xyz(r0), f4_xyz(r1), f4_xyz(r2));
    }
    if (type == SURF_QUADRIC) {
        if (cull) {
            const float4 r3 = s[k].r3;                 // centred form (c, rhs); rhs <= 0: not cullable
            if (r3.w > 0.f) {
                const bool may = pre_qo ? !quadric_miss_pre(d, r0, pre_qo[k], pre_lim[k]) : !quadric_miss(o, d, r0, r2.w, r3);
                if (!__any_sync(0xffffffffu, active && may)) return -1.0f;
            }
        }
        Quadric q = { r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z, r2.w };
        return quadric_t<E>(o, d, q, linear);
    }
    return plane_t<E>(o, d, f4_xyz(r0), r1.x);
}

// nearest hit: the loop of Scene.cpp:1710-1715 (strict <, so the first surface wins a tie)
template <bool E> RT_DEV void nearest_hit(const SceneView& sv, bool active, V3 o, V3 d, float& best_t, int& best_k, bool& best_lin,
                                          bool from_camera = false)
{
    best_t = -1.0f; best_k = -1; best_lin = false;
    const float4* pre_qo = from_camera ? sv.pre_qo : nullptr;
    for (int k = 0; k < sv.n_surf; ++k) {
        int type; bool lin;
        float t = surface_t<E>(sv.surf, k, o, d, active, sv.cull, type, lin, pre_qo, sv.pre_lim);
        if (t > RT_EPS && (t < best_t || best_t < 0.f)) { best_t = t; best_k = k; best_lin = lin; }
    }
}

// normal the reference would have stored in the winning CIntersection
template <bool E> RT_DEV V3 hit_normal(const SceneView& sv, int k, V3 o, V3 d, float t, bool lin)
{
    const float4 r0 = sv.surf[k].r0;
    const int type = __float_as_int(r0.w);
    if (type == SURF_TRIANGLE) return f4_xyz(sv.surf[k].r3);
    if (type == SURF_PLANE) return f4_xyz(r0);
    const float4 r1 = sv.surf[k].r1, r2 = sv.surf[k].r2;
    Quadric q = { r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z, r2.w };
    return quadric_normal<E>(o, d, t, q, lin);
}

// shadow filter: Scene.cpp:1842-1861.  `need` lanes accumulate; the warp leaves the scan as
// soon as every lane that needs it holds a zero filter (0 * finite stays 0).
template <bool E> RT_DEV int shadow_filter(const SceneView& sv, bool need, V3 P, V3 Ldir, float dist,
                                           float& fr, float& fg, float& fb)
{
    fr = fg = fb = 1.0f;
    int k = 0;
    for (; k < sv.n_surf; ++k) {
        int type; bool lin;
        float t = surface_t<E>(sv.surf, k, P, Ldir, need, sv.cull, type, lin);
        const bool occluded = need && t > RT_EPS && t < dist;
        if (!__any_sync(0xffffffffu, occluded)) continue;     // no filter changed: the exit test below cannot change either
        if (occluded) {
            const int m = __float_as_int(sv.surf[k].r1.w);
            const float4 m0 = sv.mat[m].m0;
            const float kt = sv.mat[m].m2.x;
            fr = FP<E>::mul(fr, FP<E>::mul(m0.x, kt));
            fg = FP<E>::mul(fg, FP<E>::mul(m0.y, kt));
            fb = FP<E>::mul(fb, FP<E>::mul(m0.z, kt));
        }
        const bool done = !need || (fr == 0.f && fg == 0.f && fb == 0.f);
        if (__all_sync(0xffffffffu, done)) { ++k; break; }
    }
    return k;                                   // surfaces tested by the lanes that needed the query
}

// warp-wide add of the per-thread counters into the spread statistics slots (once per CTA:
// the CTAs are persistent, the counters live in registers across tiles)
RT_DEV void warp_stats(unsigned long long* stats, unsigned int n_shadow, unsigned int n_refl, unsigned int n_refr,
                       unsigned int n_tests, unsigned int n_nan)
{
    n_shadow = __reduce_add_sync(0xffffffffu, n_shadow);
    n_refl   = __reduce_add_sync(0xffffffffu, n_refl);
    n_refr   = __reduce_add_sync(0xffffffffu, n_refr);
    n_tests  = __reduce_add_sync(0xffffffffu, n_tests);
    n_nan    = __reduce_add_sync(0xffffffffu, n_nan);
    if ((threadIdx.x & 31) == 0) {
        unsigned long long* st = stats + (size_t)((blockIdx.x * 8 + (threadIdx.x >> 5)) % RT_STAT_SLOTS) * ST_COUNT;
        // resident path: ST_TEST_TRI holds ALL surface tests (split by type on the host)
        if (n_shadow) atomicAdd(&st[ST_SHADOW], (unsigned long long)n_shadow);
        if (n_refl)   atomicAdd(&st[ST_REFLECT], (unsigned long long)n_refl);
        if (n_refr)   atomicAdd(&st[ST_REFRACT], (unsigned long long)n_refr);
        if (n_tests)  atomicAdd(&st[ST_TEST_TRI], (unsigned long long)n_tests);
        if (n_nan)    atomicAdd(&st[ST_NAN], (unsigned long long)n_nan);
    }
}

// the RGBA8 quantiser floor(clamp(c,0,1)*255 + 0.5), NaN -> 0 (counted)
RT_DEV uchar4 quantise_rgba8(float r, float g, float b, unsigned int& nans)
{
    float v[3] = { r, g, b };
    unsigned char q[3];
    #pragma unroll
    for (int c = 0; c < 3; ++c) {
        float w = v[c];
        if (w != w) { ++nans; w = 0.f; }
        w = fminf(fmaxf(w, 0.f), 1.f);
        q[c] = (unsigned char)__float2uint_rd(__fadd_rn(__fmul_rn(w, 255.0f), 0.5f));   // floor of a value in [0.5, 255.5]
    }
    return make_uchar4(q[0], q[1], q[2], 255);
}

// ---------------------------------------------------------------------------------------
// One wavefront level.  PERSISTENT CTAs (a few per SM): the scene is staged into shared memory
// ONCE per CTA by one bulk TMA copy, then the CTA walks 16x16 tiles of the sample lattice
// (level 0, warps are 8x4 sub-tiles so a warp's rays stay a compact bundle) or blocks of 256
// consecutive queue entries (level >= 1).  When the frame spawns no secondary rays and is not
// supersampled (p.direct), the level-0 kernel also resolves: it writes the float RGB and the
// RGBA8 framebuffers itself and the frame is ONE kernel.
// ---------------------------------------------------------------------------------------
// programmatic dependent launch (sm_90+): see LevelParams::pdl
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <bool E, bool PRIMARY>
__global__ void __launch_bounds__(RES_THREADS, RES_MIN_CTAS)
k_resident_level(const __grid_constant__ LevelParams p)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (p.pdl) pdl_launch_dependents();     // the next kernel of the chain may start its prologue now
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (tid == 0) {                         // one bulk copy stages the whole scene (never written by a kernel)
        mbar_arrive_expect_tx(&bar, (uint32_t)p.rl.bytes);
        tma_load_1d(smem, p.blob, (uint32_t)p.rl.bytes, &bar);
    }
    if (p.pdl) pdl_wait();                  // everything below may read what the previous kernel wrote
    const bool tail_mode = !PRIMARY && p.tail != nullptr;
    int in_count = p.in_count;
    if (!PRIMARY && p.in_count_dev && !tail_mode) {   // written by the level before this one, no host in between
        const unsigned int n = *p.in_count_dev;
        in_count = (int)(n < (unsigned int)p.in_cap ? n : (unsigned int)p.in_cap);
        if ((int)blockIdx.x * RES_THREADS >= in_count) {                   // whole CTA: nothing to do
            if (tid == 0) mbar_wait(&bar, 0);                              // (not before the bulk copy has landed)
            return;
        }
    }
    const FrameParams& f = p.f;
    const int tiles_x = (f.LW + 15) >> 4;
    const int n_blocks = PRIMARY ? tiles_x * ((p.chunk_rows + RES_TILE_H - 1) / RES_TILE_H) : (in_count + RES_THREADS - 1) / RES_THREADS;
    if (p.peer_ack) peer_frame_wait_ack(p.peer_ack, p.peer_ack_min, p.stats ? p.stats + ST_PEER_TIMEOUT : nullptr);      // several GPUs: the frame half is free
    mbar_wait(&bar, 0);
    SceneView sv;
    sv.surf  = reinterpret_cast<const SurfRec*>(smem);
    sv.mat   = reinterpret_cast<const MatRec*>(smem + p.rl.off_mat);
    sv.light = reinterpret_cast<const LightRec*>(smem + p.rl.off_light);
    sv.n_surf = p.rl.n_surf; sv.n_light = p.rl.n_light; sv.cull = p.cull != 0;
    sv.pre_qo = nullptr; sv.pre_lim = nullptr;
    if (PRIMARY && p.pre_table && sv.cull) {
        // every primary ray starts at the camera: the origin-dependent terms of the quadric vote, once per CTA
        float4* tq = reinterpret_cast<float4*>(smem + ((p.rl.bytes + 15) & ~15));
        float*  tl = reinterpret_cast<float*>(tq + sv.n_surf);
        const V3 cam = mk3(f.cam_pos[0], f.cam_pos[1], f.cam_pos[2]);
        for (int k = tid; k < sv.n_surf; k += RES_THREADS) {
            const float4 r0 = sv.surf[k].r0, r3 = sv.surf[k].r3;
            float4 qe = make_float4(0.f, 0.f, 0.f, 0.f); float lim = 0.f;
            if (__float_as_int(r0.w) == SURF_QUADRIC && r3.w > 0.f) quadric_miss_origin_terms(cam, r0, sv.surf[k].r2.w, r3, qe, lim);
            tq[k] = qe; tl[k] = lim;
        }
        __syncthreads();
        sv.pre_qo = tq; sv.pre_lim = tl;
    }
    unsigned int n_shadow = 0, n_tests = 0, n_refl_st = 0, n_refr_st = 0, n_nan = 0;

    // the level this launch works on -- in tail mode one after the other, from the table
    const float4* in_od0 = p.in_od0;
    const float4* in_od1 = p.in_od1;
    float4* in_colour = p.in_colour;
    float4* out_od0 = p.out_od0;
    float4* out_od1 = p.out_od1;
    int4*   links = p.links;
    unsigned int* out_count = p.out_count;
    unsigned int* link_count = p.link_count;
    int  level = p.level, out_guarded = p.out_guarded;
    unsigned int out_cap = p.out_cap;
    bool may_spawn = f.secondary != 0;
    const int n_levels = tail_mode ? p.tail_count : 1;
    for (int tl = 0; tl < n_levels; ++tl) {
    int n_blocks_l = n_blocks;
    if (tail_mode) {
        const TailLevel tv = p.tail[tl];
        in_od0 = tv.in_od0; in_od1 = tv.in_od1; in_colour = tv.in_colour;
        out_od0 = tv.out_od0; out_od1 = tv.out_od1; links = tv.links;
        level = tv.level; out_guarded = 1; out_cap = tv.out_cap; may_spawn = tv.spawn != 0;
        out_count = p.out_count + 2 * tl; link_count = p.link_count + 2 * tl;       // counters of consecutive levels
        const unsigned int n = __ldcg(p.in_count_dev + 2 * tl);                     // written by the level before
        in_count = (int)(n < (unsigned int)tv.in_cap ? n : (unsigned int)tv.in_cap);
        if (in_count == 0) break;                                                   // (the whole CTA reads the same value)
        n_blocks_l = (in_count + RES_THREADS - 1) / RES_THREADS;
    }

    // tile walk without a division per tile: (tx, ty) advance by the grid stride
    const int step_x = (int)(gridDim.x % (unsigned)tiles_x), step_y = (int)(gridDim.x / (unsigned)tiles_x);
    int tile_x = PRIMARY ? (int)(blockIdx.x % (unsigned)tiles_x) : 0, tile_y = PRIMARY ? (int)(blockIdx.x / (unsigned)tiles_x) : 0;
    for (int blk = blockIdx.x; blk < n_blocks_l; blk += gridDim.x) {
    // ---- this thread's ray ------------------------------------------------------------------
    V3 o, d;
    float energy = 1.0f, ior = 1.0f;
    bool valid;
    int  self;                              // index of this ray in its level's colour buffer
    int  pix_x = 0, pix_y = 0;
    if (PRIMARY) {
        const int tx = tile_x, ty = tile_y;
        tile_x += step_x; tile_y += step_y;
        if (tile_x >= tiles_x) { tile_x -= tiles_x; ++tile_y; }
        const int lx = (tx << 4) + ((warp & 1) << 3) + (lane & 7);
        const int cy = ty * RES_TILE_H + ((warp >> 1) << 2) + (lane >> 3);   // row inside the chunk
        const int ly = p.chunk_ly0 + cy;                                 // local lattice row
        valid = lx < f.LW && cy < p.chunk_rows && ly < p.local_LH;
        self = cy * f.LW + lx;
        pix_x = lx; pix_y = ly;
        primary_ray<E>(f, valid ? lx : 0, valid ? global_lattice_row(f, ly) : 0, o, d);
    } else {
        self = blk * RES_THREADS + tid;
        valid = self < in_count;
        const float4 a = valid ? in_od0[self] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 b = valid ? in_od1[self] : make_float4(0.f, 0.f, 0.f, 0.f);
        o = f4_xyz(a); d = f4_xyz(b); energy = a.w; ior = b.w;
    }

    // ---- nearest hit (Scene.cpp:1705-1720) -----------------------------------------------
    float t; int k; bool lin;
    nearest_hit<E>(sv, valid, o, d, t, k, lin, PRIMARY);
    const bool hit = valid && k >= 0;

    float cr = f.background[0], cg = f.background[1], cb = f.background[2];
    n_tests += valid ? (unsigned int)sv.n_surf : 0u;
    bool want_refl = false, want_refr = false;
    V3 P = mk3(0.f, 0.f, 0.f), N = mk3(0.f, 0.f, 0.f);
    int mat = 0;
    float kr = 0.f, kt = 0.f, mior = 0.f, e_refl = 0.f, e_refr = 0.f;

    if (__any_sync(0xffffffffu, hit)) {
        // ---- ObtenirCouleurSurIntersection (Scene.cpp:1740-1826) ---------------------------
        float4 m0 = make_float4(0.f, 0.f, 0.f, 0.f), m1 = m0;
        if (hit) {
            mat = __float_as_int(sv.surf[k].r1.w);
            m0 = sv.mat[mat].m0; m1 = sv.mat[mat].m1;
            const float4 m2 = sv.mat[mat].m2;
            kr = m1.w; kt = m2.x; mior = m2.y;
            N = hit_normal<E>(sv, k, o, d, t, lin);
            P = axpy3<E>(o, d, t);
            cr = FP<E>::mul(m0.x, m0.w); cg = FP<E>::mul(m0.y, m0.w); cb = FP<E>::mul(m0.z, m0.w);
        }
        // ks == 0 with a sane shininess: the Phong term is (I*0)*pow(ps, n) = 0 for every finite
        // pow, and adding lcr*0 = 0 leaves the (non-negative) colour unchanged -> no powf needed
        const bool no_phong = m1.y == 0.f && m1.z >= 0.f && m1.z <= 1.0e4f;
        for (int l = 0; l < sv.n_light; ++l) {
            const float4 l0 = sv.light[l].l0, l1 = sv.light[l].l1;
            V3 Ld = sub3<E>(f4_xyz(l0), P);
            const bool need = hit && dot3<E>(Ld, N) > 0.f;
            if (!__any_sync(0xffffffffu, need)) continue;
            // ObtenirFiltreDeSurface normalises the light ray in place (Scene.cpp:1847-1848)
            const float dist = norm3<E>(Ld);
            Ld = scale3<E>(Ld, FP<E>::rcp(dist));
            float fr, fg, fb;
            const int tested = shadow_filter<E>(sv, need, P, Ld, dist, fr, fg, fb);
            if (need) {
                ++n_shadow; n_tests += (unsigned int)tested;
                const float lcr = FP<E>::mul(l1.x, fr), lcg = FP<E>::mul(l1.y, fg), lcb = FP<E>::mul(l1.z, fb);
                // Gouraud term (Scene.cpp:1763-1765): colour * (I*kd*dot(N,L)) * lightcolour
                const float g = FP<E>::mul(FP<E>::mul(l0.w, m1.x), dot3<E>(N, Ld));
                cr = FP<E>::madd(FP<E>::mul(m0.x, g), lcr, cr);
                cg = FP<E>::madd(FP<E>::mul(m0.y, g), lcg, cg);
                cb = FP<E>::madd(FP<E>::mul(m0.z, g), lcb, cb);
                // Phong term (Scene.cpp:1768-1775)
                if (!no_phong || !(lcr == lcr && lcg == lcg && lcb == lcb && fabsf(lcr) <= 3.0e38f && fabsf(lcg) <= 3.0e38f && fabsf(lcb) <= 3.0e38f)) {
                    const V3 R = reflect3<E>(Ld, N);
                    const float ps = dot3<E>(R, d);
                    if (ps > 0.f) {
                        const float ph = FP<E>::mul(FP<E>::mul(l0.w, m1.y), powf(ps, m1.z));
                        cr = FP<E>::madd(lcr, ph, cr);
                        cg = FP<E>::madd(lcg, ph, cg);
                        cb = FP<E>::madd(lcb, ph, cb);
                    }
                }
            }
        }
        // ---- secondary rays (Scene.cpp:1779-1823) ---------------------------------------------
        if (hit && may_spawn && level < f.max_bounces) {
            e_refl = FP<E>::mul(kr, energy);
            e_refr = FP<E>::mul(kt, energy);
            want_refl = e_refl > f.min_energy;
            want_refr = e_refr > f.min_energy;
        }
    }

    // ---- this ray's own colour -------------------------------------------------------------
    if (valid) {
        if (PRIMARY) {
            if (p.direct) {                                   // k == 1, no children: resolve here
                const size_t pix = (size_t)framebuffer_row(f, pix_y) * f.W + pix_x;
                if (p.fb_f32) { float* c = p.fb_f32 + 3 * pix; c[0] = cr; c[1] = cg; c[2] = cb; }
                if (p.fb_rgba8) p.fb_rgba8[pix] = quantise_rgba8(cr, cg, cb, n_nan);
            } else {
                float* c = p.c0 + 3 * (size_t)self;
                c[0] = cr; c[1] = cg; c[2] = cb;
            }
        } else {
            in_colour[self] = make_float4(cr, cg, cb, 0.f);
        }
    }

    // ---- warp-aggregated append of the children + link -------------------------------------
    const unsigned int m_refl = __ballot_sync(0xffffffffu, want_refl);
    const unsigned int m_refr = __ballot_sync(0xffffffffu, want_refr);
    if (m_refl | m_refr) {
        const unsigned int n_refl = __popc(m_refl), n_refr = __popc(m_refr);
        const unsigned int m_any = m_refl | m_refr;
        unsigned int base = 0, lbase = 0;
        if (lane == 0) {
            base  = atomicAdd(out_count, n_refl + n_refr);
            lbase = atomicAdd(link_count, (unsigned int)__popc(m_any));
        }
        base  = __shfl_sync(0xffffffffu, base, 0);
        lbase = __shfl_sync(0xffffffffu, lbase, 0);
        const unsigned int below = (1u << lane) - 1u;
        int s_refl = -1, s_refr = -1;
        // guessed capacity exhausted: nothing is written, the link carries no children (so that the
        // fold stays in bounds) and the host is told to redo the frame with exact capacities
        const bool room = !out_guarded || base + n_refl + n_refr <= out_cap;
        if (!room && lane == 0) *p.overflow = 1u;
        if (room && want_refl) {
            s_refl = (int)(base + __popc(m_refl & below));
            const V3 rd = reflect3<E>(d, N);
            out_od0[s_refl] = make_float4(P.x, P.y, P.z, e_refl);
            out_od1[s_refl] = make_float4(rd.x, rd.y, rd.z, 0.0f);            // CRayon default ior, Rayon.cpp:16
            ++n_refl_st;
        }
        if (room && want_refr) {
            s_refr = (int)(base + n_refl + __popc(m_refr & below));
            V3 n2 = N; float ratio, nior;
            if (ior == mior) { nior = f.scene_ior; ratio = FP<E>::div(mior, f.scene_ior); n2 = neg3(N); }
            else             { nior = mior;        ratio = FP<E>::div(f.scene_ior, mior); }
            const V3 rd = refract3<E>(d, n2, ratio);
            out_od0[s_refr] = make_float4(P.x, P.y, P.z, e_refr);
            out_od1[s_refr] = make_float4(rd.x, rd.y, rd.z, nior);
            ++n_refr_st;
        }
        if (want_refl || want_refr)
            links[lbase + __popc(m_any & below)] = make_int4(self, s_refl, s_refr, mat);
    }
    }   // tiles of this CTA
    if (tail_mode) __syncthreads();         // the next level reads what this one appended (same CTA: visible after the barrier)
    }   // levels of a tail-mode launch

    // ---- statistics ---------------------------------------------------------------------------
    warp_stats(p.stats, n_shadow, n_refl_st, n_refr_st, n_tests, n_nan);

    // ---- several GPUs: completion signal to rank 0, from the kernel that made the pixels ------
    if (p.peer_flag) peer_frame_signal(p.peer_done, p.peer_flag, p.peer_seq);
}

// ---------------------------------------------------------------------------------------
// Bottom-up fold: parent += refl_child * kr, then parent += refr_child * kt
// (Scene.cpp:1789 and :1822, in that order).  One thread per link: deterministic.
// ---------------------------------------------------------------------------------------
template <bool E>
__global__ void __launch_bounds__(256) k_combine(const CombineParams p)
{
    if (p.pdl) { pdl_launch_dependents(); pdl_wait(); }
    int n_links = p.n_links;
    if (p.n_links_dev) { const unsigned int n = *p.n_links_dev; n_links = (int)(n < (unsigned int)p.n_links ? n : (unsigned int)p.n_links); }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_links; i += gridDim.x * blockDim.x) {
    const int4 lk = p.links[i];
    const float kr = p.mat[lk.w].m1.w, kt = p.mat[lk.w].m2.x;
    float r, g, b;
    if (p.parent_is_level0) { const float* c = p.parent_c0 + 3 * (size_t)lk.x; r = c[0]; g = c[1]; b = c[2]; }
    else { const float4 c = p.parent_colour[lk.x]; r = c.x; g = c.y; b = c.z; }
    if (lk.y >= 0) {
        const float4 c = p.child_colour[lk.y];
        r = FP<E>::madd(c.x, kr, r); g = FP<E>::madd(c.y, kr, g); b = FP<E>::madd(c.z, kr, b);
    }
    if (lk.z >= 0) {
        const float4 c = p.child_colour[lk.z];
        r = FP<E>::madd(c.x, kt, r); g = FP<E>::madd(c.y, kt, g); b = FP<E>::madd(c.z, kt, b);
    }
    if (p.parent_is_level0) { float* c = p.parent_c0 + 3 * (size_t)lk.x; c[0] = r; c[1] = g; c[2] = b; }
    else p.parent_colour[lk.x] = make_float4(r, g, b, 0.f);
    }
}

// the same fold for a run of deep (small) levels, by ONE CTA: level l's parents are level l - 1's children,
// so a block barrier separates the levels (global writes of a block are visible to it after the barrier)
template <bool E>
__global__ void __launch_bounds__(1024) k_tail_fold(const __grid_constant__ TailFoldParams p)
{
    if (p.pdl) { pdl_launch_dependents(); pdl_wait(); }
    for (int l = p.last; l >= p.first; --l) {
        const int s = l - p.first;
        const unsigned int nd = p.counters[2 * l + 1];
        const int n = (int)(nd < (unsigned int)p.cap[s] ? nd : (unsigned int)p.cap[s]);
        const int4* __restrict__ links = p.links[s];
        float4* parent = p.colour[s];
        const float4* child = p.colour[s + 1];
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const int4 lk = links[i];
            const float kr = p.mat[lk.w].m1.w, kt = p.mat[lk.w].m2.x;
            const float4 c0 = parent[lk.x];
            float r = c0.x, g = c0.y, b = c0.z;
            if (lk.y >= 0) {
                const float4 c = child[lk.y];
                r = FP<E>::madd(c.x, kr, r); g = FP<E>::madd(c.y, kr, g); b = FP<E>::madd(c.z, kr, b);
            }
            if (lk.z >= 0) {
                const float4 c = child[lk.z];
                r = FP<E>::madd(c.x, kt, r); g = FP<E>::madd(c.y, kt, g); b = FP<E>::madd(c.z, kt, b);
            }
            parent[lk.x] = make_float4(r, g, b, 0.f);
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// Resolve: k x k box filter of the sample lattice (j outer, i inner, times 1/(k*k); k = 1 is
// a plain copy) + the RGBA8 quantiser floor(clamp(c,0,1)*255 + 0.5), NaN -> 0 (counted).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_resolve(const ResolveParams p)
{
    if (p.pdl) { pdl_launch_dependents(); pdl_wait(); }
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int yr = blockIdx.y * 8 + (threadIdx.x >> 5);          // pixel row inside the chunk
    unsigned int nans = 0;
    if (x < p.f.W && yr < p.chunk_prows) {
        const int k = p.f.k;
        float r, g, b;
        if (k == 1) {
            const float* c = p.c0 + 3 * ((size_t)yr * p.f.LW + x);
            r = c[0]; g = c[1]; b = c[2];
        } else {
            r = g = b = 0.f;
            for (int j = 0; j < k; ++j)
                for (int i = 0; i < k; ++i) {
                    const float* c = p.c0 + 3 * ((size_t)(yr * k + j) * p.f.LW + (x * k + i));
                    r = __fadd_rn(r, c[0]); g = __fadd_rn(g, c[1]); b = __fadd_rn(b, c[2]);
                }
            const float sc = __fdiv_rn(1.0f, (float)(k * k));
            r = __fmul_rn(r, sc); g = __fmul_rn(g, sc); b = __fmul_rn(b, sc);
        }
        const size_t pix = (size_t)framebuffer_row(p.f, p.chunk_py0 + yr) * p.f.W + x;
        if (p.fb_f32) { float* o = p.fb_f32 + 3 * pix; o[0] = r; o[1] = g; o[2] = b; }
        if (p.fb_rgba8) p.fb_rgba8[pix] = quantise_rgba8(r, g, b, nans);
    }
    nans = __reduce_add_sync(0xffffffffu, nans);
    if ((threadIdx.x & 31) == 0 && nans)
        atomicAdd(&p.stats[((blockIdx.x + blockIdx.y) % RT_STAT_SLOTS) * ST_COUNT + ST_NAN], (unsigned long long)nans);
}

// ---------------------------------------------------------------------------------------
// Ray-level queries over the unified list in global memory (parity-test entry points).
// ---------------------------------------------------------------------------------------
template <bool E>
__global__ void __launch_bounds__(128) k_query_nearest(DeviceScene s, int n, const float* o3, const float* d3,
                                                       int* out_order, float* out_t, float* out_n3, bool cull)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    SceneView sv; sv.surf = s.surf; sv.mat = s.mat; sv.light = s.light; sv.n_surf = s.rl.n_surf; sv.n_light = s.rl.n_light; sv.cull = cull; sv.pre_qo = nullptr; sv.pre_lim = nullptr;
    V3 o = valid ? ld3(o3 + 3 * (size_t)i) : mk3(0.f, 0.f, 0.f);
    V3 d = valid ? ld3(d3 + 3 * (size_t)i) : mk3(0.f, 0.f, 0.f);
    float t; int k; bool lin;
    nearest_hit<E>(sv, valid, o, d, t, k, lin);
    if (!valid) return;
    V3 N = k >= 0 ? hit_normal<E>(sv, k, o, d, t, lin) : mk3(0.f, 0.f, 0.f);
    out_order[i] = k; out_t[i] = t;
    out_n3[3 * (size_t)i] = N.x; out_n3[3 * (size_t)i + 1] = N.y; out_n3[3 * (size_t)i + 2] = N.z;
}

template <bool E>
__global__ void __launch_bounds__(128) k_query_shadow(DeviceScene s, int n, const float* p3, const float* l3, float* out3, bool cull)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    SceneView sv; sv.surf = s.surf; sv.mat = s.mat; sv.light = s.light; sv.n_surf = s.rl.n_surf; sv.n_light = s.rl.n_light; sv.cull = cull; sv.pre_qo = nullptr; sv.pre_lim = nullptr;
    V3 P = valid ? ld3(p3 + 3 * (size_t)i) : mk3(0.f, 0.f, 0.f);
    V3 L = valid ? ld3(l3 + 3 * (size_t)i) : mk3(1.f, 0.f, 0.f);
    V3 Ld = sub3<E>(L, P);
    const float dist = norm3<E>(Ld);
    Ld = scale3<E>(Ld, FP<E>::rcp(dist));
    float fr, fg, fb;
    (void)shadow_filter<E>(sv, valid, P, Ld, dist, fr, fg, fb);
    if (!valid) return;
    out3[3 * (size_t)i] = fr; out3[3 * (size_t)i + 1] = fg; out3[3 * (size_t)i + 2] = fb;
}

// ---------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------
size_t resident_smem_bytes(const ResidentLayout& rl) { return (size_t)rl.bytes; }
// blob + the per-CTA table of primary-ray culling terms (20 bytes per surface)
static size_t resident_smem_with_table(const ResidentLayout& rl) { return (((size_t)rl.bytes + 15) & ~(size_t)15) + 20 * (size_t)rl.n_surf + 16; }

cudaError_t resident_prepare(size_t max_smem_bytes)
{
    cudaError_t e;
    const int b = (int)max_smem_bytes;
    if ((e = cudaFuncSetAttribute(k_resident_level<false, true>,  cudaFuncAttributeMaxDynamicSharedMemorySize, b)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_resident_level<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, b)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_resident_level<true, true>,   cudaFuncAttributeMaxDynamicSharedMemorySize, b)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_resident_level<true, false>,  cudaFuncAttributeMaxDynamicSharedMemorySize, b)) != cudaSuccess) return e;
    return cudaSuccess;
}

// occupancy of the level kernels, cached per (instantiation, dynamic shared-memory bytes): it depends on the scene
// blob size and differs between the primary and the secondary instantiation.  (A property of the kernel image and of
// sm_100 alone, so one small process-wide table serves every context; guarded for multi-threaded callers.)
static int resident_ctas_per_sm(int inst, size_t smem)
{
    static std::mutex mu;
    static std::map<std::pair<int, size_t>, int> cache;
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find({inst, smem});
    if (it != cache.end()) return it->second;
    int n = 0;
    cudaError_t e;
    switch (inst) {
    case 0:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_resident_level<true, true>, RES_THREADS, smem); break;
    case 1:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_resident_level<false, true>, RES_THREADS, smem); break;
    case 2:  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_resident_level<true, false>, RES_THREADS, smem); break;
    default: e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_resident_level<false, false>, RES_THREADS, smem); break;
    }
    if (e != cudaSuccess || n <= 0) { (void)cudaGetLastError(); n = 1; }
    cache[{inst, smem}] = n;
    return n;
}

// plain launch, or (pdl) one that may start before the kernel ahead of it in the stream has finished
template <class P>
static cudaError_t launch_chain(void (*kernel)(P), dim3 grid, unsigned block, size_t smem, cudaStream_t st, bool pdl, const P& p)
{
    cudaLaunchConfig_t cfg = cudaLaunchConfig_t();
    cfg.gridDim = grid; cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, p);
}

cudaError_t launch_resident_level(const LevelParams& p_in, bool exact, int sm_count, cudaStream_t st)
{
    LevelParams p = p_in;
    // the table rides along while the scene is small (it must not cost a resident CTA)
    p.pre_table = (p.level == 0 && p.rl.bytes <= 16 * 1024) ? 1 : 0;
    const size_t smem = p.pre_table ? resident_smem_with_table(p.rl) : resident_smem_bytes(p.rl);
    const int max_ctas = resident_ctas_per_sm((exact ? 0 : 1) + (p.level == 0 ? 0 : 2), smem) * sm_count;
    const bool pdl = p.pdl == 2;
    if (p.level == 0) {
        const int tiles_x = (p.f.LW + 15) / 16, tiles_y = (p.chunk_rows + RES_TILE_H - 1) / RES_TILE_H;
        const dim3 grid((unsigned)std::min(tiles_x * tiles_y, max_ctas));
        return launch_chain(exact ? k_resident_level<true, true> : k_resident_level<false, true>, grid, RES_THREADS, smem, st, pdl, p);
    }
    const int n_max = p.in_count_dev ? p.in_cap : p.in_count;
    const dim3 grid(p.tail ? 1u : (unsigned)std::min((n_max + RES_THREADS - 1) / RES_THREADS, max_ctas));
    return launch_chain(exact ? k_resident_level<true, false> : k_resident_level<false, false>, grid, RES_THREADS, smem, st, pdl, p);
}

cudaError_t launch_combine(const CombineParams& p, bool exact, cudaStream_t st)
{
    // device-side count: the grid covers the capacity up to 8 blocks per SM, threads stride beyond
    const unsigned want = (unsigned)((p.n_links + 255) / 256);
    const dim3 grid(p.n_links_dev ? std::min(want, 1184u) : want);
    return launch_chain(exact ? k_combine<true> : k_combine<false>, grid, 256, 0, st, p.pdl == 2, p);
}

cudaError_t launch_tail_fold(const TailFoldParams& p, bool exact, cudaStream_t st)
{
    return launch_chain(exact ? k_tail_fold<true> : k_tail_fold<false>, dim3(1), 1024, 0, st, p.pdl == 2, p);
}

cudaError_t launch_resolve(const ResolveParams& p, cudaStream_t st)
{
    const dim3 grid((unsigned)((p.f.W + 31) / 32), (unsigned)((p.chunk_prows + 7) / 8));
    return launch_chain(k_resolve, grid, 256, 0, st, p.pdl == 2, p);
}

cudaError_t launch_query_nearest(const DeviceScene& s, int n, const float* o3, const float* d3,
                                 int* out_order, float* out_t, float* out_n3, bool exact, bool cull, cudaStream_t st)
{
    const dim3 grid((unsigned)((n + 127) / 128));
    if (exact) k_query_nearest<true><<<grid, 128, 0, st>>>(s, n, o3, d3, out_order, out_t, out_n3, cull);
    else       k_query_nearest<false><<<grid, 128, 0, st>>>(s, n, o3, d3, out_order, out_t, out_n3, cull);
    return cudaGetLastError();
}

cudaError_t launch_query_shadow(const DeviceScene& s, int n, const float* p3, const float* l3,
                                float* out3, bool exact, bool cull, cudaStream_t st)
{
    const dim3 grid((unsigned)((n + 127) / 128));
    if (exact) k_query_shadow<true><<<grid, 128, 0, st>>>(s, n, p3, l3, out3, cull);
    else       k_query_shadow<false><<<grid, 128, 0, st>>>(s, n, p3, l3, out3, cull);
    return cudaGetLastError();
}

} // namespace rtb
