// rt_bvh.cu -- GPU-built LBVH over the scene's triangles and its traversal kernels.
//
// The reference has no acceleration structure: nearest hit and shadow filter are linear scans
// (Scene.cpp:1710-1715, 1850-1858).  This file is the north-star's "optional GPU-built LBVH
// for large synthetic meshes that must return the same nearest hit as the brute-force loop":
//
//   build  (all on the device, no host round trip except one 24-byte bounds read):
//     scene bounds -> 63-bit Morton code of every triangle centroid -> radix sort (key, index)
//     -> Karras 2012 binary radix tree, one thread per internal node -> bottom-up refit with
//     one atomic counter per node.  Subtrees of <= BVH_LEAF_MAX triangles become leaves
//     (contiguous ranges of the Morton-sorted triangle array).
//   exactness: a leaf test IS the reference's rounding sequence (triangle_t<true>), so the BVH
//     only has to guarantee that it never culls a triangle the linear scan would accept.  The
//     accepted hit of CTriangle::Intersection can lie outside the geometric triangle by the
//     rounding error of u, v and t, which grows like 1/|det| up to the reference's own
//     threshold |det| >= EPSILON (Triangle.cpp:142).  Every triangle's box is therefore padded by
//         pad = PADC * 2^-24 / EPSILON * R * |e1| * |e2|      (R = largest |origin - p0| possible)
//     (first-order bound of the displacement) and nodes are culled only when the
//     PADDED box lies strictly beyond the current best distance.  Ties resolve on (t, order)
//     exactly like Scene.cpp:1713.
//   traversal: one ray per thread, stack in local memory, nearer child first, both child boxes
//     in one 64-byte node fetch; primary rays are generated on the fly in 8x4-pixel warps.
//   bounded quadrics: a quadric rtb200_set_scene could write in centred form
//     sum q_i (x_i - c_i)^2 = rhs with every q_i > 0 and no mixed terms (spheres, axis-aligned
//     ellipsoids) is a primitive of the same tree; its box is the box of the INFLATED quadric
//     sum q_i (x_i - c_i)^2 <= 1.05 rhs + 2e-5 M(R), the region outside of which the reference's
//     float discriminant cannot come out positive (rt_resident.cu: quadric_miss, same margin, with
//     M bounded for every origin within R).  Its leaf test is the reference's Quadrique.cpp:160
//     sequence with the ten original coefficients.  Cylinders, cones, paraboloids, rotated quadrics
//     and planes stay in the linear side list.
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "rt_bvh.cuh"
#include "rt_context.h"
#include "rt_cull.cuh"

namespace rtb {

#define FULL_MASK 0xffffffffu
#define BVH_SLACK 1.0e-5f          // relative slack of the slab test (rounding of the test itself)
#define BVH_SHADOW_LIST 8          // transparent occluders kept per shadow ray for in-order products

struct BvhState {
    DevBuf nodes, tri_sorted, keys, keys_out, idx, idx_out, cub_tmp, parent, leaf_parent, gamma, flags, bounds;
    DevBuf tree_quad, side_quad;                   // bounded quadrics as tree records / the rest as QuadRec
    int    n_tri = 0, n_nodes = 0, root = 0;
    int    n_prims = 0, n_tree_quad = 0, n_side_quad = 0;
    bool   built = false;
    float  lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};   // vertex bounds of the triangles
    float  radius = 0.f;                           // R the boxes are padded for
    float  wmax = 0.f;                             // largest |e1| |e2| of a triangle (far rays: rt_scene.cuh, DeviceScene::bvh_grow_tri)
    float  kq = 0.f, cmax = 0.f;                   // bounded quadrics in the tree: max (sum q / min q), max |centre coordinate|
    float  padc = 1.0f;                            // PADC; RTB200_BVH_PADC overrides
    double build_ms = 0.0, refit_ms = 0.0;
    unsigned refits = 0;
};

static BvhState* bvh_state(rtb200_ctx* c)
{
    if (!c->bvh_state) c->bvh_state = new BvhState();
    return static_cast<BvhState*>(c->bvh_state);
}

// ---------------------------------------------------------------------------------------
// build kernels
// ---------------------------------------------------------------------------------------
RT_DEV int float_order(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
static inline float order_float(int i) { const int b = i >= 0 ? i : i ^ 0x7fffffff; float f; memcpy(&f, &b, 4); return f; }

// primitive i of the build: a triangle of the scene array or a bounded quadric record
RT_DEV TriRec bvh_prim(const TriRec* __restrict__ tri, int n_tri, const TriRec* __restrict__ tq, int i)
{
    return i < n_tri ? tri[i] : tq[i - n_tri];
}
RT_DEV bool rec_is_quadric(const float4& c) { return c.w != 0.0f; }

// box of a bounded quadric record; R < 0: the quadric itself, R >= 0: inflated for ray origins within R
RT_DEV void quadric_box(const float4& a, const float4& b, const float4& c, float R, float* lo, float* hi)
{
    const float q[3] = {a.x, a.y, a.z}, l[3] = {b.x, b.y, b.z};
    float ctr[3], M = 0.f;
    #pragma unroll
    for (int k = 0; k < 3; ++k) {
        ctr[k] = -l[k] / (2.0f * q[k]);
        const float far = fabsf(ctr[k]) + fmaxf(R, 0.f);
        M += q[k] * (far * far + ctr[k] * ctr[k]);
    }
    const float lim = R < 0.f ? c.y : fmaf(c.y, 1.05f, 2.0e-5f * M);
    #pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float h = sqrtf(lim / q[k]) * 1.00001f + 1.0e-6f * fabsf(ctr[k]);
        lo[k] = ctr[k] - h; hi[k] = ctr[k] + h;
    }
}

__global__ void k_bvh_bounds(const TriRec* __restrict__ tri, int n_tri, const TriRec* __restrict__ tq, int n, int* __restrict__ bounds)
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}, wmax = 0.f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const TriRec r = bvh_prim(tri, n_tri, tq, i);
        const float4 a = r.a, b = r.b, c = r.c;
        if (rec_is_quadric(c)) {
            float ql[3], qh[3];
            quadric_box(a, b, c, -1.0f, ql, qh);
            #pragma unroll
            for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], ql[k]); hi[k] = fmaxf(hi[k], qh[k]); }
            continue;
        }
        const float p[3] = {a.x, a.y, a.z}, e1[3] = {b.x, b.y, b.z}, e2[3] = {c.x, c.y, c.z};
        #pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float v1 = p[k] + e1[k], v2 = p[k] + e2[k];
            lo[k] = fminf(lo[k], fminf(p[k], fminf(v1, v2)));
            hi[k] = fmaxf(hi[k], fmaxf(p[k], fmaxf(v1, v2)));
        }
        const float w = sqrtf(b.x * b.x + b.y * b.y + b.z * b.z) * sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
        if (w <= 3.0e38f) wmax = fmaxf(wmax, w);                // (the l1 * l2 of k_bvh_refit's pad)
    }
    for (int off = 16; off > 0; off >>= 1) wmax = fmaxf(wmax, __shfl_xor_sync(FULL_MASK, wmax, off));
    if ((threadIdx.x & 31) == 0) atomicMax(&bounds[6], __float_as_int(wmax));     // non-negative floats order like ints
    #pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int off = 16; off > 0; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(FULL_MASK, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(FULL_MASK, hi[k], off));
        }
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(&bounds[k], float_order(lo[k]));
            atomicMax(&bounds[3 + k], float_order(hi[k]));
        }
    }
}

RT_DEV unsigned long long expand21(unsigned int v)
{
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

struct Bounds6 { float lo[3], inv[3]; };

__global__ void k_bvh_morton(const TriRec* __restrict__ tri, int n_tri, const TriRec* __restrict__ tq, int n, Bounds6 b,
                             unsigned long long* __restrict__ keys, unsigned int* __restrict__ idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const TriRec r = bvh_prim(tri, n_tri, tq, i);
    const float4 a = r.a, e1 = r.b, e2 = r.c;
    float c[3] = {a.x + (e1.x + e2.x) * (1.0f / 3.0f), a.y + (e1.y + e2.y) * (1.0f / 3.0f), a.z + (e1.z + e2.z) * (1.0f / 3.0f)};
    if (rec_is_quadric(e2)) { c[0] = -e1.x / (2.0f * a.x); c[1] = -e1.y / (2.0f * a.y); c[2] = -e1.z / (2.0f * a.z); }
    unsigned int q[3];
    #pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float u = fminf(fmaxf((c[k] - b.lo[k]) * b.inv[k], 0.0f), 1.0f);
        q[k] = min((unsigned int)(u * 2097152.0f), 2097151u);
    }
    keys[i] = expand21(q[0]) << 2 | expand21(q[1]) << 1 | expand21(q[2]);
    idx[i] = (unsigned int)i;
}

__global__ void k_bvh_gather(const TriRec* __restrict__ tri, int n_tri, const TriRec* __restrict__ tq,
                             const unsigned int* __restrict__ idx, int n, TriRec* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = bvh_prim(tri, n_tri, tq, (int)idx[i]);
}

// common-prefix length of keys i and j (Karras 2012, sec. 4); equal keys fall back to the index
RT_DEV int bvh_delta(const unsigned long long* __restrict__ k, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = k[i], b = k[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

RT_DEV int bvh_leaf_ref(int first, int count) { return ~(first | ((count - 1) << BVH_LEAF_SHIFT)); }

__global__ void k_bvh_hierarchy(const unsigned long long* __restrict__ keys, int n, BvhNode* __restrict__ nodes,
                                int* __restrict__ parent, int* __restrict__ leaf_parent, int* __restrict__ gamma)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (bvh_delta(keys, n, i, i + 1) - bvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = bvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (bvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (bvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = bvh_delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (bvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int g = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    gamma[i] = g;
    // Karras children: left = node/leaf g, right = node/leaf g + 1
    if (first == g) leaf_parent[g] = i; else parent[g] = i;
    if (last == g + 1) leaf_parent[g + 1] = i; else parent[g + 1] = i;
    if (i == 0) parent[0] = -1;
    // traversal references: small subtrees collapse into leaves over [first, g] / [g + 1, last]
    const int nl = g - first + 1, nr = last - g;
    nodes[i].c = make_int4(nl <= BVH_LEAF_MAX ? bvh_leaf_ref(first, nl) : g,
                           nr <= BVH_LEAF_MAX ? bvh_leaf_ref(g + 1, nr) : g + 1, first, last);
}

// bottom-up refit: every triangle carries its padded box up; the second thread to reach a node
// merges the two child boxes and continues
__global__ void k_bvh_refit(const TriRec* __restrict__ tri, int n, BvhNode* nodes, const int* __restrict__ parent,
                            const int* __restrict__ leaf_parent, const int* __restrict__ gamma, int* flags,
                            float pad_scale, float pad_floor, float radius)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = tri[i].a, b = tri[i].b, c = tri[i].c;
    const float l1 = sqrtf(b.x * b.x + b.y * b.y + b.z * b.z), l2 = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
    const float pad = fmaxf(pad_scale * l1 * l2, pad_floor);
    float lo[3], hi[3];
    if (rec_is_quadric(c)) {
        quadric_box(a, b, c, radius, lo, hi);
        #pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] -= pad_floor; hi[k] += pad_floor; }
    } else {
        const float p[3] = {a.x, a.y, a.z}, e1[3] = {b.x, b.y, b.z}, e2[3] = {c.x, c.y, c.z};
        #pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float v1 = p[k] + e1[k], v2 = p[k] + e2[k];
            lo[k] = fminf(p[k], fminf(v1, v2)) - pad;
            hi[k] = fmaxf(p[k], fmaxf(v1, v2)) + pad;
        }
    }
    int child = i, p = leaf_parent[i];
    while (p >= 0) {
        float* f = reinterpret_cast<float*>(&nodes[p]);
        const int slot = (child == gamma[p]) ? 0 : 1;
        float* mine = f + 6 * slot;
        mine[0] = lo[0]; mine[1] = lo[1]; mine[2] = lo[2]; mine[3] = hi[0]; mine[4] = hi[1]; mine[5] = hi[2];
        __threadfence();
        if (atomicAdd(&flags[p], 1) == 0) return;          // the sibling subtree is not finished yet
        __threadfence();
        const volatile float* other = f + 6 * (1 - slot);
        lo[0] = fminf(lo[0], other[0]); lo[1] = fminf(lo[1], other[1]); lo[2] = fminf(lo[2], other[2]);
        hi[0] = fmaxf(hi[0], other[3]); hi[1] = fmaxf(hi[1], other[4]); hi[2] = fmaxf(hi[2], other[5]);
        child = p;
        p = parent[p];
    }
}

// ---------------------------------------------------------------------------------------
// traversal
// ---------------------------------------------------------------------------------------
struct RayBox { float ix, iy, iz, nx, ny, nz; };           // 1/d and -(o/d)

RT_DEV float safe_rcp(float d)
{
    return fabsf(d) > 1.0e-30f ? 1.0f / d : copysignf(1.0e30f, d);
}

RT_DEV RayBox make_raybox(V3 o, V3 d)
{
    RayBox r;
    r.ix = safe_rcp(d.x); r.iy = safe_rcp(d.y); r.iz = safe_rcp(d.z);
    r.nx = -(o.x * r.ix); r.ny = -(o.y * r.iy); r.nz = -(o.z * r.iz);
#ifndef BVH_REMAT
    // keep the three products in registers: without this ptxas recomputes them at every node
    asm volatile("" : "+f"(r.nx), "+f"(r.ny), "+f"(r.nz));
#endif
    return r;
}

// does the ray meet the (already padded) box inside [0, limit]?  tn = entry distance
RT_DEV bool slab(const RayBox& r, float lx, float ly, float lz, float hx, float hy, float hz, float limit, float& tn)
{
    const float ax = fmaf(lx, r.ix, r.nx), bx = fmaf(hx, r.ix, r.nx);
    const float ay = fmaf(ly, r.iy, r.ny), by = fmaf(hy, r.iy, r.ny);
    const float az = fmaf(lz, r.iz, r.nz), bz = fmaf(hz, r.iz, r.nz);
    tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), limit));
    return tn <= tf * (1.0f + 2.0f * BVH_SLACK);
}
// The same test with the ray's OCTANT known at compile time (bit k of OCT set: direction component k is negative): which of
// a box's two planes per axis is entered first is then a matter of naming registers, not of six FMNMX per box -- the
// values of tn and tf are the ones slab() computes (fma is monotone: lo <= hi  =>  fma(lo, i, n) <= fma(hi, i, n) for
// i >= 0), so the walk, its counters and the frame are unchanged.  12 instead of 18 instructions per box.
template <int OCT>
RT_DEV bool slab_oct(const RayBox& r, float lx, float ly, float lz, float hx, float hy, float hz, float limit, float& tn)
{
    const float ax = fmaf((OCT & 1) ? hx : lx, r.ix, r.nx), bx = fmaf((OCT & 1) ? lx : hx, r.ix, r.nx);
    const float ay = fmaf((OCT & 2) ? hy : ly, r.iy, r.ny), by = fmaf((OCT & 2) ? ly : hy, r.iy, r.ny);
    const float az = fmaf((OCT & 4) ? hz : lz, r.iz, r.nz), bz = fmaf((OCT & 4) ? lz : hz, r.iz, r.nz);
    tn = fmaxf(fmaxf(ax, ay), fmaxf(az, 0.0f));
    const float tf = fminf(fminf(bx, by), fminf(bz, limit));
    return tn <= tf * (1.0f + 2.0f * BVH_SLACK);
}
// octant of a direction for slab_oct, or 8 (the generic test) when a component is so small that safe_rcp replaces its
// reciprocal, or not finite -- there the products can overflow and only slab()'s min / max ignore the NaNs
RT_DEV int ray_octant(V3 d)
{
    const bool ok = fabsf(d.x) > 1.0e-30f && fabsf(d.y) > 1.0e-30f && fabsf(d.z) > 1.0e-30f &&
                    fabsf(d.x) < 1.0e30f && fabsf(d.y) < 1.0e30f && fabsf(d.z) < 1.0e30f;
    return ok ? ((d.x < 0.f ? 1 : 0) | (d.y < 0.f ? 2 : 0) | (d.z < 0.f ? 4 : 0)) : 8;
}
// the same for the box grown by g on every side (gx = g |1/dx| ...: the slab distances move by that much)
RT_DEV bool slab_grown(const RayBox& r, float gx, float gy, float gz, float lx, float ly, float lz, float hx, float hy, float hz, float limit, float& tn)
{
    const float ax = fmaf(lx, r.ix, r.nx), bx = fmaf(hx, r.ix, r.nx);
    const float ay = fmaf(ly, r.iy, r.ny), by = fmaf(hy, r.iy, r.ny);
    const float az = fmaf(lz, r.iz, r.nz), bz = fmaf(hz, r.iz, r.nz);
    tn = fmaxf(fmaxf(fminf(ax, bx) - gx, fminf(ay, by) - gy), fmaxf(fminf(az, bz) - gz, 0.0f));
    const float tf = fminf(fminf(fmaxf(ax, bx) + gx, fmaxf(ay, by) + gy), fminf(fmaxf(az, bz) + gz, limit));
    return tn <= tf * (1.0f + 2.0f * BVH_SLACK);
}

// Far rays.  The boxes are padded for origins whose farthest corner of the primitives' bounds lies within bvh_radius
// (bvh_refit: pad = PADC 2^-24 / EPSILON R |e1| |e2|, floor 32 ulp of R; quadrics inflated for R).  `reach` is that
// distance for one origin; a ray with reach > R needs every pad scaled by reach / R: its walk grows every box by the
// pad of the LARGEST triangle at its reach (>= what any box lacks), and by the growth of the quadrics' inflated boxes,
//     sqrt(lim(D)/q) - sqrt(lim(R)/q) <= sqrt((lim(D) - lim(R)) / q_min),  lim(D) - lim(R) = 2e-5 (M(D) - M(R)),
//     M(D) - M(R) = sum q_i ((|c_i| + D)^2 - (|c_i| + R)^2) <= (D - R)(2 c_max + D + R) sum q_i.
RT_DEV float bvh_far_grow(const DeviceScene& s, float reach)
{
    float g = s.bvh_grow_tri * reach;
    if (s.bvh_kq > 0.f) {
        const float R = s.bvh_radius;
        g = fmaxf(g, sqrtf(2.1e-5f * s.bvh_kq * fmaxf(reach - R, 0.f) * (2.0f * s.bvh_cmax + reach + R)) * 1.01f);
    }
    return g;
}

// One triangle of a leaf.  MODE 0 / 2: the conservative fused filter of rt_cull.cuh (phase 1: det, u; phase 2:
// v, t -- numerator space, error-scaled margins), survivors re-evaluated in the reference's rounding sequence
// (MODE 0) or fused (MODE 2); MODE 1 (RTB200_F_EXACT_FP): the reference sequence for every test.
template <int MODE>
RT_DEV float triangle_leaf(V3 o, V3 d, V3 p0, V3 e1, V3 e2, float tmax)
{
    if (MODE == 1) return triangle_t<true>(o, d, p0, e1, e2);
    const TriMargins mg = tri_margins(e1, e2, max_abs3(d));
    if (!triangle_phase1(o, d, p0, e1, e2, mg)) return -1.0f;
    return triangle_survivor<MODE>(o, d, p0, e1, e2, tmax, mg);
}

// Generic "while-while" traversal: the inner loop walks internal nodes only, so the lanes of a
// warp leave it together and run the leaf code converged.  `Leaf` is called as
// leaf(first, count, limit) and returns the new distance limit (negative = stop the traversal).
// SHRINK: the limit can come down during the walk (nearest hit), so postponed subtrees carry their
// entry distance and are dropped when they lie beyond it; an occlusion query (fixed limit) keeps
// references only.  The bottom stack entry is a sentinel, so a pop never tests for an empty stack.
#define BVH_DONE INT_MIN              // not a leaf reference: first + count <= n < 2^28 (bvh_build)

template <bool SHRINK, bool GROW = false, int OCT = 8, class Leaf>
RT_DEV void bvh_traverse(const DeviceScene& s, V3 o, V3 d, float limit, Leaf&& leaf, unsigned& n_nodes, unsigned& overflow, float grow = 0.f)
{
    const RayBox rb = make_raybox(o, d);
    const float gx = GROW ? grow * fabsf(rb.ix) : 0.f, gy = GROW ? grow * fabsf(rb.iy) : 0.f, gz = GROW ? grow * fabsf(rb.iz) : 0.f;
    const BvhNode* __restrict__ nodes = s.bvh_nodes;
    int   stack_ref[BVH_STACK + 1];
    float stack_tn[SHRINK ? BVH_STACK + 1 : 1];
    stack_ref[0] = BVH_DONE;
    stack_tn[0] = -1.0f;
    int sp = 1;
    int ref = s.bvh_root;
    for (;;) {
        while (ref >= 0) {                                   // ---- internal nodes
            const BvhNode* __restrict__ nd = nodes + ref;
            const float4 b0 = __ldg(&nd->b0), b1 = __ldg(&nd->b1), b2 = __ldg(&nd->b2);
            const int2 c = __ldg(reinterpret_cast<const int2*>(&nd->c));
            ++n_nodes;
            float t0, t1;
            bool h0, h1;
            if (GROW) {
                h0 = slab_grown(rb, gx, gy, gz, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, limit, t0);
                h1 = slab_grown(rb, gx, gy, gz, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w, limit, t1);
            } else if (OCT < 8) {
                h0 = slab_oct<OCT & 7>(rb, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, limit, t0);
                h1 = slab_oct<OCT & 7>(rb, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w, limit, t1);
            } else {
                h0 = slab(rb, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, limit, t0);
                h1 = slab(rb, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w, limit, t1);
            }
            if (h0 && h1) {
                const bool swap = t1 < t0;
                if (sp <= BVH_STACK) {
                    stack_ref[sp] = swap ? c.x : c.y;
                    if (SHRINK) stack_tn[sp] = swap ? t0 : t1;
                    ++sp;
                } else overflow = 1;
                ref = swap ? c.y : c.x;
            } else if (h0 || h1) ref = h0 ? c.x : c.y;
            else if (SHRINK) {                               // next postponed subtree that can still matter
                do { --sp; } while (stack_tn[sp] > limit * (1.0f + 2.0f * BVH_SLACK));
                ref = stack_ref[sp];
            } else ref = stack_ref[--sp];
        }
        if (ref == BVH_DONE) return;
        const int code = ~ref;                               // ---- leaf
        RT_CHECK((code & ((1 << BVH_LEAF_SHIFT) - 1)) + (code >> BVH_LEAF_SHIFT) + 1 <= s.bvh_n_prims && sp >= 1 && sp <= BVH_STACK + 1);
        limit = leaf(code & ((1 << BVH_LEAF_SHIFT) - 1), (code >> BVH_LEAF_SHIFT) + 1, limit);
        if (limit < 0.f) return;
        if (SHRINK) {
            do { --sp; } while (stack_tn[sp] > limit * (1.0f + 2.0f * BVH_SLACK));
            ref = stack_ref[sp];
        } else ref = stack_ref[--sp];
    }
}

// the walk specialised for the ray's octant (nine copies of the loop; the lanes of a warp mostly share theirs).  Used by
// the nearest-hit kernel (config 5: 100 -> 90 ms, config 4: 1.80 -> 1.69 ms); the occlusion kernel keeps the generic
// walk -- measured with the variants: no gain (config 5 shadow 325 -> 324 ms) and + 5 % on config 4 (56 registers:
// spills, and nine copies of its larger leaf code).
template <bool SHRINK, class Leaf>
RT_DEV void bvh_traverse_oct(const DeviceScene& s, V3 o, V3 d, float limit, Leaf&& leaf, unsigned& n_nodes, unsigned& overflow)
{
    // ... only when every lane that walks now shares the octant: a warp of mixed octants (the eight lights of one hit
    // point in adjacent lanes) would serialise nine copies of the loop (measured: config 4 shadow 6.6 -> 14.1 ms).
    // Any choice is correct -- the variants compute the same values -- so the vote only has to be cheap.
    int oct = ray_octant(d);
    const unsigned int active = __activemask();
    if (!__all_sync(active, oct == __shfl_sync(active, oct, __ffs((int)active) - 1))) oct = 8;
    switch (oct) {
    case 0: bvh_traverse<SHRINK, false, 0>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 1: bvh_traverse<SHRINK, false, 1>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 2: bvh_traverse<SHRINK, false, 2>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 3: bvh_traverse<SHRINK, false, 3>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 4: bvh_traverse<SHRINK, false, 4>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 5: bvh_traverse<SHRINK, false, 5>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 6: bvh_traverse<SHRINK, false, 6>(s, o, d, limit, leaf, n_nodes, overflow); break;
    case 7: bvh_traverse<SHRINK, false, 7>(s, o, d, limit, leaf, n_nodes, overflow); break;
    default: bvh_traverse<SHRINK, false, 8>(s, o, d, limit, leaf, n_nodes, overflow); break;
    }
}

template <int MODE, bool PRIMARY, bool QUADS>    // QUADS: the tree holds bounded quadrics besides triangles
__global__ void __launch_bounds__(BVH_THREADS)
k_bvh_nearest(const __grid_constant__ TraceParams p)
{
    const DeviceScene& s = p.s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    V3 o, d;
    int slot = -1;
    if (PRIMARY) {                                            // CTA = 16x8 lattice tile, warp = 8x4
        const int tiles_x = (p.f.LW + 15) >> 4;
        const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
        const int lx = (tx << 4) + ((warp & 1) << 3) + (lane & 7);
        const int cy = (ty << 3) + ((warp >> 1) << 2) + (lane >> 3);
        const int ly = p.chunk_ly0 + cy;
        if (lx < p.f.LW && cy < p.chunk_rows && ly < p.local_LH) {
            slot = cy * p.f.LW + lx;
            primary_ray<(MODE != 2)>(p.f, lx, global_lattice_row(p.f, ly), o, d);
        }
    } else {
        const int i = blockIdx.x * BVH_THREADS + threadIdx.x;
        if (i < p.n_rays) { slot = i; o = f4_xyz(p.in_od0[i]); d = f4_xyz(p.in_od1[i]); }
        if (slot >= 0 && p.far_count && s.bvh_n_prims > 0 && bvh_is_far(s, o)) {       // beyond the padded radius: k_bvh_far_nearest
            const unsigned int k = atomicAdd(p.far_count, 1u);
            if ((int)k < p.far_cap) { p.far_list[k] = slot; slot = -1; }
        }
    }
    unsigned n_nodes = 0, n_tests = 0, n_qtests = 0, overflow = 0;
    if (slot >= 0) {
        float best_t = -1.0f;
        int best_id = -1;
        if (s.bvh_n_prims > 0) {
            const TriRec* __restrict__ tri = s.bvh_tri;
            bvh_traverse_oct<true>(s, o, d, 3.0e38f, [&](int first, int count, float limit) -> float {
                for (int k = 0; k < count; ++k) {
                    const float4 ra = __ldg(&tri[first + k].a), rb = __ldg(&tri[first + k].b), rc = __ldg(&tri[first + k].c);
                    float tt;
                    if (QUADS && rec_is_quadric(rc)) {      // bounded quadric: the reference sequence, original coefficients
                        const Quadric Q = { ra.x, ra.y, ra.z, 0.f, 0.f, 0.f, rb.x, rb.y, rb.z, rc.x };
                        bool lin;
                        tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                        ++n_qtests; --n_tests;
                    } else
                    tt = triangle_leaf<MODE>(o, d, f4_xyz(ra), f4_xyz(rb), f4_xyz(rc),
                                                         best_t < 0.f ? 3.0e38f : fmaf(best_t, CULL_TREL, CULL_TABS));
                    const int ord = __float_as_int(ra.w);
                    if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) {
                        best_t = tt; best_id = ord; limit = tt;
                    }
                }
                n_tests += (unsigned)count;
                return limit;
            }, n_nodes, overflow);
        }
        // quadrics outside the tree and planes: the brute-force side list (reference sequence; ties -> lower order)
        n_qtests += (unsigned)s.bvh_n_quad;
        for (int q = 0; q < s.bvh_n_quad; ++q) {
            const float4 q0 = __ldg(&s.bvh_quad[q].q0), q1 = __ldg(&s.bvh_quad[q].q1), q2 = __ldg(&s.bvh_quad[q].q2);
            const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
            const int ord = __float_as_int(q0.w);
            bool lin;
            const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
            if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
        }
        for (int q = 0; q < s.n_plane; ++q) {
            const float4 nc = __ldg(&s.plane[q].nc);
            const int ord = s.plane[q].order;
            const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
            if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
        }
        p.hit[slot] = make_float2(best_t, __int_as_float(best_id));
    }
    // statistics: one atomic per warp and counter
    n_nodes = __reduce_add_sync(FULL_MASK, n_nodes);         // one REDUX each instead of five shuffle rounds
    n_tests = __reduce_add_sync(FULL_MASK, n_tests);
    n_qtests = __reduce_add_sync(FULL_MASK, n_qtests);
    overflow = __reduce_or_sync(FULL_MASK, overflow);
    if (lane == 0 && p.stats) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (n_nodes) atomicAdd(&st[ST_BVH_NODES], (unsigned long long)n_nodes);
        if (n_tests) atomicAdd(&st[ST_TEST_TRI], (unsigned long long)n_tests);
        if (n_qtests) atomicAdd(&st[ST_TEST_QUAD], (unsigned long long)n_qtests);
        if (overflow) atomicAdd(&st[ST_BVH_OVERFLOW], 1ull);
    }
}

// Shadow filter through the BVH: every triangle hit strictly inside (EPSILON, dist) multiplies
// the filter by colour*kt (Scene.cpp:1856-1857).  An opaque occluder makes the filter exactly 0
// (0 * finite = 0), so the traversal may stop there; transparent occluders are collected and
// multiplied in insertion order, the order of the linear scan.
template <int MODE, bool QUADS>
#ifndef BVH_SHADOW_CTAS
#define BVH_SHADOW_CTAS 9
#endif
__global__ void __launch_bounds__(BVH_THREADS, BVH_SHADOW_CTAS)          // 56 registers: 9 CTAs per SM (64 registers: +6 % time)
k_bvh_shadow(const __grid_constant__ ShadowParams p)
{
    const DeviceScene& s = p.s;
    const int lane = threadIdx.x & 31;
    int i = blockIdx.x * BVH_THREADS + threadIdx.x;
    if (p.by_light) {
        // Queries arrive ray-major (the facing lights of one hit are neighbours).  Where many rays start on the
        // same primitive (supersampling, few large triangles) a warp walks the tree more coherently when its
        // lanes also share the LIGHT: within a block of 256 queries warp w takes every 8th one (config 5:
        // -4 % on this kernel; config 4, 8 hits per triangle: +20 %, hence the switch in stream_render).
        const int blk = i & ~255;
        if (blk + 256 <= p.n_queries) i = blk | ((i & 31) << 3) | ((i >> 5) & 7);
    }
    unsigned n_nodes = 0, n_tests = 0, overflow = 0, n_side_q = 0, n_side_p = 0;
    bool mine = i < p.n_queries;
    V3 o = mk3(0.f, 0.f, 0.f), d = o;
    float dist = 0.f;
    if (mine) {
        const bool far = shadow_query<(MODE != 2)>(p, i, o, d, dist);
        if (far && p.far_count) {                                             // beyond the padded radius (shade-A marked it): k_bvh_far_shadow
            const unsigned int k = atomicAdd(p.far_count, 1u);
            if ((int)k < p.far_cap) { p.far_list[k] = i; mine = false; }
        }
    }
    if (mine) {
        const float tlim = fmaf(dist, CULL_TREL, CULL_TABS);
        float fr = 1.0f, fg = 1.0f, fb = 1.0f;
        bool zero = false;
        int l_ord[BVH_SHADOW_LIST], l_mat[BVH_SHADOW_LIST];
        int n_list = 0;
        // the product below is taken over the tree's occluders in insertion order, then the side list's: a query with
        // three or more factors (two commute exactly) that were not met in increasing order is redone by k_shadow_fixup
        bool ooo = false;
        if (s.bvh_n_prims > 0) {
            const TriRec* __restrict__ tri = s.bvh_tri;
            bvh_traverse<false>(s, o, d, dist, [&](int first, int count, float limit) -> float {
                n_tests += (unsigned)count;
                for (int k = 0; k < count; ++k) {
                    const float4 ra = __ldg(&tri[first + k].a), rb = __ldg(&tri[first + k].b), rc = __ldg(&tri[first + k].c);
                    float tt;
                    if (QUADS && rec_is_quadric(rc)) {
                        const Quadric Q = { ra.x, ra.y, ra.z, 0.f, 0.f, 0.f, rb.x, rb.y, rb.z, rc.x };
                        bool lin;
                        tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                        ++n_side_q; --n_tests;
                    } else
                    tt = triangle_leaf<MODE>(o, d, f4_xyz(ra), f4_xyz(rb), f4_xyz(rc), tlim);
                    if (tt > RT_EPS && tt < dist) {
                        const int m = __float_as_int(rb.w);
                        const float4 m0 = __ldg(&s.mat[m].m0);
                        const float kt = __ldg(&s.mat[m].m2).x;
                        const float cr = __fmul_rn(m0.x, kt), cg = __fmul_rn(m0.y, kt), cb = __fmul_rn(m0.z, kt);
                        if (cr == 0.f && cg == 0.f && cb == 0.f) { zero = true; return -1.0f; }
                        if (n_list < BVH_SHADOW_LIST) { l_ord[n_list] = __float_as_int(ra.w); l_mat[n_list] = m; ++n_list; }
                        else { fr = __fmul_rn(fr, cr); fg = __fmul_rn(fg, cg); fb = __fmul_rn(fb, cb); ooo = true; }   // > 8 transparent layers: redone in order (fix-up list)
                    }
                }
                return limit;
            }, n_nodes, overflow);
        }
        if (zero) { fr = fg = fb = 0.0f; }
        else {
            for (int x = 1; x < n_list; ++x) {                 // insertion sort by order
                const int ko = l_ord[x], km = l_mat[x];
                int y = x - 1;
                while (y >= 0 && l_ord[y] > ko) { l_ord[y + 1] = l_ord[y]; l_mat[y + 1] = l_mat[y]; --y; }
                l_ord[y + 1] = ko; l_mat[y + 1] = km;
            }
            for (int x = 0; x < n_list; ++x) {
                const float4 m0 = __ldg(&s.mat[l_mat[x]].m0);
                const float kt = __ldg(&s.mat[l_mat[x]].m2).x;
                fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
            }
        }
        int n_occ = n_list + (ooo ? 1 : 0), max_ord = n_list ? l_ord[n_list - 1] : -1;
        if (!(fr == 0.f && fg == 0.f && fb == 0.f)) {
            n_side_q += (unsigned)s.bvh_n_quad; n_side_p = (unsigned)s.n_plane;
            for (int q = 0; q < s.bvh_n_quad; ++q) {
                const float4 q0 = __ldg(&s.bvh_quad[q].q0), q1 = __ldg(&s.bvh_quad[q].q1), q2 = __ldg(&s.bvh_quad[q].q2);
                const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
                bool lin;
                const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                if (tt > RT_EPS && tt < dist) {
                    const int m = __float_as_int(q1.w), ord = __float_as_int(q0.w);
                    ooo = ooo || ord < max_ord; max_ord = max(max_ord, ord); ++n_occ;
                    const float4 m0 = __ldg(&s.mat[m].m0);
                    const float kt = __ldg(&s.mat[m].m2).x;
                    fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
                }
            }
            for (int q = 0; q < s.n_plane; ++q) {
                const float4 nc = __ldg(&s.plane[q].nc);
                const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
                if (tt > RT_EPS && tt < dist) {
                    const int m = s.plane[q].material, ord = s.plane[q].order;
                    ooo = ooo || ord < max_ord; max_ord = max(max_ord, ord); ++n_occ;
                    const float4 m0 = __ldg(&s.mat[m].m0);
                    const float kt = __ldg(&s.mat[m].m2).x;
                    fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
                }
            }
        }
        p.filter[i] = make_float4(fr, fg, fb, 0.f);
        if (ooo && n_occ >= 3 && !(fr == 0.f && fg == 0.f && fb == 0.f) && p.fix_count) {
            const unsigned int k = atomicAdd(p.fix_count, 1u);
            if ((int)k < p.fix_cap) p.fix_list[k] = i;
            atomicAdd(&p.stats[(size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT + ST_SHADOW_FIXUP], 1ull);
        }
    }
    n_nodes = __reduce_add_sync(FULL_MASK, n_nodes);
    n_tests = __reduce_add_sync(FULL_MASK, n_tests);
    n_side_q = __reduce_add_sync(FULL_MASK, n_side_q);
    n_side_p = __reduce_add_sync(FULL_MASK, n_side_p);
    overflow = __reduce_or_sync(FULL_MASK, overflow);
    if (lane == 0 && p.stats) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (n_nodes) atomicAdd(&st[ST_BVH_NODES_SHADOW], (unsigned long long)n_nodes);
        if (n_tests) atomicAdd(&st[ST_TEST_TRI_SHADOW], (unsigned long long)n_tests);
        if (n_side_q) atomicAdd(&st[ST_TEST_QUAD], (unsigned long long)n_side_q);
        if (n_side_p) atomicAdd(&st[ST_TEST_PLANE], (unsigned long long)n_side_p);
        if (overflow) atomicAdd(&st[ST_BVH_OVERFLOW], 1ull);
    }
}

// ---------------------------------------------------------------------------------------
// Far rays (see bvh_reach): few, so one plain kernel per kind -- the same leaf tests and winner rule, every box grown
// by what this ray's own origin needs.  Shadow: the occluders are multiplied as they are met; three or more go to the
// fix-up list like everywhere else (k_shadow_fixup redoes the product in insertion order).
// ---------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128) k_bvh_far_nearest(const __grid_constant__ TraceParams p)
{
    const DeviceScene& s = p.s;
    const int n = min((int)*p.far_count, p.far_cap);
    unsigned n_nodes = 0, overflow = 0, n_far = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < n; w += gridDim.x * blockDim.x) {
        const int slot = p.far_list[w];
        RT_CHECK(slot >= 0 && slot < p.n_rays);
        const V3 o = f4_xyz(p.in_od0[slot]), d = f4_xyz(p.in_od1[slot]);
        float best_t = -1.0f;
        int best_id = -1;
        ++n_far;
        const TriRec* __restrict__ tri = s.bvh_tri;
        bvh_traverse<true, true>(s, o, d, 3.0e38f, [&](int first, int count, float limit) -> float {
            for (int k = 0; k < count; ++k) {
                const float4 ra = __ldg(&tri[first + k].a), rb = __ldg(&tri[first + k].b), rc = __ldg(&tri[first + k].c);
                float tt;
                if (rec_is_quadric(rc)) {
                    const Quadric Q = { ra.x, ra.y, ra.z, 0.f, 0.f, 0.f, rb.x, rb.y, rb.z, rc.x };
                    bool lin;
                    tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                } else
                    tt = triangle_leaf<MODE>(o, d, f4_xyz(ra), f4_xyz(rb), f4_xyz(rc), best_t < 0.f ? 3.0e38f : fmaf(best_t, CULL_TREL, CULL_TABS));
                const int ord = __float_as_int(ra.w);
                if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; limit = tt; }
            }
            return limit;
        }, n_nodes, overflow, bvh_far_grow(s, bvh_reach(s, o)));
        for (int q = 0; q < s.bvh_n_quad; ++q) {
            const float4 q0 = __ldg(&s.bvh_quad[q].q0), q1 = __ldg(&s.bvh_quad[q].q1), q2 = __ldg(&s.bvh_quad[q].q2);
            const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
            const int ord = __float_as_int(q0.w);
            bool lin;
            const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
            if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
        }
        for (int q = 0; q < s.n_plane; ++q) {
            const float4 nc = __ldg(&s.plane[q].nc);
            const int ord = s.plane[q].order;
            const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
            if (tt > RT_EPS && (tt < best_t || best_t < 0.f || (tt == best_t && ord < best_id))) { best_t = tt; best_id = ord; }
        }
        p.hit[slot] = make_float2(best_t, __int_as_float(best_id));
    }
    if (p.stats && (n_far | overflow)) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (n_far) atomicAdd(&st[ST_BVH_OUT_OF_RADIUS], (unsigned long long)n_far);
        if (n_nodes) atomicAdd(&st[ST_BVH_NODES], (unsigned long long)n_nodes);
        if (overflow) atomicAdd(&st[ST_BVH_OVERFLOW], 1ull);
    }
}

template <int MODE>
__global__ void __launch_bounds__(128) k_bvh_far_shadow(const __grid_constant__ ShadowParams p)
{
    const DeviceScene& s = p.s;
    const int n = min((int)*p.far_count, p.far_cap);
    unsigned n_nodes = 0, overflow = 0, n_far = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < n; w += gridDim.x * blockDim.x) {
        const int i = p.far_list[w];
        V3 o, d;
        float dist;
        shadow_query<(MODE != 2)>(p, i, o, d, dist);
        const float tlim = fmaf(dist, CULL_TREL, CULL_TABS);
        float fr = 1.0f, fg = 1.0f, fb = 1.0f;
        int n_occ = 0;
        ++n_far;
        auto occluder = [&](int m) {
            const float4 m0 = __ldg(&s.mat[m].m0);
            const float kt = __ldg(&s.mat[m].m2).x;
            fr = __fmul_rn(fr, __fmul_rn(m0.x, kt)); fg = __fmul_rn(fg, __fmul_rn(m0.y, kt)); fb = __fmul_rn(fb, __fmul_rn(m0.z, kt));
            ++n_occ;
        };
        const TriRec* __restrict__ tri = s.bvh_tri;
        bvh_traverse<false, true>(s, o, d, dist, [&](int first, int count, float limit) -> float {
            for (int k = 0; k < count; ++k) {
                const float4 ra = __ldg(&tri[first + k].a), rb = __ldg(&tri[first + k].b), rc = __ldg(&tri[first + k].c);
                float tt;
                if (rec_is_quadric(rc)) {
                    const Quadric Q = { ra.x, ra.y, ra.z, 0.f, 0.f, 0.f, rb.x, rb.y, rb.z, rc.x };
                    bool lin;
                    tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
                } else tt = triangle_leaf<MODE>(o, d, f4_xyz(ra), f4_xyz(rb), f4_xyz(rc), tlim);
                if (tt > RT_EPS && tt < dist) occluder(__float_as_int(rb.w));
            }
            return limit;
        }, n_nodes, overflow, bvh_far_grow(s, bvh_reach(s, o)));
        for (int q = 0; q < s.bvh_n_quad; ++q) {
            const float4 q0 = __ldg(&s.bvh_quad[q].q0), q1 = __ldg(&s.bvh_quad[q].q1), q2 = __ldg(&s.bvh_quad[q].q2);
            const Quadric Q = { q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q2.w };
            bool lin;
            const float tt = MODE == 2 ? quadric_t<false>(o, d, Q, lin) : quadric_t<true>(o, d, Q, lin);
            if (tt > RT_EPS && tt < dist) occluder(__float_as_int(q1.w));
        }
        for (int q = 0; q < s.n_plane; ++q) {
            const float4 nc = __ldg(&s.plane[q].nc);
            const float tt = MODE == 2 ? plane_t<false>(o, d, f4_xyz(nc), nc.w) : plane_t<true>(o, d, f4_xyz(nc), nc.w);
            if (tt > RT_EPS && tt < dist) occluder(s.plane[q].material);
        }
        p.filter[i] = make_float4(fr, fg, fb, 0.f);
        if (n_occ >= 3 && !(fr == 0.f && fg == 0.f && fb == 0.f) && p.fix_count) {
            const unsigned int k = atomicAdd(p.fix_count, 1u);
            if ((int)k < p.fix_cap) p.fix_list[k] = i;
            if (p.stats) atomicAdd(&p.stats[(size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT + ST_SHADOW_FIXUP], 1ull);
        }
    }
    if (p.stats && (n_far | overflow)) {
        unsigned long long* st = p.stats + (size_t)(blockIdx.x % RT_STAT_SLOTS) * ST_COUNT;
        if (n_far) atomicAdd(&st[ST_BVH_OUT_OF_RADIUS], (unsigned long long)n_far);
        if (n_nodes) atomicAdd(&st[ST_BVH_NODES_SHADOW], (unsigned long long)n_nodes);
        if (overflow) atomicAdd(&st[ST_BVH_OVERFLOW], 1ull);
    }
}

cudaError_t launch_bvh_far_nearest(const TraceParams& p, int sm_count, cudaStream_t st)
{
    if (!p.far_count || p.primary || p.s.bvh_n_prims <= 0) return cudaSuccess;
    if (p.mode == 2)      k_bvh_far_nearest<2><<<sm_count * 2, 128, 0, st>>>(p);
    else if (p.mode == 1) k_bvh_far_nearest<1><<<sm_count * 2, 128, 0, st>>>(p);
    else                  k_bvh_far_nearest<0><<<sm_count * 2, 128, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_bvh_far_shadow(const ShadowParams& p, int sm_count, cudaStream_t st)
{
    if (!p.far_count || p.s.bvh_n_prims <= 0) return cudaSuccess;
    if (p.mode == 2)      k_bvh_far_shadow<2><<<sm_count * 2, 128, 0, st>>>(p);
    else if (p.mode == 1) k_bvh_far_shadow<1><<<sm_count * 2, 128, 0, st>>>(p);
    else                  k_bvh_far_shadow<0><<<sm_count * 2, 128, 0, st>>>(p);
    return cudaGetLastError();
}

template <bool QUADS>
static void launch_nearest_t(const TraceParams& p, int blocks, cudaStream_t st)
{
    if (p.primary) {
        if (p.mode == 2)      k_bvh_nearest<2, true, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
        else if (p.mode == 1) k_bvh_nearest<1, true, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
        else                  k_bvh_nearest<0, true, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
    } else {
        if (p.mode == 2)      k_bvh_nearest<2, false, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
        else if (p.mode == 1) k_bvh_nearest<1, false, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
        else                  k_bvh_nearest<0, false, QUADS><<<blocks, BVH_THREADS, 0, st>>>(p);
    }
}

cudaError_t launch_bvh_nearest(const TraceParams& p, cudaStream_t st)
{
    int blocks;
    if (p.primary) blocks = ((p.f.LW + 15) / 16) * ((p.chunk_rows + 7) / 8);
    else blocks = (p.n_rays + BVH_THREADS - 1) / BVH_THREADS;
    if (blocks <= 0) return cudaSuccess;
    if (p.s.bvh_n_prims > p.s.n_tri) launch_nearest_t<true>(p, blocks, st);     // triangle-only trees keep the lean leaf code
    else                             launch_nearest_t<false>(p, blocks, st);
    return cudaGetLastError();
}

cudaError_t launch_bvh_shadow(const ShadowParams& p, cudaStream_t st)
{
    const int blocks = (p.n_queries + BVH_THREADS - 1) / BVH_THREADS;
    if (blocks <= 0) return cudaSuccess;
    if (p.s.bvh_n_prims > p.s.n_tri) {
        if (p.mode == 2)      k_bvh_shadow<2, true><<<blocks, BVH_THREADS, 0, st>>>(p);
        else if (p.mode == 1) k_bvh_shadow<1, true><<<blocks, BVH_THREADS, 0, st>>>(p);
        else                  k_bvh_shadow<0, true><<<blocks, BVH_THREADS, 0, st>>>(p);
    } else {
        if (p.mode == 2)      k_bvh_shadow<2, false><<<blocks, BVH_THREADS, 0, st>>>(p);
        else if (p.mode == 1) k_bvh_shadow<1, false><<<blocks, BVH_THREADS, 0, st>>>(p);
        else                  k_bvh_shadow<0, false><<<blocks, BVH_THREADS, 0, st>>>(p);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
#define BCU(call)                                                       \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return set_cuda_error(c, e__, #call);   \
    } while (0)

void bvh_release(rtb200_ctx* c)
{
    if (!c->bvh_state) return;
    BvhState* b = static_cast<BvhState*>(c->bvh_state);
    b->nodes.release(); b->tri_sorted.release(); b->keys.release(); b->keys_out.release(); b->idx.release(); b->idx_out.release();
    b->cub_tmp.release(); b->tree_quad.release(); b->side_quad.release(); b->parent.release(); b->leaf_parent.release(); b->gamma.release(); b->flags.release(); b->bounds.release();
    delete b;
    c->bvh_state = nullptr;
}

void bvh_invalidate(rtb200_ctx* c)
{
    if (c->bvh_state) static_cast<BvhState*>(c->bvh_state)->built = false;
    c->ds.bvh_nodes = nullptr; c->ds.bvh_tri = nullptr; c->ds.bvh_root = 0;
    c->ds.bvh_n_prims = 0; c->ds.bvh_quad = nullptr; c->ds.bvh_n_quad = 0;
}

bool bvh_ready(const rtb200_ctx* c)
{
    return c->bvh_state && static_cast<const BvhState*>(c->bvh_state)->built;
}

static int bvh_refit(rtb200_ctx* c, BvhState* b, float radius)
{
    const int n = b->n_prims;
    b->radius = radius;
    for (int k = 0; k < 3; ++k) { c->ds.bvh_lo[k] = b->lo[k]; c->ds.bvh_hi[k] = b->hi[k]; }
    c->ds.bvh_radius = radius;
    c->ds.bvh_grow_tri = std::max(b->padc * 5.9604645e-8f / RT_EPS * b->wmax, 32.0f * 1.1920929e-7f) * 1.01f;
    c->ds.bvh_kq = b->kq; c->ds.bvh_cmax = b->cmax;
    if (n <= BVH_LEAF_MAX) return RTB200_OK;                  // a single leaf: no boxes at all
    // pad = PADC * 2^-24 / EPSILON * R * |e1| |e2|, floor 32 ulp of R (rounding of the slab test)
    const float pad_scale = b->padc * 5.9604645e-8f / RT_EPS * radius;
    const float pad_floor = 32.0f * 1.1920929e-7f * radius;
    BCU(cudaMemsetAsync(b->flags.p, 0, sizeof(int) * (size_t)(n - 1), c->stream));
    k_bvh_refit<<<(n + 255) / 256, 256, 0, c->stream>>>(b->tri_sorted.as<TriRec>(), n, b->nodes.as<BvhNode>(), b->parent.as<int>(),
                                                        b->leaf_parent.as<int>(), b->gamma.as<int>(), b->flags.as<int>(),
                                                        pad_scale, pad_floor, radius);
    BCU(cudaGetLastError());
    return RTB200_OK;
}

// Which quadrics become primitives of the tree: centred form known (rtb200_set_scene: r3.w = rhs > 0, no mixed
// terms), every quadratic coefficient positive (a cylinder has one zero: unbounded) and mixed terms that are +0
// bit for bit (the leaf test passes literal zeros for them).
static bool quadric_is_bounded(const QuadRec& q, const float4& r3)
{
    if (!(r3.w > 0.f) || !(q.q0.x > 0.f && q.q0.y > 0.f && q.q0.z > 0.f)) return false;
    const float mixed[3] = {q.q1.x, q.q1.y, q.q1.z};
    unsigned int bits[3];
    memcpy(bits, mixed, sizeof bits);
    if (bits[0] | bits[1] | bits[2]) return false;
    const float ratio = std::max(q.q0.x, std::max(q.q0.y, q.q0.z)) / std::min(q.q0.x, std::min(q.q0.y, q.q0.z));
    return std::isfinite(ratio) && ratio < 1.0e6f && std::isfinite(r3.w);
}

int bvh_build(rtb200_ctx* c)
{
    BvhState* b = bvh_state(c);
    b->built = false;
    const int T = c->ds.n_tri;
    if (const char* e = getenv("RTB200_BVH_PADC")) { const float v = (float)atof(e); if (v >= 0.f) b->padc = v; }
    cudaError_t e;

    // ---- quadrics: bounded ones join the tree, the others stay in the linear side list -----------------------
    std::vector<TriRec> tree_q;
    std::vector<QuadRec> side_q;
    float kq = 0.f, cmax = 0.f;
    const bool quads_in_tree = getenv("RTB200_BVH_NO_QUADRICS") == nullptr;
    for (size_t i = 0; i < c->h_quad.size(); ++i) {
        const QuadRec& q = c->h_quad[i];
        const float4 r3 = i < c->h_quad_r3.size() ? c->h_quad_r3[i] : make_float4(0.f, 0.f, 0.f, -1.f);
        if (quads_in_tree && quadric_is_bounded(q, r3)) {
            TriRec r;
            r.a = q.q0;                                                   // A B C | order
            r.b = make_float4(q.q2.x, q.q2.y, q.q2.z, q.q1.w);            // lx ly lz | material
            r.c = make_float4(q.q2.w, r3.w, 0.f, 1.0f);                   // K, rhs, -, tag
            tree_q.push_back(r);
            const float qmin = std::min(q.q0.x, std::min(q.q0.y, q.q0.z));
            kq = std::max(kq, (q.q0.x + q.q0.y + q.q0.z) / qmin);
            for (int k = 0; k < 3; ++k) cmax = std::max(cmax, std::fabs((&q.q2.x)[k] / (2.0f * (&q.q0.x)[k])));
        } else side_q.push_back(q);
    }
    b->kq = kq; b->cmax = cmax;
    const int Qb = (int)tree_q.size(), Qs = (int)side_q.size();
    const long long n_ll = (long long)T + Qb;
    if (n_ll >= (1ll << BVH_LEAF_SHIFT)) { c->last_error = "too many primitives for the LBVH leaf encoding"; return RTB200_E_ARG; }
    const int n = (int)n_ll;
    if ((e = b->tree_quad.ensure(sizeof(TriRec) * (size_t)std::max(Qb, 1))) != cudaSuccess ||
        (e = b->side_quad.ensure(sizeof(QuadRec) * (size_t)std::max(Qs, 1))) != cudaSuccess) return set_cuda_error(c, e, "bvh quadric buffers");
    // (pageable sources: the copies are staged before the calls return)
    if (Qb) BCU(cudaMemcpyAsync(b->tree_quad.p, tree_q.data(), sizeof(TriRec) * (size_t)Qb, cudaMemcpyHostToDevice, c->stream));
    if (Qs) BCU(cudaMemcpyAsync(b->side_quad.p, side_q.data(), sizeof(QuadRec) * (size_t)Qs, cudaMemcpyHostToDevice, c->stream));
    const TriRec* tq = b->tree_quad.as<TriRec>();

    b->n_tri = T; b->n_prims = n; b->n_tree_quad = Qb; b->n_side_quad = Qs;
    b->n_nodes = n > BVH_LEAF_MAX ? n - 1 : 0;
    b->root = n > BVH_LEAF_MAX ? 0 : ~(0 | ((std::max(n, 1) - 1) << BVH_LEAF_SHIFT));
    cudaEvent_t ev0 = c->ev[2], ev1 = c->ev[3];
    BCU(cudaEventRecord(ev0, c->stream));
    if (n > 0) {
        if ((e = b->tri_sorted.ensure(sizeof(TriRec) * (size_t)n)) != cudaSuccess || (e = b->bounds.ensure(sizeof(int) * 8)) != cudaSuccess)
            return set_cuda_error(c, e, "bvh buffers");
        int h_init[7] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN, 0};
        BCU(cudaMemcpyAsync(b->bounds.p, h_init, sizeof h_init, cudaMemcpyHostToDevice, c->stream));
        k_bvh_bounds<<<std::min((n + 255) / 256, 4 * c->sm_count), 256, 0, c->stream>>>(c->ds.tri, T, tq, n, b->bounds.as<int>());
        BCU(cudaGetLastError());
        int h_b[7];
        BCU(cudaMemcpyAsync(h_b, b->bounds.p, sizeof h_b, cudaMemcpyDeviceToHost, c->stream));
        BCU(cudaStreamSynchronize(c->stream));
        for (int k = 0; k < 3; ++k) { b->lo[k] = order_float(h_b[k]); b->hi[k] = order_float(h_b[3 + k]); }
        memcpy(&b->wmax, &h_b[6], 4);
    } else {
        for (int k = 0; k < 3; ++k) b->lo[k] = b->hi[k] = 0.f;
        b->wmax = 0.f;
    }
    if (n > BVH_LEAF_MAX) {
        const size_t nn = (size_t)n;
        if ((e = b->nodes.ensure(sizeof(BvhNode) * (nn - 1))) != cudaSuccess ||
            (e = b->keys.ensure(8 * nn)) != cudaSuccess || (e = b->keys_out.ensure(8 * nn)) != cudaSuccess ||
            (e = b->idx.ensure(4 * nn)) != cudaSuccess || (e = b->idx_out.ensure(4 * nn)) != cudaSuccess ||
            (e = b->parent.ensure(4 * nn)) != cudaSuccess || (e = b->leaf_parent.ensure(4 * nn)) != cudaSuccess ||
            (e = b->gamma.ensure(4 * nn)) != cudaSuccess || (e = b->flags.ensure(4 * nn)) != cudaSuccess)
            return set_cuda_error(c, e, "bvh buffers");
        // Morton cells are CUBES (one scale for the three axes, the largest extent): with a scale per axis a
        // flat scene (a terrain: 2000 x 120 x 2000) is cut into horizontal sheets at every third level, and
        // rays cross those sheets all the time (config 4: 27 -> 18 nodes per ray with cubes).
        // RTB200_BVH_ANISO=1 restores the per-axis scale for A/B measurements.
        Bounds6 bb;
        const bool aniso = getenv("RTB200_BVH_ANISO") != nullptr;
        const float ext_max = std::max(b->hi[0] - b->lo[0], std::max(b->hi[1] - b->lo[1], b->hi[2] - b->lo[2]));
        for (int k = 0; k < 3; ++k) {
            bb.lo[k] = b->lo[k];
            const float ext = aniso ? b->hi[k] - b->lo[k] : ext_max;
            bb.inv[k] = ext > 0.f && std::isfinite(ext) ? 1.0f / ext : 0.0f;
        }
        const int blocks = (n + 255) / 256;
        k_bvh_morton<<<blocks, 256, 0, c->stream>>>(c->ds.tri, T, tq, n, bb, b->keys.as<unsigned long long>(), b->idx.as<unsigned int>());
        BCU(cudaGetLastError());
        size_t tmp_bytes = 0;
        BCU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, b->keys.as<unsigned long long>(), b->keys_out.as<unsigned long long>(),
                                            b->idx.as<unsigned int>(), b->idx_out.as<unsigned int>(), n, 0, 63, c->stream));
        if ((e = b->cub_tmp.ensure(tmp_bytes)) != cudaSuccess) return set_cuda_error(c, e, "sort scratch");
        BCU(cub::DeviceRadixSort::SortPairs(b->cub_tmp.p, tmp_bytes, b->keys.as<unsigned long long>(), b->keys_out.as<unsigned long long>(),
                                            b->idx.as<unsigned int>(), b->idx_out.as<unsigned int>(), n, 0, 63, c->stream));
        k_bvh_gather<<<blocks, 256, 0, c->stream>>>(c->ds.tri, T, tq, b->idx_out.as<unsigned int>(), n, b->tri_sorted.as<TriRec>());
        BCU(cudaGetLastError());
        k_bvh_hierarchy<<<blocks, 256, 0, c->stream>>>(b->keys_out.as<unsigned long long>(), n, b->nodes.as<BvhNode>(), b->parent.as<int>(),
                                                       b->leaf_parent.as<int>(), b->gamma.as<int>());
        BCU(cudaGetLastError());
    } else if (n > 0) {
        if (T) BCU(cudaMemcpyAsync(b->tri_sorted.p, c->ds.tri, sizeof(TriRec) * (size_t)T, cudaMemcpyDeviceToDevice, c->stream));
        if (Qb) BCU(cudaMemcpyAsync(b->tri_sorted.as<TriRec>() + T, tq, sizeof(TriRec) * (size_t)Qb, cudaMemcpyDeviceToDevice, c->stream));
    }
    // default radius: origins anywhere inside the primitives' bounds (secondary and shadow rays start
    // on surfaces); a camera farther out triggers one refit (bvh_prepare)
    const float dx = b->hi[0] - b->lo[0], dy = b->hi[1] - b->lo[1], dz = b->hi[2] - b->lo[2];
    const float diag = n > 0 ? sqrtf(dx * dx + dy * dy + dz * dz) : 0.f;
    int rc = bvh_refit(c, b, diag);
    if (rc != RTB200_OK) return rc;
    BCU(cudaEventRecord(ev1, c->stream));
    BCU(cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    BCU(cudaEventElapsedTime(&ms, ev0, ev1));
    b->build_ms = ms;
    b->built = true;
    c->ds.bvh_nodes = b->nodes.as<BvhNode>();
    c->ds.bvh_tri = b->tri_sorted.as<TriRec>();
    c->ds.bvh_root = b->root;
    c->ds.bvh_n_prims = n;
    c->ds.bvh_quad = b->side_quad.as<QuadRec>();
    c->ds.bvh_n_quad = Qs;
    return RTB200_OK;
}

// Make the padded boxes valid for rays starting at the given points (host array of n xyz):
// R must bound |origin - p0| for every triangle.  Refits (no rebuild) when the current pad is
// too small.
int bvh_prepare(rtb200_ctx* c, const float* origins3, int n)
{
    if (!bvh_ready(c)) { c->last_error = "RTB200_F_USE_BVH without rtb200_build_bvh"; return RTB200_E_STATE; }
    BvhState* b = static_cast<BvhState*>(c->bvh_state);
    if (b->n_prims == 0) return RTB200_OK;
    float need = 0.f;
    for (int i = 0; i < n; ++i) {
        float d2 = 0.f;
        for (int k = 0; k < 3; ++k) {
            const float o = origins3[3 * (size_t)i + k];
            const float far = std::max(std::fabs(o - b->lo[k]), std::fabs(o - b->hi[k]));
            d2 += far * far;
        }
        if (std::isfinite(d2)) need = std::max(need, std::sqrt(d2));
    }
    if (need <= b->radius) return RTB200_OK;
    cudaEvent_t ev0 = c->ev[2], ev1 = c->ev[3];
    BCU(cudaEventRecord(ev0, c->stream));
    int rc = bvh_refit(c, b, need * 1.1f);
    if (rc != RTB200_OK) return rc;
    BCU(cudaEventRecord(ev1, c->stream));
    BCU(cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    BCU(cudaEventElapsedTime(&ms, ev0, ev1));
    b->refit_ms = ms; ++b->refits;
    return RTB200_OK;
}

} // namespace rtb

using namespace rtb;

extern "C" int rtb200_build_bvh(rtb200_ctx* ctx)
{
    if (!ctx) return RTB200_E_ARG;
    if (!ctx->have_scene) { ctx->last_error = "rtb200_build_bvh before rtb200_set_scene"; return RTB200_E_STATE; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) return set_cuda_error(ctx, cudaGetLastError(), "cudaSetDevice");
    return bvh_build(ctx);
}

extern "C" int rtb200_bvh_info(rtb200_ctx* ctx, int* n_nodes, int* n_triangles, double* build_ms, double* pad_radius, double* node_bytes)
{
    if (!ctx) return RTB200_E_ARG;
    if (!bvh_ready(ctx)) { ctx->last_error = "no LBVH built"; return RTB200_E_STATE; }
    const BvhState* b = static_cast<const BvhState*>(ctx->bvh_state);
    if (n_nodes) *n_nodes = b->n_nodes;
    if (n_triangles) *n_triangles = b->n_tri;
    if (build_ms) *build_ms = b->build_ms;
    if (pad_radius) *pad_radius = b->radius;
    if (node_bytes) *node_bytes = (double)sizeof(BvhNode);
    return RTB200_OK;
}
