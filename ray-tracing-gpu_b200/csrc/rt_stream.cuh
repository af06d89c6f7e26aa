// rt_stream.cuh -- parameter blocks of the streamed (tiled brute-force) wavefront kernels.
#pragma once
#include "rt_kernels.cuh"

namespace rtb {

#define ST_THREADS   256           // threads per CTA
#define ST_RPT       4             // rays per thread (ILP + amortised shared-memory reads)
#define ST_BLOCK_RAYS (ST_THREADS * ST_RPT)
#define ST_TILE      256           // triangles per shared-memory tile (12 KB)
#define ST_STAGES    2
#define ST_SIDE_MAX  8192          // bytes of quadric + plane records kept in shared memory (170 quadrics)

// nearest hit of every ray of one wavefront level against ALL surfaces (tiled brute force)
struct TraceParams {
    FrameParams f;
    DeviceScene s;
    int   primary;                 // 1: rays generated from the sample lattice (level 0)
    int   chunk_ly0, chunk_rows, local_LH;
    const float4* in_od0;          // level >= 1 / query rays: o.xyz | energy
    const float4* in_od1;          //                          d.xyz | ior
    int   n_rays;
    float2* hit;                   // t | order (int bits, -1 = miss), indexed like the rays
    int   mode;                    // 0: FMA culling filter + reference-rounded survivors (default)
                                   // 1: no filter, reference sequence for every test (EXACT_FP)
                                   // 2: filter + fused survivors (FUSED_FP)
    unsigned long long* stats;     // LBVH kernels: nodes visited / triangle tests (may be null)
    int   cull_bundles;            // (culled scan only)
    int*          far_list;        // LBVH, secondary rays: rays that start beyond the padded radius are listed here
    unsigned int* far_count;       // (null: every ray takes the normal walk) and traced by k_bvh_far_nearest
    int           far_cap;
};

// shade-A: hit point, normal, shadow queries of the facing lights, children
struct ShadeAParams {
    FrameParams f;
    DeviceScene s;
    int   level;
    int   chunk_ly0, chunk_rows, local_LH;
    const float4* in_od0;
    const float4* in_od1;
    int   n_rays;
    const float2* hit;
    float4* hitP;                  // P.xyz | t
    float4* hitN;                  // N.xyz | order (int bits, -1 = miss)
    int*    q_base;                // first shadow query of the ray
    int2*   sq_src;                // shadow query = (ray of this level, light): the shadow kernels derive P, dist
                                   // and direction from hitP and the light (8 B per query instead of 32)
    unsigned int* sq_count;
    float4* out_od0;
    float4* out_od1;
    unsigned int* out_count;
    int4*   links;
    unsigned int* link_count;
    unsigned long long* stats;
    int   exact;                   // rounding of everything that is kept (1 unless FUSED_FP)
    int   tile_rays;               // level 0: a CTA takes 16x16 lattice samples instead of 256 consecutive ones
    // culled scan, levels >= 1 (rt_stream.cu): sort keys.  Rays that start close to each other and shadow queries
    // that go from neighbouring points to the same light are compact bundles whatever their queue positions are;
    // the host sorts (key, index) pairs and hands the trace kernels the permutation.  All four may be null.
    unsigned int* child_key;       // per child slot: 30-bit Morton code of its origin (cells of key_inv^-1 from key_lo)
    int*          child_idx;       // per child slot: the slot itself (values of the pair sort)
    unsigned int* sq_key;          // per shadow query: light << 27 | 27-bit Morton code of the hit point
    int*          sq_idx;
    float         key_lo[3], key_inv;
    int           mark_far;        // LBVH frames: set SQ_FAR_BIT on the queries of hit points beyond the padded radius
    int*  sq_dst;                  // null: queries in (ray, light) order, filter slot = query slot.  Otherwise the
                                   // CTA's queries are queued light by light (<= 8 lights) and sq_dst[query] is the
                                   // filter slot shade-B reads (q_base[ray] + rank of the light among the facing ones)
};

struct ShadowParams {
    DeviceScene s;
    const float4* sq0;             // explicit queries (rtb200_trace_shadow): P.xyz | dist
    const float4* sq1;             //                                         L.xyz (normalised) | 0
    const int2*   sq_src;          // or, inside a frame: (ray, light) pairs over hitP (shade-A), sq0 / sq1 unused
    const int*    sq_dst;          // filter slot of every query (null: the query's own index)
    const float4* hitP;
    int   by_light;                // LBVH: lanes of a warp share the light rather than the hit (see k_bvh_shadow)
    int   n_queries;
    float4* filter;                // r g b | 0 per query
    unsigned long long* stats;
    int   mode;
    int   cull_bundles;            // (culled scan only)
    int*          fix_list;        // queries whose filter product has to be redone in insertion order (rare) ...
    unsigned int* fix_count;       // ... how many (null: nothing is recorded), and the room there is
    int           fix_cap;
    int*          far_list;        // LBVH: queries whose origin lies beyond the padded radius (see TraceParams::far_list)
    unsigned int* far_count;
    int           far_cap;
};

// shade-B: ambient + per-light diffuse/Phong with the shadow filters, in light order
struct ShadeBParams {
    FrameParams f;
    DeviceScene s;
    int   level;
    int   chunk_ly0, chunk_rows, local_LH;
    const float4* in_od1;          // ray directions (level >= 1)
    int   n_rays;
    const float4* hitP;
    const float4* hitN;
    const int*    q_base;
    const float4* filter;
    float*  c0;                    // level 0 colours (3 floats per sample)
    float4* colour;                // level >= 1 colours
    int   exact;
};

// One shadow query: origin, unit direction towards the light, distance to it.  The derived form repeats
// the arithmetic shade-A / shade-B use for the same pair (Scene.cpp:1847-1848), in the same rounding mode.
#define SQ_FAR_BIT 0x40000000      // in sq_src[].y next to the light index: the hit point lies beyond the radius the LBVH boxes
                                   // are padded for (shade-A knows the point; the shadow kernel only tests the bit)
template <bool E>
RT_DEV bool shadow_query(const ShadowParams& p, int i, V3& o, V3& d, float& dist)
{
    if (p.sq_src) {
        int2 src = p.sq_src[i];
        const bool far = (src.y & SQ_FAR_BIT) != 0;
        src.y &= ~SQ_FAR_BIT;
        o = f4_xyz(p.hitP[src.x]);
        V3 Ld = sub3<E>(f4_xyz(p.s.light[src.y].l0), o);
        dist = norm3<E>(Ld);
        d = scale3<E>(Ld, FP<E>::rcp(dist));
        return far;
    }
    const float4 a = p.sq0[i], b = p.sq1[i];
    o = f4_xyz(a); d = f4_xyz(b); dist = a.w;
    return false;
}

// LBVH: distance from a ray origin to the farthest corner of the primitives' bounds -- what the padded boxes must cover
RT_DEV float bvh_reach(const DeviceScene& s, V3 o)
{
    const float fx = fmaxf(fabsf(o.x - s.bvh_lo[0]), fabsf(o.x - s.bvh_hi[0]));
    const float fy = fmaxf(fabsf(o.y - s.bvh_lo[1]), fabsf(o.y - s.bvh_hi[1]));
    const float fz = fmaxf(fabsf(o.z - s.bvh_lo[2]), fabsf(o.z - s.bvh_hi[2]));
    return sqrtf(fmaf(fx, fx, fmaf(fy, fy, fz * fz))) * 1.00001f;
}
RT_DEV bool bvh_is_far(const DeviceScene& s, V3 o) { return !(bvh_reach(s, o) <= s.bvh_radius); }

RT_DEV unsigned int expand10(unsigned int v)         // bit i of a 10-bit value -> bit 3 i
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// 30-bit Morton code of a point in cubic cells (clamped to the 1024^3 grid; NaN -> cell 0)
RT_DEV unsigned int morton30(float x, float y, float z, float lx, float ly, float lz, float inv)
{
    const float fx = fminf(fmaxf((x - lx) * inv, 0.f), 1023.f), fy = fminf(fmaxf((y - ly) * inv, 0.f), 1023.f), fz = fminf(fmaxf((z - lz) * inv, 0.f), 1023.f);
    const unsigned int ix = fx == fx ? (unsigned int)fx : 0u, iy = fy == fy ? (unsigned int)fy : 0u, iz = fz == fz ? (unsigned int)fz : 0u;
    return (expand10(ix) << 2) | (expand10(iy) << 1) | expand10(iz);
}

cudaError_t launch_trace_nearest(const TraceParams& p, int sm_count, cudaStream_t st);
cudaError_t launch_shade_a(const ShadeAParams& p, cudaStream_t st);
cudaError_t launch_trace_shadow(const ShadowParams& p, int sm_count, cudaStream_t st);
cudaError_t launch_shade_b(const ShadeBParams& p, cudaStream_t st);
cudaError_t launch_query_pack_rays(int n, const float* o3, const float* d3, float4* od0, float4* od1, cudaStream_t st);
cudaError_t launch_query_finish_nearest(const DeviceScene& s, int n, const float4* od0, const float4* od1, const float2* hit,
                                        int* out_order, float* out_t, float* out_n3, bool exact, cudaStream_t st);
cudaError_t launch_query_pack_shadow(int n, const float* p3, const float* l3, float4* sq0, float4* sq1, bool exact, cudaStream_t st);
cudaError_t launch_query_unpack_filter(int n, const float4* filter, float* out3, cudaStream_t st);

} // namespace rtb
