// rt_cullscan.cuh -- the culled scan of the streamed path: every warp walks a 32-ary hierarchy of bounding spheres
// over the Morton-sorted triangles with the cone(s) of its 128 rays and runs the filter + reference sequence only on
// the triangles that survive (rt_cullscan.cu).  Same answers as the plain tiled scan of rt_trace.cu, bit for bit.
#pragma once
#include "rt_stream.cuh"

namespace rtb {

#define CS_LEVELS 4

// Level 0: one bounding sphere per triangle (Morton order); level l >= 1: one sphere per block of 32 level-(l-1)
// entries, 3 levels above the triangles.  n(l-1) == 32 n(l): the arrays are padded from the top with empty entries
// (w == 0, never kept), n(3) = n3 = ceil(T / 32768).  All levels live in ONE array: [level 3][level 2][level 1][level 0].
struct CullScene {
    const TriRec* tri;               // triangles in Morton order, 32768 * n3 records (padding: zero edges)
    const float4* sph;               // centre | radius
    const float*  w;                 // |e1| |e2| bound of the triangle / the largest below the block; 0 = empty
    int n3;
    int top_blocks;                  // rounds of 32 at level 3 (1 up to a million triangles)
    int n_tri;                       // real triangles
    float key_lo[3], key_inv;        // grid of the sort keys of secondary rays / queries (ShadeAParams::child_key)
};

struct CullTraceParams {
    TraceParams t;
    CullScene   cs;
    unsigned int* task_counter;      // zeroed before the launch: warps take tasks of 128 rays from it
    int           task_shift;        // log2 of the rays per task: 7 (four per lane), or 5 (one per lane) for levels with few
                                     // rays -- more tasks than warps, tighter cones, and a bundle nothing can cull costs 1/4
    const int*    perm;              // null, or: the i-th ray of the launch is queue entry perm[i] (rays sorted by the
                                     // place they start at, see ShadeAParams::child_key)
    int*          lane_list;         // tasks no cone fits: listed by the kernel proper, traced ray by ray (8 lanes each) by a second
    unsigned int* lane_count;        // launch (null: such tasks keep every triangle and scan them all)
    unsigned int* lane_task_counter;
    int           lane_cap;
    float         lane_sin;          // a bundle with a cone wider than this (sine of the half angle) also goes to the lanes
};

struct CullShadowParams {
    ShadowParams s;
    CullScene    cs;
    unsigned int* task_counter;
    int           task_shift;        // as in CullTraceParams
    const int*    perm;              // null, or: the i-th query of the launch is queue entry perm[i] (sorted by light, place)
    int*          lane_list;         // as in CullTraceParams
    unsigned int* lane_count;
    unsigned int* lane_task_counter;
    int           lane_cap;
    float         lane_sin;          // a bundle with a cone wider than this (sine of the half angle) also goes to the lanes
};                                   // (fix-up list: ShadowParams::fix_list / fix_count / fix_cap)

cudaError_t launch_cull_nearest(const CullTraceParams& p, int sm_count, cudaStream_t st);
cudaError_t launch_cull_shadow(const CullShadowParams& p, int sm_count, cudaStream_t st);
cudaError_t launch_shadow_fixup(const CullShadowParams& p, int sm_count, cudaStream_t st);
// stable sort of n (key, index) pairs on the low `bits` bits of the key (cub radix sort); *tmp grows as needed
struct DevBuf;
cudaError_t cull_sort_pairs(DevBuf* tmp, const unsigned int* keys_in, unsigned int* keys_out, const int* idx_in, int* idx_out,
                            int n, int bits, cudaStream_t st);

} // namespace rtb
