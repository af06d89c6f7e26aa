// rt_bvh.cuh -- GPU-built LBVH over the triangles of the scene (no
// counterpart in the reference, whose nearest hit is a linear scan, Scene.cpp:1710-1715).
// Contract: the SAME (t, order) winner and the SAME shadow filter as the brute-force kernels.
#pragma once
#include "rt_stream.cuh"

namespace rtb {

#ifndef BVH_LEAF_MAX
#define BVH_LEAF_MAX   6          // primitives per leaf (contiguous in Morton order); config 4 (cubic Morton cells): 9.63 / 9.53 / 9.69 ms with 4 / 6 / 8
#endif
#define BVH_STACK      64         // traversal stack entries per ray
#ifndef BVH_THREADS
#define BVH_THREADS    128
#endif
#define BVH_LEAF_SHIFT 28         // leaf reference = ~(first | (count - 1) << 28)

cudaError_t launch_bvh_nearest(const TraceParams& p, cudaStream_t st);
cudaError_t launch_bvh_shadow(const ShadowParams& p, cudaStream_t st);
// the rays / queries the launches above put on their far lists (origins beyond the padded radius): same answers as the
// linear scan for them too -- every box grown by the pad their own origin needs.  No-ops without a far list.
cudaError_t launch_bvh_far_nearest(const TraceParams& p, int sm_count, cudaStream_t st);
cudaError_t launch_bvh_far_shadow(const ShadowParams& p, int sm_count, cudaStream_t st);

} // namespace rtb
