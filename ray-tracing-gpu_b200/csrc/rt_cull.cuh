// rt_cull.cuh -- the conservative two-phase triangle culling filter shared by the tiled
// brute-force kernels (rt_trace.cu) and the LBVH leaves (rt_bvh.cu).
//
// Phase 1 computes det and u in FUSED arithmetic with one MUFU.RCP (no division); phase 2
// finishes v and t.  Both reject only with margins 10^3 x the fused/unfused differences
// (DESIGN.md 5.2), and every ray that passes is re-evaluated with the reference's own rounding
// sequence (triangle_t<true>), whose result is the one that is kept.
#pragma once
#include "rt_math.cuh"

namespace rtb {

// culling margins (see DESIGN.md "conservative filter"): relative 1e-3, far above the
// ~1e-6 relative differences between fused and unfused evaluation
#define CULL_DET   0.0099f         // |det| >= EPSILON, with 1 % slack
#define CULL_UV    0.001f
#define CULL_TREL  1.001f
#define CULL_TABS  1.0e-3f

RT_DEV float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Phase-2 of the filter + exact evaluation for one (ray, triangle); returns the reference
// distance or -1.  `tmin`/`tmax` bound the distances that could still matter to the caller.
template <int MODE>
RT_DEV float triangle_survivor(V3 o, V3 d, V3 p0, V3 e1, V3 e2, float tmax)
{
    // fused re-evaluation of the whole test with margins
    const V3 P = cross3<false>(d, e2);
    const float det = dot3<false>(e1, P);
    const float rd = rcp_approx(det);
    const V3 S = sub3<false>(o, p0);
    const float u = dot3<false>(S, P) * rd;
    const V3 Q = cross3<false>(S, e1);
    const float v = dot3<false>(d, Q) * rd;
    const float t = dot3<false>(e2, Q) * rd;
    if (!(v >= -CULL_UV && u + v <= 1.0f + CULL_UV && t > RT_EPS * 0.99f && t < tmax)) return -1.0f;
    return MODE == 2 ? triangle_t<false>(o, d, p0, e1, e2) : triangle_t<true>(o, d, p0, e1, e2);
}


// Phase 1 for one (ray, triangle): can this ray still hit?  (det and u only)
RT_DEV bool triangle_candidate(V3 o, V3 d, V3 p0, V3 e1, V3 e2)
{
    const V3 P = cross3<false>(d, e2);
    const float det = dot3<false>(e1, P);
    const V3 S = sub3<false>(o, p0);
    const float u = dot3<false>(S, P) * rcp_approx(det);
    return fabsf(det) >= CULL_DET && fabsf(u - 0.5f) <= 0.5f + CULL_UV;
}

} // namespace rtb
