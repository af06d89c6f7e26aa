// rt_cull.cuh -- the conservative two-phase triangle culling filter shared by the tiled
// brute-force kernels (rt_trace.cu), the LBVH leaves (rt_bvh.cu) and the warp votes of the
// resident kernels (rt_resident.cu).
//
// The reference's test (CTriangle::Intersection, Triangle.cpp:127-172, restated as triangle_t<true>)
// rounds after every operation.  The filter evaluates the same quantities in FUSED arithmetic, without
// any division, and rejects a (ray, triangle) pair only where the reference cannot accept it: every
// comparison is made in NUMERATOR space against a margin that bounds the difference between the fused
// and the reference evaluation of that numerator (eps = 2^-24, a_k = max_i |e_k,i|, dm = max_i |d_i|,
// S1 = |S|_1):
//     |det_f - det_r|            <= 60 eps a1 a2 dm             (cross product 24 + two dot products 36)
//     |S.P_f - S.P_r|            <= 20 eps S1 a2 dm
//     |d.Q_f - d.Q_r|            <= 60 eps S1 a1 dm
//     |e2.Q_f - e2.Q_r|          <= 60 eps S1 a1 a2
// and the reference's own divisions add 2 eps relative.  With the constants below (10-30 % above the
// bounds) and the 1e-3 relative slack kept from the first version:
//   phase 1   |det_f| + E_det >= EPSILON (1 - 1 %)   and   |S.P_f - det_f / 2| <= (1/2 + 1e-3) |det_f| + E_num + E_det
//   phase 2   v, u + v, t > EPSILON and t < tmax the same way, once the sign of det is certain (|det_f| > 2 E_det)
// A fixed margin (the first version: u in [-1e-3, 1 + 1e-3]) is NOT conservative for a triangle seen edge-on:
// at |det| = 0.17, |S| = 956 the reference's u carries an error of 4 %.  tests/test_filter_margins.py
// restates both evaluations in numpy and checks 10^7 adversarial pairs: no accepted hit is rejected.
// Every ray that passes is re-evaluated with the reference's own rounding sequence, whose result is the
// one that is kept.
#pragma once
#include "rt_math.cuh"

namespace rtb {

#define CULL_DET   0.0099f         // |det| >= EPSILON, with 1 % slack
#define CULL_UV    0.001f          // relative slack on |det| in the numerator-space comparisons
#define CULL_TREL  1.001f          // slack of the distance bounds handed to the filter (tmax = t * TREL + TABS)
#define CULL_TABS  1.0e-3f
#define CULL_EPS24 5.9604645e-8f
#define CULL_KU    (26.0f * CULL_EPS24)     // S.P      (bound 20 eps)
#define CULL_KD    (70.0f * CULL_EPS24)     // det      (bound 60 eps)
#define CULL_KV    (66.0f * CULL_EPS24)     // d.Q      (bound 60 eps)
#define CULL_KT    (70.0f * CULL_EPS24)     // e2.Q     (bound 60 eps)

RT_DEV float max_abs3(V3 a) { return fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fabsf(a.z)); }

// margins of one triangle for rays whose direction components are bounded by dmax (shared by the rays of a thread)
struct TriMargins { float k_num, e_det, k_v, k_t; };

RT_DEV TriMargins tri_margins(V3 e1, V3 e2, float dmax)
{
    const float a1 = max_abs3(e1), a2 = max_abs3(e2);
    const float a2d = a2 * dmax;
    TriMargins m;
    m.k_num = CULL_KU * a2d;
    m.e_det = CULL_KD * a1 * a2d;
    m.k_v = CULL_KV * a1 * dmax;
    m.k_t = CULL_KT * a1 * a2;
    return m;
}

// Phase 1 for one (ray, triangle): can the reference still accept?  (det and u only)
RT_DEV bool triangle_phase1(V3 o, V3 d, V3 p0, V3 e1, V3 e2, const TriMargins& m)
{
    const V3 P = cross3<false>(d, e2);
    const float det = dot3<false>(e1, P);
    const V3 S = sub3<false>(o, p0);
    const float un = dot3<false>(S, P);
    const float s1 = fabsf(S.x) + fabsf(S.y) + fabsf(S.z);
    const float ad = fabsf(det);
    return ad + m.e_det >= CULL_DET && fabsf(fmaf(-0.5f, det, un)) <= fmaf(0.5f + CULL_UV, ad, fmaf(s1, m.k_num, m.e_det));
}

// Phase 2 of the filter + exact evaluation for one (ray, triangle) that passed phase 1; returns the reference
// distance or -1.  `tmax` bounds the distances that can still matter to the caller (already widened by
// CULL_TREL / CULL_TABS).
template <int MODE>
RT_DEV float triangle_survivor(V3 o, V3 d, V3 p0, V3 e1, V3 e2, float tmax, const TriMargins& m)
{
    const V3 P = cross3<false>(d, e2);
    const float det = dot3<false>(e1, P);
    const float ad = fabsf(det);
    if (ad > 2.0f * m.e_det) {                       // the sign of det is certain
        const float sg = det < 0.f ? -1.0f : 1.0f;
        const V3 S = sub3<false>(o, p0);
        const float s1 = fabsf(S.x) + fabsf(S.y) + fabsf(S.z);
        const float un = dot3<false>(S, P) * sg;
        const V3 Q = cross3<false>(S, e1);
        const float vn = dot3<false>(d, Q) * sg;
        const float tn = dot3<false>(e2, Q) * sg;
        const float e_num = s1 * m.k_num, e_v = s1 * m.k_v, e_t = s1 * m.k_t;
        const bool pass = vn >= -e_v
                       && un + vn <= fmaf(ad, 1.0f + CULL_UV, m.e_det + e_num + e_v)
                       && tn >= fmaf(RT_EPS * 0.99f, ad, -fmaf(RT_EPS, m.e_det, e_t))
                       && tn <= fmaf(tmax, ad + m.e_det, e_t);
        if (!pass) return -1.0f;
    }
    return MODE == 2 ? triangle_t<false>(o, d, p0, e1, e2) : triangle_t<true>(o, d, p0, e1, e2);
}

// Phase 1 with the margins computed on the spot (one ray per thread: resident kernels' warp votes)
RT_DEV bool triangle_candidate(V3 o, V3 d, V3 p0, V3 e1, V3 e2)
{
    return triangle_phase1(o, d, p0, e1, e2, tri_margins(e1, e2, max_abs3(d)));
}

} // namespace rtb
