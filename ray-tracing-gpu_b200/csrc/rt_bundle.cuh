// rt_bundle.cuh -- warp bundles: the cone(s) that contain the rays a warp holds, and the conservative
// "no ray of the bundle can be reported to hit anything inside this inflated sphere" test.  Shared by the tiled
// kernels (rt_trace.cu) and the hierarchical culled scan (rt_cullscan.cu).
#pragma once
#include "rt_stream.cuh"
#include "rt_cull.cuh"

namespace rtb {

#ifndef FULL_MASK
#define FULL_MASK 0xffffffffu
#endif

// ---------------------------------------------------------------------------------------
// Warp bundles: whole triangles are culled per WARP before any per-ray arithmetic.
//
// The 128 rays of a warp (32 lanes x 4) are bounded by a cone -- apex, unit axis, half angle -- that
// contains every point a ray of the warp can report a hit at:
//   * nearest hit: origins within rho of their mean m, directions within theta of their mean u
//     => all rays lie in the cone of half angle theta around u whose apex sits rho / sin(theta)
//     behind m (the ball of origins is inscribed in it); primary rays: rho = 0, apex = camera;
//   * shadow queries of one light: the segments hit point -> light lie in the cone with the LIGHT
//     as apex around the mean direction light -> hit points, and within `far` = the largest distance.
// In a round of the tile loop lane j takes triangle j of a group of 32, builds its bounding sphere
// (centroid, largest vertex distance), inflates it and tests it against the cone(s); the ballot of
// the survivors is the list of triangles the warp then tests ray by ray with the unchanged filter +
// reference-sequence code.  A culled triangle is one no ray of the warp can be reported to hit, so
// every kept value is still the reference's (same contract as the LBVH):
// CTriangle::Intersection accepts hits up to delta ~ c 2^-24 |S| |e1| |e2| |d| / |det| outside the
// geometric triangle, |det| >= EPSILON (Triangle.cpp:142), hence the same pad as the LBVH boxes,
//     pad = PADC 2^-24 / EPSILON  D |e1| |e2| max|d|,   D >= |origin - p0|, floor 32 ulp of D,
// on top of explicit slack for the rounding of the test itself.  Incoherent warps (theta >= 44 deg,
// directions in opposite half spaces, non-finite rays) do not cull at all.  Off in RTB200_F_EXACT_FP.
// ---------------------------------------------------------------------------------------
#define BUNDLE_PADK  (1.0f * 5.9604645e-8f / RT_EPS)
#define BUNDLE_MAX_LIGHTS 8
#define BUNDLE_SPLIT_SIN 0.02f    // a light's cone wider than this (1.1 degrees) is split in two

struct Bundle {
    float ax, ay, az;              // apex
    float ux, uy, uz;              // unit axis
    float sin_t, cos2;             // widened half angle
    float far;                     // hits lie within `far` of the apex (3e38: unbounded)
    float dadd;                    // |origin - x| <= |x - apex| + dadd for every origin of the bundle
    float padk;                    // BUNDLE_PADK * max(1, max |d|)
    float slack;                   // rounding of apex and centres: 4e-7 * |coordinates|
    int   state;                   // 0: no rays, 1: cone valid, 2: wide (cannot cull)
};

RT_DEV float warp_sum(float v)
{
    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL_MASK, v, off);
    return v;
}
RT_DEV float warp_max_nonneg(float v)          // v >= 0, not NaN: binary32 orders like uint32
{
    return __uint_as_float(__reduce_max_sync(FULL_MASK, __float_as_uint(v)));
}
RT_DEV bool finite3(V3 a) { return fabsf(a.x) <= 3.0e38f && fabsf(a.y) <= 3.0e38f && fabsf(a.z) <= 3.0e38f; }

// bounding sphere of a triangle record: centroid, largest vertex distance, |e1| |e2|
RT_DEV void triangle_sphere(V3 p0, V3 e1, V3 e2, V3& c, float& r, float& w)
{
    const V3 a0 = mk3((e1.x + e2.x) * (-1.0f / 3.0f), (e1.y + e2.y) * (-1.0f / 3.0f), (e1.z + e2.z) * (-1.0f / 3.0f));   // p0 - c
    c = mk3(p0.x - a0.x, p0.y - a0.y, p0.z - a0.z);
    const V3 a1 = mk3(a0.x + e1.x, a0.y + e1.y, a0.z + e1.z), a2 = mk3(a0.x + e2.x, a0.y + e2.y, a0.z + e2.z);
    const float r2 = fmaxf(dot3<false>(a0, a0), fmaxf(dot3<false>(a1, a1), dot3<false>(a2, a2)));
    r = sqrtf(r2) * 1.00001f;
    w = sqrtf(dot3<false>(e1, e1) * dot3<false>(e2, e2)) * 1.00001f;
}

// true: no point within the inflated sphere lies in the bundle's cone (and range)
RT_DEV bool bundle_culls(const Bundle& b, V3 c, float r, float w)
{
    const float vx = c.x - b.ax, vy = c.y - b.ay, vz = c.z - b.az;
    const float s = fmaf(vx, b.ux, fmaf(vy, b.uy, vz * b.uz));
    const float vv = fmaf(vx, vx, fmaf(vy, vy, vz * vz));
    const float dist = vv * rsqrtf(fmaxf(vv, 1.0e-30f)) * 1.00001f;
    const float D = dist + b.dadd + r;
    const float pad = fmaxf(b.padk * D * w, 3.8e-6f * D);
    const float rp = fmaf(r + pad, 1.0001f, 1.0e-5f + b.slack);
    // distance of the centre to the cone's boundary line in the (axis, perpendicular) half plane
    const float cx = fmaf(vy, b.uz, -vz * b.uy), cy = fmaf(vz, b.ux, -vx * b.uz), cz = fmaf(vx, b.uy, -vy * b.ux);
    const float perp2 = fmaf(cx, cx, fmaf(cy, cy, cz * cz));
    const float h = fmaf(s, b.sin_t, rp) + 2.0e-6f * dist;
    bool cull = h < 0.f || perp2 * b.cos2 > h * h;
    const float fr = b.far + rp;
    cull = cull || vv > fr * fr;
    return cull;
}

// the same for a cone whose apex is a light (moderate distances: the perpendicular distance comes from
// |v|^2 - s^2, its cancellation error is covered by an explicit slack)
RT_DEV bool light_cone_culls(const Bundle& b, float4 sph, float w)
{
    const float vx = sph.x - b.ax, vy = sph.y - b.ay, vz = sph.z - b.az;
    const float s = fmaf(vx, b.ux, fmaf(vy, b.uy, vz * b.uz));
    const float vv = fmaf(vx, vx, fmaf(vy, vy, vz * vz));
    const float dist = vv * rsqrtf(fmaxf(vv, 1.0e-30f)) * 1.00001f;
    const float D = dist + b.dadd + sph.w;
    const float pad = fmaxf(b.padk * D * w, 3.8e-6f * D);
    const float rp = fmaf(sph.w + pad, 1.0001f, 1.0e-5f + b.slack);
    const float perp2 = fmaf(-s, s, vv) - 4.0e-6f * vv;           // |axis| = 1 within 3e-7, products rounded
    const float h = fmaf(s, b.sin_t, rp);
    const float fr = b.far + rp;
    return h < 0.f || perp2 * b.cos2 > h * h || vv > fr * fr;
}

// cone of a set of rays held 4 per lane; `use[r]`: the ray takes part
RT_DEV Bundle bundle_of_rays(const V3* o, const V3* d, const bool* use)
{
    Bundle b;
    b.ax = b.ay = b.az = 0.f; b.ux = b.uy = b.uz = 0.f; b.sin_t = 0.f; b.cos2 = 1.f;
    b.far = 3.0e38f; b.dadd = 0.f; b.padk = BUNDLE_PADK; b.slack = 0.f; b.state = 2;
    float so[3] = {0.f, 0.f, 0.f}, sd[3] = {0.f, 0.f, 0.f}, n = 0.f, dmax2 = 0.f;
    V3 dn[ST_RPT];
    bool fin = true;
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) {
        dn[r] = mk3(0.f, 0.f, 0.f);
        if (use[r]) {
            const float l2 = dot3<false>(d[r], d[r]);
            const bool ok = finite3(o[r]) && finite3(d[r]) && l2 > 1.0e-20f && l2 < 1.0e20f;
            fin = fin && ok;
            if (ok) {
                const float il = rsqrtf(l2);
                dn[r] = mk3(d[r].x * il, d[r].y * il, d[r].z * il);
                so[0] += o[r].x; so[1] += o[r].y; so[2] += o[r].z;
                sd[0] += dn[r].x; sd[1] += dn[r].y; sd[2] += dn[r].z;
                dmax2 = fmaxf(dmax2, l2);
                n += 1.0f;
            }
        }
    }
    n = warp_sum(n);
    fin = __all_sync(FULL_MASK, fin);
    if (n == 0.f && fin) { b.state = 0; return b; }
    if (!fin) return b;
    #pragma unroll
    for (int k = 0; k < 3; ++k) { so[k] = warp_sum(so[k]); sd[k] = warp_sum(sd[k]); }
    dmax2 = warp_max_nonneg(dmax2);
    const float inv_n = 1.0f / n;
    const V3 m = mk3(so[0] * inv_n, so[1] * inv_n, so[2] * inv_n);
    const float len2 = fmaf(sd[0], sd[0], fmaf(sd[1], sd[1], sd[2] * sd[2]));
    if (!(len2 > 0.25f * n * n)) return b;                      // mean direction shorter than 0.5: incoherent
    const float il = 1.0f / sqrtf(len2);
    const V3 u = mk3(sd[0] * il, sd[1] * il, sd[2] * il);
    float rho2 = 0.f, sin2 = 0.f;
    bool fwd = true;
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r)
        if (use[r]) {
            const V3 q = mk3(o[r].x - m.x, o[r].y - m.y, o[r].z - m.z);
            rho2 = fmaxf(rho2, dot3<false>(q, q));
            const V3 x = cross3<false>(u, dn[r]);
            sin2 = fmaxf(sin2, dot3<false>(x, x));
            fwd = fwd && dot3<false>(u, dn[r]) > 0.5f;
        }
    rho2 = warp_max_nonneg(rho2); sin2 = warp_max_nonneg(sin2);
    fwd = __all_sync(FULL_MASK, fwd);
    if (!fwd) return b;
    const float sin_e = fmaxf(fmaf(sqrtf(sin2), 1.0001f, 2.0e-6f), 1.0e-4f);
    if (!(sin_e < 0.7f)) return b;
    const float m1 = fabsf(m.x) + fabsf(m.y) + fabsf(m.z);
    const float rho_e = fmaf(sqrtf(rho2), 1.0001f, 2.0e-6f * m1 + 1.0e-6f);
    const float shift = rho_e / sin_e * 1.00001f;
    b.ax = m.x - u.x * shift; b.ay = m.y - u.y * shift; b.az = m.z - u.z * shift;
    b.ux = u.x; b.uy = u.y; b.uz = u.z;
    b.sin_t = sin_e; b.cos2 = 1.0f - sin_e * sin_e;
    b.slack = 4.0e-7f * (m1 + shift);
    b.dadd = (shift + rho_e) * 1.0001f + b.slack;
    b.padk = BUNDLE_PADK * fmaxf(1.0f, sqrtf(dmax2) * 1.00001f);
    b.state = 1;
    return b;
}

// direction (one of 4 per lane) of the member with the largest key >= 0 over the warp; unspecified without members
RT_DEV V3 warp_pick(const V3* v, const float* key, const bool* in)
{
    float best = -1.0f;
    V3 mine = mk3(0.f, 0.f, 0.f);
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r)
        if (in[r] && key[r] > best) { best = key[r]; mine = v[r]; }
    const float kb = fmaxf(best, 0.f);
    const float m = warp_max_nonneg(kb);
    const unsigned int who = __ballot_sync(FULL_MASK, best >= 0.f && kb == m);
    const int src = who ? __ffs((int)who) - 1 : 0;
    return mk3(__shfl_sync(FULL_MASK, mine.x, src), __shfl_sync(FULL_MASK, mine.y, src), __shfl_sync(FULL_MASK, mine.z, src));
}

// members `in` (unit directions dn, 4 per lane) split around their two most distant directions
RT_DEV void split_members(const V3* dn, const bool* in, bool* in_a, bool* in_b)
{
    float key[ST_RPT];
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) key[r] = 1.0f;
    const V3 seed = warp_pick(dn, key, in);
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) key[r] = fmaxf(1.0f - dot3<false>(dn[r], seed), 0.f);
    const V3 pa = warp_pick(dn, key, in);
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) key[r] = fmaxf(1.0f - dot3<false>(dn[r], pa), 0.f);
    const V3 pb = warp_pick(dn, key, in);
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r) {
        const bool near_a = dot3<false>(dn[r], pa) >= dot3<false>(dn[r], pb);
        in_a[r] = in[r] && near_a; in_b[r] = in[r] && !near_a;
    }
}

// cone of the shadow queries of light `l` held 4 per lane (apex = the light); d = unit directions TOWARDS the light
RT_DEV Bundle bundle_of_light(V3 L, const V3* d, const float* dist, const bool* in)
{
    Bundle b;
    b.ax = L.x; b.ay = L.y; b.az = L.z; b.ux = b.uy = b.uz = 0.f; b.sin_t = 0.f; b.cos2 = 1.f;
    b.far = 3.0e38f; b.dadd = 0.f; b.padk = BUNDLE_PADK * 1.0001f; b.state = 2;
    b.slack = 4.0e-7f * (fabsf(L.x) + fabsf(L.y) + fabsf(L.z));
    float sd[3] = {0.f, 0.f, 0.f}, n = 0.f, far = 0.f;
    bool fin = true;
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r)
        if (in[r]) {
            const bool ok = finite3(d[r]) && dist[r] >= 0.f && dist[r] <= 3.0e38f;
            fin = fin && ok;
            if (ok) { sd[0] -= d[r].x; sd[1] -= d[r].y; sd[2] -= d[r].z; far = fmaxf(far, dist[r]); n += 1.0f; }
        }
    n = warp_sum(n);
    fin = __all_sync(FULL_MASK, fin) && finite3(L);
    if (n == 0.f && fin) { b.state = 0; return b; }
    if (!fin) return b;
    #pragma unroll
    for (int k = 0; k < 3; ++k) sd[k] = warp_sum(sd[k]);
    far = warp_max_nonneg(far);
    const float len2 = fmaf(sd[0], sd[0], fmaf(sd[1], sd[1], sd[2] * sd[2]));
    if (!(len2 > 0.25f * n * n)) return b;
    const float il = 1.0f / sqrtf(len2);
    const V3 u = mk3(sd[0] * il, sd[1] * il, sd[2] * il);
    float sin2 = 0.f;
    bool fwd = true;
    #pragma unroll
    for (int r = 0; r < ST_RPT; ++r)
        if (in[r]) {
            const V3 x = cross3<false>(u, d[r]);
            sin2 = fmaxf(sin2, dot3<false>(x, x));
            fwd = fwd && -dot3<false>(u, d[r]) > 0.5f;
        }
    sin2 = warp_max_nonneg(sin2);
    fwd = __all_sync(FULL_MASK, fwd);
    if (!fwd) return b;
    const float sin_e = fmaf(sqrtf(sin2), 1.0002f, 4.0e-6f);       // |d| = 1 within a few ulp (normalised by the query)
    if (!(sin_e < 0.7f)) return b;
    b.ux = u.x; b.uy = u.y; b.uz = u.z;
    b.sin_t = sin_e; b.cos2 = 1.0f - sin_e * sin_e;
    b.far = fmaf(far, 1.0001f, 1.0e-4f);
    b.dadd = b.far + b.slack;
    b.state = 1;
    return b;
}

} // namespace rtb
