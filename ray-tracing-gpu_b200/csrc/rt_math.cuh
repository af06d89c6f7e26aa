// rt_math.cuh -- binary32 arithmetic of the render path with an explicit rounding policy.
//
// The reference computes everything in `float` (MathUtils.h:23) with one rounding per
// operation (x86-64 SSE, no contraction).  Every geometric expression below is written with
// rounding-explicit intrinsics so that the sequence of roundings is OURS, not the
// compiler's, and is the same in every kernel that uses it (resident, streamed, LBVH):
//   FP<true>  ("exact"): mul/add kept separate -> the reference's rounding sequence
//                        (bit-for-bit against the CPU oracle, except powf).
//   FP<false> ("fast") : each `a*b + c` step is ONE fused multiply-add, accumulation order
//                        unchanged.  This is the default and the path the roofline is
//                        measured on; divisions and square roots stay IEEE-rounded.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace rtb {

#define RT_EPS 1.0e-2f    // EPSILON, MathUtils.h:35: self-hit offset AND degenerate threshold

#define RT_DEV __device__ __forceinline__

// Device-side checks of the CHECKED build (make checked -> ab/checked/librtb200.so; compute-sanitizer is closed on the
// GPU pool this was developed on, so the bounds the kernels rely on are asserted by the kernels themselves there):
// a violated check prints where and traps, the call returns RTB200_E_CUDA.  Compiled out of the product library.
#ifdef RTB200_CHECKED
#define RT_CHECK(cond)                                                                                        \
    do {                                                                                                      \
        if (!(cond)) { printf("RT_CHECK failed: %s (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, \
                              (int)blockIdx.x, (int)threadIdx.x); __trap(); }                                 \
    } while (0)
#else
#define RT_CHECK(cond) do { } while (0)
#endif

template <bool EXACT>
struct FP {
    static RT_DEV float mul(float a, float b) { return __fmul_rn(a, b); }
    static RT_DEV float add(float a, float b) { return __fadd_rn(a, b); }
    static RT_DEV float sub(float a, float b) { return __fsub_rn(a, b); }
    static RT_DEV float div(float a, float b) { return __fdiv_rn(a, b); }
    static RT_DEV float rcp(float a) { return __frcp_rn(a); }      // correctly rounded 1/a == the reference's 1.0f / a
    static RT_DEV float sqrt(float a) { return __fsqrt_rn(a); }
    // a*b + c
    static RT_DEV float madd(float a, float b, float c)
    {
        if (EXACT) return __fadd_rn(__fmul_rn(a, b), c);
        return __fmaf_rn(a, b, c);
    }
    // a*b - c*d
    static RT_DEV float mmsub(float a, float b, float c, float d)
    {
        if (EXACT) return __fsub_rn(__fmul_rn(a, b), __fmul_rn(c, d));
        return __fmaf_rn(a, b, -__fmul_rn(c, d));
    }
};

struct V3 { float x, y, z; };

RT_DEV V3 mk3(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
RT_DEV V3 neg3(V3 a) { return mk3(-a.x, -a.y, -a.z); }

template <bool E> RT_DEV V3 sub3(V3 a, V3 b) { return mk3(FP<E>::sub(a.x, b.x), FP<E>::sub(a.y, b.y), FP<E>::sub(a.z, b.z)); }
template <bool E> RT_DEV V3 scale3(V3 a, float s) { return mk3(FP<E>::mul(a.x, s), FP<E>::mul(a.y, s), FP<E>::mul(a.z, s)); }
// a + b*s   (reference: Origine + Distance * Direction)
template <bool E> RT_DEV V3 axpy3(V3 a, V3 b, float s) { return mk3(FP<E>::madd(b.x, s, a.x), FP<E>::madd(b.y, s, a.y), FP<E>::madd(b.z, s, a.z)); }
// a - b*s   (reference: V - (s * N)); fast form folds the negation into the fma
template <bool E> RT_DEV V3 axmy3(V3 a, V3 b, float s)
{
    if (E) return mk3(__fsub_rn(a.x, __fmul_rn(b.x, s)), __fsub_rn(a.y, __fmul_rn(b.y, s)), __fsub_rn(a.z, __fmul_rn(b.z, s)));
    return mk3(__fmaf_rn(-b.x, s, a.x), __fmaf_rn(-b.y, s, a.y), __fmaf_rn(-b.z, s, a.z));
}
// ProdScal, Vecteur3.h:454-457: (x*x' + y*y') + z*z'
template <bool E> RT_DEV float dot3(V3 a, V3 b)
{
    return FP<E>::madd(a.z, b.z, FP<E>::madd(a.y, b.y, FP<E>::mul(a.x, b.x)));
}
// ProdVect, Vecteur3.h:471-476
template <bool E> RT_DEV V3 cross3(V3 a, V3 b)
{
    return mk3(FP<E>::mmsub(a.y, b.z, a.z, b.y), FP<E>::mmsub(a.z, b.x, a.x, b.z), FP<E>::mmsub(a.x, b.y, a.y, b.x));
}
// Norme, Vecteur3.h:489-492
template <bool E> RT_DEV float norm3(V3 a) { return FP<E>::sqrt(dot3<E>(a, a)); }
// Normaliser, Vecteur3.h:522-534: ZERO when the length is <= EPSILON
template <bool E> RT_DEV V3 normalise3(V3 a)
{
    float len = norm3<E>(a);
    if (len > RT_EPS) return scale3<E>(a, FP<E>::rcp(len));
    return mk3(0.f, 0.f, 0.f);
}
// Reflect, Vecteur3.h:548-551: V - ((2*dot(V,N)) * N)
template <bool E> RT_DEV V3 reflect3(V3 v, V3 n)
{
    return axmy3<E>(v, n, FP<E>::mul(2.0f, dot3<E>(v, n)));
}
// Refract, Vecteur3.h:566-581.  pow(float,int) promotes to double: the root is a binary64
// sqrt of 1 - |Z|^2 narrowed to binary32; total internal reflection shows up as NaN and
// falls back to Reflect through the failed `< 0` compare.
template <bool E> RT_DEV V3 refract3(V3 v, V3 n, float ratio)
{
    V3 z = scale3<E>(axmy3<E>(v, n, dot3<E>(v, n)), ratio);
    double nz = (double)norm3<E>(z);
    float s = (float)__dsqrt_rn(__dsub_rn(1.0, __dmul_rn(nz, nz)));
    V3 t = axmy3<E>(z, n, s);
    if (dot3<E>(t, n) < 0.f) return t;
    return reflect3<E>(v, n);
}

// ---------------------------------------------------------------------------------------
// Surface tests.  All return the reference's "distance" (-1 = the default CIntersection,
// i.e. miss; the caller applies `t > EPSILON`).
// ---------------------------------------------------------------------------------------

// CTriangle::Intersection, Triangle.cpp:127-172 (Moller-Trumbore, two-sided).
// e1 = p1 - p0 and e2 = p2 - p0 are precomputed at scene upload: a binary32 subtraction of
// the same operands, hence the same bits as the per-test subtraction of the reference.
template <bool E> RT_DEV float triangle_t(V3 o, V3 d, V3 p0, V3 e1, V3 e2)
{
    V3 P = cross3<E>(d, e2);
    float det = dot3<E>(e1, P);
    if (fabsf(det) < RT_EPS) return -1.0f;
    float inv = FP<E>::rcp(det);
    V3 S = sub3<E>(o, p0);
    float u = FP<E>::mul(dot3<E>(S, P), inv);
    if (u < 0.f || u > 1.f) return -1.0f;
    V3 Q = cross3<E>(S, e1);
    float v = FP<E>::mul(dot3<E>(d, Q), inv);
    if (v < 0.f || FP<E>::add(u, v) > 1.f) return -1.0f;
    return FP<E>::mul(dot3<E>(e2, Q), inv);
}

// CPlan::Intersection, Plan.cpp:128-144
template <bool E> RT_DEV float plane_t(V3 o, V3 d, V3 n, float c)
{
    float vd = dot3<E>(n, d);
    if (fabsf(vd) > RT_EPS) return FP<E>::div(-FP<E>::add(dot3<E>(n, o), c), vd);
    return -1.0f;
}

struct Quadric { float qx, qy, qz, mx, my, mz, lx, ly, lz, K; };   // mx,my,mz = mixed yz,xz,xy

// CQuadrique::Intersection, Quadrique.cpp:160-249, distance only.  `linear` reports the
// A == 0 branch (unconditional "hit" at -0.5*(C/B), normal = normalise(linear terms)).
template <bool E> RT_DEV float quadric_t(V3 o, V3 d, const Quadric& q, bool& linear)
{
    typedef FP<E> F;
    linear = false;
    // A = dx*(qx*dx + mz*dy + my*dz) + dy*(qy*dy + mx*dz) + dz*(qz*dz)
    float a1 = F::mul(d.x, F::madd(q.my, d.z, F::madd(q.mz, d.y, F::mul(q.qx, d.x))));
    float a2 = F::mul(d.y, F::madd(q.mx, d.z, F::mul(q.qy, d.y)));
    float a3 = F::mul(d.z, F::mul(q.qz, d.z));
    float A = F::add(F::add(a1, a2), a3);
    // B = dx*(qx*ox + 0.5*(mz*oy + my*oz + lx)) + dy*(...) + dz*(...)
    float b1 = F::mul(d.x, F::madd(0.5f, F::add(F::madd(q.my, o.z, F::mul(q.mz, o.y)), q.lx), F::mul(q.qx, o.x)));
    float b2 = F::mul(d.y, F::madd(0.5f, F::add(F::madd(q.mx, o.z, F::mul(q.mz, o.x)), q.ly), F::mul(q.qy, o.y)));
    float b3 = F::mul(d.z, F::madd(0.5f, F::add(F::madd(q.mx, o.y, F::mul(q.my, o.x)), q.lz), F::mul(q.qz, o.z)));
    float B = F::add(F::add(b1, b2), b3);
    // C = ox*(qx*ox + mz*oy + my*oz + lx) + oy*(qy*oy + mx*oz + ly) + oz*(qz*oz + lz) + K
    float c1 = F::mul(o.x, F::add(F::madd(q.my, o.z, F::madd(q.mz, o.y, F::mul(q.qx, o.x))), q.lx));
    float c2 = F::mul(o.y, F::add(F::madd(q.mx, o.z, F::mul(q.qy, o.y)), q.ly));
    float c3 = F::mul(o.z, F::madd(q.qz, o.z, q.lz));
    float C = F::add(F::add(F::add(c1, c2), c3), q.K);

    if (A != 0.0f) {
        // Division-free miss test.  delta = fl(fl(ka*ka) - kb) with ka = fl(-B/A), kb = fl(C/A) has the
        // sign of (B^2/A^2)(1+e1)^2(1+e3) - (C/A)(1+e2), |e| <= 2^-24; it is <= 0 (the reference returns
        // "no intersection", Quadrique.cpp:205) whenever B^2 (1 + 2^-18) < A*C, evaluated here with two
        // rounded products (relative slack 61 ulp >> the 5 ulp needed; the 1e-30 floor keeps both
        // products out of the subnormal range).  Saves two IEEE divisions per missed quadric.
        {
            const float ac = __fmul_rn(A, C);
            if (ac > 1.0e-30f && __fmul_rn(__fmul_rn(B, B), 1.00000381f) < ac) return -1.0f;
        }
        float ka = F::div(-B, A);
        float kb = F::div(C, A);
        float delta = F::mmsub(ka, ka, kb, 1.0f);   // ka*ka - kb  (kb*1 is exact)
        if (delta > 0.f) {
            delta = F::sqrt(delta);
            float t0 = F::sub(ka, delta), t1 = F::add(ka, delta);
            float dist = t0 < t1 ? t0 : t1;
            if (dist < RT_EPS) dist = t0 > t1 ? t0 : t1;
            if (!(dist < 0.f)) return dist;
        }
        return -1.0f;
    }
    linear = true;
    return F::mul(-0.5f, F::div(C, B));
}

// Normal of a quadric hit, Quadrique.cpp:219-237 / :244
template <bool E> RT_DEV V3 quadric_normal(V3 o, V3 d, float t, const Quadric& q, bool linear)
{
    typedef FP<E> F;
    if (linear) return normalise3<E>(mk3(q.lx, q.ly, q.lz));
    V3 h = axpy3<E>(o, d, t);
    V3 n;
    // 2*qx*hx + my*hz + mz*hy + lx
    n.x = F::add(F::madd(q.mz, h.y, F::madd(q.my, h.z, F::mul(F::mul(2.0f, q.qx), h.x))), q.lx);
    n.y = F::add(F::madd(q.mz, h.x, F::madd(q.mx, h.z, F::mul(F::mul(2.0f, q.qy), h.y))), q.ly);
    n.z = F::add(F::madd(q.my, h.x, F::madd(q.mx, h.y, F::mul(F::mul(2.0f, q.qz), h.z))), q.lz);
    return normalise3<E>(n);
}

} // namespace rtb
