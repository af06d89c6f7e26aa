// rt_pack.cu -- scene packing on the device (large scenes).
//
// rtb200_set_scene turns the caller's arrays into (1) the insertion-ordered unified surface list and
// (2) the per-type tile arrays of the streamed / LBVH kernels (rt_scene.cuh).  For the bundled scenes
// (a handful of surfaces) the host writes both into one staging buffer.  For a mesh of a million
// triangles that loop, the 128 MB it fills and their upload were most of the end-to-end time of
// BASELINE config 4, so from RTB200_PACK_ON_DEVICE_MIN triangles on the RAW arrays are uploaded
// (56 B per triangle instead of 128) and this kernel writes the device layout: the edges are the same
// two binary32 subtractions per component the host loop does (and CTriangle::Intersection does per
// test, Triangle.cpp:132-133), so the packed scene is bit-identical either way
// (tests/test_gpu_parity.py::test_device_packing_equals_host_packing).
#include "rt_kernels.cuh"

namespace rtb {

__global__ void __launch_bounds__(256) k_pack_scene(const PackParams p)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.T) {
        const float* v = p.tri_points + 9 * (size_t)i;
        const float* n = p.tri_normal + 3 * (size_t)i;
        const int ord = p.tri_order[i], mat = p.tri_material[i];
        const float p0x = v[0], p0y = v[1], p0z = v[2];
        const float e1x = __fsub_rn(v[3], p0x), e1y = __fsub_rn(v[4], p0y), e1z = __fsub_rn(v[5], p0z);
        const float e2x = __fsub_rn(v[6], p0x), e2y = __fsub_rn(v[7], p0y), e2z = __fsub_rn(v[8], p0z);
        const float4 nn = make_float4(n[0], n[1], n[2], 0.f);
        SurfRec r;
        r.r0 = make_float4(p0x, p0y, p0z, __int_as_float(SURF_TRIANGLE));
        r.r1 = make_float4(e1x, e1y, e1z, __int_as_float(mat));
        r.r2 = make_float4(e2x, e2y, e2z, 0.f);
        r.r3 = nn;
        p.surf[ord] = r;
        TriRec t;
        t.a = make_float4(p0x, p0y, p0z, __int_as_float(ord));
        t.b = r.r1; t.c = r.r2;
        p.tri[i] = t;
        p.trn[i] = nn;
    } else if (i < p.Tpad) {                 // zero edges: det = 0, never hit
        TriRec t;
        t.a = make_float4(0.f, 0.f, 0.f, __int_as_float(0x7fffffff));
        t.b = t.c = make_float4(0.f, 0.f, 0.f, 0.f);
        p.tri[i] = t;
        p.trn[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (i < p.Q) {                           // quad[] arrived packed; scatter into the unified list
        const QuadRec q = p.quad[i];
        SurfRec r;
        r.r0 = make_float4(q.q0.x, q.q0.y, q.q0.z, __int_as_float(SURF_QUADRIC));
        r.r1 = q.q1; r.r2 = q.q2; r.r3 = p.quad_r3[i];
        p.surf[__float_as_int(q.q0.w)] = r;
    }
    if (i < p.P) {
        const PlaneRec q = p.plane[i];
        SurfRec r;
        r.r0 = make_float4(q.nc.x, q.nc.y, q.nc.z, __int_as_float(SURF_PLANE));
        r.r1 = make_float4(q.nc.w, 0.f, 0.f, __int_as_float(q.material));
        r.r2 = r.r3 = make_float4(0.f, 0.f, 0.f, 0.f);
        p.surf[q.order] = r;
    }
}

cudaError_t launch_pack_scene(const PackParams& p, cudaStream_t st)
{
    const int n = p.Tpad > p.Q ? (p.Tpad > p.P ? p.Tpad : p.P) : (p.Q > p.P ? p.Q : p.P);
    if (n <= 0) return cudaSuccess;
    k_pack_scene<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p);
    return cudaGetLastError();
}

} // namespace rtb
