// rt_api.cu -- C ABI of librtb200.so (include/rtb200.h): context, scene packing/upload and
// the frame loop.  Replaces CNuanceurCalculProg + the UBO packer and glDispatchCompute of
// CScene::LancerRayons (Scene.cpp:692-1312).  No CPU rendering path exists in this library.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

#include "rt_context.h"

using namespace rtb;

#define TILE_TRIS 256                      // streamed path: triangles per shared-memory tile

namespace rtb {
int set_cuda_error(rtb200_ctx* c, cudaError_t e, const char* where)
{
    if (c) {
        char buf[512];
        snprintf(buf, sizeof buf, "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
        c->last_error = buf;
    }
    (void)cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? RTB200_E_OOM : RTB200_E_CUDA;
}
}

#define CU(call)                                                         \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) return set_cuda_error(ctx, e__, #call);  \
    } while (0)

static std::string g_create_error;      // reason of the last failed rtb200_create (no context yet)

static int create_fail(rtb200_ctx* ctx, const char* where)
{
    cudaError_t e = cudaGetLastError();
    g_create_error = std::string(where) + ": " + cudaGetErrorString(e);
    if (ctx) rtb200_destroy(ctx);
    return RTB200_E_CUDA;
}

static int fail(rtb200_ctx* ctx, int code, const char* msg)
{
    if (ctx) ctx->last_error = msg;
    return code;
}

extern "C" {

int rtb200_abi_version(void) { return RTB200_ABI_VERSION; }

const char* rtb200_error_string(int s)
{
    switch (s) {
    case RTB200_OK: return "ok";
    case RTB200_E_ARG: return "bad argument";
    case RTB200_E_CUDA: return "CUDA error";
    case RTB200_E_NCCL: return "NCCL error";
    case RTB200_E_OOM: return "out of memory";
    case RTB200_E_CAPACITY: return "ray queue capacity exceeded";
    case RTB200_E_STATE: return "invalid call order / state";
    default: return "unknown status";
    }
}

const char* rtb200_last_error(const rtb200_ctx* ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }

int rtb200_create(int device_id, rtb200_ctx** out)
{
    if (!out) return RTB200_E_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {            // no GPU: fail loudly, there is no fallback
        g_create_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        (void)cudaGetLastError();
        return RTB200_E_CUDA;
    }
    if (device_id < 0 || device_id >= n) return RTB200_E_ARG;
    rtb200_ctx* ctx = new (std::nothrow) rtb200_ctx();
    if (!ctx) return RTB200_E_OOM;
    ctx->device = device_id;
    cudaDeviceProp prop;
    if (cudaSetDevice(device_id) != cudaSuccess) return create_fail(ctx, "cudaSetDevice");
    if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) return create_fail(ctx, "cudaGetDeviceProperties");
    ctx->sm_count = prop.multiProcessorCount;
    ctx->l2_bytes = prop.l2CacheSize;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    snprintf(ctx->name, sizeof ctx->name, "%.100s", prop.name);
    if (prop.major < 10) {                       // the library carries sm_100a SASS only
        g_create_error = "device is not sm_100 class";
        rtb200_destroy(ctx);
        return RTB200_E_CUDA;
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) return create_fail(ctx, "cudaStreamCreate");
    for (auto& ev : ctx->ev) if (cudaEventCreate(&ev) != cudaSuccess) return create_fail(ctx, "cudaEventCreate");
    for (auto& ev : ctx->timer) if (cudaEventCreate(&ev) != cudaSuccess) return create_fail(ctx, "cudaEventCreate");
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return create_fail(ctx, "cudaStreamCreate");
    for (int i = 0; i < 4; ++i)
        if (cudaEventCreateWithFlags(&ctx->copy_ev[i], i < 2 ? cudaEventDisableTiming : cudaEventDefault) != cudaSuccess) return create_fail(ctx, "cudaEventCreate");
    if (resident_prepare(ctx->smem_optin - 2048) != cudaSuccess) return create_fail(ctx, "cudaFuncSetAttribute(max dynamic smem)");
    *out = ctx;
    return RTB200_OK;
}

void rtb200_destroy(rtb200_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    comm_destroy(ctx);
    stream_release(ctx);
    cull_release(ctx);
    bvh_release(ctx);
    ctx->scene_buf.release(); ctx->stage.release(); ctx->raw_buf.release();
    ctx->c0.release(); ctx->fb_f32.release(); ctx->fb_rgba8.release(); ctx->stats.release(); ctx->counters.release();
    ctx->full_f32.release(); ctx->full_rgba8.release(); ctx->h_small.release();
    ctx->seq_fb[0].release(); ctx->seq_fb[1].release(); ctx->seq_stats.release(); ctx->seq_host.release(); ctx->h_levels.release(); ctx->tail_dev.release(); ctx->tail_host.release();
    for (auto& l : ctx->levels) { l.od0.release(); l.od1.release(); l.colour.release(); l.links.release(); }
    ctx->q_in0.release(); ctx->q_in1.release(); ctx->q_out0.release(); ctx->q_out1.release(); ctx->q_out2.release();
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->timer) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->copy_ev) if (ev) cudaEventDestroy(ev);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int rtb200_device_count(int* n)
{
    if (!n) return RTB200_E_ARG;
    *n = 0;
    if (cudaGetDeviceCount(n) != cudaSuccess) { (void)cudaGetLastError(); *n = 0; return RTB200_E_CUDA; }
    return RTB200_OK;
}

int rtb200_host_alloc(size_t bytes, void** out)
{
    if (!out) return RTB200_E_ARG;
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
    if (e != cudaSuccess) { (void)cudaGetLastError(); *out = nullptr; return e == cudaErrorMemoryAllocation ? RTB200_E_OOM : RTB200_E_CUDA; }
    return RTB200_OK;
}

void rtb200_host_free(void* p) { if (p) cudaFreeHost(p); }

int rtb200_host_register(void* p, size_t bytes)
{
    if (!p || !bytes) return RTB200_E_ARG;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return e == cudaErrorMemoryAllocation ? RTB200_E_OOM : RTB200_E_CUDA; }
    return RTB200_OK;
}

int rtb200_host_unregister(void* p)
{
    if (!p) return RTB200_E_ARG;
    if (cudaHostUnregister(p) != cudaSuccess) { (void)cudaGetLastError(); return RTB200_E_CUDA; }
    return RTB200_OK;
}

int rtb200_set_sharding(rtb200_ctx* ctx, int rank, int n_ranks)
{
    if (!ctx || n_ranks < 1 || rank < 0 || rank >= n_ranks) return RTB200_E_ARG;
    if (ctx->comm) return fail(ctx, RTB200_E_STATE, "rtb200_set_sharding on a context that has a communicator");
    ctx->rank = rank; ctx->n_ranks = n_ranks;
    ctx->level_guess.valid = false;
    return RTB200_OK;
}

int rtb200_device_info(rtb200_ctx* ctx, int* sm_count, int* l2_bytes, char* name, int name_len)
{
    if (!ctx) return RTB200_E_ARG;
    if (sm_count) *sm_count = ctx->sm_count;
    if (l2_bytes) *l2_bytes = ctx->l2_bytes;
    if (name && name_len > 0) snprintf(name, (size_t)name_len, "%s", ctx->name);
    return RTB200_OK;
}

int rtb200_timer_mark(rtb200_ctx* ctx, int slot)
{
    if (!ctx || slot < 0 || slot >= 4) return RTB200_E_ARG;
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventRecord(ctx->timer[slot], ctx->stream));
    return RTB200_OK;
}

int rtb200_timer_elapsed(rtb200_ctx* ctx, int a, int b, double* ms)
{
    if (!ctx || !ms || a < 0 || a >= 4 || b < 0 || b >= 4) return RTB200_E_ARG;
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventSynchronize(ctx->timer[b]));
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, ctx->timer[a], ctx->timer[b]));
    *ms = t;
    return RTB200_OK;
}

int rtb200_measure_fp32_peak(rtb200_ctx* ctx, double* tflops, double* sm_mhz)
{
    if (!ctx) return RTB200_E_ARG;
    CU(cudaSetDevice(ctx->device));
    double tf = 0.0;
    CU(run_fp32_peak(ctx->sm_count, &tf, ctx->stream));
    ctx->peak_tflops = tf;
    // FFMA chain: 128 lanes x 2 flop per SM per clock -> the clock the chain actually ran at
    ctx->peak_mhz = tf * 1e12 / (2.0 * 128.0 * ctx->sm_count) / 1e6;
    if (tflops) *tflops = tf;
    if (sm_mhz) *sm_mhz = ctx->peak_mhz;
    return RTB200_OK;
}

int rtb200_fp32_peak_detail(rtb200_ctx* ctx, double* detail, int n)
{
    if (!ctx || !detail || n < RTB200_PEAK_DETAIL_COUNT) return RTB200_E_ARG;
    CU(cudaSetDevice(ctx->device));
    double tf = 0.0, d[8] = {0};
    CU(run_fp32_peak(ctx->sm_count, &tf, ctx->stream, d));
    ctx->peak_tflops = tf;
    ctx->peak_mhz = tf * 1e12 / (2.0 * 128.0 * ctx->sm_count) / 1e6;
    detail[0] = tf;
    for (int i = 0; i < 8; ++i) detail[1 + i] = d[i];
    detail[9] = (double)ctx->sm_count;
    return RTB200_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------
// Scene packing.  Replaces the std140 UBO packer of Scene.cpp:799-1269 (which walked the
// surfaces through dynamic_cast + getters and capped every type at 10): the arrays of
// rtb200_scene are repacked into (1) the insertion-ordered unified list the resident kernels
// stage into shared memory and (2) per-type tile-friendly arrays for the streamed kernels,
// in ONE pinned staging buffer, uploaded with ONE async copy.
// ---------------------------------------------------------------------------------------
static inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }
static inline float i2f(int v) { float f; memcpy(&f, &v, 4); return f; }

// Centred form of a quadric for the warp-vote culling test of the resident kernels
// (rt_resident.cu: quadric_miss):  sum_i q_i (x_i - c_i)^2 = rhs  with every q_i >= 0 (not all zero),
// no mixed terms, no linear term along an axis with q_i == 0 and rhs > 0 -- spheres, ellipsoids and
// axis-aligned cylinders.  Returns (c, rhs); rhs = -1 means "not of that form" (cones, paraboloids,
// hyperboloids, rotated quadrics, degenerate cases): never culled.
static float4 quadric_bounding_sphere(const float* q)
{
    const float4 none = make_float4(0.f, 0.f, 0.f, -1.0f);
    for (int i = 0; i < 10; ++i) if (!std::isfinite(q[i])) return none;
    if (q[3] != 0.f || q[4] != 0.f || q[5] != 0.f) return none;
    double c[3] = {0, 0, 0}, rhs = -(double)q[9];
    bool any = false;
    for (int i = 0; i < 3; ++i) {
        const double a = q[i], l = q[6 + i];
        if (a < 0) return none;
        if (a == 0) { if (l != 0) return none; continue; }
        any = true;
        c[i] = -l / (2.0 * a);
        rhs += a * c[i] * c[i];
    }
    if (!any || !(rhs > 0) || !std::isfinite(rhs) || rhs > 1e30) return none;
    return make_float4((float)c[0], (float)c[1], (float)c[2], (float)rhs);
}

// The three order arrays must each be strictly increasing and together be a permutation of 0..N-1
// (slot = insertion index of CScene::m_Surfaces); materials must exist.  O(N), no scratch memory.
static const char* validate_orders(const rtb200_scene* s, long long N)
{
    const int T = s->n_triangles, Q = s->n_quadrics, P = s->n_planes, M = s->n_materials;
    for (int i = 0; i < T; ++i) if (s->tri_material[i] < 0 || s->tri_material[i] >= M) return "tri_order/tri_material invalid";
    for (int i = 0; i < Q; ++i) if (s->quad_material[i] < 0 || s->quad_material[i] >= M) return "quad_order/quad_material invalid";
    for (int i = 0; i < P; ++i) if (s->plane_material[i] < 0 || s->plane_material[i] >= M) return "plane_order/plane_material invalid";
    int it = 0, iq = 0, ip = 0;
    for (long long k = 0; k < N; ++k) {                       // slot k must be the head of exactly one list
        if (it < T && s->tri_order[it] == k) ++it;
        else if (iq < Q && s->quad_order[iq] == k) ++iq;
        else if (ip < P && s->plane_order[ip] == k) ++ip;
        else return it < T || (iq >= Q && ip >= P) ? "tri_order/tri_material invalid" : (iq < Q ? "quad_order/quad_material invalid" : "plane_order/plane_material invalid");
    }
    return nullptr;
}

// Everything rtb200_set_scene refuses, as text (null = the scene is acceptable).  Pure host code.
static const char* scene_problem(const rtb200_scene* s)
{
    if (!s) return "null scene";
    const int T = s->n_triangles, Q = s->n_quadrics, P = s->n_planes, L = s->n_lights, M = s->n_materials;
    if (T < 0 || Q < 0 || P < 0 || L < 0 || M < 0) return "negative count";
    const long long N = (long long)T + Q + P;
    if (N > 0x7fffff00ll) return "too many surfaces";
    if ((T && (!s->tri_points || !s->tri_normal || !s->tri_material || !s->tri_order)) ||
        (Q && (!s->quad_coef || !s->quad_material || !s->quad_order)) ||
        (P && (!s->plane_coef || !s->plane_material || !s->plane_order)) ||
        (M && !s->materials) || (L && !s->lights) || (N && !M))
        return "null array with non-zero count";
    return validate_orders(s, N);
}

extern "C" int rtb200_validate_scene(const rtb200_scene* scene, const char** why)
{
    const char* w = scene_problem(scene);
    if (why) *why = w ? w : "";
    return w ? RTB200_E_ARG : RTB200_OK;
}

// memcpy of a large array into the pinned staging buffer on a few threads (one core moves ~10 GB/s)
static void parallel_copy(void* dst, const void* src, size_t bytes)
{
    const size_t min_slice = 4u << 20;
    int n = (int)std::min<size_t>(4, bytes / min_slice);
    if (n <= 1) { memcpy(dst, src, bytes); return; }
    const size_t slice = ((bytes / (size_t)n) + 63) & ~(size_t)63;
    std::vector<std::thread> th;
    for (int i = 1; i < n; ++i) {
        const size_t off = slice * (size_t)i, len = std::min(slice, bytes - std::min(bytes, off));
        if (len) th.emplace_back([=] { memcpy(static_cast<unsigned char*>(dst) + off, static_cast<const unsigned char*>(src) + off, len); });
    }
    memcpy(dst, src, std::min(slice, bytes));
    for (auto& t : th) t.join();
}

extern "C" int rtb200_set_scene(rtb200_ctx* ctx, const rtb200_scene* s)
{
    if (!ctx || !s) return RTB200_E_ARG;
    if (const char* why = scene_problem(s)) return fail(ctx, RTB200_E_ARG, why);
    const int T = s->n_triangles, Q = s->n_quadrics, P = s->n_planes, L = s->n_lights, M = s->n_materials;
    const long long N = (long long)T + Q + P;
    CU(cudaSetDevice(ctx->device));
    bvh_invalidate(ctx);                         // an LBVH belongs to the scene it was built from
    ++ctx->scene_serial;

    const int Tpad = ((T + TILE_TRIS - 1) / TILE_TRIS) * TILE_TRIS;
    ResidentLayout rl;
    rl.n_surf = (int)N; rl.n_mat = M; rl.n_light = L;
    rl.off_mat = (int)align16(sizeof(SurfRec) * (size_t)N);
    rl.off_light = (int)align16((size_t)rl.off_mat + sizeof(MatRec) * (size_t)M);
    const size_t blob_bytes = align16((size_t)rl.off_light + sizeof(LightRec) * (size_t)L);
    if (blob_bytes > 0x7fffffffull) return fail(ctx, RTB200_E_ARG, "scene too large");
    rl.bytes = (int)std::max<size_t>(blob_bytes, 16);
    const size_t off_tri = align16((size_t)rl.bytes);
    const size_t off_trn = off_tri + sizeof(TriRec) * (size_t)Tpad;
    const size_t off_quad = off_trn + sizeof(float4) * (size_t)Tpad;
    const size_t off_plane = off_quad + sizeof(QuadRec) * (size_t)Q;
    const size_t total = align16(off_plane + sizeof(PlaneRec) * (size_t)P) + 16;

    // the previous upload may still be reading the staging buffer (uploads are asynchronous)
    if (ctx->upload_pending) { CU(cudaEventSynchronize(ctx->ev[5])); ctx->upload_pending = false; }
    cudaError_t e = ctx->scene_buf.ensure(total);
    if (e != cudaSuccess) return set_cuda_error(ctx, e, "scene buffer");

    // ---- large meshes: upload the raw arrays, pack on the device (rt_pack.cu) -----------------------
    int pack_min = 2048;
    if (const char* pm = getenv("RTB200_PACK_ON_DEVICE_MIN")) pack_min = atoi(pm);
    if (T >= pack_min && T > 0) {
        // staging: [materials + lights][quad + plane records][quad centred forms][raw triangle arrays]
        const size_t ml_bytes = (size_t)rl.bytes - (size_t)rl.off_mat;
        const size_t qp_bytes = sizeof(QuadRec) * (size_t)Q + sizeof(PlaneRec) * (size_t)P;
        const size_t s_qp = align16(ml_bytes), s_raw = align16(s_qp + qp_bytes);
        const size_t r_r3 = 0, r_pts = align16(sizeof(float4) * (size_t)Q), r_nrm = r_pts + align16(36 * (size_t)T),
                     r_mat = r_nrm + align16(12 * (size_t)T), r_ord = r_mat + align16(4 * (size_t)T), raw_bytes = r_ord + align16(4 * (size_t)T);
        if ((e = ctx->stage.ensure(s_raw + raw_bytes)) != cudaSuccess) return set_cuda_error(ctx, e, "pinned staging");
        if ((e = ctx->raw_buf.ensure(raw_bytes)) != cudaSuccess) return set_cuda_error(ctx, e, "raw scene buffer");
        unsigned char* h = ctx->stage.as<unsigned char>();
        memset(h, 0, s_raw + r_pts);
        MatRec* mat = reinterpret_cast<MatRec*>(h);
        LightRec* light = reinterpret_cast<LightRec*>(h + (rl.off_light - rl.off_mat));
        bool any_refl = false, any_refr = false, any_both = false;
        for (int i = 0; i < M; ++i) {
            const rtb200_material& m = s->materials[i];
            mat[i].m0 = make_float4(m.color[0], m.color[1], m.color[2], m.ambient);
            mat[i].m1 = make_float4(m.diffuse, m.specular, m.shininess, m.reflect);
            mat[i].m2 = make_float4(m.refract, m.ior, 0.f, 0.f);
            const bool a = m.reflect > 0.f, b = m.refract > 0.f;
            any_refl |= a; any_refr |= b; any_both |= (a && b);
        }
        for (int i = 0; i < L; ++i) {
            const rtb200_light& l = s->lights[i];
            light[i].l0 = make_float4(l.position[0], l.position[1], l.position[2], l.intensity);
            light[i].l1 = make_float4(l.color[0], l.color[1], l.color[2], 0.f);
        }
        QuadRec* quad = reinterpret_cast<QuadRec*>(h + s_qp);
        float4* qr3 = reinterpret_cast<float4*>(h + s_raw + r_r3);
        for (int i = 0; i < Q; ++i) {
            const float* c = s->quad_coef + 10 * (size_t)i;
            quad[i].q0 = make_float4(c[0], c[1], c[2], i2f(s->quad_order[i]));
            quad[i].q1 = make_float4(c[3], c[4], c[5], i2f(s->quad_material[i]));
            quad[i].q2 = make_float4(c[6], c[7], c[8], c[9]);
            qr3[i] = quadric_bounding_sphere(c);
        }
        ctx->h_quad.assign(quad, quad + Q); ctx->h_quad_r3.assign(qr3, qr3 + Q);
        PlaneRec* plane = reinterpret_cast<PlaneRec*>(h + s_qp + sizeof(QuadRec) * (size_t)Q);
        for (int i = 0; i < P; ++i) {
            const float* c = s->plane_coef + 4 * (size_t)i;
            plane[i].nc = make_float4(c[0], c[1], c[2], c[3]);
            plane[i].order = s->plane_order[i]; plane[i].material = s->plane_material[i];
        }
        parallel_copy(h + s_raw + r_pts, s->tri_points, 36 * (size_t)T);
        parallel_copy(h + s_raw + r_nrm, s->tri_normal, 12 * (size_t)T);
        parallel_copy(h + s_raw + r_mat, s->tri_material, 4 * (size_t)T);
        parallel_copy(h + s_raw + r_ord, s->tri_order, 4 * (size_t)T);

        unsigned char* d = ctx->scene_buf.as<unsigned char>();
        unsigned char* draw = ctx->raw_buf.as<unsigned char>();
        CU(cudaEventRecord(ctx->ev[4], ctx->stream));
        CU(cudaMemcpyAsync(d + rl.off_mat, h, ml_bytes, cudaMemcpyHostToDevice, ctx->stream));
        if (qp_bytes) CU(cudaMemcpyAsync(d + off_quad, h + s_qp, qp_bytes, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(draw, h + s_raw, raw_bytes, cudaMemcpyHostToDevice, ctx->stream));
        PackParams pp;
        memset(&pp, 0, sizeof pp);
        pp.T = T; pp.Tpad = Tpad; pp.Q = Q; pp.P = P;
        pp.tri_points = reinterpret_cast<const float*>(draw + r_pts);
        pp.tri_normal = reinterpret_cast<const float*>(draw + r_nrm);
        pp.tri_material = reinterpret_cast<const int*>(draw + r_mat);
        pp.tri_order = reinterpret_cast<const int*>(draw + r_ord);
        pp.quad_r3 = reinterpret_cast<const float4*>(draw + r_r3);
        pp.surf = reinterpret_cast<SurfRec*>(d);
        pp.tri = reinterpret_cast<TriRec*>(d + off_tri);
        pp.trn = reinterpret_cast<float4*>(d + off_trn);
        pp.quad = reinterpret_cast<const QuadRec*>(d + off_quad);
        pp.plane = reinterpret_cast<const PlaneRec*>(d + off_plane);
        CU(launch_pack_scene(pp, ctx->stream));
        CU(cudaEventRecord(ctx->ev[5], ctx->stream));
        ctx->upload_pending = true;
        DeviceScene& ds = ctx->ds;
        ds.rl = rl;
        ds.blob = d;
        ds.surf = reinterpret_cast<const SurfRec*>(d);
        ds.mat = reinterpret_cast<const MatRec*>(d + rl.off_mat);
        ds.light = reinterpret_cast<const LightRec*>(d + rl.off_light);
        ds.n_tri = T; ds.n_quad = Q; ds.n_plane = P; ds.n_tri_padded = Tpad;
        ds.tri = reinterpret_cast<const TriRec*>(d + off_tri);
        ds.tri_normal = reinterpret_cast<const float4*>(d + off_trn);
        ds.quad = reinterpret_cast<const QuadRec*>(d + off_quad);
        ds.plane = reinterpret_cast<const PlaneRec*>(d + off_plane);
        ctx->any_reflect = any_refl; ctx->any_refract = any_refr;
        ctx->max_children = any_both ? 2 : ((any_refl || any_refr) ? 1 : 0);
        ctx->have_scene = true;
        return RTB200_OK;
    }

    e = ctx->stage.ensure(total);
    if (e != cudaSuccess) return set_cuda_error(ctx, e, "pinned staging");
    unsigned char* h = ctx->stage.as<unsigned char>();
    memset(h, 0, total);

    // ---- unified list: slot = insertion order ------------------------------------------------
    SurfRec* surf = reinterpret_cast<SurfRec*>(h);
    TriRec* tri = reinterpret_cast<TriRec*>(h + off_tri);
    float4* trn = reinterpret_cast<float4*>(h + off_trn);
    for (int i = 0; i < T; ++i) {
        const float* p = s->tri_points + 9 * (size_t)i;
        const float* n = s->tri_normal + 3 * (size_t)i;
        const int ord = s->tri_order[i], mat = s->tri_material[i];
        // edges: the same binary32 subtractions CTriangle::Intersection performs per test
        const float e1x = p[3] - p[0], e1y = p[4] - p[1], e1z = p[5] - p[2];
        const float e2x = p[6] - p[0], e2y = p[7] - p[1], e2z = p[8] - p[2];
        SurfRec& r = surf[ord];
        r.r0 = make_float4(p[0], p[1], p[2], i2f(SURF_TRIANGLE));
        r.r1 = make_float4(e1x, e1y, e1z, i2f(mat));
        r.r2 = make_float4(e2x, e2y, e2z, 0.f);
        r.r3 = make_float4(n[0], n[1], n[2], 0.f);
        tri[i].a = make_float4(p[0], p[1], p[2], i2f(ord));
        tri[i].b = make_float4(e1x, e1y, e1z, i2f(mat));
        tri[i].c = make_float4(e2x, e2y, e2z, 0.f);
        trn[i] = make_float4(n[0], n[1], n[2], 0.f);
    }
    for (int i = T; i < Tpad; ++i) tri[i].a.w = i2f(0x7fffffff);   // zero edges: det = 0, never hit
    QuadRec* quad = reinterpret_cast<QuadRec*>(h + off_quad);
    for (int i = 0; i < Q; ++i) {
        const float* c = s->quad_coef + 10 * (size_t)i;
        const int ord = s->quad_order[i], mat = s->quad_material[i];
        SurfRec& r = surf[ord];
        r.r0 = make_float4(c[0], c[1], c[2], i2f(SURF_QUADRIC));
        r.r1 = make_float4(c[3], c[4], c[5], i2f(mat));
        r.r2 = make_float4(c[6], c[7], c[8], c[9]);
        r.r3 = quadric_bounding_sphere(c);
        quad[i].q0 = make_float4(c[0], c[1], c[2], i2f(ord));
        quad[i].q1 = r.r1; quad[i].q2 = r.r2;
    }
    ctx->h_quad.assign(quad, quad + Q);
    ctx->h_quad_r3.resize((size_t)Q);
    for (int i = 0; i < Q; ++i) ctx->h_quad_r3[(size_t)i] = surf[s->quad_order[i]].r3;
    PlaneRec* plane = reinterpret_cast<PlaneRec*>(h + off_plane);
    for (int i = 0; i < P; ++i) {
        const float* c = s->plane_coef + 4 * (size_t)i;
        const int ord = s->plane_order[i], mat = s->plane_material[i];
        SurfRec& r = surf[ord];
        r.r0 = make_float4(c[0], c[1], c[2], i2f(SURF_PLANE));
        r.r1 = make_float4(c[3], 0.f, 0.f, i2f(mat));
        plane[i].nc = make_float4(c[0], c[1], c[2], c[3]);
        plane[i].order = ord; plane[i].material = mat;
    }
    MatRec* mat = reinterpret_cast<MatRec*>(h + rl.off_mat);
    bool any_refl = false, any_refr = false, any_both = false;
    for (int i = 0; i < M; ++i) {
        const rtb200_material& m = s->materials[i];
        mat[i].m0 = make_float4(m.color[0], m.color[1], m.color[2], m.ambient);
        mat[i].m1 = make_float4(m.diffuse, m.specular, m.shininess, m.reflect);
        mat[i].m2 = make_float4(m.refract, m.ior, 0.f, 0.f);
        const bool a = m.reflect > 0.f, b = m.refract > 0.f;
        any_refl |= a; any_refr |= b; any_both |= (a && b);
    }
    LightRec* light = reinterpret_cast<LightRec*>(h + rl.off_light);
    for (int i = 0; i < L; ++i) {
        const rtb200_light& l = s->lights[i];
        light[i].l0 = make_float4(l.position[0], l.position[1], l.position[2], l.intensity);
        light[i].l1 = make_float4(l.color[0], l.color[1], l.color[2], 0.f);
    }

    CU(cudaEventRecord(ctx->ev[4], ctx->stream));
    CU(cudaMemcpyAsync(ctx->scene_buf.p, h, total, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaEventRecord(ctx->ev[5], ctx->stream));
    ctx->upload_pending = true;                 // stream-ordered: every later kernel sees the new scene; the
                                                // caller's arrays were already copied into the staging buffer

    const unsigned char* d = ctx->scene_buf.as<unsigned char>();
    DeviceScene& ds = ctx->ds;
    ds.rl = rl;
    ds.blob = d;
    ds.surf = reinterpret_cast<const SurfRec*>(d);
    ds.mat = reinterpret_cast<const MatRec*>(d + rl.off_mat);
    ds.light = reinterpret_cast<const LightRec*>(d + rl.off_light);
    ds.n_tri = T; ds.n_quad = Q; ds.n_plane = P; ds.n_tri_padded = Tpad;
    ds.tri = reinterpret_cast<const TriRec*>(d + off_tri);
    ds.tri_normal = reinterpret_cast<const float4*>(d + off_trn);
    ds.quad = reinterpret_cast<const QuadRec*>(d + off_quad);
    ds.plane = reinterpret_cast<const PlaneRec*>(d + off_plane);
    ctx->any_reflect = any_refl; ctx->any_refract = any_refr;
    ctx->max_children = any_both ? 2 : ((any_refl || any_refr) ? 1 : 0);
    ctx->have_scene = true;
    return RTB200_OK;
}

// ---------------------------------------------------------------------------------------
// Frame loop
// ---------------------------------------------------------------------------------------
#define MAX_CHUNK_SAMPLES (16u << 20)
#define RESIDENT_AUTO_BYTES (64 * 1024)       // scenes up to this blob size use the fused resident kernels

namespace rtb {
int local_row_count(int H, int band_rows, int n_ranks, int rank)
{
    if (n_ranks <= 1) return H;
    int rows = 0;
    for (int y0 = rank * band_rows; y0 < H; y0 += band_rows * n_ranks) rows += std::min(band_rows, H - y0);
    return rows;
}

void fill_frame_params(const rtb200_frame* fr, int n_ranks, int rank, FrameParams& f)
{
    memset(&f, 0, sizeof f);
    f.W = fr->width; f.H = fr->height;
    f.k = fr->spp_grid > 1 ? fr->spp_grid : 1;
    f.LW = f.W * f.k; f.LH = f.H * f.k;
    f.inv_w = 1.0f / (float)f.LW;             // RENDRE_REEL(1.0) / m_ResLargeur at lattice resolution
    f.inv_h = 1.0f / (float)f.LH;
    f.half_w = fr->half_w; f.half_h = fr->half_h;
    for (int i = 0; i < 3; ++i) {
        f.cam_pos[i] = fr->cam_pos[i]; f.cam_u[i] = fr->cam_u[i]; f.cam_v[i] = fr->cam_v[i]; f.cam_n[i] = fr->cam_n[i];
        f.background[i] = fr->background[i];
    }
    f.max_bounces = fr->max_bounces;
    f.min_energy = fr->min_energy;
    f.scene_ior = fr->scene_ior;
    f.secondary = (fr->flags & RTB200_F_NO_SECONDARY) ? 0 : 1;
    const int band = fr->band_rows > 0 ? fr->band_rows : 16;
    f.band_lrows = band * f.k;
    auto log2_exact = [](int v) { int s = 0; while ((1 << s) < v) ++s; return (1 << s) == v ? s : -1; };
    f.band_lshift = log2_exact(f.band_lrows);
    f.band_pshift = log2_exact(band);
    f.n_ranks = n_ranks; f.rank = rank;
}
} // namespace rtb

// Overlapped read-back: called after the kernel that finished pixel rows [py0, py0 + prows) of the
// local frame.  The copy stream waits for that kernel only, so the transfer of chunk i runs under
// the rendering of chunk i + 1 (PCIe and the SMs work at the same time).
namespace rtb {
// Local pixel rows [py0, py0 + prows) of this rank's compact frame -> the caller's host buffers, on stream `s`.
// Plain case: the host buffers hold the same compact rows.  host_bands (several ranks, RTB200_F_HOST_BANDS): the host
// buffers are the FULL frame and local band b (band_rows rows) is image band b * n_ranks + rank -- a 2-D copy whose
// "row" is one whole band (contiguous on both sides), pitch = n_ranks bands on the host side.
static cudaError_t copy_rows_out(rtb200_ctx* ctx, int W, int py0, int prows, cudaStream_t s)
{
    if (prows <= 0) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    for (int which = 0; which < 2 && e == cudaSuccess; ++which) {
        unsigned char* host = which == 0 ? reinterpret_cast<unsigned char*>(ctx->host_f32) : ctx->host_rgba8;
        if (!host) continue;
        const unsigned char* dev = which == 0 ? ctx->fb_f32.as<unsigned char>() : ctx->fb_rgba8.as<unsigned char>();
        const size_t row_bytes = (size_t)W * (which == 0 ? 3 * sizeof(float) : 4);
        if (!ctx->host_bands) {
            e = cudaMemcpyAsync(host + row_bytes * py0, dev + row_bytes * py0, row_bytes * prows, cudaMemcpyDeviceToHost, s);
            continue;
        }
        const int band = ctx->host_band_rows, n = ctx->n_ranks, r = ctx->rank;
        int y = py0;
        const int end = py0 + prows;
        if (y % band) {                                     // (chunks start on band boundaries; belt and braces)
            const int cnt = std::min(end - y, band - y % band);
            e = cudaMemcpyAsync(host + row_bytes * band_global_row(band, n, r, y), dev + row_bytes * y, row_bytes * cnt, cudaMemcpyDeviceToHost, s);
            y += cnt;
        }
        const int full = (end - y) / band;
        if (full > 0 && e == cudaSuccess) {
            e = cudaMemcpy2DAsync(host + row_bytes * band_global_row(band, n, r, y), row_bytes * band * n, dev + row_bytes * y, row_bytes * band,
                                  row_bytes * band, (size_t)full, cudaMemcpyDeviceToHost, s);
            y += full * band;
        }
        if (y < end && e == cudaSuccess)                    // the image's last, partial band
            e = cudaMemcpyAsync(host + row_bytes * band_global_row(band, n, r, y), dev + row_bytes * y, row_bytes * (end - y), cudaMemcpyDeviceToHost, s);
    }
    return e;
}

int chunk_copy_out(rtb200_ctx* ctx, int W, int py0, int prows)
{
    if ((!ctx->host_f32 && !ctx->host_rgba8) || prows <= 0) return RTB200_OK;
    cudaEvent_t done = ctx->copy_ev[ctx->copy_chunks & 1];
    CU(cudaEventRecord(done, ctx->stream));
    CU(cudaStreamWaitEvent(ctx->copy_stream, done, 0));
    if (ctx->copy_chunks == 0) CU(cudaEventRecord(ctx->copy_ev[2], ctx->copy_stream));
    CU(copy_rows_out(ctx, W, py0, prows, ctx->copy_stream));
    ++ctx->copy_chunks;
    return RTB200_OK;
}
}

// scene-resident path with secondary rays, WITHOUT a host round trip per level.  The synchronous loop
// below reads the child count back after every level kernel to size the next launch: on the reference's
// own mirror / glass scenes (scene7-9: 20 bounces allowed, a few hundred rays left after the 8th) that is
// 21 round trips of ~30 us around ~0.1 ms of work.  Here every level kernel reads its input count from
// the counter the previous one wrote (LevelParams::in_count_dev), queues are sized by a guess -- what
// the last frame of this shape needed plus half, or a halving sequence for a first frame -- and a kernel
// that runs out of room raises a flag.  Levels, folds and the resolve of a chunk are enqueued in one go;
// the host waits once per frame, and redoes the frame through the synchronous loop if the flag is up
// (*overflowed).  Same kernels, same arithmetic, same frame.
static int render_resident_async(rtb200_ctx* ctx, const rtb200_frame* fr, const FrameParams& fp, int local_rows,
                                 bool exact, int children, int rows_per_chunk, rtb200_stats* st, uint32_t* launches,
                                 bool* overflowed)
{
    const int k = fp.k, local_LH = local_rows * k, max_b = fr->max_bounces;
    *overflowed = false;
    rtb200_ctx::LevelGuess& g = ctx->level_guess;
    const bool known = g.valid && g.scene_serial == ctx->scene_serial && g.W == fp.W && g.H == fp.H && g.k == k &&
                       g.max_bounces == max_b && g.children == children && g.rows_per_chunk == rows_per_chunk &&
                       g.local_rows == local_rows;
    // deepest level that gets a kernel.  Unknown frame: probe a bounded depth (a level costs queues of >= 64 K entries
    // and one launch per chunk even when the ray tree died long before); a tree that is deeper raises the overflow
    // flag at the probe level and the frame is redone by the synchronous loop, which records the real depth.
    int last = std::min(max_b, 8);
    if (known) {
        int used = 0;
        for (size_t l = 1; l < g.rays.size(); ++l) if (g.rays[l]) used = (int)l;
        last = std::min(max_b, used + 1);
    }
    // queue capacities per level (entries)
    const size_t chunk_samples = (size_t)std::min(rows_per_chunk, local_LH) * fp.LW;
    std::vector<size_t> cap((size_t)last + 2, 0);
    cap[0] = chunk_samples;
    for (int l = 1; l <= last; ++l) {
        const size_t worst = cap[(size_t)l - 1] * (size_t)children;
        size_t want;
        if (known && (size_t)l < g.rays.size()) want = (size_t)g.rays[(size_t)l] + g.rays[(size_t)l] / 2 + 8192;
        else want = (l == 1) ? worst : std::max<size_t>(cap[(size_t)l - 1] / 2, 65536);
        cap[(size_t)l] = std::max<size_t>(std::min(worst, want), 256);
    }
    cap[(size_t)last + 1] = 256;                            // a level that may not spawn still needs valid pointers

    cudaError_t e;
    if ((e = ctx->c0.ensure(sizeof(float) * 3 * (size_t)rows_per_chunk * fp.LW)) != cudaSuccess) return set_cuda_error(ctx, e, "c0");
    const size_t n_counters = 2 * ((size_t)max_b + 2) + 1, ovf = n_counters - 1;
    const int n_chunks = (local_LH + rows_per_chunk - 1) / rows_per_chunk;
    if ((e = ctx->counters.ensure(sizeof(unsigned) * n_counters)) != cudaSuccess) return set_cuda_error(ctx, e, "counters");
    if ((e = ctx->h_levels.ensure(sizeof(unsigned) * n_counters * (size_t)n_chunks)) != cudaSuccess) return set_cuda_error(ctx, e, "level counters (host)");
    unsigned* d_cnt = ctx->counters.as<unsigned>();
    unsigned* h_all = ctx->h_levels.as<unsigned>();
    if ((int)ctx->levels.size() < last + 2) ctx->levels.resize((size_t)last + 2);
    for (int l = 0; l <= last; ++l) {
        LevelBuf& cur = ctx->levels[(size_t)l];
        LevelBuf& nxt = ctx->levels[(size_t)l + 1];
        if (l >= max_b) break;                              // the deepest allowed level never spawns
        if ((e = nxt.od0.ensure(sizeof(float4) * cap[(size_t)l + 1])) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
        if ((e = nxt.od1.ensure(sizeof(float4) * cap[(size_t)l + 1])) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
        if ((e = nxt.colour.ensure(sizeof(float4) * cap[(size_t)l + 1])) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
        if ((e = cur.links.ensure(sizeof(int4) * cap[(size_t)l])) != cudaSuccess) return set_cuda_error(ctx, e, "links");
    }

    // Deep levels that held only a few hundred rays last time are walked by ONE single-CTA launch (tail mode of the
    // level kernel, rt_resident.cu) instead of one launch each: tail_from = first such level, 0 = none.
    int tail_from = 0;
    if (known && !getenv("RTB200_NO_TAIL_LEVELS")) {
        auto rays_at = [&](int l) { return (size_t)l < g.rays.size() ? g.rays[(size_t)l] : 0u; };   // beyond the history: empty
        int t = last + 1;
        while (t - 1 >= 2 && rays_at(t - 1) <= 512u) --t;
        if (last - t + 1 >= 2 && last - t + 1 <= RT_TAIL_FOLD_MAX) tail_from = t;
    }
    if (tail_from) {
        const size_t nt = (size_t)(last - tail_from + 1);
        if ((e = ctx->tail_dev.ensure(sizeof(TailLevel) * nt)) != cudaSuccess) return set_cuda_error(ctx, e, "tail table");
        if ((e = ctx->tail_host.ensure(sizeof(TailLevel) * nt)) != cudaSuccess) return set_cuda_error(ctx, e, "tail table (host)");
        TailLevel* tv = ctx->tail_host.as<TailLevel>();
        for (int l = tail_from; l <= last; ++l) {
            TailLevel& v = tv[l - tail_from];
            LevelBuf& cur = ctx->levels[(size_t)l];
            LevelBuf& nxt = ctx->levels[(size_t)l + 1];
            v.in_od0 = cur.od0.as<float4>(); v.in_od1 = cur.od1.as<float4>(); v.in_colour = cur.colour.as<float4>();
            v.out_od0 = nxt.od0.as<float4>(); v.out_od1 = nxt.od1.as<float4>(); v.links = cur.links.as<int4>();
            v.level = l; v.in_cap = (int)cap[(size_t)l]; v.spawn = l < max_b ? 1 : 0;
            v.out_cap = (l == last) ? 0u : (unsigned)cap[(size_t)l + 1];
        }
        CU(cudaMemcpyAsync(ctx->tail_dev.p, tv, sizeof(TailLevel) * nt, cudaMemcpyHostToDevice, ctx->stream));
    }

    // levels, folds and the resolve of a chunk form one chain of dependent launches: each kernel lets the
    // next one start its prologue early and waits for its predecessor itself (RTB200_NO_PDL=1: plain launches)
    const bool pdl = getenv("RTB200_NO_PDL") == nullptr;
    int chunk = 0;
    for (int ly0 = 0; ly0 < local_LH; ly0 += rows_per_chunk, ++chunk) {
        const int rows = std::min(rows_per_chunk, local_LH - ly0);
        CU(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned) * n_counters, ctx->stream));
        for (int level = 0; level <= last; ++level) {
            const bool tail_launch = tail_from && level == tail_from;
            const bool spawn = level < max_b;
            LevelBuf& cur = ctx->levels[(size_t)level];
            LevelBuf& nxt = ctx->levels[(size_t)level + 1];
            LevelParams p;
            memset(&p, 0, sizeof p);
            if (tail_launch) { p.tail = ctx->tail_dev.as<TailLevel>(); p.tail_count = last - tail_from + 1; }
            p.f = fp;
            if (!spawn) p.f.secondary = 0;
            p.blob = ctx->ds.blob; p.rl = ctx->ds.rl;
            p.level = level;
            p.chunk_ly0 = ly0; p.chunk_rows = rows; p.local_LH = local_LH;
            p.c0 = ctx->c0.as<float>();
            p.in_od0 = cur.od0.as<float4>(); p.in_od1 = cur.od1.as<float4>(); p.in_colour = cur.colour.as<float4>();
            if (level == 0) p.in_count = rows * fp.LW;
            else { p.in_count_dev = d_cnt + 2 * (level - 1); p.in_cap = (int)cap[(size_t)level]; p.in_count = p.in_cap; }
            p.out_od0 = nxt.od0.as<float4>(); p.out_od1 = nxt.od1.as<float4>();
            p.out_count = d_cnt + 2 * level; p.link_count = d_cnt + 2 * level + 1;
            p.links = cur.links.as<int4>();
            if (spawn) {                                    // the last launched level may only REPORT that it wanted children
                p.out_guarded = 1;
                p.out_cap = (level == last) ? 0u : (unsigned)cap[(size_t)level + 1];
                p.overflow = d_cnt + ovf;
            }
            p.stats = ctx->stats.as<unsigned long long>();
            p.cull = (fr->flags & RTB200_F_EXACT_FP) ? 0 : 1;
            p.fb_f32 = ctx->out_f32_dev; p.fb_rgba8 = ctx->out_rgba8_dev;
            p.pdl = pdl ? (level == 0 ? 1 : 2) : 0;            // the chain starts at the level-0 kernel
            const bool timed = (level == 0 && ly0 == 0);
            if (timed) CU(cudaEventRecord(ctx->ev[2], ctx->stream));
            CU(launch_resident_level(p, exact, ctx->sm_count, ctx->stream));
            if (timed) CU(cudaEventRecord(ctx->ev[3], ctx->stream));
            ++*launches;
            if (tail_launch) break;                             // that launch walked every remaining level
        }
        // bottom-up fold of the levels that spawned.  The deep levels (small queues) first, all of them in ONE
        // single-CTA launch; then one grid per large level.
        int fold_top = std::min(last, max_b - 1);
        if (fold_top == last) --fold_top;                                // (a probe level has no children to fold)
        int tail_first = fold_top + 1;
        while (tail_first - 1 >= 1 && cap[(size_t)tail_first - 1] <= 16384 && fold_top - (tail_first - 1) < RT_TAIL_FOLD_MAX) --tail_first;
        if (tail_first <= fold_top && fold_top - tail_first >= 1 && !getenv("RTB200_NO_TAIL_FOLD")) {
            TailFoldParams tf;
            memset(&tf, 0, sizeof tf);
            tf.first = tail_first; tf.last = fold_top;
            for (int l = tail_first; l <= fold_top; ++l) {
                tf.links[l - tail_first] = ctx->levels[(size_t)l].links.as<int4>();
                tf.colour[l - tail_first] = ctx->levels[(size_t)l].colour.as<float4>();
                tf.cap[l - tail_first] = (int)cap[(size_t)l];
            }
            tf.colour[fold_top - tail_first + 1] = ctx->levels[(size_t)fold_top + 1].colour.as<float4>();
            tf.counters = d_cnt; tf.mat = ctx->ds.mat; tf.pdl = pdl ? 2 : 0;
            CU(launch_tail_fold(tf, exact, ctx->stream));
            ++*launches;
            fold_top = tail_first - 1;
        }
        for (int l = fold_top; l >= 0; --l) {
            CombineParams c;
            memset(&c, 0, sizeof c);
            c.links = ctx->levels[(size_t)l].links.as<int4>();
            c.n_links = (l == 0) ? rows * fp.LW : (int)cap[(size_t)l];
            c.n_links_dev = d_cnt + 2 * l + 1;
            c.parent_is_level0 = (l == 0);
            c.parent_c0 = ctx->c0.as<float>();
            c.parent_colour = ctx->levels[(size_t)l].colour.as<float4>();
            c.child_colour = ctx->levels[(size_t)l + 1].colour.as<float4>();
            c.mat = ctx->ds.mat;
            c.pdl = pdl ? 2 : 0;
            CU(launch_combine(c, exact, ctx->stream));
            ++*launches;
        }
        ResolveParams r;
        memset(&r, 0, sizeof r);
        r.f = fp;
        r.c0 = ctx->c0.as<float>();
        r.chunk_py0 = ly0 / k;
        r.chunk_prows = std::min(rows / k, local_rows - r.chunk_py0);
        r.fb_f32 = ctx->out_f32_dev;
        r.fb_rgba8 = ctx->out_rgba8_dev;
        r.stats = ctx->stats.as<unsigned long long>();
        r.pdl = pdl ? 2 : 0;
        CU(launch_resolve(r, ctx->stream));
        ++*launches;
        CU(cudaMemcpyAsync(h_all + (size_t)chunk * n_counters, d_cnt, sizeof(unsigned) * n_counters, cudaMemcpyDeviceToHost, ctx->stream));
        const int rc2 = chunk_copy_out(ctx, fp.W, r.chunk_py0, r.chunk_prows);
        if (rc2 != RTB200_OK) return rc2;
    }
    CU(cudaStreamSynchronize(ctx->stream));                 // the one wait of the frame

    std::vector<unsigned> seen((size_t)max_b + 2, 0);
    unsigned depth = 1;
    for (int c = 0; c < n_chunks; ++c) {
        const unsigned* h = h_all + (size_t)c * n_counters;
        if (h[ovf]) *overflowed = true;
        for (int l = 1; l <= last; ++l) {
            const unsigned n = h[2 * (l - 1)];
            if (n > cap[(size_t)l]) *overflowed = true;      // (cannot happen without the flag; belt and braces)
            seen[(size_t)l] = std::max(seen[(size_t)l], n);
            if (n) depth = std::max(depth, (unsigned)l + 1);
        }
    }
    if (*overflowed) { g.valid = false; return RTB200_OK; }
    g.valid = true; g.scene_serial = ctx->scene_serial;
    g.W = fp.W; g.H = fp.H; g.k = k; g.max_bounces = max_b; g.children = children;
    g.rows_per_chunk = rows_per_chunk; g.local_rows = local_rows;
    g.rays = seen;
    st->depth_levels = depth;
    st->chunks = (uint32_t)n_chunks;
    st->path = 0;
    st->ms_primary_kernel = -1.0;
    return RTB200_OK;
}

// scene-resident path: one fused kernel per level, host reads the child count between levels
static int render_resident(rtb200_ctx* ctx, const rtb200_frame* fr, const FrameParams& fp, int local_rows,
                           bool exact, rtb200_stats* st, uint32_t* launches)
{
    const int k = fp.k;
    const int local_LH = local_rows * k;
    int children = ctx->max_children;
    if (fr->min_energy < 0.f) children = 2;                 // 0 > negative: every hit spawns both
    if (!fp.secondary || fr->max_bounces <= 0) children = 0;

    int rows_per_chunk = (int)((MAX_CHUNK_SAMPLES / (unsigned)fp.LW) / (16u * k)) * 16 * k;
    if (rows_per_chunk < 16 * k) rows_per_chunk = 16 * k;
    if (rows_per_chunk > local_LH) rows_per_chunk = ((local_LH + 16 * k - 1) / (16 * k)) * 16 * k;
    // overlapped read-back (host pointers given).  The copy engine is the slower side (config 2: 12 pixels/ns
    // over PCIe against 24 rendered), so what counts is how soon the FIRST copy starts and how few copies there
    // are: frames that are one kernel per chunk use a doubling schedule -- 64 rows, 128, 256, ... --, frames
    // with a level loop per chunk about 8 equal chunks, never smaller than 64 pixel rows.
    const bool overlap = ctx->host_f32 || ctx->host_rgba8;
    const bool doubling = overlap && children == 0;
    if (overlap && !doubling) {
        int want = (((local_rows + 7) / 8 + 15) / 16) * 16;
        if (want < 64) want = 64;
        if (want * k < rows_per_chunk) rows_per_chunk = want * k;
    }

    if (children > 0 && !getenv("RTB200_SYNC_LEVELS")) {
        // secondary rays: all levels of a chunk enqueued at once, one wait per frame
        bool overflowed = false;
        const int arc = render_resident_async(ctx, fr, fp, local_rows, exact, children, rows_per_chunk, st, launches, &overflowed);
        if (arc != RTB200_OK || !overflowed) return arc;
        // a queue guess was too small: forget the attempt (counters, launches, copies) and take the exact route
        CU(cudaMemsetAsync(ctx->stats.p, 0, sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT, ctx->stream));
        *launches = 0;
        if (ctx->copy_chunks) { CU(cudaStreamSynchronize(ctx->copy_stream)); ctx->copy_chunks = 0; }
        ++ctx->level_fallbacks;
        st->level_loop_redone = 1;
    }

    cudaError_t e;
    if ((e = ctx->c0.ensure(sizeof(float) * 3 * (size_t)rows_per_chunk * fp.LW)) != cudaSuccess) return set_cuda_error(ctx, e, "c0");
    const size_t n_counters = 2 * ((size_t)std::max(fr->max_bounces, 0) + 2);
    if ((e = ctx->counters.ensure(sizeof(unsigned) * n_counters)) != cudaSuccess) return set_cuda_error(ctx, e, "counters");
    if ((e = ctx->h_small.ensure(16384)) != cudaSuccess) return set_cuda_error(ctx, e, "pinned small");
    unsigned* h_cnt = ctx->h_small.as<unsigned>();
    unsigned* d_cnt = ctx->counters.as<unsigned>();
    unsigned max_level_seen = 0, chunks = 0;
    std::vector<unsigned> seen_rays(1, 0);

    int step_rows = doubling ? std::min(64 * k, rows_per_chunk) : rows_per_chunk;
    for (int ly0 = 0, rows = 0; ly0 < local_LH; ly0 += rows, ++chunks, step_rows = doubling ? std::min(2 * step_rows, rows_per_chunk) : step_rows) {
        rows = std::min(step_rows, local_LH - ly0);
        if (children > 0) CU(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned) * n_counters, ctx->stream));
        std::vector<int> n_rays, n_links;
        n_rays.push_back(rows * fp.LW);
        int level = 0;
        bool resolved = false;
        for (;; ++level) {
            const int n_in = n_rays[level];
            if (n_in == 0) break;
            const bool spawn = children > 0 && level < fr->max_bounces;
            if ((int)ctx->levels.size() < level + 2) ctx->levels.resize((size_t)level + 2);
            LevelBuf& cur = ctx->levels[level];
            LevelBuf& nxt = ctx->levels[level + 1];
            if (spawn) {
                const size_t cap = (size_t)n_in * children;
                if ((e = nxt.od0.ensure(sizeof(float4) * cap)) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
                if ((e = nxt.od1.ensure(sizeof(float4) * cap)) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
                if ((e = nxt.colour.ensure(sizeof(float4) * cap)) != cudaSuccess) return set_cuda_error(ctx, e, "queue");
                if ((e = cur.links.ensure(sizeof(int4) * (size_t)n_in)) != cudaSuccess) return set_cuda_error(ctx, e, "links");
            }
            LevelParams p;
            memset(&p, 0, sizeof p);
            p.f = fp;
            if (!spawn) p.f.secondary = 0;
            p.blob = ctx->ds.blob; p.rl = ctx->ds.rl;
            p.level = level;
            p.chunk_ly0 = ly0; p.chunk_rows = rows; p.local_LH = local_LH;
            p.c0 = ctx->c0.as<float>();
            p.in_od0 = cur.od0.as<float4>(); p.in_od1 = cur.od1.as<float4>(); p.in_colour = cur.colour.as<float4>();
            p.in_count = n_in;
            p.out_od0 = nxt.od0.as<float4>(); p.out_od1 = nxt.od1.as<float4>();
            p.out_count = d_cnt + 2 * level; p.link_count = d_cnt + 2 * level + 1;
            p.links = cur.links.as<int4>();
            p.stats = ctx->stats.as<unsigned long long>();
            // no secondary rays and no supersampling: the level-0 kernel is the whole frame
            const bool direct = (level == 0 && !spawn && k == 1);
            p.direct = direct ? 1 : 0;
            p.cull = (fr->flags & RTB200_F_EXACT_FP) ? 0 : 1;
            p.fb_f32 = ctx->out_f32_dev; p.fb_rgba8 = ctx->out_rgba8_dev;
            if (direct && ctx->peer_cur && ly0 + rows >= local_LH) {
                // the frame's last producing kernel carries the completion signal to rank 0 itself
                const PeerFrame* pf = static_cast<const PeerFrame*>(ctx->peer_cur);
                if (pf->flag) { p.peer_flag = pf->flag; p.peer_done = pf->done; p.peer_seq = pf->seq; ctx->peer_signalled = true; }
            }
            const bool timed = (level == 0 && ly0 == 0);       // dominant kernel, timed on its own stream
            if (timed) CU(cudaEventRecord(ctx->ev[2], ctx->stream));
            CU(launch_resident_level(p, exact, ctx->sm_count, ctx->stream));
            if (direct) resolved = true;
            if (timed) CU(cudaEventRecord(ctx->ev[3], ctx->stream));
            ++*launches;
            if (!spawn) { n_links.push_back(0); ++level; break; }
            CU(cudaMemcpyAsync(h_cnt, d_cnt + 2 * level, 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            n_rays.push_back((int)h_cnt[0]);
            n_links.push_back((int)h_cnt[1]);
        }
        for (size_t l = 1; l < n_rays.size(); ++l) {           // what this chunk needed: the next frame's guess
            if (seen_rays.size() <= l) seen_rays.resize(l + 1, 0);
            seen_rays[l] = std::max(seen_rays[l], (unsigned)n_rays[l]);
        }
        const int levels_used = (int)n_links.size();          // levels that ran a kernel
        max_level_seen = std::max<unsigned>(max_level_seen, (unsigned)levels_used);
        for (int l = levels_used - 1; l >= 0; --l) {
            if (n_links[l] <= 0) continue;
            CombineParams c;
            memset(&c, 0, sizeof c);
            c.links = ctx->levels[l].links.as<int4>();
            c.n_links = n_links[l];
            c.parent_is_level0 = (l == 0);
            c.parent_c0 = ctx->c0.as<float>();
            c.parent_colour = ctx->levels[l].colour.as<float4>();
            c.child_colour = ctx->levels[l + 1].colour.as<float4>();
            c.mat = ctx->ds.mat;
            CU(launch_combine(c, exact, ctx->stream));
            ++*launches;
        }
        if (!resolved) {
            ResolveParams r;
            memset(&r, 0, sizeof r);
            r.f = fp;
            r.c0 = ctx->c0.as<float>();
            r.chunk_py0 = ly0 / k;
            r.chunk_prows = std::min(rows / k, local_rows - r.chunk_py0);
            r.fb_f32 = ctx->out_f32_dev;
            r.fb_rgba8 = ctx->out_rgba8_dev;
            r.stats = ctx->stats.as<unsigned long long>();
            CU(launch_resolve(r, ctx->stream));
            ++*launches;
        }
        {
            const int py0 = ly0 / k;
            const int rc2 = chunk_copy_out(ctx, fp.W, py0, std::min(rows / k, local_rows - py0));
            if (rc2 != RTB200_OK) return rc2;
        }
    }
    if (children > 0) {
        rtb200_ctx::LevelGuess& g = ctx->level_guess;
        g.valid = true; g.scene_serial = ctx->scene_serial;
        g.W = fp.W; g.H = fp.H; g.k = k; g.max_bounces = fr->max_bounces; g.children = children;
        g.rows_per_chunk = rows_per_chunk; g.local_rows = local_rows;
        g.rays = seen_rays;
    }
    st->depth_levels = max_level_seen;
    st->chunks = chunks;
    st->path = 0;
    st->ms_primary_kernel = -1.0;                              // filled by the caller after the final sync
    // brute force by construction: every nearest-hit query and every shadow query that is not
    // cut short by the all-zero early-out tests every surface; counted exactly on the host
    return RTB200_OK;
}

static int read_back(rtb200_ctx* ctx, float* out_f32, uint8_t* out_rgba8, double* ms)
{
    if (!out_f32 && !out_rgba8) return RTB200_OK;
    const bool full = ctx->last_gathered;
    const size_t rows = full ? (size_t)ctx->last_H : (size_t)ctx->last_local_rows;
    const size_t px = rows * (size_t)ctx->last_W;
    CU(cudaEventRecord(ctx->ev[2], ctx->stream));
    if (ctx->host_bands && !full) {                          // this rank's bands into the caller's full frame
        ctx->host_f32 = out_f32; ctx->host_rgba8 = out_rgba8;
        const cudaError_t e = copy_rows_out(ctx, ctx->last_W, 0, ctx->last_local_rows, ctx->stream);
        ctx->host_f32 = nullptr; ctx->host_rgba8 = nullptr;
        if (e != cudaSuccess) return set_cuda_error(ctx, e, "band read-back");
    } else {
        if (out_f32)
            CU(cudaMemcpyAsync(out_f32, full ? static_cast<void*>(ctx->full_f32.as<unsigned char>() + ctx->full_off32) : ctx->fb_f32.p, px * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        if (out_rgba8)
            CU(cudaMemcpyAsync(out_rgba8, full ? static_cast<void*>(ctx->full_rgba8.as<unsigned char>() + ctx->full_off8) : ctx->fb_rgba8.p, px * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    CU(cudaEventRecord(ctx->ev[3], ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, ctx->ev[2], ctx->ev[3]));
    if (ms) *ms = t;
    return RTB200_OK;
}

extern "C" int rtb200_render(rtb200_ctx* ctx, const rtb200_frame* fr, float* out_f32, uint8_t* out_rgba8, rtb200_stats* stats)
{
    if (!ctx || !fr) return RTB200_E_ARG;
    if (!ctx->have_scene) return fail(ctx, RTB200_E_STATE, "rtb200_render before rtb200_set_scene");
    if (fr->width <= 0 || fr->height <= 0 || fr->spp_grid < 0 || fr->spp_grid > 64) return fail(ctx, RTB200_E_ARG, "bad frame size");
    if ((long long)fr->width * std::max(fr->spp_grid, 1) > (1 << 20)) return fail(ctx, RTB200_E_ARG, "frame too wide");
    if (fr->max_bounces > RTB200_MAX_BOUNCES) return fail(ctx, RTB200_E_ARG, "max_bounces above RTB200_MAX_BOUNCES");
    CU(cudaSetDevice(ctx->device));
    const bool exact = (fr->flags & RTB200_F_FUSED_FP) == 0;     // reference rounding unless fused is requested
    FrameParams fp;
    fill_frame_params(fr, ctx->n_ranks, ctx->rank, fp);
    const int band = fp.band_lrows / fp.k;
    const int local_rows = local_row_count(fp.H, band, ctx->n_ranks, ctx->rank);
    rtb200_stats st;
    memset(&st, 0, sizeof st);
    uint32_t launches = 0;

    cudaError_t e;
    const size_t lpx = (size_t)std::max(local_rows, 1) * fp.W;
    if ((e = ctx->fb_f32.ensure(lpx * 3 * sizeof(float))) != cudaSuccess) return set_cuda_error(ctx, e, "fb_f32");
    if ((e = ctx->fb_rgba8.ensure(lpx * 4)) != cudaSuccess) return set_cuda_error(ctx, e, "fb_rgba8");
    const size_t stat_bytes = sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT;
    if ((e = ctx->stats.ensure(stat_bytes)) != cudaSuccess) return set_cuda_error(ctx, e, "stats");
    if ((e = ctx->h_small.ensure(16384)) != cudaSuccess) return set_cuda_error(ctx, e, "pinned small");
    if (!ctx->stats_clean) CU(cudaMemsetAsync(ctx->stats.p, 0, stat_bytes, ctx->stream));
    ctx->stats_clean = false;

    // where the pixels go: the local band buffers, or -- several GPUs with peer access -- straight
    // into rank 0's full frame over NVLink (the producing kernels do the "gather" themselves)
    ctx->out_f32_dev = ctx->fb_f32.as<float>(); ctx->out_rgba8_dev = ctx->fb_rgba8.as<uchar4>();
    // RTB200_F_HOST_BANDS: every rank writes its own bands into the caller's FULL host frame (no gather on the device)
    const bool host_bands = ctx->n_ranks > 1 && (fr->flags & RTB200_F_HOST_BANDS) != 0;
    ctx->host_bands = host_bands; ctx->host_band_rows = band;
    const bool multi = ctx->comm && ctx->n_ranks > 1 && !host_bands;
    // (a float host pointer without RTB200_F_GATHER_F32 is simply not served on several GPUs: the decision must not
    // depend on one rank's arguments, or the ranks would disagree about taking part in the frame)
    const bool gather_f32 = (fr->flags & RTB200_F_GATHER_F32) != 0;
    if (multi && !gather_f32) out_f32 = nullptr;
    bool peer = false;
    PeerFrame pf;
    ctx->peer_cur = nullptr; ctx->peer_signalled = false;
    if (multi) {
        // rank 0 entering this frame has consumed the previous one: it tells the peers so (ack)
        const int prc = comm_peer_begin(ctx, fp.W, fp.H, gather_f32, true, false, &pf, ctx->stream);
        if (prc != RTB200_OK) return prc;
        if (pf.rgba8) {
            peer = true;
            fp.fb_global = 1;
            ctx->out_rgba8_dev = static_cast<uchar4*>(pf.rgba8);
            ctx->out_f32_dev = static_cast<float*>(pf.f32);       // null unless the float frame is gathered too
            ctx->peer_cur = &pf;
        }
    }

    // path selection
    const size_t resident_limit = ctx->smem_optin > 4096 ? ctx->smem_optin - 2048 : 0;
    bool use_stream = (size_t)ctx->ds.rl.bytes > RESIDENT_AUTO_BYTES || (fr->flags & (RTB200_F_FORCE_STREAM | RTB200_F_USE_BVH));
    if (fr->flags & RTB200_F_FORCE_RESIDENT) {
        if ((size_t)ctx->ds.rl.bytes > resident_limit) return fail(ctx, RTB200_E_ARG, "scene does not fit in shared memory");
        use_stream = false;
    }

    // one GPU and host pointers given: read finished row chunks back while the next ones render
    const bool overlap = !multi && (out_f32 || out_rgba8) && (size_t)fp.W * fp.H >= (256u << 10);   // (host bands: overlapped too)
    ctx->host_f32 = overlap ? out_f32 : nullptr;
    ctx->host_rgba8 = overlap ? out_rgba8 : nullptr;
    ctx->copy_chunks = 0;

    CU(cudaEventRecord(ctx->ev[0], ctx->stream));
    int rc = RTB200_OK;
    if (local_rows > 0) {
        if (use_stream) rc = render_streamed(ctx, fr, fp, local_rows, exact, &st, &launches);
        else            rc = render_resident(ctx, fr, fp, local_rows, exact, &st, &launches);
    }
    ctx->host_f32 = nullptr; ctx->host_rgba8 = nullptr;
    ctx->peer_cur = nullptr;
    if (rc != RTB200_OK && peer && ctx->rank != 0 && !ctx->peer_signalled) comm_peer_signal(ctx, pf, ctx->stream);   // never leave rank 0 waiting
    if (rc != RTB200_OK) { cudaStreamSynchronize(ctx->stream); cudaStreamSynchronize(ctx->copy_stream); (void)cudaGetLastError(); return rc; }
    CU(cudaEventRecord(ctx->ev[1], ctx->stream));
    const bool copied = overlap && ctx->copy_chunks > 0;
    if (copied) CU(cudaEventRecord(ctx->copy_ev[3], ctx->copy_stream));

    ctx->last_W = fp.W; ctx->last_H = fp.H; ctx->last_local_rows = local_rows; ctx->last_gathered = false;
    if (multi) {
        if (peer) {
            // ranks > 0: the completion word, unless the producing kernel wrote it itself;
            // rank 0: the stream waits for every peer's word before anything reads the frame
            if (ctx->rank != 0 && !ctx->peer_signalled) rc = comm_peer_signal(ctx, pf, ctx->stream);
            if (rc == RTB200_OK && ctx->rank == 0) {
                CU(cudaEventRecord(ctx->ev[6], ctx->stream));
                rc = comm_peer_wait(ctx, pf, ctx->stream);
                CU(cudaEventRecord(ctx->ev[7], ctx->stream));
            }
        } else rc = comm_gather_frame(ctx, fp.W, fp.H, band, gather_f32, true, &st.ms_gather);
        if (rc != RTB200_OK) return rc;
        ctx->last_gathered = (ctx->rank == 0);
    }

    // statistics
    unsigned long long* h_st = reinterpret_cast<unsigned long long*>(ctx->h_small.as<unsigned char>() + 1024);
    CU(cudaMemcpyAsync(h_st, ctx->stats.p, stat_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemsetAsync(ctx->stats.p, 0, stat_bytes, ctx->stream));     // for the NEXT frame, off its critical path
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->stats_clean = true;
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    st.ms_render = ms;
    if (peer && ctx->rank == 0) {
        float w = 0.f;
        CU(cudaEventElapsedTime(&w, ctx->ev[6], ctx->ev[7]));
        st.ms_gather = w;                                       // how long rank 0 waited for the slowest peer
    }
    if (st.ms_primary_kernel < 0.0) {                          // level-0 kernel of the first chunk
        float k0 = 0.f;
        CU(cudaEventElapsedTime(&k0, ctx->ev[2], ctx->ev[3]));
        st.ms_primary_kernel = k0;
        if (st.path == 0) st.ms_trace_nearest = k0;
    }
    unsigned long long sum[ST_COUNT] = {0};
    for (int s = 0; s < RT_STAT_SLOTS; ++s) for (int c = 0; c < ST_COUNT; ++c) sum[c] += h_st[s * ST_COUNT + c];
    st.rays_primary = (uint64_t)local_rows * fp.k * (uint64_t)fp.LW;
    if (st.path == 0) st.rays_shadow = sum[ST_SHADOW];
    st.rays_reflect = sum[ST_REFLECT]; st.rays_refract = sum[ST_REFRACT];
    st.nan_pixels = sum[ST_NAN];
    if (sum[ST_PEER_TIMEOUT]) return fail(ctx, RTB200_E_NCCL, "peer frame: rank 0 did not acknowledge the previous frame within 2 s; this frame may be torn");
    if (st.path == 0) {
        // resident kernels count surface tests in one counter; split by the scene's type mix
        const double n = std::max(1, ctx->ds.rl.n_surf);
        st.tests_triangle = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_tri / n);
        st.tests_quadric  = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_quad / n);
        st.tests_plane    = (uint64_t)llround((double)sum[ST_TEST_TRI] * ctx->ds.n_plane / n);
    } else {
        st.shadow_tests_triangle = sum[ST_TEST_TRI_SHADOW];
        st.shadow_bvh_nodes = sum[ST_BVH_NODES_SHADOW];
        st.tests_triangle_executed = sum[ST_TEST_TRI_EXEC];
        st.shadow_fixups = sum[ST_SHADOW_FIXUP];
        st.bvh_out_of_radius = sum[ST_BVH_OUT_OF_RADIUS];
        st.tests_triangle = sum[ST_TEST_TRI] + sum[ST_TEST_TRI_SHADOW] + ctx->host_tests[0];
        st.tests_quadric = sum[ST_TEST_QUAD] + ctx->host_tests[1];
        st.tests_plane = sum[ST_TEST_PLANE] + ctx->host_tests[2];
        st.bvh_nodes_visited = sum[ST_BVH_NODES] + sum[ST_BVH_NODES_SHADOW];
        if (sum[ST_BVH_OVERFLOW]) return fail(ctx, RTB200_E_CAPACITY, "LBVH traversal stack overflow");
    }
    st.kernel_launches = launches;
    if (ctx->upload_pending) {                   // the frame has completed, so has the upload before it
        float h = 0.f;
        if (cudaEventElapsedTime(&h, ctx->ev[4], ctx->ev[5]) == cudaSuccess) ctx->last_h2d_ms = h;
        ctx->upload_pending = false;
    }
    st.ms_h2d = ctx->last_h2d_ms;

    if (copied) {
        CU(cudaStreamSynchronize(ctx->copy_stream));
        float t = 0.f;
        CU(cudaEventElapsedTime(&t, ctx->copy_ev[2], ctx->copy_ev[3]));
        st.ms_d2h = t;                                          // mostly hidden under the rendering
    } else if (ctx->rank == 0 || !ctx->last_gathered) {
        // every rank when it owns what it reads (host bands, or sharding without a communicator); else rank 0 only
        const bool want = (ctx->n_ranks <= 1) || ctx->rank == 0 || ctx->host_bands || !ctx->comm;
        if (want) {
            rc = read_back(ctx, out_f32, out_rgba8, &st.ms_d2h);
            if (rc != RTB200_OK) return rc;
        }
    }
    if (stats) *stats = st;
    return RTB200_OK;
}

extern "C" int rtb200_read_frame(rtb200_ctx* ctx, float* out_f32, uint8_t* out_rgba8)
{
    if (!ctx) return RTB200_E_ARG;
    if (ctx->last_W <= 0) return fail(ctx, RTB200_E_STATE, "no frame rendered yet");
    CU(cudaSetDevice(ctx->device));
    return read_back(ctx, out_f32, out_rgba8, nullptr);
}

// ---------------------------------------------------------------------------------------
// Ray-level queries
// ---------------------------------------------------------------------------------------
// After a query through the streamed / LBVH kernels: the device counters they wrote.  A traversal whose stack
// overflowed is an error here exactly as in rtb200_render (the hit would be wrong, silently).
static int query_counters_check(rtb200_ctx* ctx)
{
    const size_t stat_bytes = sizeof(unsigned long long) * RT_STAT_SLOTS * ST_COUNT;
    cudaError_t e;
    if ((e = ctx->h_small.ensure(16384)) != cudaSuccess) return set_cuda_error(ctx, e, "pinned small");
    unsigned long long* h_st = reinterpret_cast<unsigned long long*>(ctx->h_small.as<unsigned char>() + 1024);
    CU(cudaMemcpyAsync(h_st, ctx->stats.p, stat_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemsetAsync(ctx->stats.p, 0, stat_bytes, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->stats_clean = true;
    unsigned long long overflow = 0;
    for (int s = 0; s < RT_STAT_SLOTS; ++s) overflow += h_st[s * ST_COUNT + ST_BVH_OVERFLOW];
    if (overflow) return fail(ctx, RTB200_E_CAPACITY, "LBVH traversal stack overflow");
    return RTB200_OK;
}

extern "C" int rtb200_trace_nearest(rtb200_ctx* ctx, int n, const float* o3, const float* d3, uint32_t flags,
                                    int32_t* out_order, float* out_t, float* out_n3)
{
    if (!ctx || n < 0 || (n && (!o3 || !d3 || !out_order || !out_t || !out_n3))) return RTB200_E_ARG;
    if (!ctx->have_scene) return fail(ctx, RTB200_E_STATE, "no scene");
    if (n == 0) return RTB200_OK;
    CU(cudaSetDevice(ctx->device));
    cudaError_t e;
    const size_t b3 = sizeof(float) * 3 * (size_t)n;
    if ((e = ctx->q_in0.ensure(b3)) != cudaSuccess || (e = ctx->q_in1.ensure(b3)) != cudaSuccess ||
        (e = ctx->q_out0.ensure(sizeof(int) * (size_t)n)) != cudaSuccess || (e = ctx->q_out1.ensure(sizeof(float) * (size_t)n)) != cudaSuccess ||
        (e = ctx->q_out2.ensure(b3)) != cudaSuccess) return set_cuda_error(ctx, e, "query buffers");
    CU(cudaMemcpyAsync(ctx->q_in0.p, o3, b3, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->q_in1.p, d3, b3, cudaMemcpyHostToDevice, ctx->stream));
    const bool q_stream = ((size_t)ctx->ds.rl.bytes > RESIDENT_AUTO_BYTES || (flags & (RTB200_F_FORCE_STREAM | RTB200_F_USE_BVH))) && !(flags & RTB200_F_FORCE_RESIDENT);
    if (flags & RTB200_F_USE_BVH) { const int rc = bvh_prepare(ctx, o3, n); if (rc != RTB200_OK) return rc; }
    if (q_stream) {
        const int rc = stream_query_nearest(ctx, n, flags);
        if (rc != RTB200_OK) return rc;
    } else
    CU(launch_query_nearest(ctx->ds, n, ctx->q_in0.as<float>(), ctx->q_in1.as<float>(), ctx->q_out0.as<int>(),
                            ctx->q_out1.as<float>(), ctx->q_out2.as<float>(), (flags & RTB200_F_FUSED_FP) == 0,
                            (flags & RTB200_F_EXACT_FP) == 0, ctx->stream));
    CU(cudaMemcpyAsync(out_order, ctx->q_out0.p, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(out_t, ctx->q_out1.p, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(out_n3, ctx->q_out2.p, b3, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (q_stream) return query_counters_check(ctx);
    return RTB200_OK;
}

extern "C" int rtb200_trace_shadow(rtb200_ctx* ctx, int n, const float* p3, const float* l3, uint32_t flags, float* out3)
{
    if (!ctx || n < 0 || (n && (!p3 || !l3 || !out3))) return RTB200_E_ARG;
    if (!ctx->have_scene) return fail(ctx, RTB200_E_STATE, "no scene");
    if (n == 0) return RTB200_OK;
    CU(cudaSetDevice(ctx->device));
    cudaError_t e;
    const size_t b3 = sizeof(float) * 3 * (size_t)n;
    if ((e = ctx->q_in0.ensure(b3)) != cudaSuccess || (e = ctx->q_in1.ensure(b3)) != cudaSuccess ||
        (e = ctx->q_out2.ensure(b3)) != cudaSuccess) return set_cuda_error(ctx, e, "query buffers");
    CU(cudaMemcpyAsync(ctx->q_in0.p, p3, b3, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->q_in1.p, l3, b3, cudaMemcpyHostToDevice, ctx->stream));
    const bool q_stream = ((size_t)ctx->ds.rl.bytes > RESIDENT_AUTO_BYTES || (flags & (RTB200_F_FORCE_STREAM | RTB200_F_USE_BVH))) && !(flags & RTB200_F_FORCE_RESIDENT);
    if (flags & RTB200_F_USE_BVH) { const int rc = bvh_prepare(ctx, p3, n); if (rc != RTB200_OK) return rc; }
    if (q_stream) {
        const int rc = stream_query_shadow(ctx, n, flags);
        if (rc != RTB200_OK) return rc;
    } else
    CU(launch_query_shadow(ctx->ds, n, ctx->q_in0.as<float>(), ctx->q_in1.as<float>(), ctx->q_out2.as<float>(),
                           (flags & RTB200_F_FUSED_FP) == 0, (flags & RTB200_F_EXACT_FP) == 0, ctx->stream));
    CU(cudaMemcpyAsync(out3, ctx->q_out2.p, b3, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (q_stream) return query_counters_check(ctx);
    return RTB200_OK;
}

// Pure host arithmetic (no GPU needed): which rows of an H-row image rank `rank` of `n_ranks` renders.
extern "C" int rtb200_band_map(int height, int band_rows, int n_ranks, int rank, int32_t* out_global_rows, int capacity, int* n_local)
{
    if (height < 0 || n_ranks < 1 || rank < 0 || rank >= n_ranks || !n_local) return RTB200_E_ARG;
    const int band = band_rows > 0 ? band_rows : 16;
    const int rows = local_row_count(height, band, n_ranks, rank);
    *n_local = rows;
    if (!out_global_rows) return RTB200_OK;
    if (capacity < rows) return RTB200_E_ARG;
    for (int i = 0; i < rows; ++i) out_global_rows[i] = band_global_row(band, n_ranks, rank, i);
    return RTB200_OK;
}

extern "C" int rtb200_comm_unique_id(uint8_t id[RTB200_UNIQUE_ID_BYTES]) { return comm_unique_id(id); }
extern "C" int rtb200_comm_init(rtb200_ctx* ctx, int rank, int n_ranks, const uint8_t id[RTB200_UNIQUE_ID_BYTES])
{
    if (!ctx || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return RTB200_E_ARG;
    return comm_init(ctx, rank, n_ranks, id);
}
