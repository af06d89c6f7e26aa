/* rt_oracle.c -- TEST INFRASTRUCTURE ONLY (see rt_oracle.h).  Parity: PINNED against
 * oracle/_ref (the reference's own compiled code) and tests/golden/.
 *
 * CPU restatement of the reference's per-pixel path, one C function per reference function,
 * with the reference's exact operation order (every expression is evaluated left to right
 * in binary32, no contraction: build with -ffp-contract=off, no -ffast-math).
 * file:line citations are relative to Rodyll/Ray-Tracing-GPU, Projet-INF8702/.
 */
#include "rt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define EPS 1.0e-2f              /* MathUtils.h:35 */

typedef struct { float x, y, z; } v3;
typedef struct { float r, g, b; } col;

typedef struct {                 /* CRayon, Rayon.h:16-52 */
    v3 o, d;
    float ior, energy;
    int bounces;
} ray_t;

typedef struct {                 /* CIntersection, Intersection.h:23-51; defaults Intersection.cpp:15-19 */
    int kind, idx;               /* kind: -1 none, 0 triangle, 1 quadric, 2 plane */
    float t;
    v3 n;
} hit_t;

typedef struct { int kind, idx, order, mat; } surf_ref;

typedef struct {                 /* insertion-ordered view of the packed scene (m_Surfaces order) */
    const rtb200_scene* s;
    int n;
    surf_ref* list;
} scene_view;

static uint64_t g_counts[4];

/* ---- CVecteur3 (Vecteur3.h) ------------------------------------------------------------ */
static inline v3 v3_make(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 v3_sub(v3 a, v3 b) { return v3_make(a.x - b.x, a.y - b.y, a.z - b.z); }        /* :379-382 */
static inline v3 v3_add(v3 a, v3 b) { return v3_make(a.x + b.x, a.y + b.y, a.z + b.z); }        /* :362-365 */
static inline v3 v3_scale(v3 a, float s) { return v3_make(a.x * s, a.y * s, a.z * s); }         /* :396-416 */
static inline v3 v3_neg(v3 a) { return v3_make(-a.x, -a.y, -a.z); }
static inline float v3_dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }            /* ProdScal :454-457 */
static inline v3 v3_cross(v3 a, v3 b)                                                            /* ProdVect :471-476 */
{
    return v3_make(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float v3_norm(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }          /* Norme :489-492 */
static inline v3 v3_normalise(v3 a)                                                              /* Normaliser :522-534 */
{
    v3 r = { 0, 0, 0 };
    float len = v3_norm(a);
    if (len > EPS) {
        len = 1.0f / len;
        r = v3_scale(a, len);
    }
    return r;
}
static inline v3 v3_reflect(v3 v, v3 n)                                                          /* Reflect :548-551 */
{
    return v3_sub(v, v3_scale(n, 2.0f * v3_dot(v, n)));
}
static inline v3 v3_refract(v3 v, v3 n, float ratio)                                             /* Refract :566-581 */
{
    v3 z = v3_scale(v3_sub(v, v3_scale(n, v3_dot(v, n))), ratio);
    /* sqrt(1 - pow(Norme(Z), 2)): pow(float,int) promotes to double, the root is taken in
     * double and narrowed to REAL when it multiplies the normal */
    double nz = (double)v3_norm(z);
    float  s  = (float)sqrt(1.0 - nz * nz);
    v3 t = v3_sub(z, v3_scale(n, s));
    if (v3_dot(t, n) < 0) return t;
    return v3_reflect(v, n);
}

/* ---- surfaces ------------------------------------------------------------------------------ */
static inline float absf_ref(float a) { return a > 0 ? a : -a; }                                 /* Abs<> MathUtils.h:60-63 */

/* CTriangle::Intersection, Triangle.cpp:127-172 */
static inline void hit_triangle(const rtb200_scene* s, int i, const ray_t* r, hit_t* h)
{
    const float* p = s->tri_points + 9 * (size_t)i;
    v3 p0 = v3_make(p[0], p[1], p[2]), p1 = v3_make(p[3], p[4], p[5]), p2 = v3_make(p[6], p[7], p[8]);
    h->kind = -1; h->t = -1.0f; h->n = v3_make(0, 0, 0);
    v3 e1 = v3_sub(p1, p0), e2 = v3_sub(p2, p0);
    v3 P = v3_cross(r->d, e2);
    float det = v3_dot(e1, P);
    if (absf_ref(det) < EPS) return;
    float inv = 1.0f / det;
    v3 S = v3_sub(r->o, p0);
    float u = v3_dot(S, P) * inv;
    if (u < 0 || u > 1) return;
    v3 Q = v3_cross(S, e1);
    float v = v3_dot(r->d, Q) * inv;
    if (v < 0 || u + v > 1) return;
    h->kind = 0; h->idx = i;
    h->t = v3_dot(e2, Q) * inv;
    h->n = v3_make(s->tri_normal[3 * (size_t)i], s->tri_normal[3 * (size_t)i + 1], s->tri_normal[3 * (size_t)i + 2]);
}

/* CPlan::Intersection, Plan.cpp:128-144 */
static inline void hit_plane(const rtb200_scene* s, int i, const ray_t* r, hit_t* h)
{
    const float* c = s->plane_coef + 4 * (size_t)i;
    v3 n = v3_make(c[0], c[1], c[2]);
    h->kind = -1; h->t = -1.0f; h->n = v3_make(0, 0, 0);
    float vd = v3_dot(n, r->d);
    if (absf_ref(vd) > EPS) {
        h->kind = 2; h->idx = i;
        h->t = -(v3_dot(n, r->o) + c[3]) / vd;
        h->n = n;
    }
}

/* CQuadrique::Intersection, Quadrique.cpp:160-249 */
static inline void hit_quadric(const rtb200_scene* s, int i, const ray_t* r, hit_t* h)
{
    const float* c = s->quad_coef + 10 * (size_t)i;
    const float qx = c[0], qy = c[1], qz = c[2];     /* m_Quadratique */
    const float mx = c[3], my = c[4], mz = c[5];     /* m_Mixte (yz, xz, xy) */
    const float lx = c[6], ly = c[7], lz = c[8];     /* m_Lineaire */
    const float K = c[9];
    const v3 d = r->d, o = r->o;
    h->kind = -1; h->t = -1.0f; h->n = v3_make(0, 0, 0);

    const float A = d.x * (qx * d.x + mz * d.y + my * d.z) +
                    d.y * (qy * d.y + mx * d.z) +
                    d.z * (qz * d.z);
    const float B = d.x * (qx * o.x + 0.5f * (mz * o.y + my * o.z + lx)) +
                    d.y * (qy * o.y + 0.5f * (mz * o.x + mx * o.z + ly)) +
                    d.z * (qz * o.z + 0.5f * (my * o.x + mx * o.y + lz));
    const float C = o.x * (qx * o.x + mz * o.y + my * o.z + lx) +
                    o.y * (qy * o.y + mx * o.z + ly) +
                    o.z * (qz * o.z + lz) +
                    K;
    if (A != 0.0f) {
        float ka = -B / A;
        float kb = C / A;
        float delta = ka * ka - kb;
        if (delta > 0) {
            delta = sqrtf(delta);
            float t0 = ka - delta, t1 = ka + delta;
            float dist = t0 < t1 ? t0 : t1;                      /* Min<> */
            if (dist < EPS) dist = t0 > t1 ? t0 : t1;            /* Max<> */
            if (!(dist < 0)) {
                h->kind = 1; h->idx = i; h->t = dist;
                v3 hp = v3_add(o, v3_scale(d, dist));
                v3 n;
                n.x = 2.0f * qx * hp.x + my * hp.z + mz * hp.y + lx;
                n.y = 2.0f * qy * hp.y + mx * hp.z + mz * hp.x + ly;
                n.z = 2.0f * qz * hp.z + mx * hp.y + my * hp.x + lz;
                h->n = v3_normalise(n);
            }
        }
    } else {
        h->kind = 1; h->idx = i;
        h->t = -0.5f * (C / B);
        h->n = v3_normalise(v3_make(lx, ly, lz));
    }
}

static inline void hit_surface(const scene_view* sv, int k, const ray_t* r, hit_t* h)
{
    const surf_ref* sr = &sv->list[k];
    if (sr->kind == 0) hit_triangle(sv->s, sr->idx, r, h);
    else if (sr->kind == 1) hit_quadric(sv->s, sr->idx, r, h);
    else hit_plane(sv->s, sr->idx, r, h);
}

/* ---- scene view ---------------------------------------------------------------------------- */
static int cmp_order(const void* a, const void* b)
{
    return ((const surf_ref*)a)->order - ((const surf_ref*)b)->order;
}
static void view_init(scene_view* sv, const rtb200_scene* s)
{
    int n = s->n_triangles + s->n_quadrics + s->n_planes, k = 0;
    sv->s = s; sv->n = n;
    sv->list = (surf_ref*)malloc(sizeof(surf_ref) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < s->n_triangles; ++i, ++k) { surf_ref e = { 0, i, s->tri_order[i], s->tri_material[i] }; sv->list[k] = e; }
    for (int i = 0; i < s->n_quadrics; ++i, ++k)  { surf_ref e = { 1, i, s->quad_order[i], s->quad_material[i] }; sv->list[k] = e; }
    for (int i = 0; i < s->n_planes; ++i, ++k)    { surf_ref e = { 2, i, s->plane_order[i], s->plane_material[i] }; sv->list[k] = e; }
    qsort(sv->list, (size_t)n, sizeof(surf_ref), cmp_order);
}
static void view_free(scene_view* sv) { free(sv->list); sv->list = NULL; }

/* ---- tracer -------------------------------------------------------------------------------- */
typedef struct {
    const scene_view* sv;
    const rtb200_frame* f;
    uint64_t counts[4];          /* primary, shadow, reflected, refracted */
    uint64_t sig;                /* running signature (rto_signatures) */
    int want_sig;
} trace_ctx;

static inline void sig_mix(trace_ctx* tc, uint64_t v)
{
    if (!tc->want_sig) return;
    uint64_t h = tc->sig ^ (v + 0x9E3779B97F4A7C15ull + (tc->sig << 6) + (tc->sig >> 2));
    h *= 0xD6E8FEB86659FD93ull; h ^= h >> 32;
    tc->sig = h;
}

/* the nearest-hit loop of CScene::ObtenirCouleur, Scene.cpp:1707-1715; returns list slot or -1 */
static inline int nearest_hit(const scene_view* sv, const ray_t* r, hit_t* best)
{
    hit_t tmp;
    int best_k = -1;
    best->kind = -1; best->t = -1.0f; best->n = v3_make(0, 0, 0);
    for (int k = 0; k < sv->n; ++k) {
        hit_surface(sv, k, r, &tmp);
        if (tmp.t > EPS && (tmp.t < best->t || best->t < 0)) { *best = tmp; best_k = k; }
    }
    return best_k;
}

/* CScene::ObtenirFiltreDeSurface, Scene.cpp:1842-1861 (normalises the light ray in place) */
static col shadow_filter(trace_ctx* tc, ray_t* lr)
{
    const scene_view* sv = tc->sv;
    col filter = { 1.0f, 1.0f, 1.0f };
    hit_t h;
    float dist = v3_norm(lr->d);
    float inv = 1.0f / dist;                      /* operator/ : Vecteur3.h:328-335 */
    lr->d = v3_make(lr->d.x * inv, lr->d.y * inv, lr->d.z * inv);
    tc->counts[1]++;
    for (int k = 0; k < sv->n; ++k) {
        hit_surface(sv, k, lr, &h);
        if (h.t > EPS && h.t < dist) {
            const rtb200_material* m = &sv->s->materials[sv->list[k].mat];
            filter.r *= m->color[0] * m->refract;
            filter.g *= m->color[1] * m->refract;
            filter.b *= m->color[2] * m->refract;
        }
    }
    return filter;
}

static col trace_colour(trace_ctx* tc, const ray_t* r);

/* CScene::ObtenirCouleurSurIntersection, Scene.cpp:1740-1826 (reflection/refraction block of
 * :1779-1823 enabled, as in the recursion-restored reference build) */
static col shade_hit(trace_ctx* tc, const ray_t* r, const hit_t* h, int slot)
{
    const rtb200_scene* s = tc->sv->s;
    const rtb200_frame* f = tc->f;
    const rtb200_material* m = &s->materials[tc->sv->list[slot].mat];
    col res = { m->color[0] * m->ambient, m->color[1] * m->ambient, m->color[2] * m->ambient };
    v3 P = v3_add(r->o, v3_scale(r->d, h->t));

    for (int l = 0; l < s->n_lights; ++l) {
        const rtb200_light* L = &s->lights[l];
        ray_t lr;
        lr.o = P;
        lr.d = v3_sub(v3_make(L->position[0], L->position[1], L->position[2]), P);
        lr.energy = 1; lr.ior = 1; lr.bounces = 0;
        if (v3_dot(lr.d, h->n) > 0) {
            col filter = shadow_filter(tc, &lr);
            sig_mix(tc, 0x100u + (uint64_t)l);
            sig_mix(tc, ((uint64_t)(filter.r != 0) << 2) | ((uint64_t)(filter.g != 0) << 1) | (uint64_t)(filter.b != 0));
            if (tc->want_sig) {
                uint32_t qa = (uint32_t)(filter.r * 1024.0f), qb = (uint32_t)(filter.g * 1024.0f), qc = (uint32_t)(filter.b * 1024.0f);
                sig_mix(tc, ((uint64_t)qa << 40) ^ ((uint64_t)qb << 20) ^ qc);
            }
            col lc = { L->color[0] * filter.r, L->color[1] * filter.g, L->color[2] * filter.b };
            float gouraud = L->intensity * m->diffuse * v3_dot(h->n, lr.d);
            res.r += (m->color[0] * gouraud) * lc.r;
            res.g += (m->color[1] * gouraud) * lc.g;
            res.b += (m->color[2] * gouraud) * lc.b;
            v3 R = v3_reflect(lr.d, h->n);
            float ps = v3_dot(R, r->d);
            if (ps > 0) {
                float phong = L->intensity * m->specular * powf(ps, m->shininess);
                res.r += lc.r * phong;
                res.g += lc.g * phong;
                res.b += lc.b * phong;
            }
        }
    }

    if (f->flags & RTB200_F_NO_SECONDARY) return res;

    float e_refl = m->reflect * r->energy;
    if (e_refl > f->min_energy && r->bounces < f->max_bounces) {
        ray_t rr;
        rr.d = v3_reflect(r->d, h->n);
        rr.o = P;
        rr.energy = e_refl;
        rr.bounces = r->bounces + 1;
        rr.ior = 0.0f;                                       /* CRayon default, Rayon.cpp:16 */
        tc->counts[2]++;
        sig_mix(tc, 0x200u);
        col c = trace_colour(tc, &rr);
        res.r += c.r * m->reflect; res.g += c.g * m->reflect; res.b += c.b * m->reflect;
    }
    float e_refr = m->refract * r->energy;
    if (e_refr > f->min_energy && r->bounces < f->max_bounces) {
        ray_t rr;
        float ratio;
        v3 n = h->n;
        if (r->ior == m->ior) {                              /* inside -> outside */
            rr.ior = f->scene_ior;
            ratio = m->ior / f->scene_ior;
            n = v3_neg(n);
        } else {                                             /* outside -> inside */
            rr.ior = m->ior;
            ratio = f->scene_ior / m->ior;
        }
        rr.o = P;
        rr.energy = e_refr;
        rr.bounces = r->bounces + 1;
        rr.d = v3_refract(r->d, n, ratio);
        tc->counts[3]++;
        sig_mix(tc, 0x300u);
        col c = trace_colour(tc, &rr);
        res.r += c.r * m->refract; res.g += c.g * m->refract; res.b += c.b * m->refract;
    }
    return res;
}

/* CScene::ObtenirCouleur, Scene.cpp:1705-1720 */
static col trace_colour(trace_ctx* tc, const ray_t* r)
{
    hit_t best;
    int slot = nearest_hit(tc->sv, r, &best);
    if (best.t < 0) {
        sig_mix(tc, 0xFFFFFFFFull);
        col bg = { tc->f->background[0], tc->f->background[1], tc->f->background[2] };
        return bg;
    }
    sig_mix(tc, (uint64_t)tc->sv->list[slot].order);
    return shade_hit(tc, r, &best, slot);
}

/* primary ray of lattice sample (sx, sy): Scene.cpp:1543-1552 at resolution (k*W, k*H) */
static inline ray_t primary_ray(const rtb200_frame* f, int sx, int sy, float inv_w, float inv_h)
{
    ray_t r;
    v3 v = v3_make((2 * sx * inv_w - 1) * f->half_w, (2 * sy * inv_h - 1) * f->half_h, -1);
    /* CVecteur3 * CMatrice4 with rows U,V,N and a zero translation row, Matrice4.h:297-306 */
    v3 d;
    d.x = f->cam_u[0] * v.x + f->cam_v[0] * v.y + f->cam_n[0] * v.z + 0.0f;
    d.y = f->cam_u[1] * v.x + f->cam_v[1] * v.y + f->cam_n[1] * v.z + 0.0f;
    d.z = f->cam_u[2] * v.x + f->cam_v[2] * v.y + f->cam_n[2] * v.z + 0.0f;
    r.o = v3_make(f->cam_pos[0], f->cam_pos[1], f->cam_pos[2]);
    r.d = v3_normalise(d);
    r.energy = 1; r.bounces = 0; r.ior = 1;
    return r;
}

static double now_s(void)
{
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

double rto_render_region(const rtb200_scene* s, const rtb200_frame* f, int x0, int y0, int w, int h,
                         float* out, int threads)
{
    scene_view sv; view_init(&sv, s);
    const int k = f->spp_grid > 1 ? f->spp_grid : 1;
    const float inv_w = 1.0f / (f->width * k), inv_h = 1.0f / (f->height * k);
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (threads < 1) threads = 1;
    double t0 = now_s();
    #pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+:c0,c1,c2,c3)
    for (int y = 0; y < h; ++y) {
        trace_ctx tc; memset(&tc, 0, sizeof tc); tc.sv = &sv; tc.f = f;
        for (int x = 0; x < w; ++x) {
            ray_t r = primary_ray(f, x0 + x, y0 + y, inv_w, inv_h);
            tc.counts[0]++;
            col c = trace_colour(&tc, &r);
            float* o = out + 3 * ((size_t)y * w + x);
            o[0] = c.r; o[1] = c.g; o[2] = c.b;
        }
        c0 += tc.counts[0]; c1 += tc.counts[1]; c2 += tc.counts[2]; c3 += tc.counts[3];
    }
    double t1 = now_s();
    g_counts[0] = c0; g_counts[1] = c1; g_counts[2] = c2; g_counts[3] = c3;
    view_free(&sv);
    return t1 - t0;
}

double rto_render(const rtb200_scene* s, const rtb200_frame* f, float* out, int threads)
{
    const int k = f->spp_grid > 1 ? f->spp_grid : 1;
    const int W = f->width, H = f->height;
    if (k == 1) return rto_render_region(s, f, 0, 0, W, H, out, threads);
    /* k*k regular grid, pixel = mean of its samples summed j-outer / i-inner, times 1/(k*k) */
    scene_view sv; view_init(&sv, s);
    const float inv_w = 1.0f / (W * k), inv_h = 1.0f / (H * k);
    const float scale = 1.0f / (float)(k * k);
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (threads < 1) threads = 1;
    double t0 = now_s();
    #pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+:c0,c1,c2,c3)
    for (int y = 0; y < H; ++y) {
        trace_ctx tc; memset(&tc, 0, sizeof tc); tc.sv = &sv; tc.f = f;
        for (int x = 0; x < W; ++x) {
            col acc = { 0, 0, 0 };
            for (int j = 0; j < k; ++j)
                for (int i = 0; i < k; ++i) {
                    ray_t r = primary_ray(f, x * k + i, y * k + j, inv_w, inv_h);
                    tc.counts[0]++;
                    col c = trace_colour(&tc, &r);
                    acc.r += c.r; acc.g += c.g; acc.b += c.b;
                }
            float* o = out + 3 * ((size_t)y * W + x);
            o[0] = acc.r * scale; o[1] = acc.g * scale; o[2] = acc.b * scale;
        }
        c0 += tc.counts[0]; c1 += tc.counts[1]; c2 += tc.counts[2]; c3 += tc.counts[3];
    }
    double t1 = now_s();
    g_counts[0] = c0; g_counts[1] = c1; g_counts[2] = c2; g_counts[3] = c3;
    view_free(&sv);
    return t1 - t0;
}

void rto_signatures(const rtb200_scene* s, const rtb200_frame* f, uint64_t* sig, int32_t* primary_hit, int threads)
{
    scene_view sv; view_init(&sv, s);
    const int W = f->width, H = f->height;
    const float inv_w = 1.0f / W, inv_h = 1.0f / H;
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int y = 0; y < H; ++y) {
        trace_ctx tc; memset(&tc, 0, sizeof tc); tc.sv = &sv; tc.f = f; tc.want_sig = 1;
        for (int x = 0; x < W; ++x) {
            ray_t r = primary_ray(f, x, y, inv_w, inv_h);
            tc.sig = 0x1234567ull;
            hit_t best;
            int slot = nearest_hit(&sv, &r, &best);
            if (primary_hit) primary_hit[(size_t)y * W + x] = slot < 0 ? -1 : sv.list[slot].order;
            (void)trace_colour(&tc, &r);
            sig[(size_t)y * W + x] = tc.sig;
        }
    }
    view_free(&sv);
}

void rto_nearest(const rtb200_scene* s, int n, const float* o3, const float* d3,
                 int32_t* out_order, float* out_t, float* out_n3, int threads)
{
    scene_view sv; view_init(&sv, s);
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
    for (int i = 0; i < n; ++i) {
        ray_t r; memset(&r, 0, sizeof r);
        r.o = v3_make(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]);
        r.d = v3_make(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]);
        hit_t best;
        int slot = nearest_hit(&sv, &r, &best);
        out_order[i] = slot < 0 ? -1 : sv.list[slot].order;
        out_t[i] = best.t;
        out_n3[3 * i] = best.n.x; out_n3[3 * i + 1] = best.n.y; out_n3[3 * i + 2] = best.n.z;
    }
    view_free(&sv);
}

void rto_shadow(const rtb200_scene* s, int n, const float* p3, const float* l3, float* out3, int threads)
{
    scene_view sv; view_init(&sv, s);
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
    for (int i = 0; i < n; ++i) {
        trace_ctx tc; memset(&tc, 0, sizeof tc); tc.sv = &sv;
        ray_t lr; memset(&lr, 0, sizeof lr);
        lr.o = v3_make(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]);
        lr.d = v3_sub(v3_make(l3[3 * i], l3[3 * i + 1], l3[3 * i + 2]), lr.o);
        col c = shadow_filter(&tc, &lr);
        out3[3 * i] = c.r; out3[3 * i + 1] = c.g; out3[3 * i + 2] = c.b;
    }
    view_free(&sv);
}

void rto_trace(const rtb200_scene* s, const rtb200_frame* f, int n, const float* o3, const float* d3,
               float* out3, int threads)
{
    scene_view sv; view_init(&sv, s);
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
    for (int i = 0; i < n; ++i) {
        trace_ctx tc; memset(&tc, 0, sizeof tc); tc.sv = &sv; tc.f = f;
        ray_t r;
        r.o = v3_make(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]);
        r.d = v3_make(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]);
        r.energy = 1; r.bounces = 0; r.ior = 1;
        col c = trace_colour(&tc, &r);
        out3[3 * i] = c.r; out3[3 * i + 1] = c.g; out3[3 * i + 2] = c.b;
    }
    view_free(&sv);
}

uint64_t rto_quantise(const float* rgb, size_t n_pixels, uint8_t* rgba8)
{
    uint64_t nans = 0;
    for (size_t i = 0; i < n_pixels; ++i) {
        for (int c = 0; c < 3; ++c) {
            float v = rgb[3 * i + c];
            if (v != v) { nans++; v = 0.0f; }
            v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
            rgba8[4 * i + c] = (uint8_t)floorf(v * 255.0f + 0.5f);
        }
        rgba8[4 * i + 3] = 255;
    }
    return nans;
}

void rto_last_ray_counts(uint64_t out[4]) { memcpy(out, g_counts, sizeof g_counts); }

int rto_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
