// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Headless C-callable driver around the UNMODIFIED reference renderer
// (/root/reference/Projet-INF8702, compiled in place by oracle/Makefile into
// oracle/_ref/).  It is the parity pin for everything in this repo:
//   * full frames through the reference's own CScene::LancerRayons() CPU branch
//     (Scene.cpp:672, :1535-1563), read back from the glTexImage2D call at
//     Scene.cpp:1562 (the stub below *is* the framebuffer read-back);
//   * per-ray / crop / sub-pixel access through the private, const, re-entrant
//     CScene::ObtenirCouleur (Scene.cpp:1705; this TU is built with
//     -fno-access-control), optionally OpenMP-parallel over rows for the
//     "all host cores" baseline;
//   * dumps of the post-Pretraitement surface coefficients and camera basis,
//     which pin the fresh host classes under ray-tracing-gpu_b200/host/;
//   * synthetic scenes built through the reference's public API
//     (AjouterSurface / AjouterLumiere, Scene.h:41-43).
// Nothing here restates reference arithmetic except the six-line primary-ray
// set-up of Scene.cpp:1543-1552, needed for crops and sub-pixel samples; the
// test-suite checks it bit-for-bit against LancerRayons() full frames.
//
// GL: only link-time no-op symbols (this is not a graphics stack).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "Scene.h"
#include "Var.h"
#include "Triangle.h"
#include "Plan.h"
#include "Quadrique.h"
#include "NuanceurCalculProg.h"

using namespace Scene;

// ---------------------------------------------------------------- GL no-ops
static const void* g_tex_pixels = nullptr;
static int g_tex_w = 0, g_tex_h = 0;

extern "C" {
void glBindTexture(GLenum, GLuint) {}
void glDeleteTextures(GLsizei, const GLuint*) {}
void glGenTextures(GLsizei n, GLuint* t) { for (int i = 0; i < n; ++i) t[i] = 1; }
GLenum glGetError(void) { return 0; }
void glTexParameteri(GLenum, GLenum, GLint) {}
void glTexImage2D(GLenum, GLint, GLint, GLsizei w, GLsizei h, GLint, GLenum, GLenum, const void* pixels)
{
    g_tex_pixels = pixels; g_tex_w = w; g_tex_h = h;
}
static void noop_active_texture(GLenum) {}
PFNGLACTIVETEXTUREPROC __glewActiveTexture = noop_active_texture;
PFNGLBINDBUFFERPROC __glewBindBuffer = nullptr;
PFNGLBINDBUFFERBASEPROC __glewBindBufferBase = nullptr;
PFNGLBINDIMAGETEXTUREPROC __glewBindImageTexture = nullptr;
PFNGLBUFFERDATAPROC __glewBufferData = nullptr;
PFNGLDISPATCHCOMPUTEPROC __glewDispatchCompute = nullptr;
PFNGLGENBUFFERSPROC __glewGenBuffers = nullptr;
PFNGLGETACTIVEUNIFORMBLOCKIVPROC __glewGetActiveUniformBlockiv = nullptr;
PFNGLGETACTIVEUNIFORMSIVPROC __glewGetActiveUniformsiv = nullptr;
PFNGLGETBUFFERSUBDATAPROC __glewGetBufferSubData = nullptr;
PFNGLGETUNIFORMBLOCKINDEXPROC __glewGetUniformBlockIndex = nullptr;
PFNGLGETUNIFORMINDICESPROC __glewGetUniformIndices = nullptr;
PFNGLMEMORYBARRIERPROC __glewMemoryBarrier = nullptr;
}
// The GLSL wrapper is referenced (never called) by the GPU branch of LancerRayons.
void CNuanceurCalculProg::activer() {}
GLuint CNuanceurCalculProg::getProg() { return 0; }

// ---------------------------------------------------------------- state
static CScene* g_scene = nullptr;
static bool    g_prepared = false;
static float   g_halfw = 0, g_halfh = 0;

static void drop_scene()
{
    if (g_scene) {
        CScene::LibererInstance();   // ~CScene -> Liberer()
        g_scene = nullptr;
    }
    g_prepared = false;
    g_tex_pixels = nullptr;
}

static void apply_material(ISurface* s, const float* m)
{
    // m = r g b ka kd ks shininess kr kt ior ; colour written through the public
    // float members so that no clamping/scaling is applied by the driver.
    CCouleur c; c.r = m[0]; c.g = m[1]; c.b = m[2];
    s->AjusterCouleur(c);
    s->AjusterCoeffAmbiant(m[3]);
    s->AjusterCoeffDiffus(m[4]);
    s->AjusterCoeffSpeculaire(m[5]);
    s->AjusterCoeffBrillance(m[6]);
    s->AjusterCoeffReflexion(m[7]);
    s->AjusterCoeffRefraction(m[8]);
    s->AjusterIndiceRefraction(m[9]);
}

extern "C" {

const char* ref_flavour(void)
{
#ifdef REF_GPU_BINDING
    return "recursion-restored, GPU branch = librtb200 binding";
#elif defined(REF_RECURSION_RESTORED)
    return "recursion-restored";
#else
    return "as-shipped";
#endif
}

int ref_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ref_free_scene(void) { drop_scene(); }

// Which branch of CScene::LancerRayons the next ref_render_full takes (Scene.cpp:692).  Only the libref_gpu.so flavour
// has something behind the GPU branch (oracle/ref_gpu_binding.inc -> librtb200.so); in the other flavours that branch is
// the reference's GL code, which cannot run here, so they refuse.  ref_new_scene / ref_load_scene reset it to the CPU branch.
int ref_use_gpu_branch(int on)
{
#ifdef REF_GPU_BINDING
    CVar::g_ComputerShadersON = on != 0;
    return 0;
#else
    (void)on;
    return -1;
#endif
}

void ref_new_scene(void)
{
    drop_scene();
    CVar::g_ComputerShadersON = false;
    g_scene = CScene::ObtenirInstance();
}

int ref_load_scene(const char* path)
{
    ref_new_scene();
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    fclose(f);
    // the parser prints one line per object; silence it
    fflush(stdout);
    FILE* saved = stdout;
    std::streambuf* sb = std::cout.rdbuf(nullptr);
    g_scene->TraiterFichierDeScene(path);
    std::cout.rdbuf(sb);
    stdout = saved;
    return (int)g_scene->m_Surfaces.size();
}

void ref_set_camera(const float* pos, const float* look, const float* up)
{
    g_scene->AjusterPositionCamera(CVecteur3(pos[0], pos[1], pos[2]));
    g_scene->AjusterPointViseCamera(CVecteur3(look[0], look[1], look[2]));
    g_scene->AjusterVecteurUpCamera(CVecteur3(up[0], up[1], up[2]));
}

void ref_set_background(const float* rgb)
{
    CCouleur c; c.r = rgb[0]; c.g = rgb[1]; c.b = rgb[2];
    g_scene->AjusterCouleurArrierePlan(c);
}

// max_bounces < 0 keeps the reference default (20, Scene.cpp:68)
void ref_set_params(int max_bounces, float min_energy, float scene_ior)
{
    if (max_bounces >= 0) g_scene->AjusterNbRebondsMax(max_bounces);
    if (min_energy >= 0)  g_scene->AjusterEnergieMinimale(min_energy);
    if (scene_ior >= 0)   g_scene->AjusterIndiceRefraction(scene_ior);
}

int ref_add_triangle(const float* p9, const float* mat10)
{
    CTriangle* t = new CTriangle();
    t->AjusterPoints(CVecteur3(p9[0], p9[1], p9[2]), CVecteur3(p9[3], p9[4], p9[5]), CVecteur3(p9[6], p9[7], p9[8]));
    apply_material(t, mat10);
    g_scene->AjouterSurface(t);
    return (int)g_scene->m_Surfaces.size() - 1;
}

// coef10 = quadratic A B C | mixed yz xz xy | linear x y z | constant
int ref_add_quadric(const float* c, const float* mat10)
{
    CQuadrique* q = new CQuadrique();
    q->AjusterQuadratique(CVecteur3(c[0], c[1], c[2]));
    q->AjusterMixte(CVecteur3(c[3], c[4], c[5]));
    q->AjusterLineaire(CVecteur3(c[6], c[7], c[8]));
    q->AjusterConstante(c[9]);
    apply_material(q, mat10);
    g_scene->AjouterSurface(q);
    return (int)g_scene->m_Surfaces.size() - 1;
}

int ref_add_plane(const float* c4, const float* mat10)
{
    CPlan* p = new CPlan();
    p->AjusterNormale(CVecteur3(c4[0], c4[1], c4[2]));
    p->AjusterConstante(c4[3]);
    apply_material(p, mat10);
    g_scene->AjouterSurface(p);
    return (int)g_scene->m_Surfaces.size() - 1;
}

// Bulk versions so that ctypes does not pay one call per primitive.
void ref_add_triangles(int n, const float* p9, const float* mats10, const int* mat_idx)
{
    for (int i = 0; i < n; ++i) ref_add_triangle(p9 + 9 * (size_t)i, mats10 + 10 * (size_t)mat_idx[i]);
}
void ref_add_quadrics(int n, const float* c10, const float* mats10, const int* mat_idx)
{
    for (int i = 0; i < n; ++i) ref_add_quadric(c10 + 10 * (size_t)i, mats10 + 10 * (size_t)mat_idx[i]);
}

int ref_add_light(const float* pos, const float* rgb, float intensity)
{
    CLumiere* l = new CLumiere();
    l->SetPosition(CVecteur3(pos[0], pos[1], pos[2]));
    CCouleur c; c.r = rgb[0]; c.g = rgb[1]; c.b = rgb[2];
    l->AjusterCouleur(c);
    l->SetIntensity(intensity);
    g_scene->AjouterLumiere(l);
    return (int)g_scene->m_Lumieres.size() - 1;
}

int ref_num_surfaces(void) { return (int)g_scene->m_Surfaces.size(); }
int ref_num_lights(void)   { return (int)g_scene->m_Lumieres.size(); }

// Canonical reference render: CScene::LancerRayons() CPU branch, single thread.
// Returns seconds spent inside LancerRayons (includes Initialiser(), as the
// reference's own timer at Main.cpp:169-198 does).  May be called once per scene
// (Pretraitement is not idempotent, Scene.cpp:140-147).
double ref_render_full(int W, int H, float* out_rgb)
{
    if (g_prepared) return -1.0;
    g_scene->AjusterResolution(W, H);
    auto t0 = std::chrono::steady_clock::now();
    g_scene->LancerRayons();
    auto t1 = std::chrono::steady_clock::now();
    g_prepared = true;
    if (!g_tex_pixels || g_tex_w != W || g_tex_h != H) return -2.0;
    if (out_rgb) memcpy(out_rgb, g_tex_pixels, sizeof(float) * 3 * (size_t)W * H);
    // keep the per-ray entry usable after a full render
    g_halfh = tan(Deg2Rad<REAL>(g_scene->m_Camera.Angle * RENDRE_REEL(0.5)));
    g_halfw = (RENDRE_REEL(W) / H) * g_halfh;
    return std::chrono::duration<double>(t1 - t0).count();
}

// Pre-processing only (camera + Pretraitement), for per-ray access and dumps.
// W,H are the resolution the aspect ratio is taken from (Scene.cpp:676-677).
int ref_prepare(int W, int H)
{
    if (g_prepared) return -1;
    g_scene->AjusterResolution(W, H);
    g_scene->InitialiserCamera();
    for (auto* s : g_scene->m_Surfaces) s->Pretraitement();
    g_prepared = true;
    g_halfh = tan(Deg2Rad<REAL>(g_scene->m_Camera.Angle * RENDRE_REEL(0.5)));
    g_halfw = (RENDRE_REEL(W) / H) * g_halfh;
    return 0;
}

// type: 0 triangle, 1 quadric, 2 plane.  geom16 layout:
//  triangle: p0 p1 p2 normal (12) ; quadric: A B C | yz xz xy | lx ly lz | K (10) ; plane: n c (4)
void ref_dump_surface(int i, int* type, float* geom16, float* mat10)
{
    ISurface* s = g_scene->m_Surfaces[i];
    memset(geom16, 0, 16 * sizeof(float));
    if (CTriangle* t = dynamic_cast<CTriangle*>(s)) {
        *type = 0;
        for (int k = 0; k < 3; ++k) { CVecteur3 p = t->ObtenirPoint(k); geom16[3*k] = p.x; geom16[3*k+1] = p.y; geom16[3*k+2] = p.z; }
        CVecteur3 n = t->ObtenirNormale(); geom16[9] = n.x; geom16[10] = n.y; geom16[11] = n.z;
    } else if (CQuadrique* q = dynamic_cast<CQuadrique*>(s)) {
        *type = 1;
        CVecteur3 a = q->ObtenirQuadratique(), m = q->ObtenirMixte(), l = q->ObtenirLineaire();
        geom16[0] = a.x; geom16[1] = a.y; geom16[2] = a.z;
        geom16[3] = m.x; geom16[4] = m.y; geom16[5] = m.z;
        geom16[6] = l.x; geom16[7] = l.y; geom16[8] = l.z;
        geom16[9] = q->ObtenirConstante();
    } else if (CPlan* p = dynamic_cast<CPlan*>(s)) {
        *type = 2;
        CVecteur3 n = p->ObtenirNormale();
        geom16[0] = n.x; geom16[1] = n.y; geom16[2] = n.z; geom16[3] = p->ObtenirConstante();
    } else *type = -1;
    CCouleur c = s->ObtenirCouleur();
    mat10[0] = c.r; mat10[1] = c.g; mat10[2] = c.b;
    mat10[3] = s->ObtenirCoeffAmbiant();   mat10[4] = s->ObtenirCoeffDiffus();
    mat10[5] = s->ObtenirCoeffSpeculaire(); mat10[6] = s->ObtenirCoeffBrillance();
    mat10[7] = s->ObtenirCoeffReflexion();  mat10[8] = s->ObtenirCoeffRefraction();
    mat10[9] = s->ObtenirIndiceRefraction();
}

void ref_dump_light(int i, float* out7)
{
    CLumiere* l = g_scene->m_Lumieres[i];
    CVecteur3 p = l->GetPosition(); CCouleur c = l->ObtenirCouleur();
    out7[0] = p.x; out7[1] = p.y; out7[2] = p.z; out7[3] = c.r; out7[4] = c.g; out7[5] = c.b; out7[6] = l->GetIntensity();
}

// out: pos3 U3 V3 N3 angle halfW halfH | background3 | max_bounces min_energy scene_ior  (21 floats)
void ref_dump_camera(float* out)
{
    const CMatrice4& M = g_scene->m_Camera.Orientation;
    out[0] = g_scene->m_Camera.Position.x; out[1] = g_scene->m_Camera.Position.y; out[2] = g_scene->m_Camera.Position.z;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) out[3 + 3*r + c] = M.m[r][c];
    out[12] = g_scene->m_Camera.Angle; out[13] = g_halfw; out[14] = g_halfh;
    out[15] = g_scene->m_CouleurArrierePlan.r; out[16] = g_scene->m_CouleurArrierePlan.g; out[17] = g_scene->m_CouleurArrierePlan.b;
    out[18] = (float)g_scene->m_NbRebondsMax; out[19] = g_scene->m_EnergieMinRayon; out[20] = g_scene->m_IndiceRefractionScene;
}

// Primary ray of sub-sample (sx, sy) on the (k*W) x (k*H) lattice, exactly as
// Scene.cpp:1543-1552 builds it at that resolution.
static inline CRayon primary_ray(int sx, int sy, float inv_w, float inv_h)
{
    CRayon r;
    r.AjusterOrigine(g_scene->m_Camera.Position);
    r.AjusterDirection(CVecteur3((2 * sx * inv_w - 1) * g_halfw, (2 * sy * inv_h - 1) * g_halfh, -1));
    r.AjusterDirection(CVecteur3::Normaliser(r.ObtenirDirection() * g_scene->m_Camera.Orientation));
    r.AjusterEnergie(1);
    r.AjusterNbRebonds(0);
    r.AjusterIndiceRefraction(1);
    return r;
}

// Renders sub-sample rows/cols [x0,x0+w) x [y0,y0+h) of the (k*W) x (k*H) lattice
// (k = 1: plain pixels).  out = w*h*3 floats, row-major, row 0 = lowest y.
// threads <= 1: serial.  Returns seconds.
double ref_render_region(int W, int H, int k, int x0, int y0, int w, int h, float* out, int threads)
{
    const float inv_w = RENDRE_REEL(1.0) / (W * k);
    const float inv_h = RENDRE_REEL(1.0) / (H * k);
    const CScene* sc = g_scene;
    auto t0 = std::chrono::steady_clock::now();
#ifdef _OPENMP
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            CRayon r = primary_ray(x0 + x, y0 + y, inv_w, inv_h);
            CCouleur c = sc->ObtenirCouleur(r);
            float* o = out + 3 * ((size_t)y * w + x);
            o[0] = c.r; o[1] = c.g; o[2] = c.b;
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// Arbitrary rays (energy 1, bounces 0, ior 1 -- i.e. treated as primaries).
void ref_trace_rays(int n, const float* org3, const float* dir3, float* rgb3, int threads)
{
    const CScene* sc = g_scene;
#ifdef _OPENMP
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (int i = 0; i < n; ++i) {
        CRayon r;
        r.AjusterOrigine(CVecteur3(org3[3*i], org3[3*i+1], org3[3*i+2]));
        r.AjusterDirection(CVecteur3(dir3[3*i], dir3[3*i+1], dir3[3*i+2]));
        r.AjusterEnergie(1); r.AjusterNbRebonds(0); r.AjusterIndiceRefraction(1);
        CCouleur c = sc->ObtenirCouleur(r);
        rgb3[3*i] = c.r; rgb3[3*i+1] = c.g; rgb3[3*i+2] = c.b;
    }
}

// Nearest hit only (the loop of Scene.cpp:1710-1715): surface index (-1 = miss),
// distance and normal per ray.  Pins the brute-force kernels and the LBVH.
void ref_nearest_hits(int n, const float* org3, const float* dir3, int* idx, float* t, float* nrm3, int threads)
{
    const CScene* sc = g_scene;
#ifdef _OPENMP
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (int i = 0; i < n; ++i) {
        CRayon r;
        r.AjusterOrigine(CVecteur3(org3[3*i], org3[3*i+1], org3[3*i+2]));
        r.AjusterDirection(CVecteur3(dir3[3*i], dir3[3*i+1], dir3[3*i+2]));
        CIntersection best, tmp;
        int best_i = -1, k = 0;
        for (auto it = sc->m_Surfaces.begin(); it != sc->m_Surfaces.end(); ++it, ++k) {
            tmp = (*it)->Intersection(r);
            if (tmp.ObtenirDistance() > EPSILON && (tmp.ObtenirDistance() < best.ObtenirDistance() || best.ObtenirDistance() < 0)) {
                best = tmp; best_i = k;
            }
        }
        idx[i] = best_i; t[i] = best.ObtenirDistance();
        CVecteur3 nn = best.ObtenirNormale();
        nrm3[3*i] = nn.x; nrm3[3*i+1] = nn.y; nrm3[3*i+2] = nn.z;
    }
}

// Shadow filter (Scene.cpp:1842-1861) for (point, light position) pairs.
void ref_shadow_filters(int n, const float* pt3, const float* light3, float* rgb3, int threads)
{
    const CScene* sc = g_scene;
#ifdef _OPENMP
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (int i = 0; i < n; ++i) {
        CRayon r;
        CVecteur3 p(pt3[3*i], pt3[3*i+1], pt3[3*i+2]);
        r.AjusterOrigine(p);
        r.AjusterDirection(CVecteur3(light3[3*i], light3[3*i+1], light3[3*i+2]) - p);
        r.AjusterEnergie(1); r.AjusterIndiceRefraction(1);
        CCouleur c = sc->ObtenirFiltreDeSurface(r);
        rgb3[3*i] = c.r; rgb3[3*i+1] = c.g; rgb3[3*i+2] = c.b;
    }
}

} // extern "C"
