// rtb_scene.cpp -- CScene, the scene-file parser, the SoA packer and the CUDA dispatch.
#include "rtb_scene.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <thread>

using namespace Scene;

int CVar::g_FenetreID = 0;
int CVar::g_LargeurFenetre = 512;
int CVar::g_HauteurFenetre = 256;
int CVar::g_LargeurViewport = 0;
int CVar::g_HauteurViewport = 0;
CScene* CVar::g_GestionnaireDeScene = nullptr;
bool CVar::g_ComputerShadersON = true;
static char g_chemin_nuanceur[] = "shaders/rayTracing.glsl";
char* CVar::g_ComputeShaderPath = g_chemin_nuanceur;
CNuanceurCalculProg* CVar::g_ComputeShader = nullptr;
int CVar::g_Peripherique = 0;
int CVar::g_NbPeripheriques = 1;

const REAL CScene::DIM_FILM_CAM = 0.024f;

// ---- CStringUtils ---------------------------------------------------------------------------
std::string CStringUtils::TrimRight(const std::string& s, const std::string& t) { std::string r = s; return r.erase(r.find_last_not_of(t) + 1); }
std::string CStringUtils::TrimLeft(const std::string& s, const std::string& t) { std::string r = s; return r.erase(0, s.find_first_not_of(t)); }
std::string CStringUtils::Trim(const std::string& s, const std::string& t) { return TrimLeft(TrimRight(s, t), t); }

// ---- CNuanceurCalculProg ----------------------------------------------------------------------
CNuanceurCalculProg::CNuanceurCalculProg() {}
CNuanceurCalculProg::CNuanceurCalculProg(char*, bool compilerImmediatement)
{
    estVide_ = false;
    if (compilerImmediatement) compilerEtLier();
}
CNuanceurCalculProg::~CNuanceurCalculProg() { if (ctx_) rtb200_destroy(ctx_); }

void CNuanceurCalculProg::compilerEtLier()
{
    if (ctx_) return;
    const int rc = rtb200_create(CVar::g_Peripherique, &ctx_);
    if (rc != RTB200_OK) {
        // the reference's convention: print and exit (NuanceurCalculProg.cpp:300, Scene.cpp:708)
        std::cerr << "[ERREUR]: rtb200_create(" << CVar::g_Peripherique << ") : " << rtb200_error_string(rc) << " -- "
                  << rtb200_last_error(nullptr) << " (no CPU fallback in this build)" << std::endl;
        exit(EXIT_FAILURE);
    }
    estCompileEtLie_ = true;
}
void CNuanceurCalculProg::activer() { compilerEtLier(); }
void CNuanceurCalculProg::uniform1(char*, int) {}
void CNuanceurCalculProg::uniform1(char*, float) {}
void CNuanceurCalculProg::enregistrerUniformFloat(const char*, const float&) {}
void CNuanceurCalculProg::enregistrerUniformInteger(const char*, const int&) {}
GLuint CNuanceurCalculProg::getProg() { return ctx_ ? 1u : 0u; }

// ---- SceneEmballee ----------------------------------------------------------------------------
rtb200_scene SceneEmballee::vue() const
{
    rtb200_scene s;
    s.n_triangles = (int32_t)tri_order.size(); s.n_quadrics = (int32_t)quad_order.size(); s.n_planes = (int32_t)plane_order.size();
    s.n_lights = (int32_t)lights.size(); s.n_materials = (int32_t)materials.size();
    s.tri_points = tri_points.data(); s.tri_normal = tri_normal.data(); s.tri_material = tri_material.data(); s.tri_order = tri_order.data();
    s.quad_coef = quad_coef.data(); s.quad_material = quad_material.data(); s.quad_order = quad_order.data();
    s.plane_coef = plane_coef.data(); s.plane_material = plane_material.data(); s.plane_order = plane_order.data();
    s.materials = materials.data(); s.lights = lights.data();
    return s;
}

// ---- CScene: construction and setters ----------------------------------------------------------
CScene::CScene()
{
    m_Camera.Position = CVecteur3::ZERO; m_Camera.PointVise = CVecteur3::ZERO; m_Camera.Up = CVecteur3::UNIT_Y;
    m_Camera.Orientation = CMatrice4::IDENTITE;
    m_Camera.Focale = RENDRE_REEL(50.0); m_Camera.Angle = 0;     // Scene.cpp:71-77
}
CScene::~CScene() { Liberer(); }

void CScene::AjouterSurface(ISurface* const s) { m_Surfaces.push_back(s); }
void CScene::AjouterLumiere(CLumiere* const l) { m_Lumieres.push_back(l); }
void CScene::AjusterPositionCamera(const CVecteur3& p) { m_Camera.Position = p; }
void CScene::AjusterPointViseCamera(const CVecteur3& p) { m_Camera.PointVise = p; }
void CScene::AjusterVecteurUpCamera(const CVecteur3& u) { m_Camera.Up = u; }
void CScene::AjusterResolution(const int w, const int h) { m_ResLargeur = w; m_ResHauteur = h; }
void CScene::AjusterCouleurArrierePlan(const CCouleur& c) { m_CouleurArrierePlan = c; }
void CScene::AjusterNbRebondsMax(const int n) { m_NbRebondsMax = n; }
void CScene::AjusterEnergieMinimale(const REAL e) { m_EnergieMinRayon = e; }
void CScene::AjusterIndiceRefraction(const REAL i) { m_IndiceRefractionScene = i; }
GLuint CScene::ObtenirTextureGL() { return 0; }

int CScene::ChargerMaillageOBJ(const char* fichier, const CTriangle& modele)
{
    std::ifstream in(fichier);
    if (!in.is_open()) return -1;
    std::vector<CVecteur3> sommets;
    std::string ligne;
    int ajoutes = 0;
    while (std::getline(in, ligne)) {
        if (ligne.size() < 2) continue;
        if (ligne[0] == 'v' && (ligne[1] == ' ' || ligne[1] == '\t')) {
            float x = 0.f, y = 0.f, z = 0.f;
            if (sscanf(ligne.c_str() + 1, "%f %f %f", &x, &y, &z) == 3) sommets.push_back(CVecteur3(x, y, z));
        } else if (ligne[0] == 'f' && (ligne[1] == ' ' || ligne[1] == '\t')) {
            std::vector<int> idx;                              // "f a b c ..." with a, a/t, a//n or a/t/n
            const char* p = ligne.c_str() + 1;
            while (*p) {
                while (*p == ' ' || *p == '\t') ++p;
                if (!*p || *p == '\r' || *p == '#') break;
                char* fin = nullptr;
                long v = strtol(p, &fin, 10);
                if (fin == p) break;
                const long n = (long)sommets.size();
                if (v < 0) v = n + v + 1;                      // relative index
                if (v < 1 || v > n) { idx.clear(); break; }   // dangling reference: drop the face
                idx.push_back((int)v - 1);
                p = fin;
                while (*p && *p != ' ' && *p != '\t') ++p;      // skip /t/n
            }
            for (size_t k = 2; k < idx.size(); ++k) {          // fan triangulation
                CTriangle* t = modele.Copier();
                t->AjusterPoints(sommets[idx[0]], sommets[idx[k - 1]], sommets[idx[k]]);
                AjouterSurface(t);
                ++ajoutes;
            }
        }
    }
    m_Prepare = false;
    return ajoutes;
}

void CScene::Liberer()
{
    for (ISurface* s : m_Surfaces) delete s;
    m_Surfaces.clear();
    for (CLumiere* l : m_Lumieres) delete l;
    m_Lumieres.clear();
    delete[] m_InfoPixel;
    m_InfoPixel = nullptr;
    m_ImageRGBA8.clear();
    m_Prepare = false;
}

// ---- scene file ---------------------------------------------------------------------------------
// Same grammar and quirks as CScene::TraiterFichierDeScene (Scene.cpp:231-501): lines are NOT
// trimmed (the reference discards Trim's result, :255), only a first-byte '*' is a comment,
// keywords are found by substring search, values by sscanf, a header flushes the previous
// object, unknown keys are ignored.  Hardening beyond the reference: lines of any length (the
// reference loops forever on >= 80 bytes), a trailing CR is dropped, an unopenable file is
// reported (and, like the reference, leaves an empty scene).
void CScene::TraiterFichierDeScene(const char* fichier)
{
    std::ifstream in(fichier);
    if (!in.is_open()) {
        std::cerr << "[CScene::TraiterFichierDeScene()] : Incapable d'ouvrir " << fichier << "." << std::endl;
        return;
    }
    auto has = [](const std::string& s, const char* key) { return s.find(key) != std::string::npos; };
    EtatTraitementScene etat = TRAITEMENT_SCENE;
    CLumiere* lumiere = nullptr;
    ISurface* surface = nullptr;
    char mot[256];
    float v0 = 0, v1 = 0, v2 = 0;
    int R = 0, G = 0, B = 0;
    std::string ligne;
    while (std::getline(in, ligne)) {
        if (!ligne.empty() && ligne.back() == '\r') ligne.pop_back();
        if (ligne.empty() || ligne[0] == '*') continue;
        const char* txt = ligne.c_str();
        EtatTraitementScene nouveau = etat;
        bool entete = true;
        if (has(ligne, "Lumiere:")) nouveau = TRAITEMENT_LUMIERE;
        else if (has(ligne, "Poly:")) nouveau = TRAITEMENT_TRIANGLE;
        else if (has(ligne, "Plane:")) nouveau = TRAITEMENT_PLAN;
        else if (has(ligne, "Quad:")) nouveau = TRAITEMENT_QUADRIQUE;
        else entete = false;

        if (entete) {
            if (etat != TRAITEMENT_SCENE) {
                if (etat == TRAITEMENT_LUMIERE) AjouterLumiere(lumiere); else AjouterSurface(surface);
                surface = nullptr; lumiere = nullptr;
            }
            etat = nouveau;
            switch (etat) {
            case TRAITEMENT_LUMIERE:   lumiere = new CLumiere(); break;
            case TRAITEMENT_TRIANGLE:  surface = new CTriangle(); break;
            case TRAITEMENT_PLAN:      surface = new CPlan(); break;
            case TRAITEMENT_QUADRIQUE: surface = new CQuadrique(); break;
            default: break;
            }
            continue;
        }
        if (surface) {                                  // material / transform keys of any surface
            bool generique = true;
            if (has(ligne, "color:")) { sscanf(txt, "%255s %i %i %i", mot, &R, &G, &B); surface->AjusterCouleur(CCouleur(R, G, B)); }
            else if (has(ligne, "ambient:")) { sscanf(txt, "%255s %f", mot, &v0); surface->AjusterCoeffAmbiant(v0); }
            else if (has(ligne, "diffus:")) { sscanf(txt, "%255s %f", mot, &v0); surface->AjusterCoeffDiffus(v0); }
            else if (has(ligne, "specular:")) { sscanf(txt, "%255s %f %f", mot, &v0, &v1); surface->AjusterCoeffSpeculaire(v0); surface->AjusterCoeffBrillance(v1); }
            else if (has(ligne, "reflect:")) { sscanf(txt, "%255s %f", mot, &v0); surface->AjusterCoeffReflexion(v0); }
            else if (has(ligne, "refract:")) { sscanf(txt, "%255s %f %f", mot, &v0, &v1); surface->AjusterCoeffRefraction(v0); surface->AjusterIndiceRefraction(v1); }
            else if (has(ligne, "rotate:")) {           // X then Y then Z, each post-multiplied (Scene.cpp:356-364)
                sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2);
                CMatrice4 t = surface->ObtenirTransformation();
                t.RotationAutourDesX(Deg2Rad<REAL>(v0)); t.RotationAutourDesY(Deg2Rad<REAL>(v1)); t.RotationAutourDesZ(Deg2Rad<REAL>(v2));
                surface->AjusterTransformation(t);
            }
            else if (has(ligne, "translate:")) {
                sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2);
                CMatrice4 t = surface->ObtenirTransformation(); t.Translation(v0, v1, v2); surface->AjusterTransformation(t);
            }
            else if (has(ligne, "scale:")) {
                sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2);
                CMatrice4 t = surface->ObtenirTransformation(); t.MiseAEchelle(v0, v1, v2); surface->AjusterTransformation(t);
            }
            else generique = false;
            if (generique) continue;
        }
        switch (etat) {
        case TRAITEMENT_SCENE:
            if (has(ligne, "background:")) { sscanf(txt, "%255s %i %i %i", mot, &R, &G, &B); AjusterCouleurArrierePlan(CCouleur(R, G, B)); }
            else if (has(ligne, "origin:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); AjusterPositionCamera(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "eye:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); AjusterPointViseCamera(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "up:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); AjusterVecteurUpCamera(CVecteur3(v0, v1, v2)); }
            break;
        case TRAITEMENT_LUMIERE:
            if (has(ligne, "position:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); lumiere->SetPosition(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "intens:")) { sscanf(txt, "%255s %f", mot, &v0); lumiere->SetIntensity(v0); }
            else if (has(ligne, "color:")) { sscanf(txt, "%255s %i %i %i", mot, &R, &G, &B); lumiere->AjusterCouleur(CCouleur(R, G, B)); }
            break;
        case TRAITEMENT_TRIANGLE:
            if (has(ligne, "point:")) {
                int idx = 0;
                sscanf(txt, "%255s %i %f %f %f", mot, &idx, &v0, &v1, &v2);
                static_cast<CTriangle*>(surface)->AjusterPoint(idx, CVecteur3(v0, v1, v2));
            }
            break;
        case TRAITEMENT_PLAN:
            if (has(ligne, "v_linear:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); static_cast<CPlan*>(surface)->AjusterNormale(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "v_const:")) { sscanf(txt, "%255s %f", mot, &v0); static_cast<CPlan*>(surface)->AjusterConstante(v0); }
            break;
        case TRAITEMENT_QUADRIQUE:
            if (has(ligne, "v_quad:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); static_cast<CQuadrique*>(surface)->AjusterQuadratique(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "v_mixte:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); static_cast<CQuadrique*>(surface)->AjusterMixte(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "v_linear:")) { sscanf(txt, "%255s %f %f %f", mot, &v0, &v1, &v2); static_cast<CQuadrique*>(surface)->AjusterLineaire(CVecteur3(v0, v1, v2)); }
            else if (has(ligne, "v_const:")) { sscanf(txt, "%255s %f", mot, &v0); static_cast<CQuadrique*>(surface)->AjusterConstante(v0); }
            break;
        }
    }
    if (surface) AjouterSurface(surface);
    if (lumiere) AjouterLumiere(lumiere);
}

// ---- camera and pre-processing ------------------------------------------------------------------
// Scene.cpp:624-660: thin-lens field of view from the camera-to-target distance, then the
// orthonormal basis U,V,N stored as the ROWS of the orientation matrix.
void CScene::InitialiserCamera()
{
    const REAL d2 = CVecteur3::Distance(m_Camera.Position, m_Camera.PointVise);
    const REAL y2 = (d2 / (m_Camera.Focale * RENDRE_REEL(0.001)) - 1) * DIM_FILM_CAM;
    m_Camera.Angle = (360 * std::atan2(y2 * RENDRE_REEL(0.5), d2)) / RENDRE_REEL(PI);
    const CVecteur3 N = CVecteur3::Normaliser(m_Camera.Position - m_Camera.PointVise);
    const CVecteur3 V = CVecteur3::Normaliser(m_Camera.Up - N * CVecteur3::ProdScal(m_Camera.Up, N));
    const CVecteur3 U = CVecteur3::ProdVect(V, N);
    m_Camera.Orientation = CMatrice4(U.x, U.y, U.z, 0.0f, V.x, V.y, V.z, 0.0f, N.x, N.y, N.z, 0.0f, 0.0f, 0.0f, 0.0f, 1.0f);
}

void CScene::Initialiser()                               // Scene.cpp:140-147
{
    InitialiserCamera();
    for (ISurface* s : m_Surfaces) s->Pretraitement();
    delete[] m_InfoPixel;
    m_InfoPixel = new float[(size_t)m_ResLargeur * m_ResHauteur * 3];
}

// The reference re-runs Pretraitement on every LancerRayons (transforms applied twice on a
// second call, Scene.cpp:140-147 has no guard); here it runs once per loaded scene.
void CScene::Preparer()
{
    if (m_Prepare) {                                     // resolution may have changed since the last frame
        delete[] m_InfoPixel;
        m_InfoPixel = new float[(size_t)m_ResLargeur * m_ResHauteur * 3];
    } else {
        Initialiser();
        m_Prepare = true;
    }
    // ---- SoA packer: replaces the std140 UBO packer of Scene.cpp:799-1269 ------------------------
    SceneEmballee& e = m_Emballee;
    e = SceneEmballee();
    int ordre = 0;
    for (ISurface* s : m_Surfaces) {
        rtb200_material m;
        const CCouleur c = s->ObtenirCouleur();
        m.color[0] = c.r; m.color[1] = c.g; m.color[2] = c.b;
        m.ambient = s->ObtenirCoeffAmbiant(); m.diffuse = s->ObtenirCoeffDiffus();
        m.specular = s->ObtenirCoeffSpeculaire(); m.shininess = s->ObtenirCoeffBrillance();
        m.reflect = s->ObtenirCoeffReflexion(); m.refract = s->ObtenirCoeffRefraction(); m.ior = s->ObtenirIndiceRefraction();
        const int mat = (int)e.materials.size();
        e.materials.push_back(m);
        if (const CTriangle* t = dynamic_cast<const CTriangle*>(s)) {
            for (int k = 0; k < 3; ++k) { const CVecteur3 p = t->ObtenirPoint(k); e.tri_points.insert(e.tri_points.end(), { p.x, p.y, p.z }); }
            const CVecteur3 n = t->ObtenirNormale();
            e.tri_normal.insert(e.tri_normal.end(), { n.x, n.y, n.z });
            e.tri_material.push_back(mat); e.tri_order.push_back(ordre);
        } else if (const CQuadrique* q = dynamic_cast<const CQuadrique*>(s)) {
            const CVecteur3 a = q->ObtenirQuadratique(), mx = q->ObtenirMixte(), l = q->ObtenirLineaire();
            e.quad_coef.insert(e.quad_coef.end(), { a.x, a.y, a.z, mx.x, mx.y, mx.z, l.x, l.y, l.z, q->ObtenirConstante() });
            e.quad_material.push_back(mat); e.quad_order.push_back(ordre);
        } else if (const CPlan* p = dynamic_cast<const CPlan*>(s)) {
            const CVecteur3 n = p->ObtenirNormale();
            e.plane_coef.insert(e.plane_coef.end(), { n.x, n.y, n.z, p->ObtenirConstante() });
            e.plane_material.push_back(mat); e.plane_order.push_back(ordre);
        }
        ++ordre;
    }
    for (const CLumiere* l : m_Lumieres) {
        rtb200_light g;
        const CVecteur3 p = l->GetPosition(); const CCouleur c = l->ObtenirCouleur();
        g.position[0] = p.x; g.position[1] = p.y; g.position[2] = p.z;
        g.color[0] = c.r; g.color[1] = c.g; g.color[2] = c.b; g.intensity = l->GetIntensity();
        e.lights.push_back(g);
    }
    RemplirCadre(e.cadre);
}

// ---- frame block: replaces the "General" UBO (Scene.cpp:704-783) and Scene.cpp:676-679 -----------
void CScene::RemplirCadre(rtb200_frame& f) const
{
    memset(&f, 0, sizeof f);
    f.width = m_ResLargeur; f.height = m_ResHauteur; f.spp_grid = m_GrilleSpp;
    f.max_bounces = m_NbRebondsMax; f.min_energy = m_EnergieMinRayon; f.scene_ior = m_IndiceRefractionScene;
    f.background[0] = m_CouleurArrierePlan.r; f.background[1] = m_CouleurArrierePlan.g; f.background[2] = m_CouleurArrierePlan.b;
    f.cam_pos[0] = m_Camera.Position.x; f.cam_pos[1] = m_Camera.Position.y; f.cam_pos[2] = m_Camera.Position.z;
    for (int c = 0; c < 3; ++c) { f.cam_u[c] = m_Camera.Orientation.m[0][c]; f.cam_v[c] = m_Camera.Orientation.m[1][c]; f.cam_n[c] = m_Camera.Orientation.m[2][c]; }
    f.half_h = std::tan(Deg2Rad<REAL>(m_Camera.Angle * RENDRE_REEL(0.5)));
    f.half_w = (RENDRE_REEL(m_ResLargeur) / m_ResHauteur) * f.half_h;
    f.flags = m_Drapeaux;
}

const SceneEmballee& CScene::ObtenirSceneEmballee()
{
    Preparer();
    return m_Emballee;
}

// ---- render ---------------------------------------------------------------------------------------
void CScene::LancerRayons()
{
    Preparer();
    if (!CVar::g_ComputerShadersON) {
        // The reference's C++ per-pixel loop (Scene.cpp:1535-1563) is deliberately NOT part of this
        // build: the only renderer is the CUDA one.  Fail loudly, like every other fatal error here.
        std::cerr << "[ERREUR]: CVar::g_ComputerShadersON == false: this build has no CPU renderer." << std::endl;
        exit(EXIT_FAILURE);
    }
    if (!CVar::g_ComputeShader) CVar::g_ComputeShader = new CNuanceurCalculProg(CVar::g_ComputeShaderPath, true);
    CVar::g_ComputeShader->activer();
    rtb200_ctx* ctx = CVar::g_ComputeShader->contexte();
    const rtb200_scene vue = m_Emballee.vue();
    if (CVar::g_NbPeripheriques > 1) { LancerRayonsMultiGPU(ctx, vue); return; }
    int rc = rtb200_set_scene(ctx, &vue);                // was: SceneData UBO glBufferData (Scene.cpp:1268)
    if (rc == RTB200_OK && (m_Drapeaux & RTB200_F_USE_BVH)) rc = rtb200_build_bvh(ctx);
    if (rc == RTB200_OK) {
        m_ImageRGBA8.resize((size_t)m_ResLargeur * m_ResHauteur * 4);
        rc = rtb200_render(ctx, &m_Emballee.cadre, m_InfoPixel, m_ImageRGBA8.data(), &m_Stats);   // was: glDispatchCompute (Scene.cpp:1308)
    }
    if (rc != RTB200_OK) {
        std::cerr << "[ERREUR]: rendu CUDA : " << rtb200_error_string(rc) << " -- " << rtb200_last_error(ctx) << std::endl;
        exit(EXIT_FAILURE);
    }
}

// One frame over g GPUs from ONE process (rt_cli --gpus g; SURVEY.md 8b's single-process form): one context and one host
// thread per GPU, band sharding without a communicator (rtb200_set_sharding), scene replicated, and every context
// writes its own interleaved 16-row bands straight into this scene's host frame (RTB200_F_HOST_BANDS) over its own
// PCIe link.  Nothing is exchanged between the GPUs.  With fewer devices than g the contexts share devices (same
// frame, no speed-up).  `premier` = the context CVar::g_ComputeShader owns (rank 0).
void CScene::LancerRayonsMultiGPU(rtb200_ctx* premier, const rtb200_scene& vue)
{
    static std::vector<rtb200_ctx*> autres;              // ranks 1 .. g-1, kept for the life of the process
    const int g = CVar::g_NbPeripheriques;
    int n_dev = 0;
    if (rtb200_device_count(&n_dev) != RTB200_OK || n_dev < 1) { std::cerr << "[ERREUR]: aucun peripherique CUDA" << std::endl; exit(EXIT_FAILURE); }
    while ((int)autres.size() < g - 1) {
        rtb200_ctx* c = nullptr;
        const int dev = (CVar::g_Peripherique + (int)autres.size() + 1) % n_dev;
        const int rc = rtb200_create(dev, &c);
        if (rc != RTB200_OK) { std::cerr << "[ERREUR]: rtb200_create(" << dev << ") : " << rtb200_error_string(rc) << " -- " << rtb200_last_error(nullptr) << std::endl; exit(EXIT_FAILURE); }
        autres.push_back(c);
    }
    m_ImageRGBA8.resize((size_t)m_ResLargeur * m_ResHauteur * 4);
    rtb200_frame cadre = m_Emballee.cadre;
    cadre.flags |= RTB200_F_HOST_BANDS;
    std::vector<int> rcs((size_t)g, RTB200_OK);
    std::vector<rtb200_stats> stats((size_t)g);
    auto travail = [&](int rang) {
        rtb200_ctx* c = rang == 0 ? premier : autres[(size_t)rang - 1];
        int rc = rtb200_set_sharding(c, rang, g);
        if (rc == RTB200_OK) rc = rtb200_set_scene(c, &vue);
        if (rc == RTB200_OK && (m_Drapeaux & RTB200_F_USE_BVH)) rc = rtb200_build_bvh(c);
        if (rc == RTB200_OK) rc = rtb200_render(c, &cadre, m_InfoPixel, m_ImageRGBA8.data(), &stats[(size_t)rang]);
        rcs[(size_t)rang] = rc;
    };
    std::vector<std::thread> fils;
    for (int r = 1; r < g; ++r) fils.emplace_back(travail, r);
    travail(0);
    for (std::thread& t : fils) t.join();
    for (int r = 0; r < g; ++r)
        if (rcs[(size_t)r] != RTB200_OK) {
            rtb200_ctx* c = r == 0 ? premier : autres[(size_t)r - 1];
            std::cerr << "[ERREUR]: rendu CUDA (GPU " << r << ") : " << rtb200_error_string(rcs[(size_t)r]) << " -- " << rtb200_last_error(c) << std::endl;
            exit(EXIT_FAILURE);
        }
    m_Stats = stats[0];                                  // whole-frame counters: sums; times: the slowest GPU
    for (int r = 1; r < g; ++r) {
        const rtb200_stats& s = stats[(size_t)r];
        m_Stats.ms_render = std::max(m_Stats.ms_render, s.ms_render);
        m_Stats.rays_primary += s.rays_primary; m_Stats.rays_shadow += s.rays_shadow;
        m_Stats.rays_reflect += s.rays_reflect; m_Stats.rays_refract += s.rays_refract;
        m_Stats.tests_triangle += s.tests_triangle; m_Stats.tests_quadric += s.tests_quadric; m_Stats.tests_plane += s.tests_plane;
        m_Stats.kernel_launches += s.kernel_launches; m_Stats.nan_pixels += s.nan_pixels;
    }
}

// A camera path as ONE submission (rtb200_render_sequence): the reference draws a single frame per
// process (Main.cpp:181); with the scene resident on the device a fly-through costs one kernel per frame.
// positions[i] replaces the camera position of frame i; the point looked at, the up vector and the focal
// length stay, and every camera goes through InitialiserCamera like the one of the scene file.
double CScene::LancerTrajectoire(const std::vector<CVecteur3>& positions, unsigned char* images)
{
    Preparer();
    if (positions.empty()) return 0.0;
    if (!CVar::g_ComputerShadersON) {
        std::cerr << "[ERREUR]: CVar::g_ComputerShadersON == false: this build has no CPU renderer." << std::endl;
        exit(EXIT_FAILURE);
    }
    if (!CVar::g_ComputeShader) CVar::g_ComputeShader = new CNuanceurCalculProg(CVar::g_ComputeShaderPath, true);
    CVar::g_ComputeShader->activer();
    rtb200_ctx* ctx = CVar::g_ComputeShader->contexte();
    std::vector<rtb200_frame> cadres(positions.size());
    for (size_t i = 0; i < positions.size(); ++i) cadres[i] = CadreDepuis(positions[i]);
    m_StatsTrajectoire.assign(positions.size(), rtb200_stats{});
    const rtb200_scene vue = m_Emballee.vue();
    int rc = rtb200_set_scene(ctx, &vue);
    if (rc == RTB200_OK && (m_Drapeaux & RTB200_F_USE_BVH)) rc = rtb200_build_bvh(ctx);
    if (rc == RTB200_OK) rc = rtb200_render_sequence(ctx, cadres.data(), (int)cadres.size(), images, m_StatsTrajectoire.data());
    if (rc == RTB200_OK) {                               // the last frame is also the scene's current image (RGBA8:
        const size_t octets = (size_t)m_ResLargeur * m_ResHauteur * 4;     // a sequence delivers no float frame)
        m_ImageRGBA8.resize(octets);
        if (images) memcpy(m_ImageRGBA8.data(), images + (positions.size() - 1) * octets, octets);
        else rc = rtb200_read_frame(ctx, nullptr, m_ImageRGBA8.data());
    }
    if (rc != RTB200_OK) {
        std::cerr << "[ERREUR]: rendu CUDA : " << rtb200_error_string(rc) << " -- " << rtb200_last_error(ctx) << std::endl;
        exit(EXIT_FAILURE);
    }
    m_Stats = m_StatsTrajectoire.back();
    double ms = 0.0;
    for (const rtb200_stats& st : m_StatsTrajectoire) ms += st.ms_render;
    return ms;
}

// The frame block of the scene's camera moved to `position` (the scene's own camera is left as it was).
rtb200_frame CScene::CadreDepuis(const CVecteur3& position)
{
    const CVecteur3 depart = m_Camera.Position;
    m_Camera.Position = position;
    InitialiserCamera();
    rtb200_frame f;
    RemplirCadre(f);
    m_Camera.Position = depart;
    InitialiserCamera();
    return f;
}

// n camera positions on a circle around the vertical axis through the point looked at, starting at
// the scene's own camera (frame 0 is the frame LancerRayons draws).
std::vector<CVecteur3> CScene::OrbiteCamera(const int n) const
{
    std::vector<CVecteur3> pos;
    const CVecteur3 d = m_Camera.Position - m_Camera.PointVise;
    for (int i = 0; i < n; ++i) {
        const REAL a = RENDRE_REEL(2.0 * PI) * i / n, c = std::cos(a), s = std::sin(a);
        pos.push_back(i == 0 ? m_Camera.Position
                             : m_Camera.PointVise + CVecteur3(c * d.x + s * d.z, d.y, -s * d.x + c * d.z));
    }
    return pos;
}

bool CScene::EcrireImagePPM(const char* fichier) const
{
    if (m_ImageRGBA8.empty()) return false;
    return EcrirePPM(fichier, m_ImageRGBA8.data(), m_ResLargeur, m_ResHauteur);
}

bool CScene::EcrirePPM(const char* fichier, const unsigned char* rgba8, int largeur, int hauteur)
{
    FILE* f = fopen(fichier, "wb");
    if (!f) return false;
    fprintf(f, "P6\n%d %d\n255\n", largeur, hauteur);
    std::vector<unsigned char> row((size_t)largeur * 3);
    for (int y = hauteur - 1; y >= 0; --y) {             // buffer row 0 is the bottom scanline
        const unsigned char* src = rgba8 + (size_t)y * largeur * 4;
        for (int x = 0; x < largeur; ++x) { row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2]; }
        fwrite(row.data(), 1, row.size(), f);
    }
    fclose(f);
    return true;
}

// ---- C helpers for the Python test harness (ctypes) -------------------------------------------------
extern "C" {

int rtscene_load(const char* path, int max_bounces)
{
    CScene::LibererInstance();
    CScene* s = CScene::ObtenirInstance();
    CVar::g_GestionnaireDeScene = s;
    std::ifstream probe(path);
    if (!probe.is_open()) return -1;
    probe.close();
    s->TraiterFichierDeScene(path);
    if (max_bounces >= 0) s->AjusterNbRebondsMax(max_bounces);
    return (int)s->NbSurfaces();
}

// Appends an OBJ mesh to the current scene: colour r g b, reflection coefficient, uniform scale and a
// translation (applied through the prototype's transformation, i.e. by CTriangle::Pretraitement).
int rtscene_add_obj(const char* path, const float* rgb, float reflect, float scale, const float* translate)
{
    CScene* s = CScene::ObtenirInstance();
    CTriangle modele;
    modele.AjusterCouleur(CCouleur(rgb[0], rgb[1], rgb[2]));
    modele.AjusterCoeffReflexion(reflect);
    CMatrice4 m(CMatrice4::IDENTITE);                    // scale then translate, post-multiplied like the parser
    m.MiseAEchelle(scale, scale, scale);
    m.Translation(translate[0], translate[1], translate[2]);
    modele.AjusterTransformation(m);
    return s->ChargerMaillageOBJ(path, modele);
}

// Packs the current scene for resolution W x H; returns pointers into the scene's own vectors.
int rtscene_pack(int W, int H, rtb200_scene* out_scene, rtb200_frame* out_frame)
{
    CScene* s = CScene::ObtenirInstance();
    s->AjusterResolution(W, H);
    const SceneEmballee& e = s->ObtenirSceneEmballee();
    *out_scene = e.vue();
    *out_frame = e.cadre;
    return 0;
}

// Full drop-in path: CScene::LancerRayons() -> CUDA.  Copies the float and RGBA8 frames out.
int rtscene_render(int W, int H, int spp_grid, unsigned int flags, float* out_rgb, unsigned char* out_rgba8, rtb200_stats* st)
{
    CScene* s = CScene::ObtenirInstance();
    s->AjusterResolution(W, H);
    s->AjusterSurechantillonnage(spp_grid);
    s->AjusterDrapeauxRendu(flags);
    CVar::g_ComputerShadersON = true;
    s->LancerRayons();
    if (out_rgb) memcpy(out_rgb, s->ObtenirInfoPixel(), sizeof(float) * 3 * (size_t)W * H);
    if (out_rgba8) memcpy(out_rgba8, s->ObtenirImageRGBA8(), (size_t)4 * W * H);
    if (st) *st = s->ObtenirStatistiques();
    return 0;
}

// Camera block (basis, field of view) the host classes derive for a camera given through the
// public setters -- used by the synthetic-scene generators so that every camera goes through
// CScene::InitialiserCamera.
int rtscene_camera(const float* pos, const float* look, const float* up, int W, int H, rtb200_frame* out)
{
    CScene::LibererInstance();
    CScene* s = CScene::ObtenirInstance();
    s->AjusterPositionCamera(CVecteur3(pos[0], pos[1], pos[2]));
    s->AjusterPointViseCamera(CVecteur3(look[0], look[1], look[2]));
    s->AjusterVecteurUpCamera(CVecteur3(up[0], up[1], up[2]));
    s->AjusterResolution(W, H);
    *out = s->ObtenirSceneEmballee().cadre;
    CScene::LibererInstance();
    return 0;
}

// CScene::LancerTrajectoire over CScene::OrbiteCamera(n): n frames (n * W*H*4 bytes) and the camera
// blocks they were drawn with (so that a test can render the same frames one by one).
int rtscene_orbit(int W, int H, unsigned int flags, int n, unsigned char* out_rgba8, rtb200_frame* out_frames, rtb200_stats* st)
{
    CScene* s = CScene::ObtenirInstance();
    s->AjusterResolution(W, H);
    s->AjusterSurechantillonnage(1);
    s->AjusterDrapeauxRendu(flags);
    CVar::g_ComputerShadersON = true;
    s->Preparer();
    const std::vector<CVecteur3> pos = s->OrbiteCamera(n);
    s->LancerTrajectoire(pos, out_rgba8);
    for (int i = 0; i < n; ++i) {
        if (st) st[i] = s->ObtenirStatistiquesTrajectoire()[i];
        if (out_frames) out_frames[i] = s->CadreDepuis(pos[i]);
    }
    return 0;
}

// The camera blocks alone (no device needed): what LancerTrajectoire would submit for OrbiteCamera(n).
int rtscene_orbit_frames(int W, int H, int n, rtb200_frame* out_frames)
{
    CScene* s = CScene::ObtenirInstance();
    s->AjusterResolution(W, H);
    s->Preparer();
    const std::vector<CVecteur3> pos = s->OrbiteCamera(n);
    for (int i = 0; i < n; ++i) out_frames[i] = s->CadreDepuis(pos[i]);
    return 0;
}

void rtscene_free(void)
{
    CScene::LibererInstance();
    delete CVar::g_ComputeShader;
    CVar::g_ComputeShader = nullptr;
}

} // extern "C"
