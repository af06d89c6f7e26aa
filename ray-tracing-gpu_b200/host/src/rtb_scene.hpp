// rtb_scene.hpp -- CScene (singleton scene manager), CVar (global switches), CStringUtils and
// CNuanceurCalculProg, with the reference's public interface (Scene.h:39-70, Var.h:29-57,
// StringUtils.h, NuanceurCalculProg.h:73-132, Singleton.h).  What changed underneath:
//   * CNuanceurCalculProg no longer compiles a GLSL compute shader; it is the RAII owner of an
//     rtb200_ctx (librtb200.so, hand-written sm_100a CUDA).
//   * CScene::LancerRayons() keeps its signature and its `if (CVar::g_ComputerShadersON)`
//     plug-in point (Scene.cpp:692) but the GPU branch packs the scene as structure-of-arrays
//     and calls rtb200_set_scene / rtb200_render instead of the std140 UBO uploads and
//     glDispatchCompute.  There is no CPU rendering branch in this build.
#pragma once
#include <string>
#include <vector>

#include "rtb200.h"
#include "rtb_surfaces.hpp"

typedef unsigned int GLuint;                           // the reference leaks GL types in its API

template <class T>
class Singleton {                                      // Singleton.h: lazily created, not thread-safe
public:
    static T* ObtenirInstance() { if (!instance_) instance_ = new T(); return instance_; }
    static void LibererInstance() { delete instance_; instance_ = nullptr; }
protected:
    Singleton() {}
    ~Singleton() {}
private:
    static T* instance_;
    Singleton(const Singleton&);
    Singleton& operator=(const Singleton&);
};
template <typename T> T* Singleton<T>::instance_ = nullptr;

class CStringUtils {
public:
    static std::string TrimRight(const std::string& s, const std::string& t = " ");
    static std::string TrimLeft(const std::string& s, const std::string& t = " ");
    static std::string Trim(const std::string& s, const std::string& t = " ");
};

// Owner of the CUDA render context; stands where the GLSL program object stood.
class CNuanceurCalculProg {
public:
    CNuanceurCalculProg();
    CNuanceurCalculProg(char* nuanceurCalculsStr, bool compilerImmediatement);   // the path is ignored: kernels are in librtb200.so
    ~CNuanceurCalculProg();
    void compilerEtLier();                             // creates the rtb200 context on CVar::g_Peripherique
    void activer();                                    // makes sure the context exists (was glUseProgram)
    void uniform1(char* nom, int v);                   // kept for source compatibility: no-ops
    void uniform1(char* nom, float v);
    void enregistrerUniformFloat(const char* nom, const float& val);
    void enregistrerUniformInteger(const char* nom, const int& val);
    GLuint getProg();                                  // 1 when the context is live, 0 otherwise
    rtb200_ctx* contexte() { return ctx_; }
private:
    rtb200_ctx* ctx_ = nullptr;
    bool estVide_ = true, estCompileEtLie_ = false;
};

namespace Scene {

// Structure-of-arrays image of the scene after Pretraitement (what rtb200_set_scene consumes).
struct SceneEmballee {
    std::vector<float> tri_points, tri_normal, quad_coef, plane_coef;
    std::vector<int32_t> tri_material, tri_order, quad_material, quad_order, plane_material, plane_order;
    std::vector<rtb200_material> materials;
    std::vector<rtb200_light> lights;
    rtb200_frame cadre;                                // camera basis, field of view, background, knobs
    rtb200_scene vue() const;                          // pointers into the vectors above
};

class CScene : public Singleton<CScene> {
    friend class Singleton<CScene>;
public:
    void AjouterSurface(ISurface* const surface);
    void AjouterLumiere(CLumiere* const lumiere);
    void LancerRayons();

    void AjusterPositionCamera(const CVecteur3& position);
    void AjusterPointViseCamera(const CVecteur3& pointVise);
    void AjusterVecteurUpCamera(const CVecteur3& up);
    void AjusterResolution(const int resX, const int resY);
    void AjusterCouleurArrierePlan(const CCouleur& couleur);
    void AjusterNbRebondsMax(const int nbRebondsMax);
    void AjusterEnergieMinimale(const REAL energieMinimale);
    void AjusterIndiceRefraction(const REAL indiceRefraction);

    GLuint ObtenirTextureGL();                         // no GL here: always 0; use the accessors below
    void TraiterFichierDeScene(const char* fichier);
    void Liberer();

    // ---- additions of the headless B200 build -----------------------------------------------
    void AjusterSurechantillonnage(const int grille) { m_GrilleSpp = grille; }       // k x k samples per pixel
    // Triangle-mesh loader (the roadmap remark of Main.cpp:13-16): every face of a Wavefront OBJ file
    // (v / f records, polygons fan-triangulated, negative indices allowed) becomes one CTriangle that
    // copies material AND transformation from `modele`, appended in file order through AjouterSurface.
    // Returns the number of triangles added, -1 when the file cannot be read.
    int  ChargerMaillageOBJ(const char* fichier, const CTriangle& modele);
    void AjusterDrapeauxRendu(const unsigned int drapeaux) { m_Drapeaux = drapeaux; } // RTB200_F_*
    const float* ObtenirInfoPixel() const { return m_InfoPixel; }                     // W*H*3 floats, row 0 = bottom
    const unsigned char* ObtenirImageRGBA8() const { return m_ImageRGBA8.data(); }    // W*H*4 bytes
    const rtb200_stats& ObtenirStatistiques() const { return m_Stats; }
    void LancerRayonsMultiGPU(rtb200_ctx* premier, const rtb200_scene& vue);   // CVar::g_NbPeripheriques > 1 (rt_cli --gpus g)
    int ObtenirLargeur() const { return m_ResLargeur; }
    int ObtenirHauteur() const { return m_ResHauteur; }
    size_t NbSurfaces() const { return m_Surfaces.size(); }
    size_t NbLumieres() const { return m_Lumieres.size(); }
    // Camera + Pretraitement (once) and the packed scene; LancerRayons() calls these itself.
    void Preparer();
    const SceneEmballee& ObtenirSceneEmballee();
    bool EcrireImagePPM(const char* fichier) const;   // P6, rows flipped so that the top scanline comes first
    // Camera path in one pipelined submission (rtb200_render_sequence): frame i is drawn from positions[i]
    // with the scene's look-at point and up vector.  `images` (n * W*H*4 bytes, may be NULL) receives every
    // frame; the last one is also left in ObtenirImageRGBA8 / ObtenirInfoPixel.  Returns the GPU time (ms).
    double LancerTrajectoire(const std::vector<CVecteur3>& positions, unsigned char* images);
    rtb200_frame CadreDepuis(const CVecteur3& position);      // frame block of the camera moved to `position`
    std::vector<CVecteur3> OrbiteCamera(const int n) const;   // n positions on a circle around the look-at point
    const std::vector<rtb200_stats>& ObtenirStatistiquesTrajectoire() const { return m_StatsTrajectoire; }
    static bool EcrirePPM(const char* fichier, const unsigned char* rgba8, int largeur, int hauteur);

private:
    CScene();
    ~CScene();
    CScene(const CScene&);
    CScene& operator=(const CScene&);

    enum EtatTraitementScene { TRAITEMENT_SCENE, TRAITEMENT_LUMIERE, TRAITEMENT_TRIANGLE, TRAITEMENT_PLAN, TRAITEMENT_QUADRIQUE };
    static const REAL DIM_FILM_CAM;
    struct CameraDeScene { CVecteur3 Position, PointVise, Up; CMatrice4 Orientation; REAL Focale, Angle; };

    void Initialiser();
    void InitialiserCamera();

    int m_ResLargeur = 0, m_ResHauteur = 0;
    float* m_InfoPixel = nullptr;
    std::vector<unsigned char> m_ImageRGBA8;
    CameraDeScene m_Camera;
    CCouleur m_CouleurArrierePlan;
    std::vector<ISurface*> m_Surfaces;
    std::vector<CLumiere*> m_Lumieres;
    REAL m_IndiceRefractionScene = RENDRE_REEL(1.0);
    int  m_NbRebondsMax = 20;                          // Scene.cpp:68
    REAL m_EnergieMinRayon = RENDRE_REEL(0.01);        // Scene.cpp:69
    int  m_GrilleSpp = 1;
    unsigned int m_Drapeaux = 0;
    bool m_Prepare = false;
    SceneEmballee m_Emballee;
    rtb200_stats m_Stats{};
    std::vector<rtb200_stats> m_StatsTrajectoire;
    void RemplirCadre(rtb200_frame& cadre) const;
};

} // namespace Scene

class CVar : public Singleton<CVar> {
    friend class Singleton<CVar>;
    CVar() {}
    ~CVar() {}
public:
    static int g_FenetreID, g_LargeurFenetre, g_HauteurFenetre, g_LargeurViewport, g_HauteurViewport;
    static Scene::CScene* g_GestionnaireDeScene;
    static bool g_ComputerShadersON;                   // true = render on the GPU (now: CUDA on a B200)
    static CNuanceurCalculProg* g_ComputeShader;
    static char* g_ComputeShaderPath;                  // kept; unused (no GLSL file is loaded)
    static int g_Peripherique;                         // addition: CUDA device index
    static int g_NbPeripheriques;                      // addition: GPUs one frame is spread over (rt_cli --gpus g); 1 = the reference's single device
};
