// rt_cli.cpp -- headless command line of the renderer: the reference's `exe scene.dat -x W -y H`
// (Main.cpp:51-124) without GLUT/GLEW: the frame goes to a PPM file instead of a window.
//   rt_cli scene.dat [-x W] [-y H] [--depth d] [--spp k] [--device n] [--gpus g] [--fused] [--exact] [--bvh]
//          [--frames n] [--out image.ppm] [--obj mesh.obj] [--orbit n]
// --gpus g: one frame over g GPUs (one context + thread per GPU, interleaved 16-row bands, every GPU writes its bands
// into the host frame itself); not combined with --orbit.
// --orbit n: a camera path of n frames on a circle around the point looked at, as one pipelined
// submission (CScene::LancerTrajectoire); with --out f.ppm the frames go to f_000.ppm, f_001.ppm, ...
// Unlike Main.cpp:79 there is no power-of-two / multiple-of-16 restriction on W and H.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "Scene.h"
#include "Var.h"
#include "NuanceurCalculProg.h"

using namespace Scene;

int main(int argc, char** argv)
{
    if (argc < 2) {
        std::cerr << "usage: " << argv[0] << " scene.dat [-x W] [-y H] [--depth d] [--spp k] [--device n] [--gpus g] [--fused] [--exact] [--bvh] [--frames n] [--out f.ppm] [--obj mesh.obj] [--orbit n]" << std::endl;
        return EXIT_FAILURE;
    }
    int W = CVar::g_LargeurFenetre, H = CVar::g_HauteurFenetre, depth = -1, spp = 1, frames = 1, orbit = 0;
    unsigned int flags = 0;
    const char* out = nullptr;
    const char* obj = nullptr;
    for (int i = 2; i < argc; ++i) {
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { std::cerr << "[ERREUR]: " << what << " needs a value" << std::endl; exit(EXIT_FAILURE); }
            return argv[++i];
        };
        if (!strcmp(argv[i], "-x")) W = atoi(next("-x"));
        else if (!strcmp(argv[i], "-y")) H = atoi(next("-y"));
        else if (!strcmp(argv[i], "--depth")) depth = atoi(next("--depth"));
        else if (!strcmp(argv[i], "--spp")) spp = atoi(next("--spp"));
        else if (!strcmp(argv[i], "--device")) CVar::g_Peripherique = atoi(next("--device"));
        else if (!strcmp(argv[i], "--gpus")) CVar::g_NbPeripheriques = atoi(next("--gpus"));
        else if (!strcmp(argv[i], "--frames")) frames = atoi(next("--frames"));
        else if (!strcmp(argv[i], "--orbit")) orbit = atoi(next("--orbit"));
        else if (!strcmp(argv[i], "--out")) out = next("--out");
        else if (!strcmp(argv[i], "--obj")) obj = next("--obj");
        else if (!strcmp(argv[i], "--fused")) flags |= RTB200_F_FUSED_FP;
        else if (!strcmp(argv[i], "--exact")) flags |= RTB200_F_EXACT_FP;
        else if (!strcmp(argv[i], "--bvh")) flags |= RTB200_F_USE_BVH;
        else { std::cerr << "[ERREUR]: unknown option " << argv[i] << std::endl; return EXIT_FAILURE; }
    }
    if (W <= 0 || H <= 0) { std::cerr << "[ERREUR]: bad resolution" << std::endl; return EXIT_FAILURE; }
    if (CVar::g_NbPeripheriques < 1 || CVar::g_NbPeripheriques > 64 || (CVar::g_NbPeripheriques > 1 && orbit > 0)) {
        std::cerr << "[ERREUR]: --gpus wants 1..64 and no --orbit" << std::endl; return EXIT_FAILURE;
    }

    CVar::g_GestionnaireDeScene = CScene::ObtenirInstance();
    CScene* scene = CVar::g_GestionnaireDeScene;
    scene->AjusterResolution(W, H);
    std::cout << "[ETAT]: Traitement du fichier de scene " << argv[1] << std::endl;
    scene->TraiterFichierDeScene(argv[1]);
    if (obj) {                                             // triangle mesh appended to the scene file's surfaces
        CTriangle modele;
        modele.AjusterCouleur(CCouleur(0.8f, 0.8f, 0.8f));
        const int n = scene->ChargerMaillageOBJ(obj, modele);
        if (n < 0) { std::cerr << "[ERREUR]: cannot read " << obj << std::endl; return EXIT_FAILURE; }
        std::cout << "[ETAT]: " << n << " triangles charges de " << obj << std::endl;
        if (n > 64) flags |= RTB200_F_USE_BVH;             // meshes go through the LBVH
    }
    if (scene->NbSurfaces() == 0) { std::cerr << "[ERREUR]: empty scene (file missing or without surfaces)" << std::endl; return EXIT_FAILURE; }
    if (depth >= 0) scene->AjusterNbRebondsMax(depth);
    scene->AjusterSurechantillonnage(spp);
    scene->AjusterDrapeauxRendu(flags);
    CVar::g_ComputerShadersON = true;
    CVar::g_ComputeShader = new CNuanceurCalculProg(CVar::g_ComputeShaderPath, true);

    std::cout << "[ETAT]: Lancer de rayons..." << std::endl;
    if (orbit > 0) {
        const std::vector<CVecteur3> positions = scene->OrbiteCamera(orbit);
        std::vector<unsigned char> images(out ? (size_t)orbit * W * H * 4 : 0);
        const auto t0 = std::chrono::steady_clock::now();
        const double ms = scene->LancerTrajectoire(positions, out ? images.data() : nullptr);
        const auto t1 = std::chrono::steady_clock::now();
        double rays = 0.0;
        for (const rtb200_stats& st : scene->ObtenirStatistiquesTrajectoire())
            rays += (double)(st.rays_primary + st.rays_shadow + st.rays_reflect + st.rays_refract);
        std::cout << "[ETAT]: Trajectoire de " << orbit << " images terminee --> " << std::chrono::duration<double>(t1 - t0).count() << " secondes"
                  << " (GPU " << ms << " ms, " << ms / orbit << " ms/image, " << rays / (ms * 1e3) << " Mrays/s)" << std::endl;
        if (out) {
            std::string base(out);
            const size_t dot = base.rfind(".ppm");
            if (dot != std::string::npos && dot + 4 == base.size()) base.erase(dot);
            for (int f = 0; f < orbit; ++f) {
                char suffixe[32];
                snprintf(suffixe, sizeof suffixe, "_%03d.ppm", f);
                const std::string nom = base + suffixe;
                if (!CScene::EcrirePPM(nom.c_str(), images.data() + (size_t)f * W * H * 4, W, H)) { std::cerr << "[ERREUR]: cannot write " << nom << std::endl; return EXIT_FAILURE; }
            }
            std::cout << "[ETAT]: " << orbit << " images ecrites dans " << base << "_NNN.ppm" << std::endl;
        }
        CScene::LibererInstance();
        delete CVar::g_ComputeShader;
        return EXIT_SUCCESS;
    }
    for (int f = 0; f < frames; ++f) {
        const auto t0 = std::chrono::steady_clock::now();
        scene->LancerRayons();
        const auto t1 = std::chrono::steady_clock::now();
        const rtb200_stats& st = scene->ObtenirStatistiques();
        const double rays = (double)(st.rays_primary + st.rays_shadow + st.rays_reflect + st.rays_refract);
        std::cout << "[ETAT]: Termine! --> Temps total de rendu : " << std::chrono::duration<double>(t1 - t0).count() << " secondes"
                  << " (GPU " << st.ms_render << " ms, " << rays / 1e6 << " Mrays, " << rays / (st.ms_render * 1e3) << " Mrays/s)" << std::endl;
    }
    if (out) {
        if (!scene->EcrireImagePPM(out)) { std::cerr << "[ERREUR]: cannot write " << out << std::endl; return EXIT_FAILURE; }
        std::cout << "[ETAT]: image ecrite dans " << out << std::endl;
    }
    CScene::LibererInstance();
    delete CVar::g_ComputeShader;
    return EXIT_SUCCESS;
}
