// rtb_surfaces.hpp -- the scene objects of the drop-in API: CRayon, CIntersection, CLumiere,
// ISurface and its three concrete surfaces.  Same public names, signatures and defaults as the
// reference (Rayon.h, Intersection.h, Lumiere.h, ISurface.h, Triangle.h, Plan.h, Quadrique.h).
// What the renderer consumes is the state AFTER Pretraitement(), read through the getters by
// the packer in rtb_scene.cpp.  Intersection() is kept because it is part of the public
// ISurface contract; the render path never calls it (rendering happens on the GPU only).
#pragma once
#include "rtb_colour.hpp"

using namespace Math3D;

class CRayon {                                        // defaults: Rayon.cpp:13-19 (ior 0, energy 0)
    CVecteur3 m_Origine, m_Direction;
    REAL m_IndiceDeRefraction = RENDRE_REEL(0.0), m_Energie = RENDRE_REEL(0.0);
    int  m_NbRebonds = 0;
public:
    CRayon() = default;
    void AjusterOrigine(const CVecteur3& o) { m_Origine = o; }
    void AjusterDirection(const CVecteur3& d) { m_Direction = d; }
    void AjusterIndiceRefraction(const REAL i) { m_IndiceDeRefraction = i; }
    void AjusterEnergie(const REAL e) { m_Energie = e; }
    void AjusterNbRebonds(const int n) { m_NbRebonds = n; }
    CVecteur3 ObtenirOrigine() const { return m_Origine; }
    CVecteur3 ObtenirDirection() const { return m_Direction; }
    REAL ObtenirIndiceRefraction() const { return m_IndiceDeRefraction; }
    REAL ObtenirEnergie() const { return m_Energie; }
    int  ObtenirNbRebonds() const { return m_NbRebonds; }
};

namespace Scene {

class ISurface;

class CIntersection {                                 // defaults: Intersection.cpp:15-19 (distance -1)
    ISurface* m_Surface = nullptr;
    REAL m_Distance = RENDRE_REEL(-1);
    CVecteur3 m_Normale;
public:
    CIntersection() = default;
    void AjusterSurface(ISurface* s) { m_Surface = s; }
    void AjusterDistance(const REAL d) { m_Distance = d; }
    void AjusterNormale(const CVecteur3& n) { m_Normale = n; }
    ISurface* ObtenirSurface() const { return m_Surface; }
    REAL ObtenirDistance() const { return m_Distance; }
    CVecteur3 ObtenirNormale() const { return m_Normale; }
};

class CLumiere {                                      // defaults: Lumiere.cpp:15-19 (white, intensity 0)
    CVecteur3 m_Position;
    CCouleur  m_Couleur = CCouleur(1.0f);
    REAL      m_Intensite = RENDRE_REEL(0);
public:
    CLumiere() = default;
    void SetPosition(const CVecteur3& p) { m_Position = p; }
    void AjusterCouleur(const CCouleur& c) { m_Couleur = c; }
    void SetIntensity(const REAL i) { m_Intensite = i; }
    CVecteur3 GetPosition() const { return m_Position; }
    CCouleur  ObtenirCouleur() const { return m_Couleur; }
    REAL      GetIntensity() const { return m_Intensite; }
};

class ISurface {                                      // material defaults: ISurface.cpp:15-25
protected:
    CCouleur  m_Couleur;                              // black
    REAL m_CoeffAmbient = RENDRE_REEL(0.2), m_CoeffDiffus = RENDRE_REEL(0.8);
    REAL m_CoeffSpeculaire = 0, m_CoeffShininess = 0, m_CoeffReflexion = 0, m_CoeffRefraction = 0, m_IndiceDeRefraction = 0;
    CMatrice4 m_Transformation = CMatrice4::IDENTITE;
    virtual std::ostream& AfficherInfoDebug(std::ostream& out) const = 0;
public:
    ISurface() = default;
    ISurface(const ISurface&) = default;
    ISurface& operator=(const ISurface&) = default;
    virtual ~ISurface() {}

    virtual void          Pretraitement() = 0;
    virtual CIntersection Intersection(const CRayon& rayon) = 0;
    virtual ISurface*     Copier() const = 0;

    void AjusterCouleur(const CCouleur& c) { m_Couleur = c; }
    void AjusterCoeffAmbiant(const REAL v) { m_CoeffAmbient = v; }
    void AjusterCoeffDiffus(const REAL v) { m_CoeffDiffus = v; }
    void AjusterCoeffSpeculaire(const REAL v) { m_CoeffSpeculaire = v; }
    void AjusterCoeffBrillance(const REAL v) { m_CoeffShininess = v; }
    void AjusterCoeffReflexion(const REAL v) { m_CoeffReflexion = v; }
    void AjusterCoeffRefraction(const REAL v) { m_CoeffRefraction = v; }
    void AjusterIndiceRefraction(const REAL v) { m_IndiceDeRefraction = v; }
    void AjusterTransformation(const CMatrice4& t) { m_Transformation = t; }
    CCouleur  ObtenirCouleur() const { return m_Couleur; }
    REAL ObtenirCoeffAmbiant() const { return m_CoeffAmbient; }
    REAL ObtenirCoeffDiffus() const { return m_CoeffDiffus; }
    REAL ObtenirCoeffSpeculaire() const { return m_CoeffSpeculaire; }
    REAL ObtenirCoeffBrillance() const { return m_CoeffShininess; }
    REAL ObtenirCoeffReflexion() const { return m_CoeffReflexion; }
    REAL ObtenirCoeffRefraction() const { return m_CoeffRefraction; }
    REAL ObtenirIndiceRefraction() const { return m_IndiceDeRefraction; }
    CMatrice4 ObtenirTransformation() const { return m_Transformation; }

    friend std::ostream& operator<<(std::ostream& out, const ISurface& s);
};

class CTriangle : public ISurface {
    CVecteur3 m_Pts[3];
    CVecteur3 m_Normale;
    void CalculerNormale();
protected:
    std::ostream& AfficherInfoDebug(std::ostream& out) const override;
public:
    CTriangle() = default;
    void AjusterPoints(const CVecteur3& a, const CVecteur3& b, const CVecteur3& c) { m_Pts[0] = a; m_Pts[1] = b; m_Pts[2] = c; }
    void AjusterPoint(const int i, const CVecteur3& p) { assert(i > -1 && i < 3); m_Pts[i] = p; }
    void AjusterNormale(const CVecteur3& n) { m_Normale = n; }
    CVecteur3 ObtenirPoint(const int i) const { assert(i > -1 && i < 3); return m_Pts[i]; }
    CVecteur3 ObtenirNormale() const { return m_Normale; }
    void          Pretraitement() override;
    CIntersection Intersection(const CRayon& rayon) override;
    CTriangle*    Copier() const override { return new CTriangle(*this); }
};

class CPlan : public ISurface {
    CVecteur3 m_Normale;
    REAL      m_Cst = RENDRE_REEL(0);
protected:
    std::ostream& AfficherInfoDebug(std::ostream& out) const override;
public:
    CPlan() = default;
    void AjusterNormale(const CVecteur3& n) { m_Normale = n; }
    void AjusterConstante(const REAL c) { m_Cst = c; }
    CVecteur3 ObtenirNormale() const { return m_Normale; }
    REAL      ObtenirConstante() const { return m_Cst; }
    void          Pretraitement() override;
    CIntersection Intersection(const CRayon& rayon) override;
    CPlan*        Copier() const override { return new CPlan(*this); }
};

class CQuadrique : public ISurface {
    CVecteur3 m_Quadratique, m_Lineaire, m_Mixte;     // A B C | G H J | mixed (yz, xz, xy)
    REAL      m_Cst = RENDRE_REEL(0);
protected:
    std::ostream& AfficherInfoDebug(std::ostream& out) const override;
public:
    CQuadrique() = default;
    void AjusterQuadratique(const CVecteur3& q) { m_Quadratique = q; }
    void AjusterLineaire(const CVecteur3& l) { m_Lineaire = l; }
    void AjusterMixte(const CVecteur3& m) { m_Mixte = m; }
    void AjusterConstante(const REAL c) { m_Cst = c; }
    CVecteur3 ObtenirQuadratique() const { return m_Quadratique; }
    CVecteur3 ObtenirLineaire() const { return m_Lineaire; }
    CVecteur3 ObtenirMixte() const { return m_Mixte; }
    REAL      ObtenirConstante() const { return m_Cst; }
    void          Pretraitement() override;
    CIntersection Intersection(const CRayon& rayon) override;
    CQuadrique*   Copier() const override { return new CQuadrique(*this); }
};

} // namespace Scene
