// rtb_surfaces.cpp -- out-of-line parts of the host math + surface classes.
#include "rtb_surfaces.hpp"

using namespace Math3D;
using namespace Scene;

const CVecteur3 CVecteur3::ZERO(0, 0, 0);
const CVecteur3 CVecteur3::UNIT_X(1, 0, 0);
const CVecteur3 CVecteur3::UNIT_Y(0, 1, 0);
const CVecteur3 CVecteur3::UNIT_Z(0, 0, 1);
const CMatrice4 CMatrice4::ZERO(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
const CMatrice4 CMatrice4::IDENTITE(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1);
const float CCouleur::UBYTE_2_FLOAT = 1.0f / 255.0f;
const CCouleur CCouleur::NOIR = CCouleur(0.0f);
const CCouleur CCouleur::BLANC = CCouleur(1.0f);

// ---- CMatrice4 ---------------------------------------------------------------------------
// Cofactor inverse.  The 2x2 minors and 3-term cofactors are evaluated in the operation order of
// Matrice4.cpp:29-88 (a*b - c*d ; p*x - q*y + r*z, left to right) so that CQuadrique::
// Pretraitement yields the reference's coefficients bit for bit.
namespace {
inline REAL minor2(REAL a, REAL b, REAL c, REAL d) { return a * b - c * d; }
inline REAL cof3(REAL p, REAL x, REAL q, REAL y, REAL r, REAL z) { return p * x - q * y + r * z; }
}

const CMatrice4 CMatrice4::Inverse() const
{
    const REAL(&a)[4][4] = m;
    REAL s[6];
    auto lower = [&](int r0, int r1) {               // minors of rows r0, r1 over column pairs
        s[0] = minor2(a[r0][0], a[r1][1], a[r0][1], a[r1][0]);
        s[1] = minor2(a[r0][0], a[r1][2], a[r0][2], a[r1][0]);
        s[2] = minor2(a[r0][0], a[r1][3], a[r0][3], a[r1][0]);
        s[3] = minor2(a[r0][1], a[r1][2], a[r0][2], a[r1][1]);
        s[4] = minor2(a[r0][1], a[r1][3], a[r0][3], a[r1][1]);
        s[5] = minor2(a[r0][2], a[r1][3], a[r0][3], a[r1][2]);
    };
    // cofactor of row `r` against the current minors: columns (1,2,3), (0,2,3), (0,1,3), (0,1,2)
    auto c0 = [&](int r) { return cof3(s[5], a[r][1], s[4], a[r][2], s[3], a[r][3]); };
    auto c1 = [&](int r) { return cof3(s[5], a[r][0], s[2], a[r][2], s[1], a[r][3]); };
    auto c2 = [&](int r) { return cof3(s[4], a[r][0], s[2], a[r][1], s[0], a[r][3]); };
    auto c3 = [&](int r) { return cof3(s[3], a[r][0], s[1], a[r][1], s[0], a[r][2]); };

    lower(2, 3);
    const REAL t00 = c0(1), t10 = -c1(1), t20 = c2(1), t30 = -c3(1);
    const REAL inv = RENDRE_REEL(1) / (t00 * a[0][0] + t10 * a[0][1] + t20 * a[0][2] + t30 * a[0][3]);
    CMatrice4 r;
    r.m[0][0] = t00 * inv; r.m[1][0] = t10 * inv; r.m[2][0] = t20 * inv; r.m[3][0] = t30 * inv;
    r.m[0][1] = -c0(0) * inv; r.m[1][1] = c1(0) * inv; r.m[2][1] = -c2(0) * inv; r.m[3][1] = c3(0) * inv;
    lower(1, 3);
    r.m[0][2] = c0(0) * inv; r.m[1][2] = -c1(0) * inv; r.m[2][2] = c2(0) * inv; r.m[3][2] = -c3(0) * inv;
    // last column: minors of rows (2,1) taken with the opposite orientation (v = a2j*a1i - a2i*a1j)
    s[0] = minor2(a[2][1], a[1][0], a[2][0], a[1][1]);
    s[1] = minor2(a[2][2], a[1][0], a[2][0], a[1][2]);
    s[2] = minor2(a[2][3], a[1][0], a[2][0], a[1][3]);
    s[3] = minor2(a[2][2], a[1][1], a[2][1], a[1][2]);
    s[4] = minor2(a[2][3], a[1][1], a[2][1], a[1][3]);
    s[5] = minor2(a[2][3], a[1][2], a[2][2], a[1][3]);
    r.m[0][3] = -c0(0) * inv; r.m[1][3] = c1(0) * inv; r.m[2][3] = -c2(0) * inv; r.m[3][3] = c3(0) * inv;
    return r;
}

const CMatrice4 CMatrice4::Transpose() const
{
    CMatrice4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = m[j][i];
    return r;
}

// ---- ISurface -------------------------------------------------------------------------------
namespace Scene {
std::ostream& operator<<(std::ostream& out, const ISurface& s)
{
    s.AfficherInfoDebug(out);
    out << "\n[DEBUG]: Surface.Couleur          = " << s.m_Couleur
        << "\n[DEBUG]: Surface.Ambiant          = " << s.m_CoeffAmbient
        << "\n[DEBUG]: Surface.Diffus           = " << s.m_CoeffDiffus
        << "\n[DEBUG]: Surface.Speculaire       = " << s.m_CoeffSpeculaire
        << "\n[DEBUG]: Surface.Shininess        = " << s.m_CoeffShininess
        << "\n[DEBUG]: Surface.Reflexion        = " << s.m_CoeffReflexion
        << "\n[DEBUG]: Surface.Refraction       = " << s.m_CoeffRefraction
        << "\n[DEBUG]: Surface.IndiceRefraction = " << s.m_IndiceDeRefraction
        << "\n[DEBUG]: Surface.Transformation   = " << s.m_Transformation;
    return out;
}
}

// ---- CTriangle ------------------------------------------------------------------------------
std::ostream& CTriangle::AfficherInfoDebug(std::ostream& out) const
{
    for (int i = 0; i < 3; ++i) out << "[DEBUG]: Triangle.Point" << (i + 1) << "         = " << m_Pts[i] << "\n";
    return out << "[DEBUG]: Triangle.Normale        = " << m_Normale;
}

void CTriangle::CalculerNormale()
{
    m_Normale = CVecteur3::Normaliser(CVecteur3::ProdVect(m_Pts[1] - m_Pts[0], m_Pts[2] - m_Pts[0]));
}

// Triangle.cpp:108-113: points through the transformation (as points), then the face normal
void CTriangle::Pretraitement()
{
    for (CVecteur3& p : m_Pts) p = p * m_Transformation;
    CalculerNormale();
}

// Moller-Trumbore, two-sided, |det| < EPSILON rejects (Triangle.cpp:127-172)
CIntersection CTriangle::Intersection(const CRayon& rayon)
{
    CIntersection hit;
    const CVecteur3 e1 = m_Pts[1] - m_Pts[0], e2 = m_Pts[2] - m_Pts[0];
    const CVecteur3 d = rayon.ObtenirDirection();
    const CVecteur3 P = CVecteur3::ProdVect(d, e2);
    const REAL det = CVecteur3::ProdScal(e1, P);
    if (Abs<REAL>(det) < EPSILON) return hit;
    const REAL inv = RENDRE_REEL(1.0) / det;
    const CVecteur3 S = rayon.ObtenirOrigine() - m_Pts[0];
    const REAL u = CVecteur3::ProdScal(S, P) * inv;
    if (u < 0 || u > 1) return hit;
    const CVecteur3 Q = CVecteur3::ProdVect(S, e1);
    const REAL v = CVecteur3::ProdScal(d, Q) * inv;
    if (v < 0 || u + v > 1) return hit;
    hit.AjusterSurface(this);
    hit.AjusterDistance(CVecteur3::ProdScal(e2, Q) * inv);
    hit.AjusterNormale(m_Normale);
    return hit;
}

// ---- CPlan ----------------------------------------------------------------------------------
std::ostream& CPlan::AfficherInfoDebug(std::ostream& out) const
{
    return out << "[DEBUG]: Plan.Normal            = " << m_Normale << "\n[DEBUG]: Plan.Cst               = " << m_Cst;
}

// Plan.cpp:101-114, quirks included: the normal goes through the matrix as a POINT (translation
// row and all) before normalisation; the anchor point uses the NEW normal with the OLD constant.
void CPlan::Pretraitement()
{
    m_Normale = CVecteur3::Normaliser(m_Normale * m_Transformation);
    CVecteur3 anchor;
    for (int i = 0; i < 3; ++i)
        if (m_Normale[i] != 0) anchor[i] = -(m_Cst / m_Normale[i]);
    anchor = anchor * m_Transformation;
    m_Cst = CVecteur3::ProdScal(-m_Normale, anchor);
}

CIntersection CPlan::Intersection(const CRayon& rayon)            // Plan.cpp:128-144
{
    CIntersection hit;
    const REAL vd = CVecteur3::ProdScal(m_Normale, rayon.ObtenirDirection());
    if (Abs<REAL>(vd) > EPSILON) {
        hit.AjusterSurface(this);
        hit.AjusterDistance(-(CVecteur3::ProdScal(m_Normale, rayon.ObtenirOrigine()) + m_Cst) / vd);
        hit.AjusterNormale(m_Normale);
    }
    return hit;
}

// ---- CQuadrique -----------------------------------------------------------------------------
std::ostream& CQuadrique::AfficherInfoDebug(std::ostream& out) const
{
    return out << "[DEBUG]: Quadric.Quadratique       = " << m_Quadratique << "\n[DEBUG]: Quadric.Lineaire          = " << m_Lineaire
               << "\n[DEBUG]: Quadric.Mixte             = " << m_Mixte << "\n[DEBUG]: Quadric.Constante         = " << m_Cst;
}

// Quadrique.cpp:110-146: Q' = T^-1 * Q * (T^-1)^T with Q the symmetric 4x4 of the 10 coefficients
void CQuadrique::Pretraitement()
{
    const REAL h = RENDRE_REEL(0.5);
    const REAL A = m_Quadratique.x, B = m_Quadratique.y, C = m_Quadratique.z;
    const REAL D = m_Mixte.z * h, E = m_Mixte.x * h, F = m_Mixte.y * h;
    const REAL G = m_Lineaire.x * h, H = m_Lineaire.y * h, J = m_Lineaire.z * h, K = m_Cst;
    CMatrice4 Q(A, D, F, G,
                D, B, E, H,
                F, E, C, J,
                G, H, J, K);
    const CMatrice4 inv = m_Transformation.Inverse();
    Q = inv * Q * inv.Transpose();
    const REAL two = RENDRE_REEL(2.0);
    m_Quadratique = CVecteur3(Q.m[0][0], Q.m[1][1], Q.m[2][2]);
    m_Cst = Q.m[3][3];
    m_Mixte = CVecteur3(Q.m[1][2] * two, Q.m[0][2] * two, Q.m[0][1] * two);
    m_Lineaire = CVecteur3(Q.m[0][3] * two, Q.m[1][3] * two, Q.m[2][3] * two);
}

CIntersection CQuadrique::Intersection(const CRayon& rayon)       // Quadrique.cpp:160-249
{
    CIntersection hit;
    const CVecteur3 d = rayon.ObtenirDirection(), o = rayon.ObtenirOrigine();
    const CVecteur3 &q = m_Quadratique, &mx = m_Mixte, &l = m_Lineaire;
    const REAL h = RENDRE_REEL(0.5);
    const REAL a = d.x * (q.x * d.x + mx.z * d.y + mx.y * d.z) + d.y * (q.y * d.y + mx.x * d.z) + d.z * (q.z * d.z);
    const REAL b = d.x * (q.x * o.x + h * (mx.z * o.y + mx.y * o.z + l.x)) +
                   d.y * (q.y * o.y + h * (mx.z * o.x + mx.x * o.z + l.y)) +
                   d.z * (q.z * o.z + h * (mx.y * o.x + mx.x * o.y + l.z));
    const REAL c = o.x * (q.x * o.x + mx.z * o.y + mx.y * o.z + l.x) + o.y * (q.y * o.y + mx.x * o.z + l.y) +
                   o.z * (q.z * o.z + l.z) + m_Cst;
    if (a != 0.0) {
        const REAL ka = -b / a, kb = c / a;
        REAL delta = ka * ka - kb;
        if (delta > 0) {
            delta = std::sqrt(delta);
            const REAL t0 = ka - delta, t1 = ka + delta;
            REAL t = Min<REAL>(t0, t1);
            if (t < EPSILON) t = Max<REAL>(t0, t1);
            if (!(t < 0)) {
                hit.AjusterDistance(t);
                hit.AjusterSurface(this);
                const CVecteur3 p = o + t * d;
                CVecteur3 n;
                n.x = RENDRE_REEL(2.0) * q.x * p.x + mx.y * p.z + mx.z * p.y + l.x;
                n.y = RENDRE_REEL(2.0) * q.y * p.y + mx.x * p.z + mx.z * p.x + l.y;
                n.z = RENDRE_REEL(2.0) * q.z * p.z + mx.x * p.y + mx.y * p.x + l.z;
                hit.AjusterNormale(CVecteur3::Normaliser(n));
            }
        }
    } else {
        hit.AjusterSurface(this);
        hit.AjusterDistance(-h * (c / b));
        hit.AjusterNormale(CVecteur3::Normaliser(l));
    }
    return hit;
}
