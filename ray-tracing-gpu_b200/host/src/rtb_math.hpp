// rtb_math.hpp -- host-side math value types of the scene API (namespace Math3D), written
// fresh with the public names, signatures and -- where a result feeds the renderer -- the
// binary32 operation order of the reference (MathUtils.h, Vecteur3.h, Matrice4.h/.cpp), so
// that camera bases and pre-processed surface coefficients come out bit-identical.
// All matrix work is pre-render host work; nothing here runs per pixel.
#pragma once
#include <cassert>
#include <cmath>
#include <cstring>
#include <iomanip>
#include <iostream>

namespace Math3D {

typedef float REAL;                                   // MathUtils.h:23
#define RENDRE_REEL(A) static_cast<Math3D::REAL>(A)
#define EPSILON RENDRE_REEL(1.0e-2)                   // MathUtils.h:35
#ifdef M_PI
#define PI M_PI
#else
#define PI 3.14159265359
#endif

template <typename T> inline const T Max(const T& a, const T& b) { return a > b ? a : b; }
template <typename T> inline const T Min(const T& a, const T& b) { return a < b ? a : b; }
template <typename T> inline const T Abs(const T& a) { return a > 0 ? a : -a; }
template <typename T> inline const T Borner(const T& v, const T& lo, const T& hi) { return v < lo ? lo : (v > hi ? hi : v); }
template <typename T> inline const T Deg2Rad(const T& a) { return (a / T(180.0)) * T(PI); }
template <typename T> inline const T Rad2Deg(const T& a) { return (a / T(PI)) * T(180.0); }

class CVecteur3 {
public:
    static const CVecteur3 ZERO, UNIT_X, UNIT_Y, UNIT_Z;
    REAL x, y, z;

    CVecteur3() : x(0), y(0), z(0) {}
    CVecteur3(const REAL X, const REAL Y, const REAL Z) : x(X), y(Y), z(Z) {}
    explicit CVecteur3(const REAL* const c) : x(c[0]), y(c[1]), z(c[2]) {}
    explicit CVecteur3(const REAL s) : x(s), y(s), z(s) {}
    CVecteur3(const CVecteur3& v) = default;
    CVecteur3& operator=(const CVecteur3& v) = default;

    REAL& operator[](const int i) { assert(i > -1 && i < 3); return *(&x + i); }
    bool operator==(const CVecteur3& v) const
    {
        return Abs<REAL>(x - v.x) < EPSILON && Abs<REAL>(y - v.y) < EPSILON && Abs<REAL>(z - v.z) < EPSILON;
    }
    bool operator!=(const CVecteur3& v) const { return !(*this == v); }
    CVecteur3& operator+=(const CVecteur3& v) { x += v.x; y += v.y; z += v.z; return *this; }
    CVecteur3& operator-=(const CVecteur3& v) { x -= v.x; y -= v.y; z -= v.z; return *this; }
    CVecteur3& operator*=(const REAL s) { x *= s; y *= s; z *= s; return *this; }
    CVecteur3& operator/=(const REAL s) { assert(std::fabs(s) > EPSILON); const REAL inv = RENDRE_REEL(1) / s; return (*this) *= inv; }
    const CVecteur3 operator-() const { return CVecteur3(-x, -y, -z); }
    const CVecteur3 operator/(const REAL s) const
    {
        assert(std::fabs(s) > EPSILON);
        const REAL inv = RENDRE_REEL(1.0) / s;          // reciprocal first, then three products
        return CVecteur3(x * inv, y * inv, z * inv);
    }
    friend const CVecteur3 operator+(const CVecteur3& a, const CVecteur3& b) { return CVecteur3(a.x + b.x, a.y + b.y, a.z + b.z); }
    friend const CVecteur3 operator-(const CVecteur3& a, const CVecteur3& b) { return CVecteur3(a.x - b.x, a.y - b.y, a.z - b.z); }
    friend const CVecteur3 operator*(const REAL s, const CVecteur3& v) { return CVecteur3(v.x * s, v.y * s, v.z * s); }
    friend const CVecteur3 operator*(const CVecteur3& v, const REAL s) { return CVecteur3(v.x * s, v.y * s, v.z * s); }
    friend std::ostream& operator<<(std::ostream& o, const CVecteur3& v)
    {
        return o << std::setprecision(3) << "[ " << std::setw(3) << v.x << ", " << v.y << ", " << v.z << " ]" << std::setw(0);
    }

    static const REAL ProdScal(const CVecteur3& a, const CVecteur3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
    static const CVecteur3 ProdVect(const CVecteur3& a, const CVecteur3& b)
    {
        return CVecteur3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
    }
    static const REAL Norme(const CVecteur3& v) { return std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z); }
    static const REAL Distance(const CVecteur3& a, const CVecteur3& b) { return Norme(a - b); }
    // zero vector when the length does not exceed EPSILON (Vecteur3.h:522-534)
    static const CVecteur3 Normaliser(const CVecteur3& v)
    {
        const REAL len = Norme(v);
        if (!(len > EPSILON)) return CVecteur3();
        return v * (RENDRE_REEL(1.0) / len);
    }
    static const CVecteur3 Reflect(const CVecteur3& v, const CVecteur3& n) { return v - (RENDRE_REEL(2.0) * ProdScal(v, n) * n); }
    // total internal reflection surfaces as a NaN that fails the `< 0` test (Vecteur3.h:566-581)
    static const CVecteur3 Refract(const CVecteur3& v, const CVecteur3& n, const REAL ratio)
    {
        const CVecteur3 z = ratio * (v - ProdScal(v, n) * n);
        const double nz = static_cast<double>(Norme(z));
        const CVecteur3 t = z - static_cast<REAL>(std::sqrt(1.0 - nz * nz)) * n;
        return ProdScal(t, n) < 0 ? t : Reflect(v, n);
    }
};

// Row-major 4x4, ROW-vector convention v' = v * M with the translation in row 3
// (Matrice4.h:297-306); transforms compose by POST-multiplication unless PRE is asked.
class CMatrice4 {
public:
    enum OrdreTransformation { PRE, POST };
    static const CMatrice4 ZERO, IDENTITE;
    REAL m[4][4];

    CMatrice4() { std::memset(m, 0, sizeof m); }
    CMatrice4(REAL a00, REAL a01, REAL a02, REAL a03, REAL a10, REAL a11, REAL a12, REAL a13,
              REAL a20, REAL a21, REAL a22, REAL a23, REAL a30, REAL a31, REAL a32, REAL a33)
    {
        const REAL v[16] = { a00, a01, a02, a03, a10, a11, a12, a13, a20, a21, a22, a23, a30, a31, a32, a33 };
        std::memcpy(m, v, sizeof m);
    }
    CMatrice4(const CMatrice4& o) = default;
    CMatrice4& operator=(const CMatrice4& o) = default;

    const REAL* operator[](const int row) { assert(row > -1 && row < 4); return m[row]; }
    const CMatrice4 operator-() const { CMatrice4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = -m[i][j]; return r; }
    CMatrice4& operator+=(const CMatrice4& o) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i][j] += o.m[i][j]; return *this; }
    CMatrice4& operator-=(const CMatrice4& o) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i][j] -= o.m[i][j]; return *this; }
    CMatrice4& operator*=(const CMatrice4& o) { *this = Concatene(o); return *this; }
    friend const CMatrice4 operator*(const CMatrice4& a, const CMatrice4& b) { return a.Concatene(b); }
    friend const CVecteur3 operator*(const CVecteur3& v, const CMatrice4& M)
    {
        CVecteur3 r;
        r.x = M.m[0][0] * v.x + M.m[1][0] * v.y + M.m[2][0] * v.z + M.m[3][0];
        r.y = M.m[0][1] * v.x + M.m[1][1] * v.y + M.m[2][1] * v.z + M.m[3][1];
        r.z = M.m[0][2] * v.x + M.m[1][2] * v.y + M.m[2][2] * v.z + M.m[3][2];
        return r;
    }
    friend std::ostream& operator<<(std::ostream& o, const CMatrice4& M)
    {
        o << std::setprecision(3) << std::setw(3);
        for (int i = 0; i < 4; ++i)
            o << "| " << M.m[i][0] << ", " << M.m[i][1] << ", " << M.m[i][2] << ", " << M.m[i][3] << " |" << (i < 3 ? "\n" : "");
        return o;
    }

    void Ajuster(const CMatrice4& o) { *this = o; }
    void RotationAutourDesX(const REAL rad, const OrdreTransformation ordre = POST)
    {
        CMatrice4 r(IDENTITE);
        r.m[1][1] = std::cos(rad); r.m[1][2] = std::sin(rad); r.m[2][2] = r.m[1][1]; r.m[2][1] = -r.m[1][2];
        Appliquer(r, ordre);
    }
    void RotationAutourDesY(const REAL rad, const OrdreTransformation ordre = POST)
    {
        CMatrice4 r(IDENTITE);
        r.m[0][0] = std::cos(rad); r.m[0][2] = -std::sin(rad); r.m[2][2] = r.m[0][0]; r.m[2][0] = -r.m[0][2];
        Appliquer(r, ordre);
    }
    void RotationAutourDesZ(const REAL rad, const OrdreTransformation ordre = POST)
    {
        CMatrice4 r(IDENTITE);
        r.m[0][0] = std::cos(rad); r.m[0][1] = std::sin(rad); r.m[1][1] = r.m[0][0]; r.m[1][0] = -r.m[0][1];
        Appliquer(r, ordre);
    }
    void Translation(const REAL x, const REAL y, const REAL z, const OrdreTransformation ordre = POST)
    {
        CMatrice4 t(IDENTITE);
        t.m[3][0] = x; t.m[3][1] = y; t.m[3][2] = z;
        Appliquer(t, ordre);
    }
    void MiseAEchelle(const REAL x, const REAL y, const REAL z, const OrdreTransformation ordre = POST)
    {
        CMatrice4 s(IDENTITE);
        s.m[0][0] = x; s.m[1][1] = y; s.m[2][2] = z;
        Appliquer(s, ordre);
    }
    const CMatrice4 Inverse() const;
    const CMatrice4 Transpose() const;
    // (this) * o, each entry accumulated left to right over k = 0..3
    const CMatrice4 Concatene(const CMatrice4& o) const
    {
        CMatrice4 r;
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j)
                r.m[i][j] = m[i][0] * o.m[0][j] + m[i][1] * o.m[1][j] + m[i][2] * o.m[2][j] + m[i][3] * o.m[3][j];
        return r;
    }

private:
    void Appliquer(const CMatrice4& t, const OrdreTransformation ordre) { *this = (ordre == PRE) ? t.Concatene(*this) : Concatene(t); }
};

} // namespace Math3D
