// rtb_colour.hpp -- CCouleur: float RGB whose int constructor scales by 1/255 and whose
// 3-float constructor, Ajuster() and binary +/- clamp to [0,1], while *, *= and += do NOT
// (Couleur.h:98-122, 245-417) -- shading therefore accumulates unclamped.
#pragma once
#include "rtb_math.hpp"

class CCouleur {
    static const float UBYTE_2_FLOAT;
public:
    static const CCouleur NOIR, BLANC;
    float r, g, b;

    CCouleur() : r(0.0f), g(0.0f), b(0.0f) {}
    CCouleur(const float R, const float G, const float B) { Ajuster(R, G, B); }
    CCouleur(const int R, const int G, const int B) : r(R * UBYTE_2_FLOAT), g(G * UBYTE_2_FLOAT), b(B * UBYTE_2_FLOAT) {}
    explicit CCouleur(const float v) { Ajuster(v, v, v); }
    CCouleur(const CCouleur&) = default;
    CCouleur& operator=(const CCouleur&) = default;

    float& operator[](const int i) { assert(i > -1 && i < 3); return *(&r + i); }
    bool operator==(const CCouleur& c) const { return r == c.r && g == c.g && b == c.b; }
    bool operator!=(const CCouleur& c) const { return !(*this == c); }
    CCouleur& operator+=(const CCouleur& c) { r += c.r; g += c.g; b += c.b; return *this; }
    CCouleur& operator-=(const CCouleur& c) { r -= c.r; g -= c.g; b -= c.b; return *this; }
    CCouleur& operator*=(const CCouleur& c) { r *= c.r; g *= c.g; b *= c.b; return *this; }
    CCouleur& operator*=(const Math3D::REAL v) { r *= v; g *= v; b *= v; return *this; }
    friend const CCouleur operator+(const CCouleur& a, const CCouleur& c) { return CCouleur(a.r + c.r, a.g + c.g, a.b + c.b); }
    friend const CCouleur operator-(const CCouleur& a, const CCouleur& c) { return CCouleur(a.r - c.r, a.g - c.g, a.b - c.b); }
    friend const CCouleur operator*(const CCouleur& a, const CCouleur& c) { CCouleur o; o.r = a.r * c.r; o.g = a.g * c.g; o.b = a.b * c.b; return o; }
    friend const CCouleur operator*(const CCouleur& a, const Math3D::REAL v) { CCouleur o; o.r = a.r * v; o.g = a.g * v; o.b = a.b * v; return o; }
    friend const CCouleur operator*(const Math3D::REAL v, const CCouleur& a) { return a * v; }
    friend std::ostream& operator<<(std::ostream& o, const CCouleur& c)
    {
        return o << std::setprecision(3) << "[ " << std::setw(3) << c.r << ", " << c.g << ", " << c.b << " ]";
    }
    void Ajuster(const float R, const float G, const float B)
    {
        r = Math3D::Borner<float>(R, 0.0f, 1.0f); g = Math3D::Borner<float>(G, 0.0f, 1.0f); b = Math3D::Borner<float>(B, 0.0f, 1.0f);
    }
};
