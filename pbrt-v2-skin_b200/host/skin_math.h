// skin_math.h -- the small slice of pbrt's geometry/transform arithmetic the host side needs.
//
// Every function states the reference routine whose fp32 operation order it keeps
// (core/geometry.h, core/transform.{h,cpp}); world-space vertices, BVH bounds, surface points
// and camera rays must come out bit-identical to the reference's so that the device path
// can be compared with the oracle on the same inputs.  Compile with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>

namespace skin {

struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return (&x)[i]; }
    float &operator[](int i) { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline float Dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// core/geometry.h:529-536 (double products)
inline V3 Cross(V3 a, V3 b) {
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return V3((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
inline float LengthSquared(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
inline float Length(V3 a) { return sqrtf(LengthSquared(a)); }
inline V3 Div(V3 a, float f) { float inv = 1.f / f; return V3(a.x * inv, a.y * inv, a.z * inv); }   // geometry.h:102-106
inline V3 Normalize(V3 a) { return Div(a, Length(a)); }
inline void CoordinateSystem(V3 v1, V3 *v2, V3 *v3) {   // geometry.h:564-574
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        *v2 = V3(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        *v2 = V3(0.f, v1.z * invLen, -v1.y * invLen);
    }
    *v3 = Cross(v1, *v2);
}

struct BBox {
    V3 pMin, pMax;
    BBox() : pMin(INFINITY, INFINITY, INFINITY), pMax(-INFINITY, -INFINITY, -INFINITY) {}
    BBox(V3 p) : pMin(p), pMax(p) {}
    BBox(V3 a, V3 b) : pMin(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)),
                       pMax(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)) {}
    float SurfaceArea() const { V3 d = pMax - pMin; return 2.f * (d.x * d.y + d.x * d.z + d.y * d.z); }
    int MaximumExtent() const {
        V3 d = pMax - pMin;
        if (d.x > d.y && d.x > d.z) return 0;
        else if (d.y > d.z) return 1;
        else return 2;
    }
};
inline BBox Union(const BBox &b, V3 p) {
    BBox r;
    r.pMin = V3(std::min(b.pMin.x, p.x), std::min(b.pMin.y, p.y), std::min(b.pMin.z, p.z));
    r.pMax = V3(std::max(b.pMax.x, p.x), std::max(b.pMax.y, p.y), std::max(b.pMax.z, p.z));
    return r;
}
inline BBox Union(const BBox &a, const BBox &b) {
    BBox r;
    r.pMin = V3(std::min(a.pMin.x, b.pMin.x), std::min(a.pMin.y, b.pMin.y), std::min(a.pMin.z, b.pMin.z));
    r.pMax = V3(std::max(a.pMax.x, b.pMax.x), std::max(a.pMax.y, b.pMax.y), std::max(a.pMax.z, b.pMax.z));
    return r;
}

struct Mat4 {
    float m[4][4];
    Mat4() { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[i][j] = i == j ? 1.f : 0.f; }
    Mat4(float a00, float a01, float a02, float a03, float a10, float a11, float a12, float a13,
         float a20, float a21, float a22, float a23, float a30, float a31, float a32, float a33) {
        m[0][0] = a00; m[0][1] = a01; m[0][2] = a02; m[0][3] = a03; m[1][0] = a10; m[1][1] = a11; m[1][2] = a12; m[1][3] = a13;
        m[2][0] = a20; m[2][1] = a21; m[2][2] = a22; m[2][3] = a23; m[3][0] = a30; m[3][1] = a31; m[3][2] = a32; m[3][3] = a33;
    }
};
inline Mat4 Mul(const Mat4 &a, const Mat4 &b) {   // transform.h:83-92
    Mat4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j] + a.m[i][3] * b.m[3][j];
    return r;
}
inline Mat4 Transpose(const Mat4 &a) { Mat4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i]; return r; }
Mat4 Inverse(const Mat4 &m);   // Gauss-Jordan with full pivoting, transform.cpp:75-133

struct Transform {
    Mat4 m, mInv;
    Transform() {}
    Transform(const Mat4 &a) : m(a), mInv(Inverse(a)) {}
    Transform(const Mat4 &a, const Mat4 &ai) : m(a), mInv(ai) {}
    Transform operator*(const Transform &t) const { return Transform(Mul(m, t.m), Mul(t.mInv, mInv)); }   // transform.cpp:282-286
    V3 Point(V3 p) const {   // transform.h:192-201
        float xp = m.m[0][0] * p.x + m.m[0][1] * p.y + m.m[0][2] * p.z + m.m[0][3];
        float yp = m.m[1][0] * p.x + m.m[1][1] * p.y + m.m[1][2] * p.z + m.m[1][3];
        float zp = m.m[2][0] * p.x + m.m[2][1] * p.y + m.m[2][2] * p.z + m.m[2][3];
        float wp = m.m[3][0] * p.x + m.m[3][1] * p.y + m.m[3][2] * p.z + m.m[3][3];
        if (wp == 1.) return V3(xp, yp, zp);
        float inv = 1.f / wp;
        return V3(inv * xp, inv * yp, inv * zp);
    }
    V3 Vector(V3 v) const {
        return V3(m.m[0][0] * v.x + m.m[0][1] * v.y + m.m[0][2] * v.z, m.m[1][0] * v.x + m.m[1][1] * v.y + m.m[1][2] * v.z,
                  m.m[2][0] * v.x + m.m[2][1] * v.y + m.m[2][2] * v.z);
    }
    V3 Normal(V3 n) const {   // transform.h:232-237
        return V3(mInv.m[0][0] * n.x + mInv.m[1][0] * n.y + mInv.m[2][0] * n.z, mInv.m[0][1] * n.x + mInv.m[1][1] * n.y + mInv.m[2][1] * n.z,
                  mInv.m[0][2] * n.x + mInv.m[1][2] * n.y + mInv.m[2][2] * n.z);
    }
    BBox Box(const BBox &b) const {   // transform.cpp:268-279
        BBox r(Point(V3(b.pMin.x, b.pMin.y, b.pMin.z)));
        r = Union(r, Point(V3(b.pMax.x, b.pMin.y, b.pMin.z)));
        r = Union(r, Point(V3(b.pMin.x, b.pMax.y, b.pMin.z)));
        r = Union(r, Point(V3(b.pMin.x, b.pMin.y, b.pMax.z)));
        r = Union(r, Point(V3(b.pMin.x, b.pMax.y, b.pMax.z)));
        r = Union(r, Point(V3(b.pMax.x, b.pMax.y, b.pMin.z)));
        r = Union(r, Point(V3(b.pMax.x, b.pMin.y, b.pMax.z)));
        r = Union(r, Point(V3(b.pMax.x, b.pMax.y, b.pMax.z)));
        return r;
    }
    bool SwapsHandedness() const {   // transform.cpp:289-300
        float det = ((m.m[0][0] * (m.m[1][1] * m.m[2][2] - m.m[1][2] * m.m[2][1])) -
                     (m.m[0][1] * (m.m[1][0] * m.m[2][2] - m.m[1][2] * m.m[2][0])) +
                     (m.m[0][2] * (m.m[1][0] * m.m[2][1] - m.m[1][1] * m.m[2][0])));
        return det < 0.f;
    }
};
inline Transform InverseT(const Transform &t) { return Transform(t.mInv, t.m); }
inline float Radians(float deg) { return ((float)M_PI / 180.f) * deg; }
Transform Translate(V3 d);
Transform Scale(float x, float y, float z);
Transform Rotate(float angle, V3 axis);
Transform LookAt(V3 pos, V3 look, V3 up);
Transform Perspective(float fov, float n, float f);

}  // namespace skin
