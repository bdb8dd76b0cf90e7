// srt_math.hpp — the handful of vector / matrix operations the scene surface needs, with the
// operation ORDER of the GLM 1.0.2 scalar code paths the reference is built on (vendored under the
// reference's library/glm-master/glm; cited per function), so that vertices, boxes and light
// positions come out bit-identical to the reference's.  fp32 throughout; build with
// -ffp-contract=off.  This is not a general math library.
#ifndef SRT_MATH_HPP_
#define SRT_MATH_HPP_

#include <cmath>

namespace srt {

struct vec2 { float x = 0.f, y = 0.f; vec2() = default; vec2(float a, float b) : x(a), y(b) {} };
struct ivec2 { int x = 0, y = 0; ivec2() = default; ivec2(int a, int b) : x(a), y(b) {} };
struct vec3 {
    float x = 0.f, y = 0.f, z = 0.f;
    vec3() = default;
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct vec4 {
    float x = 0.f, y = 0.f, z = 0.f, w = 0.f;
    vec4() = default;
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

// Column-major 4x4 like glm::mat4: m[col][row].
struct mat4 {
    vec4 c[4];
    mat4() = default;
    explicit mat4(float d) { c[0].x = d; c[1].y = d; c[2].z = d; c[3].w = d; }   // glm::mat4(s): s on the diagonal
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};

inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(const vec4& a, const vec4& b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }

// mat4 * vec4 — glm/detail/type_mat4x4.inl:562-573: (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
inline vec4 operator*(const mat4& m, const vec4& v) {
    const vec4 Mul0 = m[0] * v.x;
    const vec4 Mul1 = m[1] * v.y;
    const vec4 Add0 = Mul0 + Mul1;
    const vec4 Mul2 = m[2] * v.z;
    const vec4 Mul3 = m[3] * v.w;
    const vec4 Add1 = Mul2 + Mul3;
    return Add0 + Add1;
}

// mat4 * mat4 — glm/detail/type_mat4x4.inl:660-704 (non-SIMD mul4x4):
// column j = ((A0*b.x + A1*b.y) + A2*b.z) + A3*b.w, accumulated left to right.
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; ++j) {
        vec4 tmp = a[0] * b[j].x;
        tmp = tmp + a[1] * b[j].y;
        tmp = tmp + a[2] * b[j].z;
        tmp = tmp + a[3] * b[j].w;
        r[j] = tmp;
    }
    return r;
}

// glm::inverse(mat4) — glm/detail/func_matrix.inl:388-444 (cofactor expansion, one 1/det).
inline mat4 inverse(const mat4& m) {
    const float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    const float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    const float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    const float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    const float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    const float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    const float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    const float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    const float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    const float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    const float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    const float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    const float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

    const vec4 Fac0(Coef00, Coef00, Coef02, Coef03);
    const vec4 Fac1(Coef04, Coef04, Coef06, Coef07);
    const vec4 Fac2(Coef08, Coef08, Coef10, Coef11);
    const vec4 Fac3(Coef12, Coef12, Coef14, Coef15);
    const vec4 Fac4(Coef16, Coef16, Coef18, Coef19);
    const vec4 Fac5(Coef20, Coef20, Coef22, Coef23);

    const vec4 Vec0(m[1][0], m[0][0], m[0][0], m[0][0]);
    const vec4 Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
    const vec4 Vec2(m[1][2], m[0][2], m[0][2], m[0][2]);
    const vec4 Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);

    const vec4 Inv0((Vec1 * Fac0 - Vec2 * Fac1) + Vec3 * Fac2);
    const vec4 Inv1((Vec0 * Fac0 - Vec2 * Fac3) + Vec3 * Fac4);
    const vec4 Inv2((Vec0 * Fac1 - Vec1 * Fac3) + Vec3 * Fac5);
    const vec4 Inv3((Vec0 * Fac2 - Vec1 * Fac4) + Vec2 * Fac5);

    const vec4 SignA(+1.f, -1.f, +1.f, -1.f);
    const vec4 SignB(-1.f, +1.f, -1.f, +1.f);
    mat4 Inverse;
    Inverse[0] = Inv0 * SignA; Inverse[1] = Inv1 * SignB; Inverse[2] = Inv2 * SignA; Inverse[3] = Inv3 * SignB;

    const vec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
    const vec4 Dot0(m[0] * Row0);
    const float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    const float OneOverDeterminant = 1.0f / Dot1;
    mat4 r;
    for (int j = 0; j < 4; ++j) r[j] = Inverse[j] * OneOverDeterminant;
    return r;
}

// glm::radians(float) = degrees * 0.01745329251994329576923690768489   (glm/detail/func_trigonometric.inl)
inline float radians(float degrees) { return degrees * 0.01745329251994329576923690768489f; }

// glm::min / glm::max: (y < x) ? y : x  and  (x < y) ? y : x   (glm/detail/func_common.inl:22-35)
inline float glm_min(float x, float y) { return (y < x) ? y : x; }
inline float glm_max(float x, float y) { return (x < y) ? y : x; }

}  // namespace srt
#endif
