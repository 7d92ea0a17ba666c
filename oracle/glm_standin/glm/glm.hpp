// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Stand-in for glm 0.9.9.8, the one third-party dependency of the reference's
// search/physics sources that is not vendored under /root/reference
// (reference ext/glm.cmake:2-6 fetches it from the network).  Only the surface the
// reference's lib/billiard_physics and lib/billiard_search actually use is provided.
//
// Arithmetic follows glm 0.9.9.8's scalar (non-SIMD) code path, which is what the
// reference gets (it never defines GLM_FORCE_INTRINSICS):
//   dot(vec2 a, b)   = (a.x*b.x) + (a.y*b.y)          (detail::compute_dot<vec<2>>)
//   length(v)        = sqrt(dot(v, v))
//   normalize(v)     = v * (1 / sqrt(dot(v, v)))       (inversesqrt is 1/sqrt, NOT rsqrt)
//   reflect(I, N)    = I - N * dot(N, I) * 2
//   mat*vec, vec*mat = left-to-right sums of products
// These semantics are pinned end-to-end by the thesis cost logs (tests/test_oracle_ref.py).
#pragma once

#include <cassert>
#include <cmath>
#include <cstdint>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <type_traits>
#include <unordered_set>

namespace glm {

struct vec3;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    template <class A, class B>
    vec2(A a, B b) : x(static_cast<float>(a)), y(static_cast<float>(b)) {}
    vec2(const vec3& v);  // glm: implicit unless GLM_FORCE_EXPLICIT_CTOR
    float& operator[](int i) { return i ? y : x; }
    const float& operator[](int i) const { return i ? y : x; }
};

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    template <class A, class B, class C>
    vec3(A a, B b, C c) : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)) {}
    template <class C>
    vec3(const vec2& v, C c) : x(v.x), y(v.y), z(static_cast<float>(c)) {}
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const float& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

inline vec2::vec2(const vec3& v) : x(v.x), y(v.y) {}

struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    template <class A, class B, class C, class D>
    vec4(A a, B b, C c, D d)
        : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)), w(static_cast<float>(d)) {}
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
    const float& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
};

// ---- vec2 operators -------------------------------------------------------------------------
inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2& a, const vec2& b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator-(const vec2& a) { return vec2(-a.x, -a.y); }
inline vec2 operator*(const vec2& a, const vec2& b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(const vec2& a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, const vec2& a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator/(const vec2& a, float s) { return vec2(a.x / s, a.y / s); }
inline vec2& operator+=(vec2& a, const vec2& b) { a.x += b.x; a.y += b.y; return a; }
inline vec2& operator-=(vec2& a, const vec2& b) { a.x -= b.x; a.y -= b.y; return a; }
inline vec2& operator*=(vec2& a, float s) { a.x *= s; a.y *= s; return a; }
inline bool operator==(const vec2& a, const vec2& b) { return a.x == b.x && a.y == b.y; }
inline bool operator!=(const vec2& a, const vec2& b) { return !(a == b); }

// Free functions are templates so that the reference's own billiard::physics::normalize(vec2)
// (billiard_physics.cpp:637) stays the better overload inside its namespace, as with real glm.
template <class V, class = typename std::enable_if<std::is_same<V, vec2>::value>::type>
inline float dot(const V& a, const V& b) {
    vec2 t = a * b;
    return t.x + t.y;
}
template <class V, class = typename std::enable_if<std::is_same<V, vec2>::value>::type>
inline float length(const V& v) {
    return std::sqrt(dot(v, v));
}
template <class V, class = typename std::enable_if<std::is_same<V, vec2>::value>::type>
inline vec2 normalize(const V& v) {
    return v * (1.0f / std::sqrt(dot(v, v)));
}
template <class V, class = typename std::enable_if<std::is_same<V, vec2>::value>::type>
inline vec2 reflect(const V& I, const V& N) {
    return I - N * dot(N, I) * 2.0f;
}
template <class V, class = typename std::enable_if<std::is_same<V, vec2>::value>::type>
inline vec2 abs(const V& v) {
    return vec2(std::fabs(v.x), std::fabs(v.y));
}

// ---- matrices (column-major, as glm) --------------------------------------------------------
struct mat3 {
    vec3 c[3];
    mat3() { c[0] = vec3(1, 0, 0); c[1] = vec3(0, 1, 0); c[2] = vec3(0, 0, 1); }
    template <class A, class B, class C, class D, class E, class F, class G, class H, class I>
    mat3(A x0, B y0, C z0, D x1, E y1, F z1, G x2, H y2, I z2) {
        c[0] = vec3(x0, y0, z0);
        c[1] = vec3(x1, y1, z1);
        c[2] = vec3(x2, y2, z2);
    }
    vec3& operator[](int i) { return c[i]; }
    const vec3& operator[](int i) const { return c[i]; }
};
using mat3x3 = mat3;

inline vec3 operator*(const mat3& m, const vec3& v) {
    return vec3(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z,
                m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z,
                m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z);
}

struct mat4x4 {
    vec4 c[4];
    mat4x4() {}
    mat4x4(const vec4& a, const vec4& b, const vec4& cc, const vec4& d) { c[0] = a; c[1] = b; c[2] = cc; c[3] = d; }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
using mat4 = mat4x4;

struct mat2x4 {  // 2 columns of 4 rows
    vec4 c[2];
    mat2x4() {}
    mat2x4(const vec4& a, const vec4& b) { c[0] = a; c[1] = b; }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};

// row-vector * matrix
inline vec4 operator*(const vec4& v, const mat4x4& m) {
    return vec4(m[0][0] * v[0] + m[0][1] * v[1] + m[0][2] * v[2] + m[0][3] * v[3],
                m[1][0] * v[0] + m[1][1] * v[1] + m[1][2] * v[2] + m[1][3] * v[3],
                m[2][0] * v[0] + m[2][1] * v[1] + m[2][2] * v[2] + m[2][3] * v[3],
                m[3][0] * v[0] + m[3][1] * v[1] + m[3][2] * v[2] + m[3][3] * v[3]);
}
inline vec2 operator*(const vec4& v, const mat2x4& m) {
    return vec2(v.x * m[0][0] + v.y * m[0][1] + v.z * m[0][2] + v.w * m[0][3],
                v.x * m[1][0] + v.y * m[1][1] + v.z * m[1][2] + v.w * m[1][3]);
}
// matrix * column-vector
inline vec4 operator*(const mat4x4& m, const vec4& v) {
    return vec4(m[0][0] * v[0] + m[1][0] * v[1] + m[2][0] * v[2] + m[3][0] * v[3],
                m[0][1] * v[0] + m[1][1] * v[1] + m[2][1] * v[2] + m[3][1] * v[3],
                m[0][2] * v[0] + m[1][2] * v[1] + m[2][2] * v[2] + m[3][2] * v[3],
                m[0][3] * v[0] + m[1][3] * v[1] + m[2][3] * v[2] + m[3][3] * v[3]);
}
inline mat2x4 operator*(const mat4x4& m, const mat2x4& p) { return mat2x4(m * p[0], m * p[1]); }

}  // namespace glm
