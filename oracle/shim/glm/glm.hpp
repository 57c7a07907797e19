// TEST INFRASTRUCTURE — minimal stand-in for the GLM headers, which are not on this image.
//
// The reference (/root/reference) depends on GLM, a header-only third-party library that is neither
// vendored in the reference nor installed here.  This file supplies only the handful
// of types and functions that the reference's CPU path (src/update-cpu.cpp, include/surface_impl.h,
// include/light_impl.h, src/surface.cpp, src/light.cpp, src/scene-exception.cpp) touches, written
// from scratch, with the SAME FLOATING-POINT OPERATION ORDER as GLM 0.9.9's scalar code path:
//   dot(a,b)        = (a.x*b.x + a.y*b.y) + a.z*b.z
//   normalize(v)    = v * (1 / sqrt(dot(v,v)))          (multiply by reciprocal, not divide)
//   mat4 * vec4     = (m[0]*v.x + m[1]*v.y) + (m[2]*v.z + m[3]*v.w)
//   min(x,y)        = (y < x) ? y : x ;   max(x,y) = (x < y) ? y : x   (component-wise)
//   vec / scalar    = true division per component
//   converting constructors between element types and from vec4 to vec3 are implicit
// Only oracle/ (the parity checker) includes this header.  The product never does.
#ifndef RT_ORACLE_SHIM_GLM_HPP
#define RT_ORACLE_SHIM_GLM_HPP

#include <cmath>
#include <cstddef>

namespace glm {

enum qualifier { packed_highp, defaultp = packed_highp };
typedef int length_t;

template <length_t L, typename T, qualifier Q = defaultp>
struct vec;

template <typename T, qualifier Q>
struct vec<3, T, Q> {
    union { T x, r, s; };
    union { T y, g, t; };
    union { T z, b, p; };

    vec() = default;
    vec(T v) : x(v), y(v), z(v) {}
    vec(T a, T b_, T c) : x(a), y(b_), z(c) {}
    template <typename U, qualifier P>
    vec(const vec<3, U, P> &o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)), z(static_cast<T>(o.z)) {}
    template <typename U, qualifier P>
    vec(const vec<4, U, P> &o);

    T &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const T &operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }

    template <typename U>
    vec &operator+=(const vec<3, U, Q> &o)
    {
        x += static_cast<T>(o.x);
        y += static_cast<T>(o.y);
        z += static_cast<T>(o.z);
        return *this;
    }
    template <typename U>
    vec &operator-=(const vec<3, U, Q> &o)
    {
        x -= static_cast<T>(o.x);
        y -= static_cast<T>(o.y);
        z -= static_cast<T>(o.z);
        return *this;
    }
    vec &operator*=(T s)
    {
        x *= s;
        y *= s;
        z *= s;
        return *this;
    }
};

template <typename T, qualifier Q>
struct vec<4, T, Q> {
    union { T x, r; };
    union { T y, g; };
    union { T z, b; };
    union { T w, a; };

    vec() = default;
    vec(T v) : x(v), y(v), z(v), w(v) {}
    vec(T a_, T b_, T c, T d) : x(a_), y(b_), z(c), w(d) {}
    vec(const vec<3, T, Q> &v, T w_) : x(v.x), y(v.y), z(v.z), w(w_) {}

    T &operator[](int i) { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
    const T &operator[](int i) const { return i == 0 ? x : (i == 1 ? y : (i == 2 ? z : w)); }
};

template <typename T, qualifier Q>
template <typename U, qualifier P>
vec<3, T, Q>::vec(const vec<4, U, P> &o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)), z(static_cast<T>(o.z))
{}

typedef vec<3, float> vec3;
typedef vec<3, double> dvec3;
typedef vec<3, int> ivec3;
typedef vec<4, float> vec4;
typedef vec<4, double> dvec4;

// ---- vec3 arithmetic (component-wise) -------------------------------------------------------
template <typename T, qualifier Q>
vec<3, T, Q> operator+(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator-(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator*(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator/(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(a.x / b.x, a.y / b.y, a.z / b.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator-(const vec<3, T, Q> &a) { return vec<3, T, Q>(-a.x, -a.y, -a.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator*(T s, const vec<3, T, Q> &v) { return vec<3, T, Q>(s * v.x, s * v.y, s * v.z); }
template <typename T, qualifier Q>
vec<3, T, Q> operator*(const vec<3, T, Q> &v, T s) { return vec<3, T, Q>(v.x * s, v.y * s, v.z * s); }
template <typename T, qualifier Q>
vec<3, T, Q> operator/(const vec<3, T, Q> &v, T s) { return vec<3, T, Q>(v.x / s, v.y / s, v.z / s); }
template <typename T, qualifier Q>
vec<3, T, Q> operator+(const vec<3, T, Q> &v, T s) { return vec<3, T, Q>(v.x + s, v.y + s, v.z + s); }
template <typename T, qualifier Q>
vec<3, T, Q> operator-(const vec<3, T, Q> &v, T s) { return vec<3, T, Q>(v.x - s, v.y - s, v.z - s); }

// ---- vec4 arithmetic ------------------------------------------------------------------------
template <typename T, qualifier Q>
vec<4, T, Q> operator+(const vec<4, T, Q> &a, const vec<4, T, Q> &b) { return vec<4, T, Q>(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
template <typename T, qualifier Q>
vec<4, T, Q> operator-(const vec<4, T, Q> &a, const vec<4, T, Q> &b) { return vec<4, T, Q>(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
template <typename T, qualifier Q>
vec<4, T, Q> operator*(const vec<4, T, Q> &a, const vec<4, T, Q> &b) { return vec<4, T, Q>(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
template <typename T, qualifier Q>
vec<4, T, Q> operator*(const vec<4, T, Q> &v, T s) { return vec<4, T, Q>(v.x * s, v.y * s, v.z * s, v.w * s); }
template <typename T, qualifier Q>
vec<4, T, Q> operator*(T s, const vec<4, T, Q> &v) { return vec<4, T, Q>(s * v.x, s * v.y, s * v.z, s * v.w); }

// ---- geometric / common functions -----------------------------------------------------------
template <typename T, qualifier Q>
T dot(const vec<3, T, Q> &a, const vec<3, T, Q> &b)
{
    vec<3, T, Q> m(a * b);
    return m.x + m.y + m.z;
}
template <typename T, qualifier Q>
T dot(const vec<4, T, Q> &a, const vec<4, T, Q> &b)
{
    vec<4, T, Q> m(a * b);
    return (m.x + m.y) + (m.z + m.w);
}
template <typename T>
T inversesqrt(T v) { return static_cast<T>(1) / std::sqrt(v); }
template <typename T, qualifier Q>
vec<3, T, Q> normalize(const vec<3, T, Q> &v) { return v * inversesqrt(dot(v, v)); }
template <typename T, qualifier Q>
T length(const vec<3, T, Q> &v) { return std::sqrt(dot(v, v)); }
template <typename T, qualifier Q>
vec<3, T, Q> cross(const vec<3, T, Q> &a, const vec<3, T, Q> &b)
{
    return vec<3, T, Q>(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}

template <typename T>
T min(T a, T b) { return (b < a) ? b : a; }
template <typename T>
T max(T a, T b) { return (a < b) ? b : a; }
template <typename T, qualifier Q>
vec<3, T, Q> min(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
template <typename T, qualifier Q>
vec<3, T, Q> max(const vec<3, T, Q> &a, const vec<3, T, Q> &b) { return vec<3, T, Q>(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }

// GLM pulls std::pow into its namespace for scalars, so glm::pow(double, int) is std::pow.
using std::pow;

template <typename T>
T radians(T degrees) { return degrees * static_cast<T>(0.01745329251994329576923690768489); }

// ---- 4x4 double matrix, column-major: m[col][row] -------------------------------------------
template <typename T>
struct mat4x4_t {
    typedef vec<4, T> col_type;
    col_type cols[4];

    mat4x4_t() = default;
    explicit mat4x4_t(T d)
    {
        cols[0] = col_type(d, 0, 0, 0);
        cols[1] = col_type(0, d, 0, 0);
        cols[2] = col_type(0, 0, d, 0);
        cols[3] = col_type(0, 0, 0, d);
    }
    col_type &operator[](int i) { return cols[i]; }
    const col_type &operator[](int i) const { return cols[i]; }
};
typedef mat4x4_t<double> dmat4;
typedef mat4x4_t<float> mat4;

template <typename T>
vec<4, T> operator*(const mat4x4_t<T> &m, const vec<4, T> &v)
{
    const vec<4, T> mul0 = m[0] * vec<4, T>(v.x);
    const vec<4, T> mul1 = m[1] * vec<4, T>(v.y);
    const vec<4, T> add0 = mul0 + mul1;
    const vec<4, T> mul2 = m[2] * vec<4, T>(v.z);
    const vec<4, T> mul3 = m[3] * vec<4, T>(v.w);
    const vec<4, T> add1 = mul2 + mul3;
    return add0 + add1;
}

inline ivec3 iround(const vec3 &v) { return ivec3((int) (v.x + 0.5f), (int) (v.y + 0.5f), (int) (v.z + 0.5f)); }

} // namespace glm

#endif
