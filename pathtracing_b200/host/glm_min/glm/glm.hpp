// Minimal stand-in for the subset of glm the drop-in host API (PathTracing.hpp) touches.
// A program that already has glm 1.0 on its include path (the reference vendors it) uses the real
// library: put that directory BEFORE this one.  Column-major mat4 like glm.
#pragma once
#include <cmath>
#include <cstddef>

namespace glm {

template <typename T> struct tvec2 {
    T x{}, y{};
    constexpr tvec2() = default;
    constexpr explicit tvec2(T s) : x(s), y(s){}
    template <typename A, typename B> constexpr tvec2(A a, B b) : x(T(a)), y(T(b)){}
    template <typename U> constexpr tvec2(const tvec2<U>& o) : x(T(o.x)), y(T(o.y)){}
    T& operator[](int i){ return i == 0 ? x : y; }
    const T& operator[](int i) const{ return i == 0 ? x : y; }
};
template <typename T> struct tvec3 {
    T x{}, y{}, z{};
    constexpr tvec3() = default;
    constexpr explicit tvec3(T s) : x(s), y(s), z(s){}
    template <typename A, typename B, typename C> constexpr tvec3(A a, B b, C c) : x(T(a)), y(T(b)), z(T(c)){}
    template <typename U> constexpr tvec3(const tvec3<U>& o) : x(T(o.x)), y(T(o.y)), z(T(o.z)){}
    T& operator[](int i){ return i == 0 ? x : (i == 1 ? y : z); }
    const T& operator[](int i) const{ return i == 0 ? x : (i == 1 ? y : z); }
};
template <typename T> struct tvec4 {
    T x{}, y{}, z{}, w{};
    constexpr tvec4() = default;
    constexpr explicit tvec4(T s) : x(s), y(s), z(s), w(s){}
    template <typename A, typename B, typename C, typename D> constexpr tvec4(A a, B b, C c, D d) : x(T(a)), y(T(b)), z(T(c)), w(T(d)){}
    template <typename U> constexpr tvec4(const tvec3<U>& v, T ww) : x(T(v.x)), y(T(v.y)), z(T(v.z)), w(ww){}
    T& operator[](int i){ return (&x)[i]; }
    const T& operator[](int i) const{ return (&x)[i]; }
};
using vec2 = tvec2<float>; using dvec2 = tvec2<double>; using ivec2 = tvec2<int>; using uvec2 = tvec2<unsigned>;
using vec3 = tvec3<float>; using dvec3 = tvec3<double>; using ivec3 = tvec3<int>;
using vec4 = tvec4<float>;

template <typename T> constexpr tvec2<T> operator+(const tvec2<T>& a, const tvec2<T>& b){ return { a.x + b.x, a.y + b.y }; }
template <typename T> constexpr tvec2<T> operator-(const tvec2<T>& a, const tvec2<T>& b){ return { a.x - b.x, a.y - b.y }; }
template <typename T> constexpr tvec2<T> operator*(const tvec2<T>& a, const tvec2<T>& b){ return { a.x * b.x, a.y * b.y }; }
template <typename T> constexpr tvec2<T> operator*(const tvec2<T>& a, T s){ return { a.x * s, a.y * s }; }
template <typename T> constexpr tvec2<T> operator*(T s, const tvec2<T>& a){ return { a.x * s, a.y * s }; }
template <typename T> constexpr tvec2<T> operator/(const tvec2<T>& a, T s){ return { a.x / s, a.y / s }; }
template <typename T> constexpr bool operator==(const tvec2<T>& a, const tvec2<T>& b){ return a.x == b.x && a.y == b.y; }
template <typename T> constexpr tvec3<T> operator+(const tvec3<T>& a, const tvec3<T>& b){ return { a.x + b.x, a.y + b.y, a.z + b.z }; }
template <typename T> constexpr tvec3<T> operator-(const tvec3<T>& a, const tvec3<T>& b){ return { a.x - b.x, a.y - b.y, a.z - b.z }; }
template <typename T> constexpr tvec3<T> operator-(const tvec3<T>& a){ return { -a.x, -a.y, -a.z }; }
template <typename T> constexpr tvec3<T> operator*(const tvec3<T>& a, const tvec3<T>& b){ return { a.x * b.x, a.y * b.y, a.z * b.z }; }
template <typename T> constexpr tvec3<T> operator/(const tvec3<T>& a, const tvec3<T>& b){ return { a.x / b.x, a.y / b.y, a.z / b.z }; }
template <typename T> constexpr tvec3<T> operator*(const tvec3<T>& a, T s){ return { a.x * s, a.y * s, a.z * s }; }
template <typename T> constexpr tvec3<T> operator*(T s, const tvec3<T>& a){ return { a.x * s, a.y * s, a.z * s }; }
template <typename T> constexpr tvec3<T> operator/(const tvec3<T>& a, T s){ return { a.x / s, a.y / s, a.z / s }; }
template <typename T> constexpr tvec3<T>& operator+=(tvec3<T>& a, const tvec3<T>& b){ a.x += b.x; a.y += b.y; a.z += b.z; return a; }
template <typename T> constexpr tvec3<T>& operator*=(tvec3<T>& a, T s){ a.x *= s; a.y *= s; a.z *= s; return a; }
template <typename T> constexpr bool operator==(const tvec3<T>& a, const tvec3<T>& b){ return a.x == b.x && a.y == b.y && a.z == b.z; }

template <typename T> constexpr T dot(const tvec3<T>& a, const tvec3<T>& b){ return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> constexpr tvec3<T> cross(const tvec3<T>& a, const tvec3<T>& b){ return { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y }; }
template <typename T> inline T length(const tvec3<T>& a){ return std::sqrt(dot(a, a)); }
template <typename T> inline tvec3<T> normalize(const tvec3<T>& a){ return a * (T(1) / length(a)); }
template <typename T> constexpr T radians(T deg){ return deg * T(0.01745329251994329576923690768489); }
template <typename T> constexpr T clamp(T v, T lo, T hi){ return v < lo ? lo : (v > hi ? hi : v); }
template <typename T> constexpr T pi(){ return T(3.14159265358979323846264338327950288); }

// column-major 4x4: m[c][r]
struct mat4 {
    vec4 c[4];
    constexpr mat4() = default;
    constexpr explicit mat4(float d){ c[0].x = d; c[1].y = d; c[2].z = d; c[3].w = d; }
    vec4& operator[](int i){ return c[i]; }
    const vec4& operator[](int i) const{ return c[i]; }
};
inline mat4 operator*(const mat4& a, const mat4& b){
    mat4 r;
    for(int j = 0; j < 4; j++)for(int i = 0; i < 4; i++){
        float s = 0;
        for(int k = 0; k < 4; k++)s += a[k][i] * b[j][k];
        r[j][i] = s;
    }
    return r;
}
inline vec4 operator*(const mat4& m, const vec4& v){
    vec4 r;
    for(int i = 0; i < 4; i++)r[i] = m[0][i] * v.x + m[1][i] * v.y + m[2][i] * v.z + m[3][i] * v.w;
    return r;
}
inline const float* value_ptr(const mat4& m){ return &m.c[0].x; }

} // namespace glm
