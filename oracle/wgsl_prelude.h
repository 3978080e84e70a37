// wgsl_prelude.h — the C++ meaning of the WGSL subset oracle/transpile_wgsl.py emits (TEST INFRASTRUCTURE).
//
// The transpiler rewrites the reference's shader text token by token; this header supplies what the tokens refer to:
// vecN<T> with WGSL's component-wise operators, the built-in functions the shaders call, WGSL's value-conversion rules,
// storage arrays with robust-buffer-access behaviour, and rg32uint storage textures. Arithmetic is IEEE binary32,
// round-to-nearest-even, compiled with -ffp-contract=off (WGSL permits contraction and 2.5-ulp division; SURVEY.md
// §8(c) pins the uncontracted, correctly rounded reading — the same pin the hand-written restatement uses, so a
// bit-for-bit match between the two checks the RESTATEMENT, not the pin).
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>

namespace wgsl {

// ---- value conversions (WGSL spec "Value Constructor Built-in Functions": f32 -> u32/i32 truncates toward zero and
// saturates, NaN -> 0; integer <-> integer reinterprets the bits; integer -> f32 rounds to nearest) ----
template <class To, class From> struct Conv { static To f(From v) { return (To)v; } };
template <> struct Conv<uint32_t, float> {
    static uint32_t f(float v) {
        if (!(v == v)) return 0u;
        if (v <= 0.0f) return 0u;
        if (v >= 4294967296.0f) return 0xffffffffu;
        return (uint32_t)v;
    }
};
template <> struct Conv<int32_t, float> {
    static int32_t f(float v) {
        if (!(v == v)) return 0;
        if (v <= -2147483648.0f) return INT32_MIN;
        if (v >= 2147483648.0f) return INT32_MAX;
        return (int32_t)v;
    }
};
template <class To, class From> inline To conv(From v) { return Conv<To, From>::f(v); }
template <class F> inline uint32_t to_u32(F v) { return conv<uint32_t, F>(v); }
template <class F> inline int32_t to_i32(F v) { return conv<int32_t, F>(v); }
template <class F> inline float to_f32(F v) { return conv<float, F>(v); }

// ---- vectors ----
template <class T> struct vec2 {
    T x, y;
    vec2() : x(), y() {}
    vec2(T a, T b) : x(a), y(b) {}
    explicit vec2(T a) : x(a), y(a) {}
    template <class U> explicit vec2(const vec2<U>& o) : x(conv<T, U>(o.x)), y(conv<T, U>(o.y)) {}
    vec2& operator+=(const vec2& o) { x = x + o.x; y = y + o.y; return *this; }
    vec2& operator-=(const vec2& o) { x = x - o.x; y = y - o.y; return *this; }
    vec2& operator*=(T s) { x = x * s; y = y * s; return *this; }
};
template <class T> vec2(T, T) -> vec2<T>;
template <class T> inline vec2<T> operator+(vec2<T> a, vec2<T> b) { return {a.x + b.x, a.y + b.y}; }
template <class T> inline vec2<T> operator-(vec2<T> a, vec2<T> b) { return {a.x - b.x, a.y - b.y}; }
template <class T> inline vec2<T> operator*(vec2<T> a, vec2<T> b) { return {a.x * b.x, a.y * b.y}; }
template <class T> inline vec2<T> operator/(vec2<T> a, vec2<T> b) { return {a.x / b.x, a.y / b.y}; }
template <class T> inline vec2<T> operator+(vec2<T> a, T s) { return {a.x + s, a.y + s}; }
template <class T> inline vec2<T> operator-(vec2<T> a, T s) { return {a.x - s, a.y - s}; }
template <class T> inline vec2<T> operator*(vec2<T> a, T s) { return {a.x * s, a.y * s}; }
template <class T> inline vec2<T> operator/(vec2<T> a, T s) { return {a.x / s, a.y / s}; }
template <class T> inline vec2<T> operator*(T s, vec2<T> a) { return {s * a.x, s * a.y}; }
template <class T> inline vec2<T> operator-(vec2<T> a) { return {-a.x, -a.y}; }
struct bvec2 { bool x, y; };
template <class T> inline bvec2 operator==(vec2<T> a, vec2<T> b) { return {a.x == b.x, a.y == b.y}; }
inline bool all(bvec2 b) { return b.x && b.y; }

template <class T> struct vec3 {
    T x, y, z;
    vec2<T> xy() const { return {x, y}; }
};
template <class T> struct vec4 {
    T x, y, z, w;
    vec4() : x(), y(), z(), w() {}
    vec4(T a, T b, T c, T d) : x(a), y(b), z(c), w(d) {}
    vec4(vec2<T> a, int c, int d) : x(a.x), y(a.y), z((T)c), w((T)d) {}
    vec2<T> xy() const { return {x, y}; }
};
template <class T> vec4(vec2<T>, int, int) -> vec4<T>;

// ---- built-in functions used by the shaders ----
inline float dot(vec2<float> a, vec2<float> b) { return a.x * b.x + a.y * b.y; }
inline float length(vec2<float> a) { return sqrtf(a.x * a.x + a.y * a.y); }
inline vec2<float> normalize(vec2<float> a) { return a / length(a); }
inline float sqrt(float a) { return sqrtf(a); }
inline float pow(float a, float b) { return powf(a, b); }
inline float abs(float a) { return fabsf(a); }
inline float sign(float a) { return a > 0.0f ? 1.0f : (a < 0.0f ? -1.0f : 0.0f); }
inline float floor(float a) { return floorf(a); }
inline vec2<float> floor(vec2<float> a) { return {floorf(a.x), floorf(a.y)}; }
inline float max(float a, float b) { return a < b ? b : a; }  // WGSL: b if a < b else a
template <class T> inline T select(T f, T t, bool c) { return c ? t : f; }

// ---- storage: runtime-sized arrays with robust buffer access. WebGPU lets an out-of-bounds access be clamped to an
// in-bounds element; SURVEY.md §8(c) pins exactly that for the force-field texel (compute.wgsl:129 at uv == 1.0 on the
// last row) and "cells outside the grid are empty" for start_indices — which clamping also gives: the entry it returns
// points at a particle whose key is not the out-of-range cell id, so the walk stops at once. No shader of the path
// STORES out of bounds (keys < cells, ids < n, sort.wgsl:39-41 checks index_high). ----
template <class T> struct Storage {
    T* data = nullptr;
    uint32_t len = 0;
    T& operator[](uint32_t i) { return data[i < len ? i : len - 1u]; }
};
// rg32uint storage textures (jump_flood.wgsl): textureLoad returns (r, g, 0, 1); textureStore writes r, g
struct Texture2D {
    uint32_t* data = nullptr;  // w*h pairs
    uint32_t w = 0, h = 0;
};
template <class I> inline vec4<uint32_t> textureLoad(const Texture2D& t, vec2<I> p) {
    const uint32_t x = (uint32_t)p.x, y = (uint32_t)p.y;
    if (x >= t.w || y >= t.h) return {0u, 0u, 0u, 1u};
    const uint32_t* q = t.data + 2 * ((size_t)y * t.w + x);
    return {q[0], q[1], 0u, 1u};
}
template <class I> inline void textureStore(Texture2D& t, vec2<I> p, vec4<uint32_t> v) {
    const uint32_t x = (uint32_t)p.x, y = (uint32_t)p.y;
    if (x >= t.w || y >= t.h) return;
    uint32_t* q = t.data + 2 * ((size_t)y * t.w + x);
    q[0] = v.x; q[1] = v.y;
}

}  // namespace wgsl
