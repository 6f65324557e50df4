/* wgsl_builtins.h — TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * The small run-time a C++ translation of a WGSL compute shader needs: scalar aliases,
 * vec2/vec3/vec4 with the WGSL operators, the builtin functions the reference shader calls
 * (/root/reference/shaders/ray_tracer.wgsl: dot cross normalize length reflect mix smoothstep
 * sign pow log cos sqrt abs min max, value conversions, arrayLength, textureLoad/Store),
 * storage-buffer views that decode records at their WGSL byte offsets, and call hooks.
 *
 * oracle/make_ref.py rewrites the reference's shader TEXT (read from /root/reference at build
 * time, never copied into the repo) token by token into C++ against this header; the result
 * is oracle/_ref/librt_wgsl_ref*.so.  Nothing here knows the shader: loop bounds, RNG order,
 * branch structure, tie rule and record layouts all come from the reference's own text.
 *
 * Two arithmetic modes, chosen at compile time:
 *
 *   default ("plain")  every operator and builtin is its WGSL-spec formula evaluated in
 *       binary32 with one rounding per operation, nothing fused (-ffp-contract=off), and
 *       log / cos / pow are glibc's logf / cosf / powf.  normalize(v) = v / length(v),
 *       mix(a,b,t) = a*(1-t) + b*t, reflect(i,n) = i - 2*dot(n,i)*n, dot = x*x + y*y + z*z.
 *
 *   -DWGSL_CONTRACT    the contraction a compiler is allowed to apply (WGSL leaves fusing and
 *       the accuracy of builtins to the implementation), fixed to the ONE uniform rule set
 *       that oracle/rt_oracle.cpp and the CUDA kernels implement ("Arithmetic
 *       contract"):
 *         - dot(a,b) = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x)); cross components fma(p,q, -(r*s));
 *         - a vec3 scaled by a scalar and then added is fused into the add:
 *           v + w*s = fma(w,s,v) per component, (v*s) + (w*t) = fma(w,t, v*s);
 *         - reflect(i,n) = fma(n, -(2*dot(n,i)), i); mix(a,b,t) = fma(b,t, a*(1-t));
 *         - normalize(v) = v * (1/sqrt(dot(v,v)));
 *         - pow(x,2) = x*x, pow(x,5) = ((x*x)*(x*x))*x, any other pow / log / cos are the
 *           fixed polynomial routines of rto_math.h;
 *         - everything else (all scalar arithmetic, vec2 / vec4 arithmetic, sqrt, division)
 *           is a single correctly rounded binary32 operation in source order.
 *
 * u32(f32) / i32(f32) truncate and saturate (WGSL spec); u32 arithmetic wraps.
 */
#ifndef WGSL_BUILTINS_H
#define WGSL_BUILTINS_H
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#ifdef WGSL_CONTRACT
#include "rto_math.h"
#endif

namespace wgsl {

typedef float f32;
typedef uint32_t u32;
typedef int32_t i32;

/* ---- value conversions (WGSL 16.1) --------------------------------------------------
 * u32(f32) of a value >= 2^32: saturates (the WGSL rule; default) or, with f2u_wrap set (the "wide" seed mode of
 * the CUDA path, a labelled deviation), wraps modulo 2^32 as an out-of-range SPIR-V OpConvertFToU may. */
static int f2u_wrap = 0;
inline u32 to_u32(f32 f)
{
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) {
        if (f2u_wrap) return (f >= 18446744073709551616.0f) ? 0u : (u32)(uint64_t)f;
        return 0xFFFFFFFFu;
    }
    return (u32)f;
}
inline u32 to_u32(i32 i) { return (u32)i; }
inline u32 to_u32(u32 u) { return u; }
inline u32 to_u32(bool b) { return b ? 1u : 0u; }
inline i32 to_i32(f32 f) { if (f != f) return 0; if (f >= 2147483648.0f) return 2147483647; if (f <= -2147483648.0f) return (i32)0x80000000u; return (i32)f; }
inline i32 to_i32(u32 u) { return (i32)u; }
inline i32 to_i32(i32 i) { return i; }
inline i32 to_i32(bool b) { return b ? 1 : 0; }
inline f32 to_f32(u32 u) { return (f32)u; }            /* round to nearest even */
inline f32 to_f32(i32 i) { return (f32)i; }
inline f32 to_f32(f32 f) { return f; }
inline f32 to_f32(bool b) { return b ? 1.0f : 0.0f; }

/* ---- vectors ------------------------------------------------------------------------ */
template <class T> struct vec2 {
    T x{}, y{};
    vec2() = default;
    vec2(T x_, T y_) : x(x_), y(y_) {}
    explicit vec2(T s) : x(s), y(s) {}
    friend vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
    friend vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
    friend vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
    friend vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }
    friend vec2 operator+(vec2 a, T s) { return vec2(a.x + s, a.y + s); }
    friend vec2 operator-(vec2 a, T s) { return vec2(a.x - s, a.y - s); }
    friend vec2 operator*(vec2 a, T s) { return vec2(a.x * s, a.y * s); }
    friend vec2 operator/(vec2 a, T s) { return vec2(a.x / s, a.y / s); }
    friend vec2 operator+(T s, vec2 a) { return vec2(s + a.x, s + a.y); }
    friend vec2 operator-(T s, vec2 a) { return vec2(s - a.x, s - a.y); }
    friend vec2 operator*(T s, vec2 a) { return vec2(s * a.x, s * a.y); }
    friend vec2 operator/(T s, vec2 a) { return vec2(s / a.x, s / a.y); }
    friend vec2 operator-(vec2 a) { return vec2(-a.x, -a.y); }
    vec2& operator+=(vec2 b) { *this = *this + b; return *this; }
    vec2& operator-=(vec2 b) { *this = *this - b; return *this; }
    vec2& operator*=(vec2 b) { *this = *this * b; return *this; }
    vec2& operator*=(T s) { *this = *this * s; return *this; }
};

template <class T> struct scaled3;

template <class T> struct vec3 {
    T x{}, y{}, z{};
    vec3() = default;
    vec3(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
    explicit vec3(T s) : x(s), y(s), z(s) {}
    friend vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
    friend vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
    friend vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
    friend vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
    friend vec3 operator+(vec3 a, T s) { return vec3(a.x + s, a.y + s, a.z + s); }
    friend vec3 operator-(vec3 a, T s) { return vec3(a.x - s, a.y - s, a.z - s); }
    friend vec3 operator/(vec3 a, T s) { return vec3(a.x / s, a.y / s, a.z / s); }
    friend vec3 operator+(T s, vec3 a) { return vec3(s + a.x, s + a.y, s + a.z); }
    friend vec3 operator-(T s, vec3 a) { return vec3(s - a.x, s - a.y, s - a.z); }
    friend vec3 operator/(T s, vec3 a) { return vec3(s / a.x, s / a.y, s / a.z); }
    friend vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
    vec3& operator+=(vec3 b) { *this = *this + b; return *this; }
    vec3& operator-=(vec3 b) { *this = *this - b; return *this; }
    vec3& operator*=(vec3 b) { *this = *this * b; return *this; }
    vec3& operator*=(T s) { x = x * s; y = y * s; z = z * s; return *this; }
};

#ifdef WGSL_CONTRACT
/* vec3 * scalar: the rounded product (base class: usable as a plain vec3, `.x` etc.) plus its two
 * factors, so that a following vec3 addition can fuse it (see the header comment). */
template <class T> struct scaled3 : vec3<T> {
    vec3<T> v; T s;
    scaled3(vec3<T> v_, T s_) : vec3<T>(v_.x * s_, v_.y * s_, v_.z * s_), v(v_), s(s_) {}
};
template <class T> inline scaled3<T> operator*(vec3<T> a, T s) { return scaled3<T>(a, s); }
template <class T> inline scaled3<T> operator*(T s, vec3<T> a) { return scaled3<T>(a, s); }
inline scaled3<f32> operator*(vec3<f32> a, f32 s) { return scaled3<f32>(a, s); }
inline scaled3<f32> operator*(f32 s, vec3<f32> a) { return scaled3<f32>(a, s); }
inline scaled3<f32> operator*(const scaled3<f32>& a, f32 s) { return scaled3<f32>((const vec3<f32>&)a, s); }
inline scaled3<f32> operator*(f32 s, const scaled3<f32>& a) { return scaled3<f32>((const vec3<f32>&)a, s); }
inline vec3<f32> fma3_(vec3<f32> a, f32 s, vec3<f32> c) { return vec3<f32>(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y), fmaf(a.z, s, c.z)); }
inline vec3<f32> operator+(const scaled3<f32>& l, const scaled3<f32>& r) { return fma3_(r.v, r.s, (const vec3<f32>&)l); }
inline vec3<f32> operator+(vec3<f32> a, const scaled3<f32>& p) { return fma3_(p.v, p.s, a); }
inline vec3<f32> operator+(const scaled3<f32>& p, vec3<f32> a) { return fma3_(p.v, p.s, a); }
#else
template <class T> inline vec3<T> operator*(vec3<T> a, T s) { return vec3<T>(a.x * s, a.y * s, a.z * s); }
template <class T> inline vec3<T> operator*(T s, vec3<T> a) { return vec3<T>(s * a.x, s * a.y, s * a.z); }
inline vec3<f32> operator*(vec3<f32> a, f32 s) { return vec3<f32>(a.x * s, a.y * s, a.z * s); }
inline vec3<f32> operator*(f32 s, vec3<f32> a) { return vec3<f32>(s * a.x, s * a.y, s * a.z); }
#endif

template <class T> struct vec4 {
    T x{}, y{}, z{}, w{};
    vec4() = default;
    vec4(T x_, T y_, T z_, T w_) : x(x_), y(y_), z(z_), w(w_) {}
    explicit vec4(T s) : x(s), y(s), z(s), w(s) {}
    friend vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
    friend vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
    friend vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
    friend vec4 operator/(vec4 a, vec4 b) { return vec4(a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w); }
    friend vec4 operator+(vec4 a, T s) { return vec4(a.x + s, a.y + s, a.z + s, a.w + s); }
    friend vec4 operator-(vec4 a, T s) { return vec4(a.x - s, a.y - s, a.z - s, a.w - s); }
    friend vec4 operator*(vec4 a, T s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
    friend vec4 operator/(vec4 a, T s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
    friend vec4 operator+(T s, vec4 a) { return vec4(s + a.x, s + a.y, s + a.z, s + a.w); }
    friend vec4 operator-(T s, vec4 a) { return vec4(s - a.x, s - a.y, s - a.z, s - a.w); }
    friend vec4 operator*(T s, vec4 a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
    friend vec4 operator/(T s, vec4 a) { return vec4(s / a.x, s / a.y, s / a.z, s / a.w); }
    friend vec4 operator-(vec4 a) { return vec4(-a.x, -a.y, -a.z, -a.w); }
    vec4& operator+=(vec4 b) { *this = *this + b; return *this; }
    vec4& operator-=(vec4 b) { *this = *this - b; return *this; }
    vec4& operator*=(vec4 b) { *this = *this * b; return *this; }
    vec4& operator*=(T s) { *this = *this * s; return *this; }
};
template <class T> vec2(T, T) -> vec2<T>;
template <class T> vec3(T, T, T) -> vec3<T>;
template <class T> vec4(T, T, T, T) -> vec4<T>;

/* ---- numeric builtins (WGSL 17.5); a shader's own function of the same name wins
 * (make_ref.py defines WGSL_USER_<name> for every function the shader declares) ------------ */
#ifndef WGSL_USER_sqrt
inline f32 sqrt(f32 x) { return ::sqrtf(x); }
#endif
#ifndef WGSL_USER_abs
inline f32 abs(f32 x) { return ::fabsf(x); }
inline i32 abs(i32 x) { return x < 0 ? (i32)(0u - (u32)x) : x; }
#endif
#ifndef WGSL_USER_min
inline f32 min(f32 a, f32 b) { return ::fminf(a, b); }
inline i32 min(i32 a, i32 b) { return b < a ? b : a; }
inline u32 min(u32 a, u32 b) { return b < a ? b : a; }
#endif
#ifndef WGSL_USER_max
inline f32 max(f32 a, f32 b) { return ::fmaxf(a, b); }
inline i32 max(i32 a, i32 b) { return a < b ? b : a; }
inline u32 max(u32 a, u32 b) { return a < b ? b : a; }
#endif
#ifndef WGSL_USER_sign
inline f32 sign(f32 x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }
#endif
#ifndef WGSL_USER_log
#ifdef WGSL_CONTRACT
inline f32 log(f32 x) { return rto_logf(x); }
#else
inline f32 log(f32 x) { return ::logf(x); }
#endif
#endif
#ifndef WGSL_USER_cos
#ifdef WGSL_CONTRACT
inline f32 cos(f32 x) { return rto_cosf(x); }
#else
inline f32 cos(f32 x) { return ::cosf(x); }
#endif
#endif
#ifndef WGSL_USER_pow
#ifdef WGSL_CONTRACT
inline f32 pow(f32 x, f32 y)
{
    if (y == 2.0f) return x * x;
    if (y == 5.0f) { f32 x2 = x * x; return (x2 * x2) * x; }
    return rto_powf(x, y);
}
#else
inline f32 pow(f32 x, f32 y) { return ::powf(x, y); }
#endif
#endif
#ifndef WGSL_USER_dot
#ifdef WGSL_CONTRACT
inline f32 dot(vec3<f32> a, vec3<f32> b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
#else
inline f32 dot(vec3<f32> a, vec3<f32> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
#endif
inline f32 dot(vec2<f32> a, vec2<f32> b) { return a.x * b.x + a.y * b.y; }
inline f32 dot(vec4<f32> a, vec4<f32> b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
#endif
#ifndef WGSL_USER_cross
#ifdef WGSL_CONTRACT
inline vec3<f32> cross(vec3<f32> a, vec3<f32> b)
{
    return vec3<f32>(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
#else
inline vec3<f32> cross(vec3<f32> a, vec3<f32> b)
{
    return vec3<f32>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
#endif
#endif
#ifndef WGSL_USER_length
inline f32 length(vec3<f32> a) { return ::sqrtf(dot(a, a)); }
inline f32 length(vec2<f32> a) { return ::sqrtf(dot(a, a)); }
inline f32 length(vec4<f32> a) { return ::sqrtf(dot(a, a)); }
#endif
#ifndef WGSL_USER_normalize
#ifdef WGSL_CONTRACT
inline vec3<f32> normalize(vec3<f32> a) { f32 inv = 1.0f / ::sqrtf(dot(a, a)); return vec3<f32>(a.x * inv, a.y * inv, a.z * inv); }
#else
inline vec3<f32> normalize(vec3<f32> a) { f32 l = length(a); return vec3<f32>(a.x / l, a.y / l, a.z / l); }
#endif
#endif
#ifndef WGSL_USER_reflect
#ifdef WGSL_CONTRACT
inline vec3<f32> reflect(vec3<f32> i, vec3<f32> n) { f32 t = 2.0f * dot(n, i); return fma3_(n, -t, i); }
#else
inline vec3<f32> reflect(vec3<f32> i, vec3<f32> n) { f32 t = 2.0f * dot(n, i); return vec3<f32>(i.x - t * n.x, i.y - t * n.y, i.z - t * n.z); }
#endif
#endif
#ifndef WGSL_USER_mix
#ifdef WGSL_CONTRACT
inline f32 mix(f32 a, f32 b, f32 t) { return fmaf(b, t, a * (1.0f - t)); }
#else
inline f32 mix(f32 a, f32 b, f32 t) { return a * (1.0f - t) + b * t; }
#endif
inline vec2<f32> mix(vec2<f32> a, vec2<f32> b, f32 t) { return vec2<f32>(mix(a.x, b.x, t), mix(a.y, b.y, t)); }
inline vec3<f32> mix(vec3<f32> a, vec3<f32> b, f32 t) { return vec3<f32>(mix(a.x, b.x, t), mix(a.y, b.y, t), mix(a.z, b.z, t)); }
inline vec4<f32> mix(vec4<f32> a, vec4<f32> b, f32 t) { return vec4<f32>(mix(a.x, b.x, t), mix(a.y, b.y, t), mix(a.z, b.z, t), mix(a.w, b.w, t)); }
#endif
#ifndef WGSL_USER_smoothstep
inline f32 smoothstep(f32 e0, f32 e1, f32 x)
{
    f32 t = ::fminf(::fmaxf((x - e0) / (e1 - e0), 0.0f), 1.0f);     /* clamp((x - low) / (high - low), 0, 1) */
    return (t * t) * (3.0f - 2.0f * t);
}
#endif

/* ---- resources ---------------------------------------------------------------------- */
/* Host-shareable layout (WGSL 13.4): make_ref.py emits layout<S> for every struct the shader
 * declares, with the member offsets computed from the WGSL alignment rules, so a storage or
 * uniform buffer is decoded from the caller's raw bytes exactly as the shader would see it. */
template <class T> struct layout;
template <> struct layout<f32> { enum { size = 4, align = 4 }; static f32 load(const uint8_t* p) { f32 v; memcpy(&v, p, 4); return v; } };
template <> struct layout<u32> { enum { size = 4, align = 4 }; static u32 load(const uint8_t* p) { u32 v; memcpy(&v, p, 4); return v; } };
template <> struct layout<i32> { enum { size = 4, align = 4 }; static i32 load(const uint8_t* p) { i32 v; memcpy(&v, p, 4); return v; } };
template <class T> struct layout<vec2<T>> { enum { size = 8, align = 8 };
    static vec2<T> load(const uint8_t* p) { return vec2<T>(layout<T>::load(p), layout<T>::load(p + 4)); } };
template <class T> struct layout<vec3<T>> { enum { size = 12, align = 16 };
    static vec3<T> load(const uint8_t* p) { return vec3<T>(layout<T>::load(p), layout<T>::load(p + 4), layout<T>::load(p + 8)); } };
template <class T> struct layout<vec4<T>> { enum { size = 16, align = 16 };
    static vec4<T> load(const uint8_t* p) { return vec4<T>(layout<T>::load(p), layout<T>::load(p + 4), layout<T>::load(p + 8), layout<T>::load(p + 12)); } };

/* var<storage, read> name: array<T> — a runtime-sized array over the caller's bytes */
template <class T> struct storage_array {
    const uint8_t* base = nullptr; u32 count = 0;
    T operator[](u32 i) const { return layout<T>::load(base + (size_t)i * (size_t)layout<T>::size); }
};
template <class T> inline u32 arrayLength(const storage_array<T>* a) { return a->count; }

/* texture_storage_2d<rgba32float, read_write>: out-of-range stores are dropped, loads give 0 */
struct texture_storage_2d_rgba32float { f32* texels = nullptr; u32 width = 0, height = 0; };
inline vec4<f32> textureLoad(const texture_storage_2d_rgba32float& t, vec2<i32> c)
{
    if (c.x < 0 || c.y < 0 || (u32)c.x >= t.width || (u32)c.y >= t.height) return vec4<f32>();
    const f32* p = t.texels + ((size_t)c.y * t.width + (size_t)c.x) * 4;
    return vec4<f32>(p[0], p[1], p[2], p[3]);
}
inline void textureStore(const texture_storage_2d_rgba32float& t, vec2<i32> c, vec4<f32> v)
{
    if (c.x < 0 || c.y < 0 || (u32)c.x >= t.width || (u32)c.y >= t.height) return;
    f32* p = t.texels + ((size_t)c.y * t.width + (size_t)c.x) * 4;
    p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
}

} /* namespace wgsl */
#endif /* WGSL_BUILTINS_H */
