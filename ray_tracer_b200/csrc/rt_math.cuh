// rt_math.cuh — device-side scalar routines for the shader builtins whose results the
// reference leaves to the Vulkan driver (log, cos, pow/exp2: shaders/ray_tracer.wgsl:201,
// 202, 298, 301).  Each routine is a fixed sequence of binary32 operations (RNE, explicit
// FMAs, no MUFU approximations), so its result is a pure function of its input bits on
// any IEEE-754 machine; DESIGN.md "Arithmetic contract" lists the sequences.  Polynomial
// coefficients: single-precision Cephes minimax sets (public domain).
//
// This translation unit is compiled with -fmad=false: nvcc fuses nothing on its own, an
// FFMA is emitted only where __fmaf_rn is written.  Division and sqrtf are the IEEE
// correctly rounded forms (no --use_fast_math, -prec-div/-prec-sqrt left at true).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rt {

__device__ __forceinline__ float qnan() { return __uint_as_float(0x7fc00000u); }
__device__ __forceinline__ float pinf() { return __uint_as_float(0x7f800000u); }

// natural logarithm
__device__ __forceinline__ float spec_logf(float x)
{
    if (x != x) return x;
    if (x < 0.0f) return qnan();
    if (x == 0.0f) return -pinf();
    if (x == pinf()) return x;
    uint32_t ix = __float_as_uint(x);
    int e = 0;
    if (ix < 0x00800000u) { x = x * 8388608.0f; ix = __float_as_uint(x); e = -23; }
    e += (int)(ix >> 23) - 126;
    float m = __uint_as_float((ix & 0x007fffffu) | 0x3f000000u);
    if (m < 0.707106781186547524f) { e -= 1; m = (m + m) - 1.0f; }
    else { m = m - 1.0f; }
    float z = m * m;
    float p = 7.0376836292E-2f;
    p = __fmaf_rn(p, m, -1.1514610310E-1f);
    p = __fmaf_rn(p, m, 1.1676998740E-1f);
    p = __fmaf_rn(p, m, -1.2420140846E-1f);
    p = __fmaf_rn(p, m, 1.4249322787E-1f);
    p = __fmaf_rn(p, m, -1.6668057665E-1f);
    p = __fmaf_rn(p, m, 2.0000714765E-1f);
    p = __fmaf_rn(p, m, -2.4999993993E-1f);
    p = __fmaf_rn(p, m, 3.3333331174E-1f);
    float y = (m * z) * p;
    float fe = (float)e;
    y = __fmaf_rn(fe, -2.12194440e-4f, y);
    y = __fmaf_rn(-0.5f, z, y);
    float r = m + y;
    r = __fmaf_rn(fe, 0.693359375f, r);
    return r;
}

// 2^t
__device__ __forceinline__ float spec_exp2f(float t)
{
    if (t != t) return t;
    if (t >= 128.0f) return pinf();
    if (t < -150.0f) return 0.0f;
    float n = rintf(t);
    float r = t - n;
    float p = 1.535336188319500E-004f;
    p = __fmaf_rn(p, r, 1.339887440266574E-003f);
    p = __fmaf_rn(p, r, 9.618437357674640E-003f);
    p = __fmaf_rn(p, r, 5.550332471162809E-002f);
    p = __fmaf_rn(p, r, 2.402264791363012E-001f);
    p = __fmaf_rn(p, r, 6.931472028550421E-001f);
    p = __fmaf_rn(p, r, 1.0f);
    int ni = (int)n;
    int n1 = ni >> 1, n2 = ni - n1;
    float s1 = __uint_as_float((uint32_t)(n1 + 127) << 23);
    float s2 = __uint_as_float((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}

// pow(x, y) for x >= 0, defined as exp2(y * log2(x))
__device__ __forceinline__ float spec_powf(float x, float y)
{
    if (x == 0.0f) return (y > 0.0f) ? 0.0f : ((y == 0.0f) ? 1.0f : pinf());
    float l2 = spec_logf(x) * 1.44269504088896341f;
    return spec_exp2f(y * l2);
}

// cosine, Cody-Waite reduction by pi/2
__device__ __forceinline__ float spec_cosf(float x)
{
    if (!(fabsf(x) <= 1.0e6f)) return qnan();
    float j = rintf(x * 0.636619772367581343f);
    float r = __fmaf_rn(-j, 1.5703125f, x);
    r = __fmaf_rn(-j, 4.837512969970703125e-4f, r);
    r = __fmaf_rn(-j, 7.54978995489188216e-8f, r);
    int q = ((int)j) & 3;
    float z = r * r;
    float s = -1.9515295891E-4f;
    s = __fmaf_rn(s, z, 8.3321608736E-3f);
    s = __fmaf_rn(s, z, -1.6666654611E-1f);
    s = __fmaf_rn(s * z, r, r);
    float c = 2.443315711809948E-005f;
    c = __fmaf_rn(c, z, -1.388731625493765E-003f);
    c = __fmaf_rn(c, z, 4.166664568298827E-002f);
    c = __fmaf_rn(c * z, z, __fmaf_rn(-0.5f, z, 1.0f));
    float v = (q & 1) ? s : c;
    return (q == 1 || q == 2) ? -v : v;
}

} // namespace rt
