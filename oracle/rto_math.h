/* rto_math.h — TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Scalar f32 "builtin" functions the reference shader calls (shaders/ray_tracer.wgsl
 * uses log :201, cos :202, pow :111,:287,:298,:301, sqrt, normalize, length ...).
 *
 * In the reference these are WGSL builtins lowered by naga to SPIR-V GLSL.std.450 and
 * evaluated by whatever Vulkan driver runs the shader; their results are only bounded
 * (Vulkan: log 3 ULP outside [0.5,2], cos abs err 2^-11, pow inherited from
 * exp2(y*log2 x)), not pinned.  No reference test pins them either.
 * The oracle therefore fixes ONE concrete, portable evaluation for each builtin —
 * plain binary32 arithmetic, round-to-nearest-even, explicit fmaf() where fused, no
 * libm — so that a CPU and a GPU evaluating the same sequence agree bit for bit.
 * tests/test_oracle_math.py bounds each routine against float64 libm (<= 2 ulp on the
 * ranges the shader uses), i.e. well inside what the reference's own stack guarantees.
 *
 * Coefficients are the classic single-precision Cephes minimax sets (public domain).
 * sqrtf() and '/' are IEEE correctly rounded on both sides and are used directly.
 */
#ifndef RTO_MATH_H
#define RTO_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline uint32_t rto_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float rto_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* natural log; x==0 -> -inf, x<0 -> NaN, +inf -> +inf */
static inline float rto_logf(float x)
{
    if (x != x) return x;
    if (x < 0.0f) return rto_u2f(0x7fc00000u);
    if (x == 0.0f) return -INFINITY;
    if (x == INFINITY) return x;
    uint32_t ix = rto_f2u(x);
    int e = 0;
    if (ix < 0x00800000u) { x = x * 8388608.0f; ix = rto_f2u(x); e = -23; }
    e += (int)(ix >> 23) - 126;
    float m = rto_u2f((ix & 0x007fffffu) | 0x3f000000u); /* [0.5,1) */
    if (m < 0.707106781186547524f) { e -= 1; m = (m + m) - 1.0f; }
    else { m = m - 1.0f; }
    float z = m * m;
    float p = 7.0376836292E-2f;
    p = fmaf(p, m, -1.1514610310E-1f);
    p = fmaf(p, m, 1.1676998740E-1f);
    p = fmaf(p, m, -1.2420140846E-1f);
    p = fmaf(p, m, 1.4249322787E-1f);
    p = fmaf(p, m, -1.6668057665E-1f);
    p = fmaf(p, m, 2.0000714765E-1f);
    p = fmaf(p, m, -2.4999993993E-1f);
    p = fmaf(p, m, 3.3333331174E-1f);
    float y = (m * z) * p;
    float fe = (float)e;
    y = fmaf(fe, -2.12194440e-4f, y);
    y = fmaf(-0.5f, z, y);
    float r = m + y;
    r = fmaf(fe, 0.693359375f, r);
    return r;
}

/* 2^t */
static inline float rto_exp2f(float t)
{
    if (t != t) return t;
    if (t >= 128.0f) return INFINITY;
    if (t < -150.0f) return 0.0f;
    float n = rintf(t);
    float r = t - n; /* exact, in [-0.5,0.5] */
    float p = 1.535336188319500E-004f;
    p = fmaf(p, r, 1.339887440266574E-003f);
    p = fmaf(p, r, 9.618437357674640E-003f);
    p = fmaf(p, r, 5.550332471162809E-002f);
    p = fmaf(p, r, 2.402264791363012E-001f);
    p = fmaf(p, r, 6.931472028550421E-001f);
    p = fmaf(p, r, 1.0f);
    int ni = (int)n;
    int n1 = ni >> 1, n2 = ni - n1; /* both within the normal exponent range */
    float s1 = rto_u2f((uint32_t)(n1 + 127) << 23);
    float s2 = rto_u2f((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}

/* pow for the shader's uses (x >= 0): exp2(y * log2(x)), as GLSL.std.450 defines it */
static inline float rto_powf(float x, float y)
{
    if (x == 0.0f) return (y > 0.0f) ? 0.0f : ((y == 0.0f) ? 1.0f : INFINITY);
    float l2 = rto_logf(x) * 1.44269504088896341f;
    return rto_exp2f(y * l2);
}

/* cos; 3-term Cody-Waite reduction by pi/2 (the shader's argument is in [0, 2*pi]) */
static inline float rto_cosf(float x)
{
    /* outside the reduction's useful range (never reached by the shader): defined as NaN */
    if (!(fabsf(x) <= 1.0e6f)) return rto_u2f(0x7fc00000u);
    float j = rintf(x * 0.636619772367581343f);
    float r = fmaf(-j, 1.5703125f, x);
    r = fmaf(-j, 4.837512969970703125e-4f, r);
    r = fmaf(-j, 7.54978995489188216e-8f, r);
    int q = ((int)j) & 3;
    float z = r * r;
    float s = -1.9515295891E-4f;
    s = fmaf(s, z, 8.3321608736E-3f);
    s = fmaf(s, z, -1.6666654611E-1f);
    s = fmaf(s * z, r, r);
    float c = 2.443315711809948E-005f;
    c = fmaf(c, z, -1.388731625493765E-003f);
    c = fmaf(c, z, 4.166664568298827E-002f);
    c = fmaf(c * z, z, fmaf(-0.5f, z, 1.0f));
    float v = (q & 1) ? s : c;
    return (q == 1 || q == 2) ? -v : v;
}

#endif
