// rt_shade.cuh — the exact per-primitive tests and the per-segment shading of the reference
// shader, shared by the trace kernels.  Every value computed here can change a pixel, so the
// arithmetic follows DESIGN.md's contract (same f32 operations, same order, explicit FMAs).
//   ray_sphere wgsl:106-123, ray_triangle wgsl:125-147, reflectance/refract wgsl:284-295,
//   rand_normal_dist wgsl:199-203, get_environment_light wgsl:297-304, seed wgsl:307-308,
//   hit shading wgsl:236-273.
#pragma once
#include "rt_device.cuh"

namespace rt {

#define RT_MISS 0xFFFFFFFFu
#define RT_FLT_MAX 3.40282346638528859812e+38f

struct HitRec {
    float dst, u, v;
    uint32_t id;
};

// wgsl:106-123 on slot data. Updates `best` under the tie rule of wgsl:154 (strict '<' in
// the reference's scan order == smaller id wins on equal distance).
__device__ __forceinline__ void test_sphere(float3 o, float3 d, float4 q0, float4 q1, uint32_t id, HitRec& best)
{
    float3 oc = o - xyz(q0);
    float a = dot3(d, d);
    float b = 2.0f * dot3(oc, d);
    float c = dot3(oc, oc) - q1.y;                       // q1.y = radius*radius
    float disc = __fmaf_rn(b, b, -((4.0f * a) * c));
    if (disc >= 0.0f) {
        float dst = (-b - sqrtf(disc)) / (2.0f * a);
        if (dst >= 0.0f && (dst < best.dst || (dst == best.dst && id < best.id && dst != RT_FLT_MAX))) {
            best.dst = dst; best.id = id;
        }
    }
}

// wgsl:125-147 on slot data.  `n` = cross(e1, e2) is precomputed at upload with the very
// operations the reference performs per test (wgsl:129), so the back-face / parallel reject
// of wgsl:131,143 needs only the first 16 bytes of the record; the caller fetches the rest
// (A pre-translated, e1 = B-A, e2 = C-A) for the survivors.
__device__ __forceinline__ bool triangle_faces_ray(float3 d, float3 n, float& det)
{
    det = -dot3(d, n);
    return det >= 1e-6f;
}
__device__ __forceinline__ void test_triangle(float3 o, float3 d, float3 n, float det, float3 A, float3 e1, float3 e2, uint32_t id, HitRec& best)
{
    float3 ao = o - A;
    float3 dao = cross3(ao, d);
    float inv = 1.0f / det;
    float dst = dot3(ao, n) * inv;
    float u = dot3(e2, dao) * inv;
    float v = -(dot3(e1, dao) * inv);
    float w = (1.0f - u) - v;
    if (dst >= 0.0f && u >= 0.0f && v >= 0.0f && w >= 0.0f &&
        (dst < best.dst || (dst == best.dst && id < best.id && dst != RT_FLT_MAX))) {
        best.dst = dst; best.u = u; best.v = v; best.id = id;
    }
}

// wgsl:284-288
__device__ __forceinline__ float reflectance(float cosine, float ratio)
{
    float r0 = (1.0f - ratio) / (1.0f + ratio);
    r0 = r0 * r0;
    float x = 1.0f - cosine;
    float x2 = x * x;
    return r0 + (1.0f - r0) * ((x2 * x2) * x);
}
// wgsl:290-295 (length, not length squared — kept)
__device__ __forceinline__ float3 refract3(float3 uv, float3 n, float ratio)
{
    float cos_theta = fminf(dot3(-uv, n), 1.0f);
    float3 perp = fma3(n, cos_theta, uv) * ratio;
    float k = -sqrtf(fabsf(1.0f - length3(perp)));
    return fma3(n, k, perp);
}
// wgsl:199-203
__device__ __forceinline__ float rand_normal_dist(uint32_t& rng)
{
    float theta = 6.2831852f * rand01(rng);
    float rho = sqrtf(-2.0f * spec_logf(rand01(rng)));
    return rho * spec_cosf(theta);
}
// wgsl:297-304
__device__ __forceinline__ float4 environment_light(float3 d)
{
    float sky_t = spec_powf(smoothstep1(0.0f, 0.4f, d.y), 0.35f);
    float g2s = smoothstep1(-0.01f, 0.0f, d.y);
    float sun = spec_powf(fmaxf(0.0f, dot3(d, f3(0.1f, 1.0f, 0.1f))), 500.0f) * 0.1f;
    float s = sun * ((g2s >= 1.0f) ? 1.0f : 0.0f);
    float4 sky = make_float4(mix1(1.0f, 0.0788092f, sky_t), mix1(1.0f, 0.36480793f, sky_t), mix1(1.0f, 0.7264151f, sky_t), mix1(0.0f, 0.0f, sky_t));
    return make_float4(mix1(0.35f, sky.x, g2s) + s, mix1(0.3f, sky.y, g2s) + s, mix1(0.35f, sky.z, g2s) + s, mix1(0.0f, sky.w, g2s) + s);
}

// wgsl:307-308; strict mode saturates the f32->u32 conversion (WGSL rule)
__device__ __forceinline__ uint32_t pixel_seed(uint32_t x, uint32_t y, uint32_t w, uint32_t h, int32_t frames)
{
    float px = (float)x * (float)w;
    float py = (float)y * (float)h;
    float lin = py * (float)w + px;
    uint32_t af = (frames < 0) ? (0u - (uint32_t)frames) : (uint32_t)frames;
    return __float2uint_rz(lin) + af * 71939u;
}

// One path segment's shading (wgsl:236-273). Consumes RNG exactly as the reference does and
// replaces the ray by the bounce ray.
__device__ __forceinline__ void shade_hit(const SceneDev& sc, const HitRec& hit, float3& ro, float3& rd, uint32_t& rng,
                                          float4& incoming, float4& color)
{
    float3 hp = fma3(rd, hit.dst, ro);
    float4 k0 = __ldg(sc.c0 + hit.id);
    float3 normal;
    if (hit.id < sc.n_spheres) {
        normal = normalize3(hp - xyz(k0));
    } else {
        float4 k1 = __ldg(sc.c1 + hit.id), k2 = __ldg(sc.c2 + hit.id);
        float w = (1.0f - hit.u) - hit.v;
        normal = normalize3(fma3(xyz(k2), hit.v, fma3(xyz(k1), hit.u, xyz(k0) * w)));
    }
    const uint32_t mi = __float_as_uint(k0.w);
    float4 mcolor = __ldg(sc.m0 + mi), emitted = __ldg(sc.m1 + mi);
    const float smooth = __ldg(sc.m2 + mi).x;
    float3 unit_dir = normalize3(rd);
    float3 nd;
    if (smooth == -1.0f) {                       // glass, wgsl:240-264
        mcolor = make_float4(1.f, 1.f, 1.f, 1.f);
        bool front_face = !(dot3(rd, normal) > 0.0f);
        float ratio = front_face ? (1.0f / 1.5f) : 1.5f;
        float cos_theta = fminf(dot3(-unit_dir, normal), 1.0f);
        float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
        bool cannot_refract = ratio * sin_theta > 1.0f;
        if (cannot_refract || reflectance(cos_theta, ratio) > rand01(rng)) nd = reflect3(unit_dir, normal);
        else nd = refract3(unit_dir, normal, ratio);
    } else {                                     // wgsl:265-269
        float gx = rand_normal_dist(rng), gy = rand_normal_dist(rng), gz = rand_normal_dist(rng);
        float3 dir = normalize3(f3(gx, gy, gz));
        float3 diffuse = dir * sign1(dot3(normal, dir));
        float3 spec = reflect3(unit_dir, normal);
        nd = f3(mix1(diffuse.x, spec.x, smooth), mix1(diffuse.y, spec.y, smooth), mix1(diffuse.z, spec.z, smooth));
    }
    ro = hp; rd = nd;
    incoming = incoming + emitted * color;         // wgsl:271-272
    color = color * mcolor;                        // wgsl:273
}

} // namespace rt
