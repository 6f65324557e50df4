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
    uint32_t id;          // global primitive order index (the reference's scan order): tie-break key, primary-id AOV
    uint32_t aux;         // instanced scenes only: index of the cold records, or 0x80000000 | slot of the instance's proxy
};
#define RT_MARKER 0x7ffffffe                        // stack word: "leaving the current instance" (instanced scenes)
#define RT_NO_INST 0xFFFFFFFFu
// a lane can take a node step when `node` is an inner node index: not a leaf (< 0), not RT_MARKER, not RT_EMPTY (0x7fffffff)
__device__ __forceinline__ bool is_inner(int node) { return (uint32_t)node < (uint32_t)RT_MARKER; }

// wgsl:106-123 on slot data. Updates `best` under the tie rule of wgsl:154 (strict '<' in
// the reference's scan order == smaller id wins on equal distance).
__device__ __forceinline__ void test_sphere(float3 o, float3 d, float4 q0, float4 q1, uint32_t id, HitRec& best, uint32_t aux = 0u)
{
    float3 oc = o - xyz(q0);
    float a = dot3(d, d);
    float b = 2.0f * dot3(oc, d);
    float c = dot3(oc, oc) - q1.y;                       // q1.y = radius*radius
    float disc = b * b - (4.0f * a) * c;                  // scalar arithmetic is never fused (contract)
    if (disc >= 0.0f) {
        float dst = (-b - sqrtf(disc)) / (2.0f * a);
        if (dst >= 0.0f && (dst < best.dst || (dst == best.dst && id < best.id && dst != RT_FLT_MAX))) {
            best.dst = dst; best.id = id; best.aux = aux;
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
__device__ __forceinline__ void test_triangle(float3 o, float3 d, float3 n, float det, float3 A, float3 e1, float3 e2, uint32_t id, HitRec& best, uint32_t aux = 0u)
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
        best.dst = dst; best.u = u; best.v = v; best.id = id; best.aux = aux;
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
// wgsl:297-304.  The two pow() calls are the most expensive part of a sky miss; most rays need at most
// one of them: pow(t, 0.35) is t itself for t = 0 and t = 1 (below the horizon / above d.y = 0.4), and
// pow(x, 500) is +0 for every x <= 0.8 (away from the sun) -- identities of the contract's pow checked
// exhaustively by tests/native/sky_shortcuts_check.cpp.  Lanes that need exactly one share a single
// evaluation with their own arguments; a second one runs only where a lane needs both.
__device__ __forceinline__ float4 environment_light(float3 d)
{
    const float t = smoothstep1(0.0f, 0.4f, d.y);
    const float g2s = smoothstep1(-0.01f, 0.0f, d.y);
    const float x = fmaxf(0.0f, dot3(d, f3(0.1f, 1.0f, 0.1f)));
    const bool need1 = !(t == 0.0f || t == 1.0f), need2 = !(x <= 0.8f);
    float p = 0.0f;
    if (need1 || need2) p = spec_powf(need1 ? t : x, need1 ? 0.35f : 500.0f);
    const float sky_t = need1 ? p : t;
    float sun_p = need2 ? p : 0.0f;
    if (need1 && need2) sun_p = spec_powf(x, 500.0f);
    const float sun = sun_p * 0.1f;
    float s = sun * ((g2s >= 1.0f) ? 1.0f : 0.0f);
    float4 sky = make_float4(mix1(1.0f, 0.0788092f, sky_t), mix1(1.0f, 0.36480793f, sky_t), mix1(1.0f, 0.7264151f, sky_t), mix1(0.0f, 0.0f, sky_t));
    return make_float4(mix1(0.35f, sky.x, g2s) + s, mix1(0.3f, sky.y, g2s) + s, mix1(0.35f, sky.z, g2s) + s, mix1(0.0f, sky.w, g2s) + s);
}

// wgsl:307-308.  The f32 -> u32 conversion of a value >= 2^32 (3840x2160: every row from 518 on):
//   seed_mode 0 "strict": saturates to 0xFFFFFFFF — the WGSL rule; those pixels then share one RNG stream per dispatch;
//   seed_mode 1 "wide"  : wraps modulo 2^32 (the float is an exact integer there) — what a SPIR-V OpConvertFToU, which is
//                         undefined out of range, may equally do; identical to strict wherever the value is below 2^32.
__device__ __forceinline__ uint32_t pixel_seed(uint32_t x, uint32_t y, uint32_t w, uint32_t h, int32_t frames, int32_t seed_mode)
{
    float px = (float)x * (float)w;
    float py = (float)y * (float)h;
    float lin = py * (float)w + px;
    uint32_t af = (frames < 0) ? (0u - (uint32_t)frames) : (uint32_t)frames;
    const uint32_t base = (seed_mode == 1) ? (uint32_t)__float2ull_rz(lin) : __float2uint_rz(lin);
    return base + af * 71939u;
}

// Work item g -> (pixel, dispatch): items are numbered tile-major, dispatch-minor, 32 pixels (an 8x4
// warp tile) at a time; shard tile t belongs to rank t % tile_world.  False when the pixel lies
// outside the image (the tile grid is rounded up).
template <bool ORD>
__device__ __forceinline__ bool decode_item(const LaunchParams& lp, uint64_t g, uint32_t& px, uint32_t& py, uint32_t& fi)
{
    const uint32_t pos = (uint32_t)(g & 31ull);
    const uint64_t tf = g >> 5;
    uint32_t wt;
    if ((tf >> 32) == 0ull) { wt = lp.div_frames.div((uint32_t)tf); fi = (uint32_t)tf - wt * lp.n_frames; }
    else { fi = (uint32_t)(tf % lp.n_frames); wt = (uint32_t)(tf / lp.n_frames); }
    const uint32_t lt = (ORD && lp.tile_order) ? __ldg(lp.tile_order + (wt >> 1)) : (wt >> 1);
    const uint32_t t = lt * lp.tile_world + lp.tile_rank;
    const uint32_t ty = lp.div_tiles_x.div(t), tx = t - ty * lp.tiles_x;
    px = tx * 8u + (pos & 7u);
    py = ty * 8u + (wt & 1u) * 4u + (pos >> 3);
    return px < lp.width && py < lp.height;
}
// FOLD items: work item g -> pixel (one item per pixel of this rank's tiles, all fused dispatches), same tile numbering.
template <bool ORD>
__device__ __forceinline__ bool decode_pixel(const LaunchParams& lp, uint64_t g, uint32_t& px, uint32_t& py)
{
    const uint32_t pos = (uint32_t)(g & 31ull);
    const uint32_t wt = (uint32_t)(g >> 5);
    const uint32_t lt = (ORD && lp.tile_order) ? __ldg(lp.tile_order + (wt >> 1)) : (wt >> 1);
    const uint32_t t = lt * lp.tile_world + lp.tile_rank;
    const uint32_t ty = lp.div_tiles_x.div(t), tx = t - ty * lp.tiles_x;
    px = tx * 8u + (pos & 7u);
    py = ty * 8u + (wt & 1u) * 4u + (pos >> 3);
    return px < lp.width && py < lp.height;
}
// local index (in this rank's list) of the 8x8 tile a pixel lies in
__device__ __forceinline__ uint32_t local_tile_of(const LaunchParams& lp, uint32_t pix)
{
    const uint32_t py = lp.div_width.div(pix), px = pix - py * lp.width;
    const uint32_t t = (py >> 3) * lp.tiles_x + (px >> 3);
    return (t - lp.tile_rank) / lp.tile_world;
}
// wgsl:326 `pixel_color / f32(rpp)`: an exact power-of-two scale when it can be, else the IEEE division
__device__ __forceinline__ float4 frag_of_total(const LaunchParams& lp, float4 total)
{
    if (lp.rpp_scale != 0.0f) return make_float4(total.x * lp.rpp_scale, total.y * lp.rpp_scale, total.z * lp.rpp_scale, total.w * lp.rpp_scale);
    const float rppf = (float)lp.rpp;
    return make_float4(total.x / rppf, total.y / rppf, total.z / rppf, total.w / rppf);
}

// ---- texture extension (no reference counterpart; arithmetic specified by the oracle's
// sample_texture / apply_textures) ----------------------------------------------------------
// The four texels come through the texture path (CUDA array + texture object, point
// sampling, unnormalised coordinates); wrap and the bilinear weights are computed here with
// the contract's mix(), so the result does not depend on the texture unit's fixed-point
// filtering.
__device__ __forceinline__ int wrap_index(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }
__device__ __forceinline__ float4 fetch_texel(const TexDev& t, int ix, int iy)
{
    const uchar4 p = tex2D<uchar4>(t.obj, (float)ix + 0.5f, (float)iy + 0.5f);
    const float k = 0.003921568859368563f;           // f32(1/255)
    return make_float4((float)p.x * k, (float)p.y * k, (float)p.z * k, (float)p.w * k);
}
__device__ __forceinline__ float4 mix4(float4 a, float4 b, float t)
{
    return make_float4(mix1(a.x, b.x, t), mix1(a.y, b.y, t), mix1(a.z, b.z, t), mix1(a.w, b.w, t));
}
__device__ __forceinline__ float4 sample_texture(const TexDev& t, float u, float v)
{
    float x = __fmaf_rn(u, (float)t.width, -0.5f), y = __fmaf_rn(v, (float)t.height, -0.5f);
    float x0 = floorf(x), y0 = floorf(y);
    float fx = x - x0, fy = y - y0;
    if (!(fabsf(x0) <= 1.0e9f)) { x0 = 0.0f; fx = 0.0f; }
    if (!(fabsf(y0) <= 1.0e9f)) { y0 = 0.0f; fy = 0.0f; }
    const int ix0 = wrap_index((int)x0, (int)t.width), iy0 = wrap_index((int)y0, (int)t.height);
    const int ix1 = (ix0 + 1 == (int)t.width) ? 0 : ix0 + 1, iy1 = (iy0 + 1 == (int)t.height) ? 0 : iy0 + 1;
    const float4 top = mix4(fetch_texel(t, ix0, iy0), fetch_texel(t, ix1, iy0), fx);
    const float4 bot = mix4(fetch_texel(t, ix0, iy1), fetch_texel(t, ix1, iy1), fx);
    return mix4(top, bot, fy);
}
// diffuse: mcolor *= texel; normal map: normal = normalize(T*n.x + B*n.y + N*n.z), n = 2*rgb - 1
static __device__ __noinline__ void apply_textures(const SceneDev& sc, uint32_t id, float bu, float bv, uint32_t dtex, uint32_t ntex,
                                            float3& normal, float4& mcolor)
{
    const float4 k3 = __ldg(sc.c3 + id), k4 = __ldg(sc.c4 + id);
    const float bw = (1.0f - bu) - bv;
    const float tu = __fmaf_rn(k4.x, bv, __fmaf_rn(k3.z, bu, k3.x * bw));
    const float tv = __fmaf_rn(k4.y, bv, __fmaf_rn(k3.w, bu, k3.y * bw));
    if (ntex != 0u && ntex <= sc.n_tex) {
        const float du1 = k3.z - k3.x, dv1 = k3.w - k3.y, du2 = k4.x - k3.x, dv2 = k4.y - k3.y;
        const float det = __fmaf_rn(du1, dv2, -(du2 * dv1));
        if (det != 0.0f && det == det) {
            const float4 k5 = __ldg(sc.c5 + id);
            const float3 e1 = f3(k4.z, k4.w, k5.x), e2 = f3(k5.y, k5.z, k5.w);
            const float sg = sign1(det);
            const float3 N = normal;
            const float3 Traw = fma3(e2, -dv1, e1 * dv2) * sg;
            const float3 Braw = fma3(e1, -du2, e2 * du1) * sg;
            float3 T = fma3(N, -dot3(N, Traw), Traw);
            if (dot3(T, T) > 0.0f) {
                T = normalize3(T);
                float3 Bt = cross3(N, T);
                Bt = Bt * sign1(dot3(Bt, Braw));
                const float4 tx = sample_texture(sc.tex[ntex - 1u], tu, tv);
                const float nx = __fmaf_rn(tx.x, 2.0f, -1.0f), ny = __fmaf_rn(tx.y, 2.0f, -1.0f), nz = __fmaf_rn(tx.z, 2.0f, -1.0f);
                normal = normalize3(fma3(N, nz, fma3(Bt, ny, T * nx)));
            }
        }
    }
    if (dtex != 0u && dtex <= sc.n_tex) mcolor = mcolor * sample_texture(sc.tex[dtex - 1u], tu, tv);
}

// One path segment's shading (wgsl:236-273). Consumes RNG exactly as the reference does and
// replaces the ray by the bounce ray.
// INST (instanced scenes): the cold records are addressed through hit.aux, the material of an instanced triangle is its proxy's.
template <bool TEX, bool INST = false>
__device__ __forceinline__ void shade_hit(const SceneDev& sc, const HitRec& hit, float3& ro, float3& rd, uint32_t& rng,
                                          float4& incoming, float4& color)
{
    float3 hp = fma3(rd, hit.dst, ro);
    uint32_t cidx = hit.id, mat_inst = RT_NO_INST;
    if (INST) {
        if (hit.aux & 0x80000000u) {
            const float4 p1 = __ldg(sc.prims + 4 * (size_t)(hit.aux & 0x7fffffffu) + 1);    // (root, first global id, first cold index, material)
            cidx = __float_as_uint(p1.z) + (hit.id - __float_as_uint(p1.y));
            mat_inst = __float_as_uint(p1.w);
        } else cidx = hit.aux;
    }
    float4 k0 = __ldg(sc.c0 + cidx);
    float3 normal;
    if (hit.id < sc.n_spheres) {
        normal = normalize3(hp - xyz(k0));
    } else {
        float4 k1 = __ldg(sc.c1 + cidx), k2 = __ldg(sc.c2 + cidx);
        float w = (1.0f - hit.u) - hit.v;
        normal = normalize3(fma3(xyz(k2), hit.v, fma3(xyz(k1), hit.u, xyz(k0) * w)));
    }
    const uint32_t mi = (INST && mat_inst != RT_NO_INST) ? mat_inst : __float_as_uint(k0.w);
    float4 mcolor = __ldg(sc.m0 + mi), emitted = __ldg(sc.m1 + mi);
    const float4 mat2 = __ldg(sc.m2 + mi);
    const float smooth = mat2.x;
    if (TEX && hit.id >= sc.n_spheres) {
        const uint32_t dtex = __float_as_uint(mat2.y), ntex = __float_as_uint(mat2.z);
        if ((dtex | ntex) != 0u) apply_textures(sc, cidx, hit.u, hit.v, dtex, ntex, normal, mcolor);
    }
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
