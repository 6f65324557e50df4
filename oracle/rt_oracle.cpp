/* rt_oracle.cpp — TEST INFRASTRUCTURE. CPU oracle for the path-tracing hot path.
 *
 * A function-by-function C++ restatement of the reference's compute shader
 * /root/reference/shaders/ray_tracer.wgsl ("wgsl:" below) — the reference has no CPU
 * renderer of its own and its WGSL cannot be run here (no Rust, no WGSL tooling, no
 * graphics stack), so this file IS the checker.  It keeps the reference's brute-force
 * algorithm, loop bounds (rpp+1 samples, bounces+1 segments), seeding, RNG, AoS records
 * and running-mean accumulation exactly as written there.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (ray_tracer_b200/) never does.
 *
 * PARITY PIN STATUS: the reference ships no golden vectors, tests or fixtures for this
 * path and cannot be executed in this environment, so this oracle is
 * pinned only by self-derived known answers (RNG stream, seed formula, camera uniform,
 * struct offsets, analytic intersections) => "parity unpinned" against the reference's
 * own GPU output, as DESIGN.md states.  Its own transcription is cross-checked by a second
 * restatement of the same WGSL written independently in scalar Python
 * (tests/test_oracle_restatement.py: equal bit for bit on balls / metal / room pixels).
 *
 * EXTENSION (no reference counterpart, SURVEY.md §8(a')): diffuse / tangent-space normal-map
 * textures.  UVs ride in the two padding floats of `Vertex`, 1-based texture handles in
 * padding words of `Mesh`; all-zero padding (what the reference's constructors write) means
 * "no texture" and leaves every result unchanged.  The sampling and tangent-frame
 * arithmetic is specified here (sample_texture / apply_textures) and is what the CUDA path
 * must reproduce bit for bit; it is pinned by nothing upstream.
 *
 * Arithmetic contract (DESIGN.md §"Arithmetic contract"): binary32, RNE, denormals kept,
 * no implicit contraction (-ffp-contract=off); fused multiply-adds appear only where
 * fmaf() is written: dot3, cross, ray point o+d*t, reflect, mix, normal interpolation
 * and the polynomial kernels in rto_math.h.  WGSL permits an implementation to fuse or
 * not, so fixing one pattern is a valid reading of the reference.
 */
#include <stdint.h>
#include <stddef.h>
#include <math.h>
#include <string.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "rto_math.h"

extern "C" {

/* ---- records, byte-identical to the reference's #[repr(C)] structs ------------------ */
struct rto_params {            /* src/core/context.rs:11-21 ; wgsl:1-8 reads the first 24 B */
    uint32_t width, height;
    int32_t number_of_bounces, rays_per_pixel, toggle /* Rust: skybox */, frames;
    int32_t accumulate;        /* host-only */
};
struct rto_camera {            /* src/core/camera.rs:9-29 ; wgsl:69-80 */
    float origin[3], _p1;
    float lower_left_corner[3], _p2;
    float horizontal[3], _p3;
    float vertical[3], _p4;
    float near_, far_, _p5[2];
    float w[3], _p6;
    float u[3], _p7;
    float v[3], lens_radius;
};
struct rto_sphere {            /* src/core/scene.rs:11-21 ; wgsl:9-20 */
    float position[3], radius;
    float color[4], emission_color[4];
    float emission_strength, smoothness, _pad[2];
};
struct rto_mesh {              /* src/core/scene.rs:23-37 ; wgsl:27-33 */
    uint32_t first, triangles, offset;
    uint32_t diffuse_tex;      /* Rust `_padding`: extension, 0 = none, k = texture k-1 */
    float pos[3], _p1;
    float color[4], emission_color[4];
    float emission_strength, smoothness /* Rust: specular */;
    uint32_t normal_tex;       /* Rust `_padding3[0]`: extension, same encoding */
    uint32_t _p2;
};
struct rto_vertex {            /* src/core/scene.rs:79-86 ; wgsl:22-25 */
    float pos[3], tu;          /* Rust `_padding1`: extension, texture coordinate u */
    float normal[3], tv;       /* Rust `_padding2`: extension, texture coordinate v */
};
struct rto_texture {           /* extension: tightly packed RGBA8, row 0 first */
    uint32_t width, height;
    const uint8_t* rgba8;
};

static_assert(sizeof(rto_params) == 28, "Params");
static_assert(sizeof(rto_camera) == 128 && offsetof(rto_camera, w) == 80 &&
              offsetof(rto_camera, lens_radius) == 124, "CameraUniform");
static_assert(sizeof(rto_sphere) == 64 && offsetof(rto_sphere, smoothness) == 52, "Sphere");
static_assert(sizeof(rto_mesh) == 80 && offsetof(rto_mesh, pos) == 16 && offsetof(rto_mesh, diffuse_tex) == 12 &&
              offsetof(rto_mesh, smoothness) == 68 && offsetof(rto_mesh, normal_tex) == 72, "Mesh");
static_assert(sizeof(rto_vertex) == 32 && offsetof(rto_vertex, normal) == 16 && offsetof(rto_vertex, tu) == 12 &&
              offsetof(rto_vertex, tv) == 28, "Vertex");

} /* extern "C" */

namespace {

struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }
inline V4 v4(const float* p) { return V4{p[0], p[1], p[2], p[3]}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V4 operator*(V4 a, float s) { return V4{a.x * s, a.y * s, a.z * s, a.w * s}; }
inline V4 operator*(V4 a, V4 b) { return V4{a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline V4 operator+(V4 a, V4 b) { return V4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }

/* contract helpers ------------------------------------------------------------------- */
inline float dot3(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
inline V3 cross3(V3 a, V3 b) {
    return V3{fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x))};
}
inline V3 fma3(V3 a, float s, V3 c) { return V3{fmaf(a.x, s, c.x), fmaf(a.y, s, c.y), fmaf(a.z, s, c.z)}; }
inline float length3(V3 a) { return sqrtf(dot3(a, a)); }
inline V3 normalize3(V3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return a * inv; }
inline V3 reflect3(V3 i, V3 n) { float t = 2.0f * dot3(n, i); return fma3(n, -t, i); }
inline float mix1(float a, float b, float t) { return fmaf(b, t, a * (1.0f - t)); }
inline V3 mix3(V3 a, V3 b, float t) { return V3{mix1(a.x, b.x, t), mix1(a.y, b.y, t), mix1(a.z, b.z, t)}; }
inline V4 mix4(V4 a, V4 b, float t) { return V4{mix1(a.x, b.x, t), mix1(a.y, b.y, t), mix1(a.z, b.z, t), mix1(a.w, b.w, t)}; }
inline float sign1(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }
inline float smoothstep1(float e0, float e1, float x) {
    float t = fminf(fmaxf((x - e0) / (e1 - e0), 0.0f), 1.0f);
    return (t * t) * (3.0f - 2.0f * t);
}
/* f32 -> u32, truncating, NaN -> 0.  Values >= 2^32 (the seed at 3840x2160, rows >= 518):
 *   seed mode 0 "strict": saturate (WGSL spec);
 *   seed mode 1 "wide"  : wrap modulo 2^32 (the float is an exact integer there) — what SPIR-V's OpConvertFToU,
 *                         undefined out of range, may equally do; identical to strict below 2^32. */
static int g_seed_mode = 0;
inline uint32_t f2u_sat(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) {
        if (g_seed_mode == 1) return (f >= 18446744073709551616.0f) ? 0u : (uint32_t)(uint64_t)f;
        return 0xFFFFFFFFu;
    }
    return (uint32_t)f;
}

struct Material { V4 color, emission_color; float emission_strength, smoothness; };
struct Ray { V3 origin, dir; };
struct Hit {
    bool hit = false; float dst = 0.0f; V3 hit_point{0, 0, 0}; V3 normal{0, 0, 0};
    Material material{{0, 0, 0, 0}, {0, 0, 0, 0}, 0.0f, 0.0f};
    uint32_t prim = 0xFFFFFFFFu;   /* AOV only: global primitive order index */
    /* texture extension: barycentrics, world-space edges, per-vertex UVs, handles */
    float bu = 0.0f, bv = 0.0f, bw = 0.0f; V3 e1{0, 0, 0}, e2{0, 0, 0};
    float uva[2] = {0, 0}, uvb[2] = {0, 0}, uvc[2] = {0, 0};
    uint32_t diffuse_tex = 0, normal_tex = 0;
};

struct SceneView {
    const rto_sphere* spheres; uint32_t ns;
    const rto_vertex* vertices; uint32_t nv;
    const uint32_t* indices; uint32_t ni;
    const rto_mesh* meshes; uint32_t nm;
    const rto_texture* textures; uint32_t ntex;
};

struct Counters { uint64_t segments = 0, sphere_tests = 0, tri_tests = 0, rng_draws = 0, samples = 0; };
static thread_local Counters* g_ctr = nullptr;   /* the running row's counters (RNG draws are counted where they happen) */

/* wgsl:205-210 */
inline uint32_t next_random_number(uint32_t* seed) {
    if (g_ctr) g_ctr->rng_draws++;
    *seed = *seed * 747796405u + 2891336453u;
    uint32_t result = ((*seed >> ((*seed >> 28u) + 4u)) ^ *seed) * 277803737u;
    result = (result >> 22u) ^ result;
    return result;
}
/* wgsl:187-189 — the literal 4294967295.0 is 2^32 in f32 */
inline float rand01(uint32_t* seed) { return (float)next_random_number(seed) / 4294967296.0f; }

/* wgsl:199-203 — 2.0*3.1415926 is folded at f64 then rounded to f32 */
inline float rand_normal_dist(uint32_t* seed) {
    const float two_pi = (float)(2.0 * 3.1415926);
    float theta = two_pi * rand01(seed);
    float rho = sqrtf(-2.0f * rto_logf(rand01(seed)));
    return rho * rto_cosf(theta);
}
/* wgsl:191-197 */
inline V3 rand_unit_sphere(uint32_t* seed) {
    float x = rand_normal_dist(seed);
    float y = rand_normal_dist(seed);
    float z = rand_normal_dist(seed);
    return normalize3(V3{x, y, z});
}
/* wgsl:211-214 */
inline V3 rand_hemisphere_dir_dist(V3 normal, uint32_t* seed) {
    V3 dir = rand_unit_sphere(seed);
    return dir * sign1(dot3(normal, dir));
}
/* wgsl:216-227 — `i -= 1` makes the loop unbounded */
inline V3 rand_in_unit_disk(uint32_t* seed) {
    for (;;) {
        float r1 = rand01(seed) * 2.0f;
        float r2 = rand01(seed) * 2.0f;
        V3 p{r1 - 1.0f, r2 - 1.0f, 1.0f - 1.0f};
        if (length3(p) <= 1.0f) return p;
    }
}

/* wgsl:106-123 */
inline Hit ray_sphere(const Ray& ray, V3 pos, float radius) {
    Hit hit;
    V3 oc = ray.origin - pos;
    float a = dot3(ray.dir, ray.dir);
    float b = 2.0f * dot3(oc, ray.dir);
    float c = dot3(oc, oc) - radius * radius;          /* pow(radius,2.0) */
    float discriminant = b * b - (4.0f * a) * c;       /* scalar arithmetic is never fused */
    if (discriminant >= 0.0f) {
        float dst = (-b - sqrtf(discriminant)) / (2.0f * a);
        if (dst >= 0.0f) {
            hit.hit = true;
            hit.hit_point = fma3(ray.dir, dst, ray.origin);
            hit.dst = dst;
            hit.normal = normalize3(hit.hit_point - pos);
        }
    }
    return hit;
}

struct Vtx { V3 pos, normal; };

/* wgsl:125-147 */
inline Hit ray_triangle(const Ray& ray, const Vtx& a, const Vtx& b, const Vtx& c) {
    Hit hit;
    V3 edge_ab = b.pos - a.pos;
    V3 edge_ac = c.pos - a.pos;
    V3 normal = cross3(edge_ab, edge_ac);
    V3 ao = ray.origin - a.pos;
    V3 dao = cross3(ao, ray.dir);

    float determinant = -dot3(ray.dir, normal);
    float inverse_determinant = 1.0f / determinant;

    float dst = dot3(ao, normal) * inverse_determinant;
    float u = dot3(edge_ac, dao) * inverse_determinant;
    float v = -(dot3(edge_ab, dao) * inverse_determinant);
    float w = (1.0f - u) - v;

    hit.hit = determinant >= 1e-6f && dst >= 0.0f && u >= 0.0f && v >= 0.0f && w >= 0.0f;
    hit.hit_point = fma3(ray.dir, dst, ray.origin);
    V3 nsum = fma3(c.normal, v, fma3(b.normal, u, a.normal * w));
    hit.normal = normalize3(nsum);
    hit.dst = dst;
    hit.bu = u; hit.bv = v; hit.bw = w; hit.e1 = edge_ab; hit.e2 = edge_ac;
    return hit;
}

/* ---- texture extension --------------------------------------------------------------
 * Bilinear, repeat addressing, texel centres at (i + 0.5) / size, channel = byte * (1/255).
 * Written out so that a GPU implementation can fetch the four texels with point sampling
 * and reproduce the interpolation exactly. */
inline int wrap_index(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }
inline V4 texel(const rto_texture& t, int ix, int iy) {
    const uint8_t* p = t.rgba8 + ((size_t)iy * t.width + (size_t)ix) * 4;
    const float k = 0.003921568859368563f;          /* f32(1/255) */
    return V4{(float)p[0] * k, (float)p[1] * k, (float)p[2] * k, (float)p[3] * k};
}
inline V4 sample_texture(const rto_texture& t, float u, float v) {
    float x = fmaf(u, (float)t.width, -0.5f), y = fmaf(v, (float)t.height, -0.5f);
    float x0 = floorf(x), y0 = floorf(y);
    float fx = x - x0, fy = y - y0;
    if (!(fabsf(x0) <= 1.0e9f)) { x0 = 0.0f; fx = 0.0f; }     /* non-finite / absurd coordinates */
    if (!(fabsf(y0) <= 1.0e9f)) { y0 = 0.0f; fy = 0.0f; }
    int ix0 = wrap_index((int)x0, (int)t.width), iy0 = wrap_index((int)y0, (int)t.height);
    int ix1 = (ix0 + 1 == (int)t.width) ? 0 : ix0 + 1, iy1 = (iy0 + 1 == (int)t.height) ? 0 : iy0 + 1;
    V4 top = mix4(texel(t, ix0, iy0), texel(t, ix1, iy0), fx);
    V4 bot = mix4(texel(t, ix0, iy1), texel(t, ix1, iy1), fx);
    return mix4(top, bot, fy);
}
/* Applied to the closest triangle hit before shading: material.color *= diffuse texel;
 * normal = normalize(T*n.x + B*n.y + N*n.z) with n = 2*texel.rgb - 1 and (T, B) the
 * Gram-Schmidt tangent frame of the UV mapping around the interpolated normal N. */
inline void apply_textures(const SceneView& sc, Hit& hit) {
    if (hit.diffuse_tex == 0u && hit.normal_tex == 0u) return;
    float tu = fmaf(hit.uvc[0], hit.bv, fmaf(hit.uvb[0], hit.bu, hit.uva[0] * hit.bw));
    float tv = fmaf(hit.uvc[1], hit.bv, fmaf(hit.uvb[1], hit.bu, hit.uva[1] * hit.bw));
    if (hit.normal_tex != 0u && hit.normal_tex <= sc.ntex) {
        float du1 = hit.uvb[0] - hit.uva[0], dv1 = hit.uvb[1] - hit.uva[1];
        float du2 = hit.uvc[0] - hit.uva[0], dv2 = hit.uvc[1] - hit.uva[1];
        float det = fmaf(du1, dv2, -(du2 * dv1));
        if (det != 0.0f && det == det) {
            float sg = sign1(det);
            V3 N = hit.normal;
            V3 Traw = fma3(hit.e2, -dv1, hit.e1 * dv2) * sg;
            V3 Braw = fma3(hit.e1, -du2, hit.e2 * du1) * sg;
            V3 T = fma3(N, -dot3(N, Traw), Traw);
            if (dot3(T, T) > 0.0f) {
                T = normalize3(T);
                V3 Bt = cross3(N, T);
                Bt = Bt * sign1(dot3(Bt, Braw));
                V4 tx = sample_texture(sc.textures[hit.normal_tex - 1u], tu, tv);
                float nx = fmaf(tx.x, 2.0f, -1.0f), ny = fmaf(tx.y, 2.0f, -1.0f), nz = fmaf(tx.z, 2.0f, -1.0f);
                hit.normal = normalize3(fma3(N, nz, fma3(Bt, ny, T * nx)));
            }
        }
    }
    if (hit.diffuse_tex != 0u && hit.diffuse_tex <= sc.ntex)
        hit.material.color = hit.material.color * sample_texture(sc.textures[hit.diffuse_tex - 1u], tu, tv);
}

inline Material sphere_material(const rto_sphere& s) {
    return Material{v4(s.color), v4(s.emission_color), s.emission_strength, s.smoothness};
}
inline Material mesh_material(const rto_mesh& m) {
    return Material{v4(m.color), v4(m.emission_color), m.emission_strength, m.smoothness};
}


/* wgsl:149-185 — brute force, spheres then meshes in order, strict '<' keeps the earliest */
inline Hit calculate_ray_collions(const SceneView& sc, const Ray& ray, Counters* ctr) {
    Hit closest_hit;
    closest_hit.dst = 3.40282346638528859812e+38f;  /* 0x1.fffffep+127f */
    for (uint32_t i = 0; i < sc.ns; i++) {
        Hit hit = ray_sphere(ray, v3(sc.spheres[i].position), sc.spheres[i].radius);
        if (hit.hit && hit.dst < closest_hit.dst) {
            closest_hit = hit;
            closest_hit.material = sphere_material(sc.spheres[i]);
            closest_hit.prim = i;
        }
    }
    uint32_t tri_base = sc.ns;
    for (uint32_t mesh_index = 0; mesh_index < sc.nm; mesh_index++) {
        const rto_mesh& mesh = sc.meshes[mesh_index];
        for (uint32_t i = 0; i < mesh.triangles; i++) {
            uint32_t first = mesh.first, offset = mesh.offset;
            uint32_t index1 = sc.indices[first + i * 3u];
            uint32_t index2 = sc.indices[first + i * 3u + 1u];
            uint32_t index3 = sc.indices[first + i * 3u + 2u];
            const rto_vertex& r1 = sc.vertices[offset + index1];
            const rto_vertex& r2 = sc.vertices[offset + index2];
            const rto_vertex& r3 = sc.vertices[offset + index3];
            V3 mp = v3(mesh.pos);
            Vtx v1{v3(r1.pos) + mp, v3(r1.normal)};
            Vtx v2{v3(r2.pos) + mp, v3(r2.normal)};
            Vtx v3_{v3(r3.pos) + mp, v3(r3.normal)};
            Hit hit = ray_triangle(ray, v1, v2, v3_);
            if (hit.hit && hit.dst < closest_hit.dst) {
                closest_hit = hit;
                closest_hit.material = mesh_material(mesh);
                closest_hit.prim = tri_base + i;
                closest_hit.uva[0] = r1.tu; closest_hit.uva[1] = r1.tv; closest_hit.uvb[0] = r2.tu; closest_hit.uvb[1] = r2.tv;
                closest_hit.uvc[0] = r3.tu; closest_hit.uvc[1] = r3.tv;
                closest_hit.diffuse_tex = mesh.diffuse_tex; closest_hit.normal_tex = mesh.normal_tex;
            }
        }
        tri_base += mesh.triangles;
    }
    if (ctr) { ctr->segments++; ctr->sphere_tests += sc.ns; ctr->tri_tests += tri_base - sc.ns; }
    if (closest_hit.hit && sc.ntex != 0u) apply_textures(sc, closest_hit);   /* extension; no-op without handles */
    return closest_hit;
}

/* wgsl:284-288 ; pow(x,5.0) evaluated as ((x*x)*(x*x))*x */
inline float reflectance(float cosine, float refraction_ratio) {
    float r0 = (1.0f - refraction_ratio) / (1.0f + refraction_ratio);
    r0 = r0 * r0;
    float x = 1.0f - cosine;
    float x2 = x * x;
    return r0 + (1.0f - r0) * ((x2 * x2) * x);
}
/* wgsl:290-295 — keeps the reference's length() where RTIOW has length squared */
inline V3 refract3(V3 uv, V3 normal, float refraction_ratio) {
    float cos_theta = fminf(dot3(-uv, normal), 1.0f);
    V3 r_out_perp = fma3(normal, cos_theta, uv) * refraction_ratio;
    float k = -sqrtf(fabsf(1.0f - length3(r_out_perp)));
    return fma3(normal, k, r_out_perp);
}
/* wgsl:297-304, constants wgsl:100-104 */
inline V4 get_environment_light(const Ray& ray) {
    const V4 SKY_HORIZON{1.0f, 1.0f, 1.0f, 0.0f};
    const V4 SKY_ZENITH{0.0788092f, 0.36480793f, 0.7264151f, 0.0f};
    const V4 GROUND_COLOR{0.35f, 0.3f, 0.35f, 0.0f};
    const float SUN_INTENSITY = 0.1f, SUN_FOCUS = 500.0f;
    float sky_gradient_t = rto_powf(smoothstep1(0.0f, 0.4f, ray.dir.y), 0.35f);
    float ground_to_sky_t = smoothstep1(-0.01f, 0.0f, ray.dir.y);
    V4 sky_gradient = mix4(SKY_HORIZON, SKY_ZENITH, sky_gradient_t);
    float sun = rto_powf(fmaxf(0.0f, dot3(ray.dir, V3{0.1f, 1.0f, 0.1f})), SUN_FOCUS) * SUN_INTENSITY;
    float s = sun * ((ground_to_sky_t >= 1.0f) ? 1.0f : 0.0f);
    V4 m = mix4(GROUND_COLOR, sky_gradient, ground_to_sky_t);
    return V4{m.x + s, m.y + s, m.z + s, m.w + s};
}

/* wgsl:229-282 */
inline V4 trace(const SceneView& sc, const rto_params& params, Ray ray, uint32_t* seed,
                Counters* ctr, uint32_t* first_prim) {
    V4 ray_color{1.0f, 1.0f, 1.0f, 1.0f};
    V4 incoming_light{0.0f, 0.0f, 0.0f, 0.0f};
    for (int i = 0; i <= params.number_of_bounces; i++) {
        Hit hit = calculate_ray_collions(sc, ray, ctr);
        if (i == 0 && first_prim) *first_prim = hit.hit ? hit.prim : 0xFFFFFFFFu;
        if (hit.hit) {
            ray.origin = hit.hit_point;
            V3 unit_dir = normalize3(ray.dir);
            if (hit.material.smoothness == -1.0f) {
                hit.material.color = V4{1.0f, 1.0f, 1.0f, 1.0f};
                bool front_face = !(dot3(ray.dir, hit.normal) > 0.0f);
                float refraction_ratio = 1.5f;
                if (front_face) refraction_ratio = 1.0f / 1.5f;
                float cos_theta = fminf(dot3(-unit_dir, hit.normal), 1.0f);
                float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
                bool cannot_refract = refraction_ratio * sin_theta > 1.0f;
                if (cannot_refract || reflectance(cos_theta, refraction_ratio) > rand01(seed)) {
                    ray.dir = reflect3(unit_dir, hit.normal);
                } else {
                    ray.dir = refract3(unit_dir, hit.normal, refraction_ratio);
                }
            } else {
                V3 diffuse_dir = rand_hemisphere_dir_dist(hit.normal, seed);
                V3 specular_dir = reflect3(unit_dir, hit.normal);
                ray.dir = mix3(diffuse_dir, specular_dir, hit.material.smoothness);
            }
            V4 emitted_light = hit.material.emission_color * hit.material.emission_strength;
            incoming_light = incoming_light + emitted_light * ray_color;
            ray_color = ray_color * hit.material.color;
        } else {
            if (params.toggle != 0) {
                incoming_light = incoming_light + get_environment_light(ray) * ray_color;
            }
            break;
        }
    }
    return incoming_light;
}

/* wgsl:307-308 */
inline uint32_t pixel_seed(uint32_t x, uint32_t y, const rto_params& params) {
    float px = (float)x * (float)params.width;    /* pixel_coord = i.pos * i.size */
    float py = (float)y * (float)params.height;
    float lin = py * (float)params.width + px;
    int32_t f = params.frames;
    uint32_t af = (f < 0) ? (0u - (uint32_t)f) : (uint32_t)f;
    return f2u_sat(lin) + af * 71939u;
}

/* wgsl:306-327 */
inline V4 frag(const SceneView& sc, const rto_params& params, const rto_camera& camera,
               uint32_t x, uint32_t y, Counters* ctr, uint32_t* first_prim) {
    uint32_t rng_state = pixel_seed(x, y, params);
    V4 total{0.0f, 0.0f, 0.0f, 0.0f};
    float sx = (float)params.width, sy = (float)params.height;
    V3 cu = v3(camera.u), cv = v3(camera.v), co = v3(camera.origin);
    V3 llc = v3(camera.lower_left_corner), hor = v3(camera.horizontal), ver = v3(camera.vertical);
    for (int j = 0; j <= params.rays_per_pixel; j++) {
        float aax = rand01(&rng_state);
        float aay = rand01(&rng_state);
        float posx = ((float)x + aax) / sx;
        float posy = ((float)y + aay) / sy;
        V3 rd = rand_in_unit_disk(&rng_state) * camera.lens_radius;
        V3 offset = fma3(cv, rd.y, cu * rd.x);
        Ray ray;
        ray.origin = co + offset;
        ray.dir = fma3(ver, posy, fma3(hor, posx, llc)) - ray.origin;
        if (ctr) ctr->samples++;
        V4 t = trace(sc, params, ray, &rng_state, ctr, (j == 0) ? first_prim : nullptr);
        total = total + t;
    }
    float n = (float)params.rays_per_pixel;
    return V4{total.x / n, total.y / n, total.z / n, total.w / n};
}

} /* namespace */

static uint64_t* g_stats = nullptr;

extern "C" {

/* One dispatch of the compute entry `main` (wgsl:50-67) over the W x H image.
 *   accum      : W*H*4 f32, read-modify-written exactly like the storage texture
 *                (row 0 = bottom row); must be zero-filled by the
 *                caller before the first dispatch, as wgpu zero-initialises textures.
 *   primary_ids: optional W*H u32 AOV, global primitive order index hit by the primary
 *                ray of sample j=0 (spheres [0,S), then triangles in mesh order), or
 *                0xFFFFFFFF on a miss.
 *   counters   : optional 3 x u64 {segments, sphere tests, triangle tests}, ADDED to.
 *   x_step/y_step: >1 renders only pixels with x%x_step==0 && y%y_step==0 (bounded
 *                CPU-baseline samples); the others are left untouched.
 *   nthreads   : OpenMP threads (<=0: all).
 * Returns 0, or -1 on bad arguments.
 */
int rto_dispatch_tex(const rto_params* params, const rto_camera* camera,
                     const rto_sphere* spheres, uint32_t ns,
                     const rto_vertex* vertices, uint32_t nv,
                     const uint32_t* indices, uint32_t ni,
                     const rto_mesh* meshes, uint32_t nm,
                     const rto_texture* textures, uint32_t ntex,
                     float* accum, uint32_t* primary_ids, uint64_t* counters,
                     int x_step, int y_step, int nthreads);

int rto_dispatch(const rto_params* params, const rto_camera* camera,
                 const rto_sphere* spheres, uint32_t ns,
                 const rto_vertex* vertices, uint32_t nv,
                 const uint32_t* indices, uint32_t ni,
                 const rto_mesh* meshes, uint32_t nm,
                 float* accum, uint32_t* primary_ids, uint64_t* counters,
                 int x_step, int y_step, int nthreads)
{
    return rto_dispatch_tex(params, camera, spheres, ns, vertices, nv, indices, ni, meshes, nm, nullptr, 0,
                            accum, primary_ids, counters, x_step, y_step, nthreads);
}

/* The same dispatch with the texture extension: `textures` are addressed by the 1-based
 * handles in the Mesh records (handles above ntex are ignored). */
int rto_dispatch_tex(const rto_params* params, const rto_camera* camera,
                     const rto_sphere* spheres, uint32_t ns,
                     const rto_vertex* vertices, uint32_t nv,
                     const uint32_t* indices, uint32_t ni,
                     const rto_mesh* meshes, uint32_t nm,
                     const rto_texture* textures, uint32_t ntex,
                     float* accum, uint32_t* primary_ids, uint64_t* counters,
                     int x_step, int y_step, int nthreads)
{
    if (!params || !camera || !accum) return -1;
    if (x_step < 1) x_step = 1;
    if (y_step < 1) y_step = 1;
    const uint32_t W = params->width, H = params->height;
    SceneView sc{spheres, ns, vertices, nv, indices, ni, meshes, nm, textures, ntex};
    uint64_t tot_seg = 0, tot_sph = 0, tot_tri = 0, tot_rng = 0, tot_smp = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) reduction(+ : tot_seg, tot_sph, tot_tri, tot_rng, tot_smp)
#endif
    for (int64_t yy = 0; yy < (int64_t)H; yy += y_step) {
        uint32_t y = (uint32_t)yy;
        Counters ctr;
        g_ctr = &ctr;
        for (uint32_t x = 0; x < W; x += (uint32_t)x_step) {
            uint32_t prim = 0xFFFFFFFFu;
            V4 f = frag(sc, *params, *camera, x, y, &ctr, &prim);
            float* px = accum + ((size_t)y * W + x) * 4;
            if (params->frames >= 1) {                       /* wgsl:59-63 */
                float weight = 1.0f / (float)(params->frames + 1);
                float om = 1.0f - weight;
                px[0] = px[0] * om + f.x * weight;
                px[1] = px[1] * om + f.y * weight;
                px[2] = px[2] * om + f.z * weight;
                px[3] = px[3] * om + f.w * weight;
            } else {                                         /* wgsl:64-66 */
                px[0] = f.x; px[1] = f.y; px[2] = f.z; px[3] = f.w;
            }
            if (primary_ids) primary_ids[(size_t)y * W + x] = prim;
        }
        g_ctr = nullptr;
        tot_seg += ctr.segments; tot_sph += ctr.sphere_tests; tot_tri += ctr.tri_tests; tot_rng += ctr.rng_draws; tot_smp += ctr.samples;
    }
    if (counters) { counters[0] += tot_seg; counters[1] += tot_sph; counters[2] += tot_tri; }
    if (g_stats) { g_stats[0] += tot_rng; g_stats[1] += tot_smp; }
    return 0;
}

/* Optional extra totals for tests: when set, every dispatch ADDS {RNG draws (next_random_number calls),
 * camera paths (trace calls)} to stats[0..1].  Pass NULL to stop. */
void rto_set_stats(uint64_t* stats) { g_stats = stats; }
/* 0 = strict (default), 1 = wide: see f2u_sat.  Process-wide; returns the previous mode. */
int rto_set_seed_mode(int mode) { int old = g_seed_mode; g_seed_mode = (mode == 1) ? 1 : 0; return old; }

int rto_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---- small probes for known-answer tests ------------------------------------------- */
uint32_t rto_next_random_number(uint32_t* state) { return next_random_number(state); }
float rto_rand(uint32_t* state) { return rand01(state); }
uint32_t rto_pixel_seed(uint32_t x, uint32_t y, uint32_t w, uint32_t h, int32_t frames) {
    rto_params p{w, h, 0, 0, 0, frames, 0};
    return pixel_seed(x, y, p);
}
float rto_log(float x) { return rto_logf(x); }
float rto_cos(float x) { return rto_cosf(x); }
float rto_pow(float x, float y) { return rto_powf(x, y); }
float rto_exp2(float x) { return rto_exp2f(x); }
void rto_rand_hemisphere(const float* n, uint32_t* state, float* out) {
    V3 d = rand_hemisphere_dir_dist(v3(n), state); out[0] = d.x; out[1] = d.y; out[2] = d.z;
}
void rto_rand_in_unit_disk(uint32_t* state, float* out) {
    V3 d = rand_in_unit_disk(state); out[0] = d.x; out[1] = d.y; out[2] = d.z;
}
/* out = {hit, dst, hit_point[3], normal[3]} */
void rto_ray_sphere(const float* o, const float* d, const float* pos, float radius, float* out) {
    Ray r{v3(o), v3(d)};
    Hit h = ray_sphere(r, v3(pos), radius);
    out[0] = h.hit ? 1.0f : 0.0f; out[1] = h.dst;
    out[2] = h.hit_point.x; out[3] = h.hit_point.y; out[4] = h.hit_point.z;
    out[5] = h.normal.x; out[6] = h.normal.y; out[7] = h.normal.z;
}
/* tri = 3 x {pos[3], normal[3]} = 18 floats */
void rto_ray_triangle(const float* o, const float* d, const float* tri, float* out) {
    Ray r{v3(o), v3(d)};
    Vtx a{v3(tri), v3(tri + 3)}, b{v3(tri + 6), v3(tri + 9)}, c{v3(tri + 12), v3(tri + 15)};
    Hit h = ray_triangle(r, a, b, c);
    out[0] = h.hit ? 1.0f : 0.0f; out[1] = h.dst;
    out[2] = h.hit_point.x; out[3] = h.hit_point.y; out[4] = h.hit_point.z;
    out[5] = h.normal.x; out[6] = h.normal.y; out[7] = h.normal.z;
}
void rto_environment_light(const float* d, float* out) {
    Ray r{V3{0, 0, 0}, v3(d)};
    V4 c = get_environment_light(r); out[0] = c.x; out[1] = c.y; out[2] = c.z; out[3] = c.w;
}
void rto_refract(const float* uv, const float* n, float ratio, float* out) {
    V3 r = refract3(v3(uv), v3(n), ratio); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float rto_reflectance(float c, float ratio) { return reflectance(c, ratio); }
void rto_sample_texture(const rto_texture* t, float u, float v, float* out) {
    V4 c = sample_texture(*t, u, v); out[0] = c.x; out[1] = c.y; out[2] = c.z; out[3] = c.w;
}

} /* extern "C" */
