// rt_trace.cu — the path-tracing kernel: camera ray generation, LBVH closest hit over the
// SoA primitive layout, material bounce loop, sky, progressive accumulation.
//
// Replaces the compute entry `main` of shaders/ray_tracer.wgsl (wgsl:50-67) and everything
// it calls (frag :306-327, trace :229-282, calculate_ray_collions :149-185, ray_sphere
// :106-123, ray_triangle :125-147, the RNG :187-227, reflectance/refract :284-295,
// get_environment_light :297-304).  It is NOT a transliteration: one launch covers many
// reference dispatches, the per-pixel work is a flat state machine that refills a lane with
// its next camera path as soon as its current one ends (no lane waits for the longest path
// of its warp), closest hit goes through a BVH instead of the O(S+T) scan, and only the
// winning primitive's shading data is ever fetched.  What is kept bit for bit is the
// arithmetic contract of every value that can change a pixel.
//
// Build: -fmad=false (no implicit contraction), IEEE div/sqrt, no fast-math.
#include "rt_device.cuh"
#include "rt_kernels.h"

namespace rt {

#define RT_MISS 0xFFFFFFFFu
#define RT_FLT_MAX 3.40282346638528859812e+38f
#define RT_STACK 64

struct HitRec {
    float dst, u, v;
    uint32_t id;
};

struct RayCtx {
    float3 o, d;
    float a2, a4;        // 2*dot(d,d), 4*dot(d,d) for the sphere quadratic
};

// wgsl:106-123 on slot data. Returns through `best` under the tie rule of wgsl:154.
__device__ __forceinline__ void test_sphere(const RayCtx& r, float4 q0, float4 q1, uint32_t id, HitRec& best)
{
    float3 oc = r.o - xyz(q0);
    float b = 2.0f * dot3(oc, r.d);
    float c = dot3(oc, oc) - q1.y;                       // q1.y = radius*radius
    float disc = __fmaf_rn(b, b, -(r.a4 * c));
    if (disc >= 0.0f) {
        float dst = (-b - sqrtf(disc)) / r.a2;
        if (dst >= 0.0f && (dst < best.dst || (dst == best.dst && id < best.id && dst != RT_FLT_MAX))) {
            best.dst = dst; best.id = id;
        }
    }
}

// wgsl:125-147 on slot data (A pre-translated, e1 = B-A, e2 = C-A).
__device__ __forceinline__ void test_triangle(const RayCtx& r, float4 q0, float4 q1, float4 q2, uint32_t id, HitRec& best)
{
    float3 e1 = f3(q1.x, q1.y, q1.z);
    float3 e2 = f3(q1.w, q2.x, q2.y);
    float3 n = cross3(e1, e2);
    float det = -dot3(r.d, n);
    if (!(det >= 1e-6f)) return;
    float3 ao = r.o - xyz(q0);
    float3 dao = cross3(ao, r.d);
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

template <bool DETAIL>
__device__ __forceinline__ void test_slot(const SceneDev& sc, const RayCtx& r, uint32_t k, HitRec& best, uint32_t& n_tri, uint32_t& n_sph)
{
    float4 q0 = __ldg(sc.g0 + k);
    float4 q1 = __ldg(sc.g1 + k);
    uint32_t id = __float_as_uint(q0.w);
    if (id < sc.n_spheres) {
        if (DETAIL) n_sph++;
        test_sphere(r, q0, q1, id, best);
    } else {
        if (DETAIL) n_tri++;
        float4 q2 = __ldg(sc.g2 + k);
        test_triangle(r, q0, q1, q2, id, best);
    }
}

// Closest hit. BRUTE scans every slot (validation mode, the reference's own algorithm);
// otherwise a stack traversal of the LBVH with near-child-first ordering. Box tests are
// padded by `pad` in position space so that culling can never drop a primitive the exact
// test would accept (DESIGN.md "Conservative culling").
template <bool BRUTE, bool DETAIL>
__device__ __forceinline__ HitRec closest_hit(const SceneDev& sc, const RayCtx& r, uint32_t& n_nodes, uint32_t& n_tri, uint32_t& n_sph)
{
    HitRec best;
    best.dst = RT_FLT_MAX; best.u = 0.0f; best.v = 0.0f; best.id = RT_MISS;
    if (BRUTE) {
        for (uint32_t k = 0; k < sc.n_prims; k++) test_slot<DETAIL>(sc, r, k, best, n_tri, n_sph);
        return best;
    }
    if (sc.root < 0) return best;

    // per-ray slab setup: t = plane*inv - (o*inv +/- pad*|inv|); not parity-relevant
    float3 inv = f3(__fdividef(1.0f, r.d.x), __fdividef(1.0f, r.d.y), __fdividef(1.0f, r.d.z));
    float pad = sc.pad_scale * fmaxf(fmaxf(sc.max_abs, fabsf(r.o.x)), fmaxf(fabsf(r.o.y), fabsf(r.o.z)));
    float3 oi = f3(r.o.x * inv.x, r.o.y * inv.y, r.o.z * inv.z);
    float3 pi = f3(pad * fabsf(inv.x), pad * fabsf(inv.y), pad * fabsf(inv.z));
    float3 on = f3(-(oi.x + pi.x), -(oi.y + pi.y), -(oi.z + pi.z));   // near-plane offset
    float3 of = f3(-(oi.x - pi.x), -(oi.y - pi.y), -(oi.z - pi.z));   // far-plane offset
    const bool sx = inv.x >= 0.0f, sy = inv.y >= 0.0f, sz = inv.z >= 0.0f;

    int stack[RT_STACK];
    int sp = 0;
    int node = sc.root;
    for (;;) {
        if (node >= 0) {
            const float4* np = sc.nodes + 4 * (size_t)node;
            float4 n0 = __ldg(np + 0), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
            if (DETAIL) n_nodes++;
            // left box: min (n0.x,n0.y,n0.z) max (n0.w,n1.x,n1.y); right: min (n1.z,n1.w,n2.x) max (n2.y,n2.z,n2.w)
            float lnx = sx ? n0.x : n0.w, lfx = sx ? n0.w : n0.x;
            float lny = sy ? n0.y : n1.x, lfy = sy ? n1.x : n0.y;
            float lnz = sz ? n0.z : n1.y, lfz = sz ? n1.y : n0.z;
            float rnx = sx ? n1.z : n2.y, rfx = sx ? n2.y : n1.z;
            float rny = sy ? n1.w : n2.z, rfy = sy ? n2.z : n1.w;
            float rnz = sz ? n2.x : n2.w, rfz = sz ? n2.w : n2.x;
            float ltn = fmaxf(fmaxf(__fmaf_rn(lnx, inv.x, on.x), __fmaf_rn(lny, inv.y, on.y)), fmaxf(__fmaf_rn(lnz, inv.z, on.z), 0.0f));
            float ltf = fminf(fminf(__fmaf_rn(lfx, inv.x, of.x), __fmaf_rn(lfy, inv.y, of.y)), fminf(__fmaf_rn(lfz, inv.z, of.z), best.dst));
            float rtn = fmaxf(fmaxf(__fmaf_rn(rnx, inv.x, on.x), __fmaf_rn(rny, inv.y, on.y)), fmaxf(__fmaf_rn(rnz, inv.z, on.z), 0.0f));
            float rtf = fminf(fminf(__fmaf_rn(rfx, inv.x, of.x), __fmaf_rn(rfy, inv.y, of.y)), fminf(__fmaf_rn(rfz, inv.z, of.z), best.dst));
            bool hl = ltn <= ltf, hr = rtn <= rtf;
            int cl = __float_as_int(n3.x), cr = __float_as_int(n3.y);
            if (hl && hr) {
                bool left_first = ltn <= rtn;
                stack[sp++] = left_first ? cr : cl;
                node = left_first ? cl : cr;
                continue;
            } else if (hl) { node = cl; continue; }
            else if (hr) { node = cr; continue; }
        } else {
            uint32_t leaf = (uint32_t)(~node);
            uint32_t first = leaf >> 3, count = (leaf & 7u) + 1u;
            for (uint32_t k = first; k < first + count; k++) test_slot<DETAIL>(sc, r, k, best, n_tri, n_sph);
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
    return best;
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

// One path segment's shading (wgsl:236-273). Consumes RNG exactly as the reference does,
// updates the ray in place, returns nothing: the caller decides about path termination.
__device__ __forceinline__ void shade_hit(const SceneDev& sc, const HitRec& hit, RayCtx& ray, uint32_t& rng,
                                          float4& incoming, float4& color)
{
    float3 hp = fma3(ray.d, hit.dst, ray.o);
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
    float3 unit_dir = normalize3(ray.d);
    float3 nd;
    if (smooth == -1.0f) {                       // glass, wgsl:240-264
        mcolor = make_float4(1.f, 1.f, 1.f, 1.f);
        bool front_face = !(dot3(ray.d, normal) > 0.0f);
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
    ray.o = hp; ray.d = nd;
    incoming = incoming + emitted * color;         // wgsl:271-272
    color = color * mcolor;                        // wgsl:273
}

// ---- the trace kernel ---------------------------------------------------------------------
// Persistent warps pull work from one global queue. A work item is one (pixel, dispatch)
// pair: the rpp+1 camera paths of one pixel in one reference dispatch, whose mean/rpp
// ("frag", wgsl:326) goes to the staging plane of that dispatch.  Items are numbered
// tile-major, dispatch-minor, 32 pixels (an 8x4 tile) at a time, so the lanes of a warp stay
// on neighbouring pixels while heavy tiles are spread over every SM of the chip.  A lane
// takes its next item the moment its current one ends — no lane waits for the longest path
// of its warp, no warp waits for its block.
template <bool BRUTE, bool DETAIL>
__global__ void __launch_bounds__(RT_BLOCK_THREADS)
trace_kernel(SceneDev sc, CameraDev cam, LaunchParams lp, float4* __restrict__ staging,
             uint32_t* __restrict__ primary_ids, unsigned long long* __restrict__ counters,
             unsigned long long* __restrict__ queue)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull;
    const size_t plane = (size_t)lp.width * lp.height;
    const float sw = (float)lp.width, sh = (float)lp.height, rppf = (float)lp.rpp;
    uint32_t n_rays = 0, n_samples = 0, n_nodes = 0, n_tri = 0, n_sph = 0, n_hits = 0;

    uint64_t pool_next = 0, pool_end = 0;          // warp-uniform: items granted to this warp
    bool exhausted = false;                        // warp-uniform: the global queue is empty

    // per-lane state
    bool has_item = false, need_path = true;
    uint32_t px = 0, py = 0, fi = 0, rng = 0;
    float fx = 0.f, fy = 0.f;
    int32_t j = 0, seg = 0;
    float4 total = make_float4(0.f, 0.f, 0.f, 0.f), incoming = total, color = total;
    RayCtx ray;
    ray.o = ray.d = f3(0.f, 0.f, 0.f); ray.a2 = ray.a4 = 0.f;

    for (;;) {
        // ---- item boundary: store frag, take the next item (warp-uniform control flow) ----
        if (has_item && need_path && j > lp.rpp) {
            staging[(size_t)fi * plane + (size_t)py * lp.width + px] =
                make_float4(total.x / rppf, total.y / rppf, total.z / rppf, total.w / rppf);   // wgsl:326
            has_item = false;
        }
        unsigned want = __ballot_sync(0xffffffffu, !has_item && !exhausted);
        while (want) {
            if (pool_next == pool_end && !exhausted) {
                unsigned long long g = 0;
                if (lane == 0) g = atomicAdd(queue, (unsigned long long)RT_QUEUE_CHUNK);
                g = __shfl_sync(0xffffffffu, g, 0);
                if (g >= n_items) { exhausted = true; }
                else { pool_next = g; pool_end = (g + RT_QUEUE_CHUNK < n_items) ? g + RT_QUEUE_CHUNK : n_items; }
            }
            if (exhausted) break;
            const bool mine = (want >> lane) & 1u;
            const uint32_t rank = __popc(want & ((1u << lane) - 1u));
            const uint64_t avail = pool_end - pool_next;
            if (mine && rank < avail) {
                const uint64_t g = pool_next + rank;
                const uint32_t pos = (uint32_t)(g & 31ull);
                const uint64_t tf = g >> 5;
                fi = (uint32_t)(tf % lp.n_frames);
                const uint32_t wt = (uint32_t)(tf / lp.n_frames);            // local 8x4 warp tile
                const uint32_t t = (wt >> 1) * lp.tile_world + lp.tile_rank; // 8x8 shard tile (tile t -> rank t % world)
                px = (t % lp.tiles_x) * 8u + (pos & 7u);
                py = (t / lp.tiles_x) * 8u + (wt & 1u) * 4u + (pos >> 3);
                has_item = true;
                if (px < lp.width && py < lp.height) {
                    fx = (float)px; fy = (float)py;
                    rng = pixel_seed(px, py, lp.width, lp.height, lp.first_frame + (int32_t)(fi * lp.frame_world));
                    j = 0; need_path = true;
                    total = make_float4(0.f, 0.f, 0.f, 0.f);
                } else {
                    has_item = false;                                        // outside the image: nothing to do
                }
            }
            const uint32_t granted = (uint32_t)min((uint64_t)__popc(want), avail);
            pool_next += granted;
            want = __ballot_sync(0xffffffffu, !has_item && !exhausted);
        }
        if (__ballot_sync(0xffffffffu, has_item) == 0u) break;               // queue empty, every lane idle

        if (has_item) {
            if (need_path) {
                // camera ray, wgsl:313-321 (j <= rpp is guaranteed here)
                float aax = rand01(rng), aay = rand01(rng);
                float posx = (fx + aax) / sw, posy = (fy + aay) / sh;
                float rdx, rdy;
                for (;;) {                                   // rand_in_unit_disk, wgsl:216-227
                    float r1 = rand01(rng) * 2.0f, r2 = rand01(rng) * 2.0f;
                    rdx = r1 - 1.0f; rdy = r2 - 1.0f;
                    if (length3(f3(rdx, rdy, 0.0f)) <= 1.0f) break;
                }
                rdx = rdx * cam.lens_radius; rdy = rdy * cam.lens_radius;
                float3 offset = fma3(cam.v, rdy, cam.u * rdx);
                ray.o = cam.origin + offset;
                ray.d = fma3(cam.vertical, posy, fma3(cam.horizontal, posx, cam.llc)) - ray.o;
                incoming = make_float4(0.f, 0.f, 0.f, 0.f);
                color = make_float4(1.f, 1.f, 1.f, 1.f);
                seg = 0;
                need_path = false;
                n_samples++;
            }
            bool done;
            if (lp.bounces < 0) {
                done = true;                                                 // wgsl:233 loop never runs
            } else {
                // ---- one path segment: wgsl:234 ---------------------------------------
                float a = dot3(ray.d, ray.d);
                ray.a2 = 2.0f * a; ray.a4 = 4.0f * a;
                HitRec hit = closest_hit<BRUTE, DETAIL>(sc, ray, n_nodes, n_tri, n_sph);
                n_rays++;
                if (primary_ids != nullptr && fi == 0 && j == 0 && seg == 0) primary_ids[(size_t)py * lp.width + px] = hit.id;
                if (hit.id != RT_MISS) {
                    if (DETAIL) n_hits++;
                    shade_hit(sc, hit, ray, rng, incoming, color);
                    seg++;
                    done = seg > lp.bounces;
                } else {
                    if (lp.toggle != 0) incoming = incoming + environment_light(ray.d) * color;  // wgsl:274-279
                    done = true;
                }
            }
            if (done) {
                total = total + incoming;                      // wgsl:323
                j++;
                need_path = true;
            }
        }
    }

    // counters: one atomic per warp
    unsigned long long r = n_rays, s = n_samples, nn = n_nodes, nt = n_tri, ns = n_sph, nh = n_hits;
    for (int o = 16; o > 0; o >>= 1) {
        r += __shfl_xor_sync(0xffffffffu, r, o);
        s += __shfl_xor_sync(0xffffffffu, s, o);
        if (DETAIL) {
            nn += __shfl_xor_sync(0xffffffffu, nn, o); nt += __shfl_xor_sync(0xffffffffu, nt, o);
            ns += __shfl_xor_sync(0xffffffffu, ns, o); nh += __shfl_xor_sync(0xffffffffu, nh, o);
        }
    }
    if (lane == 0) {
        atomicAdd(counters + 0, r);
        atomicAdd(counters + 1, s);
        if (DETAIL) { atomicAdd(counters + 2, nn); atomicAdd(counters + 3, nt); atomicAdd(counters + 4, ns); atomicAdd(counters + 5, nh); }
    }
}

// The store of wgsl:59-66 for the fused dispatches, in dispatch order: one thread per pixel,
// running mean kept in registers, planes read coalesced. HBM-bound: 16 B per pixel per
// dispatch in, 16 B per pixel (+16 B when blending into an existing image) per launch.
__global__ void blend_kernel(LaunchParams lp, const float4* __restrict__ staging, float4* __restrict__ accum)
{
    const size_t plane = (size_t)lp.width * lp.height;
    // thread -> pixel of this rank's 8x8 tiles, row-major inside the tile
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t n_local = (size_t)lp.n_warp_tiles * 32u;
    if (idx >= n_local) return;
    const uint32_t lt = (uint32_t)(idx >> 6), in_tile = (uint32_t)(idx & 63u);
    const uint32_t t = lt * lp.tile_world + lp.tile_rank;
    const uint32_t px = (t % lp.tiles_x) * 8u + (in_tile & 7u), py = (t / lp.tiles_x) * 8u + (in_tile >> 3);
    if (px >= lp.width || py >= lp.height) return;
    const size_t pix = (size_t)py * lp.width + px;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    bool loaded = false;
    for (uint32_t fi = 0; fi < lp.n_frames; fi++) {
        const int32_t frames = lp.first_frame + (int32_t)(fi * lp.frame_world);
        const int32_t blend = (frames >= 0) ? (int32_t)((uint32_t)frames / lp.frame_world) : frames;
        const float4 fr = staging[(size_t)fi * plane + pix];
        if (blend >= 1) {                                                   // wgsl:59-63
            if (!loaded) { acc = accum[pix]; loaded = true; }
            float weight = 1.0f / (float)(blend + 1);
            float om = 1.0f - weight;
            acc = make_float4(acc.x * om + fr.x * weight, acc.y * om + fr.y * weight,
                              acc.z * om + fr.z * weight, acc.w * om + fr.w * weight);
        } else {                                                            // wgsl:64-66
            acc = fr; loaded = true;
        }
    }
    accum[pix] = acc;
}

int trace_blocks_per_sm(bool brute, bool detail)
{
    int n = 0;
    cudaError_t e;
    if (brute) e = detail ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<true, true>, RT_BLOCK_THREADS, 0)
                          : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<true, false>, RT_BLOCK_THREADS, 0);
    else e = detail ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<false, true>, RT_BLOCK_THREADS, 0)
                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<false, false>, RT_BLOCK_THREADS, 0);
    return (e == cudaSuccess && n > 0) ? n : 1;
}

cudaError_t launch_trace(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, float4* staging, float4* accum,
                         uint32_t* primary_ids, unsigned long long* counters, unsigned long long* queue,
                         bool brute, bool detail, int n_sms, cudaStream_t stream)
{
    if (lp.n_warp_tiles == 0 || lp.n_frames == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(queue, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    const uint64_t warps_needed = ((uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull + RT_QUEUE_CHUNK - 1) / RT_QUEUE_CHUNK;
    uint64_t blocks = (uint64_t)n_sms * trace_blocks_per_sm(brute, detail);
    const uint64_t cap = (warps_needed + RT_BLOCK_THREADS / 32 - 1) / (RT_BLOCK_THREADS / 32);
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks), block(RT_BLOCK_THREADS);
    if (brute) {
        if (detail) trace_kernel<true, true><<<grid, block, 0, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
        else trace_kernel<true, false><<<grid, block, 0, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
    } else {
        if (detail) trace_kernel<false, true><<<grid, block, 0, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
        else trace_kernel<false, false><<<grid, block, 0, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const size_t n_local = (size_t)lp.n_warp_tiles * 32u;
    blend_kernel<<<(unsigned)((n_local + 255) / 256), 256, 0, stream>>>(lp, staging, accum);
    return cudaGetLastError();
}

} // namespace rt
