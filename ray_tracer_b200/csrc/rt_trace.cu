// rt_trace.cu — the path-tracing kernel: camera ray generation, LBVH closest hit over the
// SoA primitive layout, material bounce loop, sky; plus the blend kernel that folds the
// staged dispatches into the progressive accumulation image.
//
// Replaces the compute entry `main` of shaders/ray_tracer.wgsl (wgsl:50-67) and everything
// it calls (frag :306-327, trace :229-282, calculate_ray_collions :149-185, ray_sphere
// :106-123, ray_triangle :125-147, the RNG :187-227, reflectance/refract :284-295,
// get_environment_light :297-304).  It is NOT a transliteration:
//   * persistent warps pull (pixel, dispatch) work items from one global queue, so heavy
//     pixels are spread over the whole chip and no lane waits for the longest path of its
//     warp or block;
//   * each warp alternates between a SHADE section (finish the segment, fetch work, generate
//     and set up the next ray) and a TRAVERSE section of tight while-while loops in which a
//     lane that reaches a leaf parks it and keeps descending, so primitives are tested with
//     most lanes busy; lanes keep their traversal state across the switch, and the warp only
//     returns to SHADE when enough lanes wait for it (warp votes);
//   * closest hit goes through the LBVH instead of the O(S+T) scan, and only the winning
//     primitive's shading data is ever fetched.
// What is kept bit for bit is the arithmetic contract of every value that can change a
// pixel: same f32 operations, same order, same RNG consumption.
//
// Build: -fmad=false (no implicit contraction), IEEE div/sqrt, no fast-math.
#include "rt_device.cuh"
#include "rt_kernels.h"

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

// wgsl:125-147 on slot data (A pre-translated, e1 = B-A, e2 = C-A).
__device__ __forceinline__ void test_triangle(float3 o, float3 d, float4 q0, float4 q1, float4 q2, uint32_t id, HitRec& best)
{
    float3 e1 = f3(q1.x, q1.y, q1.z);
    float3 e2 = f3(q1.w, q2.x, q2.y);
    float3 n = cross3(e1, e2);
    float det = -dot3(d, n);
    if (!(det >= 1e-6f)) return;
    float3 ao = o - xyz(q0);
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

// ---- the trace kernel ---------------------------------------------------------------------
// Work item = one (pixel, dispatch) pair: the rpp+1 camera paths of one pixel in one
// reference dispatch, whose sum/rpp ("frag", wgsl:326) goes to that dispatch's staging plane.
// Items are numbered tile-major, dispatch-minor, 32 pixels (an 8x4 tile) at a time.
// ---- the trace kernel ---------------------------------------------------------------------
// Work item = one (pixel, dispatch) pair: the rpp+1 camera paths of one pixel in one
// reference dispatch, whose sum/rpp ("frag", wgsl:326) goes to that dispatch's staging plane.
// Items are numbered tile-major, dispatch-minor, 32 pixels (an 8x4 tile) at a time.
//
// Each warp alternates between two sections:
//   SHADE    — lanes whose traversal has ended shade the hit (or the sky), finish / fetch
//              work items, generate the next ray and set up its traversal;
//   TRAVERSE — tight while-while loops over the LBVH: a node loop in which a lane that
//              reaches a leaf parks it and keeps descending, left by warp vote when too few
//              lanes can still take a node step; then a leaf loop where the parked
//              primitives are tested with most lanes busy.
// The warp goes back to SHADE when enough lanes have finished their ray (sched_s_min), the
// others keep their traversal state (node, stack, closest hit) across the switch.
#define RT_EMPTY 0x7fffffff                        // "no current node": traversal stack exhausted

template <bool BRUTE, bool DETAIL>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_MIN_BLOCKS)
trace_kernel(SceneDev sc, CameraDev cam, LaunchParams lp, float4* __restrict__ staging,
             uint32_t* __restrict__ primary_ids, unsigned long long* __restrict__ counters,
             unsigned long long* __restrict__ queue)
{
    const uint32_t lane = threadIdx.x & 31u;
    int stack[RT_STACK];                           // local memory: lane-interleaved by the hardware, only touched levels occupy L1

    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull;
    const size_t plane = (size_t)lp.width * lp.height;
    const float sw = (float)lp.width, sh = (float)lp.height, rppf = (float)lp.rpp;
    uint32_t n_rays = 0, n_samples = 0, n_nodes = 0, n_tri = 0, n_sph = 0, n_hits = 0;

    uint64_t pool_next = 0, pool_end = 0;          // warp-uniform: items granted to this warp
    bool exhausted = false;                        // warp-uniform: the global queue is empty

    // ---- per-lane state ---------------------------------------------------------------
    bool shading = true;                           // lane is between rays (else: mid-traversal)
    bool idle = false;                             // no item and the queue is empty
    bool has_item = false, traced = false;         // traced: a finished traversal awaits shading
    uint32_t pix = 0, fi = 0, rng = 0;
    int32_t j = 0, seg = 0;
    float4 total = make_float4(0.f, 0.f, 0.f, 0.f), incoming = total, color = total;
    float3 ro = f3(0.f, 0.f, 0.f), rd = ro;        // current ray
    float3 inv = ro, on = ro, of = ro;             // slab setup of the current ray
    HitRec best; best.dst = RT_FLT_MAX; best.u = best.v = 0.f; best.id = RT_MISS;
    int node = RT_EMPTY, sp = 0;                   // current node (inner >= 0, leaf < 0, or RT_EMPTY) / stack height
    uint32_t leaf_first = 0, leaf_count = 0;       // parked leaf primitives

    for (;;) {
        // ======================= SHADE section =======================
        {
            bool need_path = false;
            if (shading && has_item) {
                if (traced) {
                    traced = false;
                    n_rays++;
                    if (primary_ids != nullptr && fi == 0 && j == 0 && seg == 0) primary_ids[pix] = best.id;
                    bool done;
                    if (best.id != RT_MISS) {
                        if (DETAIL) n_hits++;
                        shade_hit(sc, best, ro, rd, rng, incoming, color);
                        seg++;
                        done = seg > lp.bounces;
                    } else {
                        if (lp.toggle != 0) incoming = incoming + environment_light(rd) * color;  // wgsl:274-279
                        done = true;
                    }
                    if (done) { total = total + incoming; j++; need_path = true; }                   // wgsl:323
                } else {
                    need_path = true;                                                              // fresh item
                }
                if (need_path && j > lp.rpp) {
                    staging[(size_t)fi * plane + pix] =
                        make_float4(total.x / rppf, total.y / rppf, total.z / rppf, total.w / rppf);   // wgsl:326
                    has_item = false; need_path = false;
                }
            }
            // ---- take the next item (warp-uniform control flow) ----
            unsigned want = __ballot_sync(0xffffffffu, shading && !idle && !has_item && !exhausted);
            while (want) {
                if (pool_next == pool_end) {
                    unsigned long long g = 0;
                    if (lane == 0) g = atomicAdd(queue, (unsigned long long)RT_QUEUE_CHUNK);
                    g = __shfl_sync(0xffffffffu, g, 0);
                    if (g >= n_items) { exhausted = true; break; }
                    pool_next = g; pool_end = (g + RT_QUEUE_CHUNK < n_items) ? g + RT_QUEUE_CHUNK : n_items;
                }
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
                    const uint32_t px = (t % lp.tiles_x) * 8u + (pos & 7u);
                    const uint32_t py = (t / lp.tiles_x) * 8u + (wt & 1u) * 4u + (pos >> 3);
                    if (px < lp.width && py < lp.height) {                       // else: outside the image, ask again
                        pix = py * lp.width + px;
                        rng = pixel_seed(px, py, lp.width, lp.height, lp.first_frame + (int32_t)(fi * lp.frame_world));
                        j = 0; has_item = true; need_path = true;
                        total = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (lp.rpp < 0) {                                        // wgsl:312 loop never runs
                            staging[(size_t)fi * plane + pix] = make_float4(0.0f / rppf, 0.0f / rppf, 0.0f / rppf, 0.0f / rppf);
                            has_item = false; need_path = false;
                        }
                    }
                }
                const uint32_t granted = (uint32_t)min((uint64_t)__popc(want), avail);
                pool_next += granted;
                want = __ballot_sync(0xffffffffu, shading && !idle && !has_item && !exhausted);
            }
            if (shading && !has_item) idle = true;                               // only reachable once the queue is empty

            if (shading && !idle) {
                if (need_path) {
                    // camera ray, wgsl:313-321
                    const uint32_t px = pix % lp.width, py = pix / lp.width;
                    float aax = rand01(rng), aay = rand01(rng);
                    float posx = ((float)px + aax) / sw, posy = ((float)py + aay) / sh;
                    float rdx, rdy;
                    for (;;) {                                   // rand_in_unit_disk, wgsl:216-227
                        float r1 = rand01(rng) * 2.0f, r2 = rand01(rng) * 2.0f;
                        rdx = r1 - 1.0f; rdy = r2 - 1.0f;
                        if (length3(f3(rdx, rdy, 0.0f)) <= 1.0f) break;
                    }
                    rdx = rdx * cam.lens_radius; rdy = rdy * cam.lens_radius;
                    float3 offset = fma3(cam.v, rdy, cam.u * rdx);
                    ro = cam.origin + offset;
                    rd = fma3(cam.vertical, posy, fma3(cam.horizontal, posx, cam.llc)) - ro;
                    incoming = make_float4(0.f, 0.f, 0.f, 0.f);
                    color = make_float4(1.f, 1.f, 1.f, 1.f);
                    seg = 0;
                    n_samples++;
                }
                // ---- start the traversal of (ro, rd): wgsl:234 ----
                best.dst = RT_FLT_MAX; best.u = 0.0f; best.v = 0.0f; best.id = RT_MISS;
                sp = 0; node = RT_EMPTY; leaf_count = 0;
                if (lp.bounces < 0) {
                    total = total + incoming; j++;                // wgsl:233 loop never runs: the path ends untraced
                    traced = false;                               // stays "shading" as a fresh lane needing a path
                } else if (BRUTE) {
                    leaf_first = 0; leaf_count = sc.n_prims;
                    traced = true; shading = (leaf_count == 0);
                } else if (sc.root < 0) {
                    traced = true;                                // empty scene: immediate miss
                } else {
                    // slab setup: t = plane*inv - (o*inv +/- pad*|inv|); not parity-relevant
                    inv = f3(__fdividef(1.0f, rd.x), __fdividef(1.0f, rd.y), __fdividef(1.0f, rd.z));
                    const float pad = sc.pad_scale * fmaxf(fmaxf(sc.max_abs, fabsf(ro.x)), fmaxf(fabsf(ro.y), fabsf(ro.z)));
                    const float3 oi = f3(ro.x * inv.x, ro.y * inv.y, ro.z * inv.z);
                    const float3 pi = f3(pad * fabsf(inv.x), pad * fabsf(inv.y), pad * fabsf(inv.z));
                    on = f3(-(oi.x + pi.x), -(oi.y + pi.y), -(oi.z + pi.z));   // near-plane offset
                    of = f3(-(oi.x - pi.x), -(oi.y - pi.y), -(oi.z - pi.z));   // far-plane offset
                    node = sc.root;
                    traced = true; shading = false;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, !idle) == 0u) break;                      // queue empty, every lane done

        // ======================= TRAVERSE section =======================
        for (;;) {
            // ---- node loop: descend; a lane reaching a leaf parks it (once) and keeps going ----
            // (left when fewer than `thr` lanes can still take a node step; thr adapts so a thin warp still advances)
            const int cnt0 = __popc(__ballot_sync(0xffffffffu, !shading && node >= 0 && node != RT_EMPTY));
            const int thr = max(1, min(lp.sched_n_stay, (cnt0 + 1) >> 1));
            while (__popc(__ballot_sync(0xffffffffu, !shading && node >= 0 && node != RT_EMPTY)) >= thr) {
                if (!shading && node >= 0 && node != RT_EMPTY) {
                    const float4* np = sc.nodes + 4 * (size_t)node;
                    const float4 n0 = __ldg(np + 0), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3);
                    if (DETAIL) n_nodes++;
                    const bool sx = inv.x >= 0.0f, sy = inv.y >= 0.0f, sz = inv.z >= 0.0f;
                    // left box: min (n0.x,n0.y,n0.z) max (n0.w,n1.x,n1.y); right: min (n1.z,n1.w,n2.x) max (n2.y,n2.z,n2.w)
                    const float lnx = sx ? n0.x : n0.w, lfx = sx ? n0.w : n0.x;
                    const float lny = sy ? n0.y : n1.x, lfy = sy ? n1.x : n0.y;
                    const float lnz = sz ? n0.z : n1.y, lfz = sz ? n1.y : n0.z;
                    const float rnx = sx ? n1.z : n2.y, rfx = sx ? n2.y : n1.z;
                    const float rny = sy ? n1.w : n2.z, rfy = sy ? n2.z : n1.w;
                    const float rnz = sz ? n2.x : n2.w, rfz = sz ? n2.w : n2.x;
                    const float ltn = fmaxf(fmaxf(__fmaf_rn(lnx, inv.x, on.x), __fmaf_rn(lny, inv.y, on.y)), fmaxf(__fmaf_rn(lnz, inv.z, on.z), 0.0f));
                    const float ltf = fminf(fminf(__fmaf_rn(lfx, inv.x, of.x), __fmaf_rn(lfy, inv.y, of.y)), fminf(__fmaf_rn(lfz, inv.z, of.z), best.dst));
                    const float rtn = fmaxf(fmaxf(__fmaf_rn(rnx, inv.x, on.x), __fmaf_rn(rny, inv.y, on.y)), fmaxf(__fmaf_rn(rnz, inv.z, on.z), 0.0f));
                    const float rtf = fminf(fminf(__fmaf_rn(rfx, inv.x, of.x), __fmaf_rn(rfy, inv.y, of.y)), fminf(__fmaf_rn(rfz, inv.z, of.z), best.dst));
                    const bool hl = ltn <= ltf, hr = rtn <= rtf;
                    const int cl = __float_as_int(n3.x), cr = __float_as_int(n3.y);
                    if (hl && hr) {
                        const bool left_first = ltn <= rtn;
                        stack[sp++] = left_first ? cr : cl;
                        node = left_first ? cl : cr;
                    } else if (hl) node = cl;
                    else if (hr) node = cr;
                    else node = (sp > 0) ? stack[--sp] : RT_EMPTY;
                    if (node < 0 && leaf_count == 0u) {          // park the leaf, keep descending
                        const uint32_t leaf = (uint32_t)(~node);
                        leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                        node = (sp > 0) ? stack[--sp] : RT_EMPTY;
                    }
                }
            }
            // ---- leaf loop: test parked primitives, one per lane per step ----
            while (__ballot_sync(0xffffffffu, !shading && leaf_count != 0u)) {
                if (!shading && leaf_count != 0u) {
                    const float4 q0 = __ldg(sc.g0 + leaf_first);
                    const float4 q1 = __ldg(sc.g1 + leaf_first);
                    const uint32_t id = __float_as_uint(q0.w);
                    if (id < sc.n_spheres) {
                        if (DETAIL) n_sph++;
                        test_sphere(ro, rd, q0, q1, id, best);
                    } else {
                        if (DETAIL) n_tri++;
                        const float4 q2 = __ldg(sc.g2 + leaf_first);
                        test_triangle(ro, rd, q0, q1, q2, id, best);
                    }
                    leaf_first++;
                    if (--leaf_count == 0u && node < 0) {        // a second leaf was waiting in `node`
                        const uint32_t leaf = (uint32_t)(~node);
                        leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                        node = (sp > 0) ? stack[--sp] : RT_EMPTY;
                    }
                }
            }
            // lanes with nothing left have their closest hit
            if (!shading && node == RT_EMPTY) shading = true;
            const int n_shade = __popc(__ballot_sync(0xffffffffu, shading && !idle));
            const int n_trav = __popc(__ballot_sync(0xffffffffu, !shading));
            if (n_trav == 0 || n_shade >= lp.sched_s_min) break;
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

static const size_t kTraceSmem = 0;

int trace_blocks_per_sm(bool brute, bool detail)
{
    int n = 0;
    cudaError_t e;
    if (brute) e = detail ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<true, true>, RT_BLOCK_THREADS, kTraceSmem)
                          : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<true, false>, RT_BLOCK_THREADS, kTraceSmem);
    else e = detail ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<false, true>, RT_BLOCK_THREADS, kTraceSmem)
                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_kernel<false, false>, RT_BLOCK_THREADS, kTraceSmem);
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
        if (detail) trace_kernel<true, true><<<grid, block, kTraceSmem, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
        else trace_kernel<true, false><<<grid, block, kTraceSmem, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
    } else {
        if (detail) trace_kernel<false, true><<<grid, block, kTraceSmem, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
        else trace_kernel<false, false><<<grid, block, kTraceSmem, stream>>>(sc, cam, lp, staging, primary_ids, counters, queue);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    const size_t n_local = (size_t)lp.n_warp_tiles * 32u;
    blend_kernel<<<(unsigned)((n_local + 255) / 256), 256, 0, stream>>>(lp, staging, accum);
    return cudaGetLastError();
}

} // namespace rt
