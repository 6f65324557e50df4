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
#include "rt_shade.cuh"
#include <mutex>

// scratch of rt_render_frames_static (end of this file): queue head in the first 256 bytes, then the staging planes
static __device__ __align__(256) unsigned char g_static_scratch[256 + RT_STATIC_SCRATCH_BYTES];

namespace rt {

#ifdef RT_TIME_PROBE
// tuning builds only: when every warp of the last launch started its last SHADE pass and when it left the kernel (globaltimer, ns)
__device__ unsigned long long g_warp_times[2 * 8192];
extern "C" int rt_debug_warp_times(unsigned long long* out, int n)
{
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(out, g_warp_times, (size_t)n * 2 * sizeof(unsigned long long));
    return e == cudaSuccess ? 0 : -1;
}
__device__ __forceinline__ unsigned long long probe_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif
#ifdef RT_LANE_STATS
// tuning builds only (scripts/build_variant.sh stats "-DRT_LANE_STATS"): where the lanes of a warp are while it runs
// the node loop / leaf loop / SHADE, summed over the launch; read with rt_debug_lane_stats (not part of the ABI).
__device__ unsigned long long g_lane_stats[16];
#define LANE_STAT(k, pred) do { const unsigned b_ = __ballot_sync(0xffffffffu, (pred)); if ((threadIdx.x & 31u) == 0) ls[k] += __popc(b_); } while (0)
#else
#define LANE_STAT(k, pred) do { } while (0)
#endif

// ---- the trace kernel ---------------------------------------------------------------------
// Work item, FOLD = true (default): one PIXEL with all the dispatches fused into the launch.  The lane
// that owns the pixel runs its dispatches in order and applies the reference's progressive store
// (wgsl:59-66: running mean with weight 1/(frames+1)) to the accumulation image itself, one
// read-modify-write of 16 bytes per dispatch that stays in L1/L2 — no staging planes, no blend
// kernel, the only DRAM traffic of a launch is the image read once and written once.  The store is
// order-dependent per pixel, which is why one lane owns the pixel.
// Work item, FOLD = false (RT_OPT_ITEM_MODE = 1, and the wavefront kernels): one (pixel, dispatch)
// pair whose sum/rpp ("frag", wgsl:326) goes to that dispatch's staging plane; blend_kernel folds the
// planes in dispatch order afterwards.  Finer items (a better tail on very short launches) for
// 32 B per pixel per dispatch of DRAM traffic.
// Items are numbered tile-major, 32 pixels (an 8x4 tile) at a time.
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
#ifndef RT_NODE_U2
#define RT_NODE_U2 1                 // lanes able to step for a second / third node step on the same warp vote
#endif
#ifndef RT_NODE_U3
#define RT_NODE_U3 12
#endif
#ifndef RT_NODE_U4
#define RT_NODE_U4 20
#endif
#define RT_EMPTY 0x7fffffff                        // "no current node": traversal stack exhausted

// wgsl:59-66 for one dispatch of one pixel: `frames >= 1` blends with weight 1/(frames+1), else overwrites.
// (frames of the fi-th fused dispatch = first_frame + fi * frame_world; its weight index counts this rank's own
// dispatches: blend0 + fi, see LaunchParams.)
__device__ __forceinline__ void fold_store(const LaunchParams& lp, float4* accum, uint32_t pix, uint32_t fi, float4 fr)
{
    const int32_t blend = lp.blend0 + (int32_t)fi;
    if (blend >= 1) {
        const float4 acc = accum[pix];
        const float weight = 1.0f / (float)(blend + 1);
        const float om = 1.0f - weight;
        accum[pix] = make_float4(acc.x * om + fr.x * weight, acc.y * om + fr.y * weight, acc.z * om + fr.z * weight, acc.w * om + fr.w * weight);
    } else {
        accum[pix] = fr;
    }
}

// INST = true (scenes with instanced mesh ranges, SceneBuffers): the hierarchy has two levels.  A top-level leaf may hold a proxy
// record (pos, 0x80000000 | k) (root, first id, first cold index, material); a lane that meets one pushes the rest of its leaf, its
// next node and RT_MARKER, shifts its slab constants into the range's object space (o' = o - pos: only -o/d changes) and descends
// into the range's own nodes.  Leaves there hold RAW vertex triples: the test forms A = a + pos, B, C, the edges and their cross
// product per instance with the reference's own operations (wgsl:172-174,127-129), so the result is that of the flattened record
// bit for bit.  Popping RT_MARKER with no leaf pending restores the world-space constants.
template <bool BRUTE, bool DETAIL, bool TEX, bool PF, bool FOLD, bool INST, bool ORD>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_MIN_BLOCKS)
trace_kernel(SceneDev sc, CameraDev cam, LaunchParams lp, float4* __restrict__ staging, float4* accum,
             uint32_t* __restrict__ primary_ids, unsigned long long* __restrict__ counters,
             unsigned long long* __restrict__ queue)
{
    const uint32_t lane = threadIdx.x & 31u;
    int stack[RT_STACK];                           // local memory: lane-interleaved by the hardware, only touched levels occupy L1
    stack[0] = RT_EMPTY;                           // sentinel under every traversal (sp starts at 1): a pop needs no emptiness test

    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * (FOLD ? 1u : lp.n_frames) * 32ull;
    const unsigned long long guide = (unsigned long long)gridDim.x * (RT_BLOCK_THREADS / 32) * lp.queue_guide;   // warps x K
    const size_t plane = (size_t)lp.width * lp.height;
    const float sw = (float)lp.width, sh = (float)lp.height, rppf = (float)lp.rpp;
    uint32_t n_rays = 0, n_samples = 0, n_nodes = 0, n_tri = 0, n_sph = 0, n_hits = 0;
#ifdef RT_LANE_STATS
    unsigned long long ls[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, ls8 = 0;
#endif

    uint64_t pool_next = 0, pool_end = 0;          // warp-uniform: items granted to this warp (kept in registers: in shared memory C3 -4.6 %)
    bool exhausted = false;                        // warp-uniform: the global queue is empty
#ifdef RT_TIME_PROBE
    unsigned long long t_exhausted = 0;
#endif

    // ---- per-lane state ---------------------------------------------------------------
    bool shading = true;                           // lane is between rays (else: mid-traversal)
    bool idle = false;                             // no item and the queue is empty
    bool has_item = false, traced = false;         // traced: a finished traversal awaits shading
    uint32_t rng = 0;
    int32_t j = 0, seg = 0;
    float4 incoming = make_float4(0.f, 0.f, 0.f, 0.f), color = incoming;
    float3 ro = f3(0.f, 0.f, 0.f), rd = ro;        // current ray
    // Packed 32-byte nodes (node_test_packed) for large flat scenes: the variants without the push-time prefetch.  Same-box A/B
    // against 64-byte float nodes: C5 (1.1 M primitives, L1 data pipe at 85 %) +8.2 %, but C3 -0.1 %, C4 -2.5 %, C2 -4.7 %,
    // C1 -1.5 % (cache-resident scenes gain nothing from the halved node traffic and pay 7 more instructions per node visit
    // and 1.2 % more visits), so the builder packs only above RT_PF_MAX_PRIMS primitives.
    constexpr bool PK = RT_PACKED_NODES && !INST && !PF && !BRUTE;
    const unsigned magic = PK ? node_magic() : 0u;  // 0x4B000000 in a register
    RaySlab slab;                                  // slab-test constants of the current ray: 1/d, -o/d, |1/d|, pad*|1/d|
    slab.inv_xy = slab.noi_xy = slab.ainv_xy = slab.pi_xy = 0ull; slab.inv_z = slab.noi_z = slab.ainv_z = slab.pi_z = 0.f;
    // item state touched once per camera path lives in shared memory (keeps the traversal loops at 64 registers)
    __shared__ float4 s_total[RT_BLOCK_THREADS];   // sum of the item's finished paths (wgsl:323)
    __shared__ uint2 s_item[RT_BLOCK_THREADS];     // (pixel index, dispatch index within the launch)
    float4& total = s_total[threadIdx.x];
    uint32_t& pix = s_item[threadIdx.x].x;
    uint32_t& fi = s_item[threadIdx.x].y;
    total = make_float4(0.f, 0.f, 0.f, 0.f); pix = 0; fi = 0;
    HitRec best; best.dst = RT_FLT_MAX; best.u = best.v = 0.f; best.id = RT_MISS; best.aux = 0u;
    uint32_t cur_inst = RT_NO_INST;                // INST: slot of the proxy whose range the lane is inside
    auto leave_instance = [&]() {                  // back to world space: -o/d again
        cur_inst = RT_NO_INST;
        slab.noi_xy = f2_pack(-(ro.x * f2_lo(slab.inv_xy)), -(ro.y * f2_hi(slab.inv_xy))); slab.noi_z = -(ro.z * slab.inv_z);
    };
    float cull = RT_FLT_MAX;                       // best.dst with slack: the distance beyond which a subtree is skipped (see RT_CULL_SLACK)
    int node = RT_EMPTY;                           // current node (inner >= 0, leaf < 0, or RT_EMPTY)
    int* sp = stack;                               // the stack top as an ADDRESS: a push / pop is one STL / LDL and one add, no index scaling (3 LEA fewer per node step: C3 +0.6 %, C4 +0.8 %)
#define RT_SP_RESET() (sp = stack + 1)
#define RT_PUSH(v) (*sp++ = (v))
#define RT_POP() (*--sp)
    uint32_t leaf_first = 0, leaf_count = 0;       // parked leaf primitives

    for (;;) {
        // ======================= SHADE section =======================
        {
            LANE_STAT(9, lane == 0); LANE_STAT(10, shading && has_item && traced); LANE_STAT(11, shading && has_item && traced && best.id != RT_MISS);
            LANE_STAT(12, !shading);
            bool need_path = false;
            if (shading && has_item) {
                if (traced) {
                    traced = false;
                    n_rays++;
                    if (primary_ids != nullptr && fi == 0 && j == 0 && seg == 0) primary_ids[pix] = best.id;
                    bool done;
                    if (best.id != RT_MISS) {
                        if (DETAIL) n_hits++;
                        shade_hit<TEX, INST>(sc, best, ro, rd, rng, incoming, color);
                        seg++;
                        done = seg > lp.bounces;
                    } else {
                        if (lp.toggle != 0) incoming = incoming + environment_light(rd) * color;  // wgsl:274-279
                        done = true;
                    }
                    if (done) {                                                                    // wgsl:323
                        total = total + incoming; j++; need_path = true;
                        // this tile's weight in the next call's order: the segments of the first dispatch's paths, every 4th pixel
                        if (ORD && lp.tile_cost != nullptr && fi == 0u && (pix & 3u) == 0u) atomicAdd(lp.tile_cost + local_tile_of(lp, pix), (uint32_t)seg + 1u);
                    }
                } else {
                    need_path = true;                                                              // fresh item
                }
                if (need_path && j > lp.rpp) {                                                     // the dispatch's frag is complete, wgsl:326
                    const float4 fr = frag_of_total(lp, total);
                    if (FOLD) {
                        fold_store(lp, accum, pix, fi, fr);                                        // wgsl:59-66
                        fi++;
                        if (fi < lp.n_frames) {                                                    // next dispatch of this pixel
                            const uint32_t py = lp.div_width.div(pix), px = pix - py * lp.width;
                            rng = pixel_seed(px, py, lp.width, lp.height, lp.first_frame + (int32_t)(fi * lp.frame_world), lp.seed_mode);
                            j = 0; total = make_float4(0.f, 0.f, 0.f, 0.f);
                        } else { has_item = false; need_path = false; }
                    } else {
                        __stcs(staging + ((size_t)fi * plane + pix), fr);        // written once, read once by the blend: streaming store (+0.3..0.5 %)
                        has_item = false; need_path = false;
                    }
                }
            }
            // ---- take the next item (warp-uniform control flow) ----
            unsigned want = __ballot_sync(0xffffffffu, shading && !idle && !has_item && !exhausted);
            while (want) {
                if (pool_next == pool_end) {
                    // guided self-scheduling: RT_QUEUE_CHUNK items per grab while plenty remain, down to one
                    // 8x4 tile of one dispatch (32 items) near the end, so that no warp is left holding a
                    // large pool of heavy pixels when the queue runs dry (a heavy 128-item pool takes one
                    // warp milliseconds: the launch's fixed cost went from 3.3 ms to under 1 ms on C3)
                    unsigned long long g = 0, chunk = 32ull;                    // FOLD: one 8x4 tile (32 pixels x all dispatches) per grab
                    if (lane == 0) {
                        if (!FOLD) {
#ifdef RT_QUEUE_READ
                            const unsigned long long cur = *(volatile unsigned long long*)queue;
#else
                            // the queue head as of this warp's previous grab: stale by what the other warps took since (at most a few percent of a
                            // launch), which the guide does not mind — and one L2 round trip per grab less than reading the head first
                            const unsigned long long cur = pool_end;
#endif
                            const unsigned long long rem = cur < n_items ? n_items - cur : 0ull;
                            chunk = (rem / guide) & ~31ull;
                            chunk = chunk < 32ull ? 32ull : (chunk > (unsigned long long)lp.queue_chunk ? (unsigned long long)lp.queue_chunk : chunk);
                        }
                        g = atomicAdd(queue, chunk);
                    }
                    g = __shfl_sync(0xffffffffu, g, 0); chunk = __shfl_sync(0xffffffffu, chunk, 0);
#ifdef RT_TIME_PROBE
                    if (g >= n_items && !exhausted) t_exhausted = probe_now();
#endif
                    if (g >= n_items) { exhausted = true; break; }
                    pool_next = g; pool_end = (g + chunk < n_items) ? g + chunk : n_items;
                }
                const bool mine = (want >> lane) & 1u;
                const uint32_t rank = __popc(want & ((1u << lane) - 1u));
                const uint64_t avail = pool_end - pool_next;
                if (mine && rank < avail) {
                    uint32_t px, py;
                    bool inside;
                    if (FOLD) { inside = decode_pixel<ORD>(lp, pool_next + rank, px, py); fi = 0; }
                    else inside = decode_item<ORD>(lp, pool_next + rank, px, py, fi);
                    if (inside) {                                                // else: outside the image, ask again
                        pix = py * lp.width + px;
                        rng = pixel_seed(px, py, lp.width, lp.height, lp.first_frame + (int32_t)(fi * lp.frame_world), lp.seed_mode);
                        j = 0; has_item = true; need_path = true;
                        total = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (lp.rpp < 0) {                                        // wgsl:312 loop never runs
                            const float4 z = make_float4(0.0f / rppf, 0.0f / rppf, 0.0f / rppf, 0.0f / rppf);
                            if (FOLD) { for (uint32_t f = 0; f < lp.n_frames; f++) fold_store(lp, accum, pix, f, z); }
                            else staging[(size_t)fi * plane + pix] = z;
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
                    const uint32_t py = lp.div_width.div(pix), px = pix - py * lp.width;
                    float aax = rand01(rng), aay = rand01(rng);
                                        float posx = ((float)px + aax) / sw, posy = ((float)py + aay) / sh;
                    float rdx, rdy;
                    for (;;) {                                   // rand_in_unit_disk, wgsl:216-227
                        float r1 = rand01(rng) * 2.0f, r2 = rand01(rng) * 2.0f;
                        rdx = r1 - 1.0f; rdy = r2 - 1.0f;
                        // wgsl:222 `length(p) <= 1.0` without the square root: for a correctly rounded sqrt, sqrt(x) <= 1 exactly when
                        // x <= 1 + 2^-23 (sqrt(1 + 2^-23) rounds to 1, sqrt(1 + 2^-22) to 1 + 2^-23; checked in tests/test_oracle_math.py).
                        // The loop runs until the last lane of the warp accepts, so its body is paid about three times per pass.
                        if (dot3(f3(rdx, rdy, 0.0f), f3(rdx, rdy, 0.0f)) <= 1.00000011920928955078125f) break;
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
                best.dst = RT_FLT_MAX; best.u = 0.0f; best.v = 0.0f; best.id = RT_MISS; cull = RT_FLT_MAX;
                if (INST) { best.aux = 0u; cur_inst = RT_NO_INST; }
                RT_SP_RESET(); node = RT_EMPTY; leaf_count = 0;
                if (lp.bounces < 0) {
                    total = total + incoming; j++;                // wgsl:233 loop never runs: the path ends untraced
                    traced = false;                               // stays "shading" as a fresh lane needing a path
                } else if (BRUTE) {
                    leaf_first = 0; leaf_count = sc.n_prims;
                    traced = true; shading = (leaf_count == 0);
                } else if (sc.root < 0) {
                    traced = true;                                // empty scene: immediate miss
                } else {
                    if (!PK) slab_setup(slab, ro, rd, sc.max_abs, sc.pad_scale);
                    else slab_setup_packed(slab, ro, rd, sc.max_abs, sc.pad_scale, sc.grid_o, sc.grid_c);
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
            const int cnt0 = __popc(__ballot_sync(0xffffffffu, !shading && is_inner(node)));
            const int thr = max(1, min(lp.sched_n_stay, (cnt0 + 1) >> 1));
            auto node_step = [&]() {
                LANE_STAT(0, lane == 0); LANE_STAT(1, !shading && is_inner(node)); LANE_STAT(2, shading && !idle); LANE_STAT(3, !shading && node < 0);
                LANE_STAT(4, !shading && node == RT_EMPTY && leaf_count != 0u); LANE_STAT(5, idle);
                if (!shading && is_inner(node)) {
                    if (DETAIL) n_nodes++;
                    float ltn, ltf, rtn, rtf; int cl, cr;
                    // the node's address from a 32-bit byte offset (rt_upload_scene bounds the node count accordingly): LEA + carry on the uniform
                    // base instead of a per-thread constant load feeding a 64-bit multiply-add in front of every node fetch (C3 +0.6 %, C5 +1.3 %)
                    if (!PK) node_test((const float4*)((const char*)sc.nodes + ((uint32_t)node << 6)), slab, cull, ltn, ltf, rtn, rtf, cl, cr);
                    else node_test_packed((const float4*)((const char*)sc.nodes + ((uint32_t)node << 5)), slab, cull, magic, ltn, ltf, rtn, rtf, cl, cr);
                    const bool hl = ltn <= ltf, hr = rtn <= rtf;
                    if (hl && hr) {
                        const bool left_first = ltn <= rtn;
                        RT_PUSH(left_first ? cr : cl);
                        node = left_first ? cl : cr;
                        if (PF) {                                // small scenes: start fetching the postponed child's node now
                            const int far_ = left_first ? cr : cl;
                            if (far_ >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"((const char*)sc.nodes + ((uint32_t)far_ << (PK ? 5 : 6))));
                        }
                    } else if (hl) node = cl;
                    else if (hr) node = cr;
                    else node = RT_POP();
                    if (INST && node == RT_MARKER && leaf_count == 0u) { leave_instance(); node = RT_POP(); }
                    if (node < 0 && leaf_count == 0u) {          // park the leaf, keep descending
                        const uint32_t leaf = (uint32_t)(~node);
                        leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                        node = RT_POP();
                    }
                }
            };
            for (;;) {
                // one vote pays for up to four node steps when most lanes can take them (a lane that runs out idles for the rest)
                const int cnt = __popc(__ballot_sync(0xffffffffu, !shading && is_inner(node)));
                if (cnt < thr) break;
                node_step();
                if (cnt >= RT_NODE_U2) node_step();
                if (cnt >= RT_NODE_U3) node_step();
                if (cnt >= RT_NODE_U4) node_step();
            }
            // ---- leaf loop: test parked primitives, one per lane per step ----
            while (__ballot_sync(0xffffffffu, !shading && leaf_count != 0u)) {
                LANE_STAT(6, lane == 0); LANE_STAT(7, !shading && leaf_count != 0u);
                if (!shading && leaf_count != 0u) {
                    if (INST && cur_inst != RT_NO_INST) {
                        // a triangle of the instanced range: raw a, b, c + the proxy's pos, then the reference's own sequence
                        const float4* pr = sc.iprims + 4 * (size_t)leaf_first;
                        float4 r0, r1, p0, p1;
                        ldg256(pr, r0, r1);
                        const float4 r2 = __ldg(pr + 2);
                        ldg256(sc.prims + 4 * (size_t)cur_inst, p0, p1);
                        const float3 pos = xyz(p0);
                        const float3 A = xyz(r0) + pos, B = f3(r0.w, r1.x, r1.y) + pos, C = f3(r1.z, r1.w, r2.x) + pos;   // wgsl:172-174
                        const float3 e1 = B - A, e2 = C - A;                                                            // wgsl:127-128
                        const float3 nrm = cross3(e1, e2);                                                              // wgsl:129
                        if (DETAIL) n_tri++;
                        float det;
                        if (triangle_faces_ray(rd, nrm, det))
                            test_triangle(ro, rd, nrm, det, A, e1, e2, __float_as_uint(p1.y) + __float_as_uint(r2.y), best, 0x80000000u | cur_inst);
                    } else {
                    const float4* pr = sc.prims + 4 * (size_t)leaf_first;
                    float4 q0, q1;
                    ldg256(pr, q0, q1);
                    const uint32_t id = __float_as_uint(q0.w);
                    if (INST && (id & 0x80000000u)) {
                        // proxy of an instance: remember what is left of this leaf and where to go on, then enter the range's hierarchy
                        // (a lane whose node is RT_EMPTY popped the sentinel, sp == 0: the sentinel itself is its continuation)
                        if (node == RT_EMPTY) RT_SP_RESET();
                        if (leaf_count > 1u) RT_PUSH(~(int)(((leaf_first + 1u) << 3) | (leaf_count - 2u)));
                        if (node != RT_EMPTY) RT_PUSH(node);
                        RT_PUSH(RT_MARKER);
                        node = __float_as_int(q1.x);
                        cur_inst = leaf_first;
                        slab.noi_xy = f2_fma(f2_pack(q0.x, q0.y), slab.inv_xy, slab.noi_xy);          // -(o - pos)/d = -o/d + pos/d
                        slab.noi_z = __fmaf_rn(q0.z, slab.inv_z, slab.noi_z);
                        leaf_count = 1u;
                    } else if (id < sc.n_spheres) {
                        if (DETAIL) n_sph++;
                        test_sphere(ro, rd, q0, q1, id, best, id);
                    } else {
                        if (DETAIL) n_tri++;
                        float det;
#ifdef RT_LANE_STATS
                        ls8++;
#endif
                        if (triangle_faces_ray(rd, xyz(q0), det)) {
                            float4 q2, q3;
                            ldg256(pr + 2, q2, q3);
                            test_triangle(ro, rd, xyz(q0), det, xyz(q1), f3(q1.w, q2.x, q2.y), f3(q2.z, q2.w, q3.x), id, best, __float_as_uint(q3.y));
                        }
                    }
                    }
                    cull = best.dst * RT_CULL_SLACK;
                    leaf_first++;
                    if (--leaf_count == 0u) {
                        if (INST && node == RT_MARKER) { leave_instance(); node = RT_POP(); }
                        if (node < 0) {                              // a leaf was waiting in `node`
                            const uint32_t leaf = (uint32_t)(~node);
                            leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                            node = RT_POP();
                        }
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

#ifdef RT_TIME_PROBE
    if (lane == 0) { const uint32_t w = blockIdx.x * (RT_BLOCK_THREADS / 32) + (threadIdx.x >> 5); if (w < 8192u) { g_warp_times[2 * w] = t_exhausted; g_warp_times[2 * w + 1] = probe_now(); } }
#endif
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
#ifdef RT_LANE_STATS
    for (int o = 16; o > 0; o >>= 1) ls8 += __shfl_xor_sync(0xffffffffu, ls8, o);
    ls[8] = ls8;
    if (lane == 0) for (int k = 0; k < 16; k++) atomicAdd(&g_lane_stats[k], ls[k]);
#endif
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

namespace {
typedef void (*TraceFn)(SceneDev, CameraDev, LaunchParams, float4*, float4*, uint32_t*, unsigned long long*, unsigned long long*);
// PF: prefetch the postponed child of a two-hit node step into L1.  Pays while the hierarchy and the records fit the
// caches (C3 +0.8 %, C4 +1.8 %, C2 +0.4 %); on C5 (1.1 M primitives, 140 MB of nodes + records, L1 data pipe at 85 %)
// the extra requests cost 1.5 %, so the launcher picks the variant by scene size.
template <bool FOLD, bool ORD>
TraceFn trace_variant_t(bool brute, bool detail, bool tex, bool pf, bool inst)
{
    static const TraceFn table[12] = {
        trace_kernel<false, false, false, false, FOLD, false, ORD>, trace_kernel<false, false, true, false, FOLD, false, ORD>, trace_kernel<false, true, false, false, FOLD, false, ORD>, trace_kernel<false, true, true, false, FOLD, false, ORD>,
        trace_kernel<false, false, false, true, FOLD, false, ORD>,  trace_kernel<false, false, true, true, FOLD, false, ORD>,  trace_kernel<false, true, false, true, FOLD, false, ORD>,  trace_kernel<false, true, true, true, FOLD, false, ORD>,
        trace_kernel<true, false, false, false, FOLD, false, ORD>,  trace_kernel<true, false, true, false, FOLD, false, ORD>,  trace_kernel<true, true, false, false, FOLD, false, ORD>,  trace_kernel<true, true, true, false, FOLD, false, ORD>};
    // instanced scenes: hierarchy traversal only (the exhaustive scan needs flattened records), no push-time prefetch
    static const TraceFn inst_table[4] = {
        trace_kernel<false, false, false, false, FOLD, true, ORD>, trace_kernel<false, false, true, false, FOLD, true, ORD>,
        trace_kernel<false, true, false, false, FOLD, true, ORD>,  trace_kernel<false, true, true, false, FOLD, true, ORD>};
    if (inst) return inst_table[(detail ? 2 : 0) | (tex ? 1 : 0)];
    return table[(brute ? 8 : (pf ? 4 : 0)) | (detail ? 2 : 0) | (tex ? 1 : 0)];
}
// ord: the variants that hand tiles out in the learned order and collect the next one's weights (rt_order.cu) — short launches only,
// so that the long ones run the very code they ran without it (compiled into one kernel it cost C3 x 128 0.4 % and C4 x 32 1.4 %)
TraceFn trace_variant(bool brute, bool detail, bool tex, bool pf, bool fold, bool inst, bool ord)
{
    if (ord) return fold ? trace_variant_t<true, true>(brute, detail, tex, pf, inst) : trace_variant_t<false, true>(brute, detail, tex, pf, inst);
    return fold ? trace_variant_t<true, false>(brute, detail, tex, pf, inst) : trace_variant_t<false, false>(brute, detail, tex, pf, inst);
}
} // namespace

int trace_blocks_per_sm(bool brute, bool detail, bool tex, bool pf, bool fold, bool inst, bool ord)
{
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, trace_variant(brute, detail, tex, pf, fold, inst, ord), RT_BLOCK_THREADS, 0);
    return (e == cudaSuccess && n > 0) ? n : 1;
}

// fold = true: `accum` is updated in place by the kernel (staging unused); fold = false: the kernel fills the staging
// planes and blend_kernel folds them into `accum`.
cudaError_t launch_trace(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, float4* staging, float4* accum,
                         uint32_t* primary_ids, unsigned long long* counters, unsigned long long* queue,
                         bool brute, bool detail, bool fold, int n_sms, cudaStream_t stream, bool with_blend)
{
    if (lp.n_warp_tiles == 0 || lp.n_frames == 0) return cudaSuccess;
    if (sc.iprims != nullptr && brute) return cudaErrorNotSupported;   // rt_api.cu refuses this combination with a message
    const bool pf = !brute && sc.iprims == nullptr && !sc.packed && sc.n_prims <= RT_PF_MAX_PRIMS;
    // float nodes are read by the PF variants (small scenes) and the INST variants (instanced scenes; also large flat scenes
    // whose bounds are not finite and so have no packed nodes); the variants in between read packed nodes
    const bool inst = sc.iprims != nullptr || (RT_PACKED_NODES && !brute && !pf && !sc.packed);
    cudaError_t e = cudaMemsetAsync(queue, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    const bool tex = sc.c3 != nullptr && sc.n_tex != 0u;       // texture code only when the scene carries handles and a table is bound
    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * (fold ? 1u : lp.n_frames) * 32ull;
    const uint64_t per_grab = fold ? 32ull : (uint64_t)lp.queue_chunk;
    const uint64_t warps_needed = (n_items + per_grab - 1) / per_grab;
    const bool ord = lp.tile_order != nullptr || lp.tile_cost != nullptr;
    uint64_t blocks = (uint64_t)n_sms * trace_blocks_per_sm(brute, detail, tex, pf, fold, inst, ord);
    const uint64_t cap = (warps_needed + RT_BLOCK_THREADS / 32 - 1) / (RT_BLOCK_THREADS / 32);
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks), block(RT_BLOCK_THREADS);
    trace_variant(brute, detail, tex, pf, fold, inst, ord)<<<grid, block, 0, stream>>>(sc, cam, lp, staging, accum, primary_ids, counters, queue);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return (fold || !with_blend) ? cudaSuccess : launch_blend(lp, staging, accum, stream);
}

#ifdef RT_LANE_STATS
extern "C" int rt_debug_lane_stats(unsigned long long* out16, int reset)
{
    unsigned long long z[16] = {0};
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess && out16) e = cudaMemcpyFromSymbol(out16, g_lane_stats, sizeof z);
    if (e == cudaSuccess && reset) e = cudaMemcpyToSymbol(g_lane_stats, z, sizeof z);
    return e == cudaSuccess ? 0 : -1;
}
#endif

cudaError_t launch_blend(const LaunchParams& lp, const float4* staging, float4* accum, cudaStream_t stream)
{
    const size_t n_local = (size_t)lp.n_warp_tiles * 32u;
    if (n_local == 0 || lp.n_frames == 0) return cudaSuccess;
    blend_kernel<<<(unsigned)((n_local + 255) / 256), 256, 0, stream>>>(lp, staging, accum);
    return cudaGetLastError();
}

// ---- rt_render_frames_static: the trace launch with its scratch in a static device array ---------------------------
// For callers that may not allocate (rt_b200.h).  The scratch of a staged trace launch is the work-queue head and one
// W*H float4 plane per fused dispatch; here both live in one file-scope array that every GPU loading this library holds
// until the process exits: 256 bytes of header (queue head in the first 8) + RT_STATIC_SCRATCH_BYTES of planes.
// In host code `g_static_scratch` is only a host-side stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
// All calls share the array and it keeps the last call's contents: the queue head is cleared in stream order before
// every launch (the planes need no clearing: the kernel writes every slot the blend reads), and two calls that overlap
// on the GPU (different streams / contexts of one device) race on it.  (The array is defined at file scope, at the top.)

namespace {
const int kStaticMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kStaticMaxDevices];          // device address of g_static_scratch, per device ordinal

// `device` must be the current device of the calling thread.
cudaError_t static_scratch(int device, unsigned char** out)
{
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p || ((uintptr_t)p & 255u) != 0) return cudaErrorInvalidSymbol;
        g_static_ptr[device] = p;
    }
    *out = (unsigned char*)g_static_ptr[device];
    return cudaSuccess;
}
} // namespace

size_t static_scratch_bytes() { return RT_STATIC_SCRATCH_BYTES; }

// (pixel, dispatch) items, planes and queue head in the static array, then the blend: the same two kernels as
// launch_trace(fold = false) on other scratch.  No allocation, no event, no synchronisation.
cudaError_t launch_trace_static(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, float4* accum,
                                uint32_t* primary_ids, unsigned long long* counters, bool brute, bool detail,
                                int n_sms, int device, cudaStream_t stream)
{
    if (lp.n_warp_tiles == 0 || lp.n_frames == 0) return cudaSuccess;
    if ((uint64_t)lp.n_frames * lp.width * lp.height * sizeof(float4) > (uint64_t)RT_STATIC_SCRATCH_BYTES) return cudaErrorInvalidValue;
    if (sc.iprims != nullptr && brute) return cudaErrorNotSupported;
    const bool pf = !brute && sc.iprims == nullptr && !sc.packed && sc.n_prims <= RT_PF_MAX_PRIMS;
    const bool inst = sc.iprims != nullptr || (RT_PACKED_NODES && !brute && !pf && !sc.packed);
    unsigned char* base = nullptr;
    cudaError_t e = static_scratch(device, &base);
    if (e != cudaSuccess) return e;
    unsigned long long* queue = (unsigned long long*)base;
    float4* staging = (float4*)(base + 256);
    if ((e = cudaMemsetAsync(queue, 0, sizeof(unsigned long long), stream)) != cudaSuccess) return e;
    const bool tex = sc.c3 != nullptr && sc.n_tex != 0u;
    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull;
    const uint64_t warps_needed = (n_items + lp.queue_chunk - 1) / lp.queue_chunk;
    uint64_t blocks = (uint64_t)n_sms * trace_blocks_per_sm(brute, detail, tex, pf, false, inst, false);
    const uint64_t cap = (warps_needed + RT_BLOCK_THREADS / 32 - 1) / (RT_BLOCK_THREADS / 32);
    if (blocks > cap) blocks = cap;
    trace_variant(brute, detail, tex, pf, false, inst, false)<<<dim3((unsigned)blocks), dim3(RT_BLOCK_THREADS), 0, stream>>>(sc, cam, lp, staging, accum, primary_ids, counters, queue);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return launch_blend(lp, staging, accum, stream);
}

} // namespace rt
