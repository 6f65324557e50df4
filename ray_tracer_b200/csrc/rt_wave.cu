// rt_wave.cu — wavefront execution of the same path tracer (RT_OPT_KERNEL = 1).
//
// The default trace kernel (rt_trace.cu) keeps a path on one lane from camera ray to the last
// bounce; its cost is SIMT efficiency: ncu shows 22 of 32 lanes in the node loop and 18 in
// shading, because finished rays wait for a warp-wide SHADE section and hit / miss / new-path
// code diverge inside it.  This file runs the identical arithmetic decoupled:
//   * a pool of N path slots lives in global memory (ray, closest hit, radiance state, item
//     bookkeeping; 128 B per slot, 32-byte groups moved with 256-bit loads / stores);
//   * wave_trace: persistent warps, a lane takes the next queued ray the moment its own ends
//     (warp-aggregated fetch), so the node / leaf loops only ever wait for each other; it writes
//     the closest hit and appends the slot to the hit queue or the miss queue;
//   * wave_hit: one thread per hit — the shading of wgsl:236-273, every lane on the same branch;
//   * wave_miss: one thread per terminated path — sky, running total, frag store, next work item
//     from the global item counter, camera ray (wgsl:306-327);
//   * the three kernels repeat until no slot is alive; kernel arguments never change, counts live
//     on the device, and the host only polls the live count every few rounds.
// Each slot consumes its RNG stream in exactly the reference's order, so images, ids and ray
// counts are bit-identical to the megakernel (tests/test_gpu_configs.py).
//
// Build: -fmad=false (no implicit contraction), IEEE div/sqrt, no fast-math.
#include "rt_device.cuh"
#include "rt_kernels.h"
#include "rt_shade.cuh"

namespace rt {

#define WV_EMPTY 0x7fffffff
#define WV_SLOT_MASK 0x0fffffffu
#define WV_KIND_MISS 0u       // ray missed: sky, then the path is finished
#define WV_KIND_DONE 1u       // bounce limit reached after a hit: the path is finished
#define WV_KIND_FRESH 2u      // empty slot: take a work item

// counters on the device
enum { WC_CUR = 0, WC_NEXT = 1, WC_HIT = 2, WC_MISS = 3, WC_FETCH = 4, WC_ITEM = 5, WC_COUNT = 8 };

__device__ __forceinline__ void stg256(float4* p, float4 a, float4 b)
{
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
}
__device__ __forceinline__ void ld256(const float4* p, float4& a, float4& b)      // coherent (the pool is rewritten every round)
{
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p) : "memory");
}

// Every queue cursor is one address in L2, and same-address atomics retire at about one per clock
// for the whole chip, so nothing here takes a cursor per ray:
//  * the trace kernel's warps take input in chunks and reserve output space in chunks (chunk_push);
//    what a warp has not filled when it exits is padded with WV_NONE, which consumers skip;
//  * the one-thread-per-entry kernels reserve exactly, once per block and tile (block_reserve).
#define WV_NONE 0xffffffffu
#ifndef WV_CHUNK
#define WV_CHUNK 64ull
#endif

// all 32 lanes call; pos / end are warp-uniform cursors into the warp's current output chunk
__device__ __forceinline__ void chunk_push(bool want, uint32_t value, uint32_t* q, unsigned long long* counter,
                                           unsigned long long& pos, unsigned long long& end)
{
    const unsigned lane = threadIdx.x & 31u;
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (m == 0u) return;
    const uint32_t k = __popc(m), rank = __popc(m & ((1u << lane) - 1u));
    const uint32_t room = (uint32_t)(end - pos);
    if (k > room) {
        if (want && rank < room) q[pos + rank] = value;
        unsigned long long g = 0;
        if (lane == 0) g = atomicAdd(counter, (unsigned long long)WV_CHUNK);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (want && rank >= room) q[g + (rank - room)] = value;
        pos = g + (k - room); end = g + WV_CHUNK;
    } else {
        if (want) q[pos + rank] = value;
        pos += k;
    }
}
__device__ __forceinline__ void chunk_close(uint32_t* q, unsigned long long pos, unsigned long long end)
{
    for (unsigned long long i = pos + (threadIdx.x & 31u); i < end; i += 32ull) q[i] = WV_NONE;
}

// all threads of a 256-thread block call; returns this thread's index in [old counter, old counter + #want)
struct BlockScratch { uint32_t cnt[8]; unsigned long long base; };
__device__ __forceinline__ unsigned long long block_reserve(bool want, unsigned long long* counter, BlockScratch& sm)
{
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (lane == 0) sm.cnt[warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; w++) { const uint32_t c = sm.cnt[w]; sm.cnt[w] = t; t += c; }
        sm.base = t ? atomicAdd(counter, (unsigned long long)t) : 0ull;
    }
    __syncthreads();
    const unsigned long long r = sm.base + sm.cnt[warp] + __popc(m & ((1u << lane) - 1u));
    __syncthreads();
    return r;
}

// ---- trace: closest hit for every queued slot -------------------------------------------------
template <bool BRUTE, bool DETAIL>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_WAVE_MIN_BLOCKS)
wave_trace(SceneDev sc, WaveDev wv, int n_stay, int refill_min)
{
    const uint32_t lane = threadIdx.x & 31u;
    int stack[RT_STACK];
    stack[0] = WV_EMPTY;                             // sentinel under every traversal (sp starts at 1)
    const unsigned long long n_queued = wv.counts[WC_CUR];
    const uint32_t* __restrict__ qcur = wv.q_cur;
    uint32_t n_rays = 0, n_nodes = 0, n_tri = 0, n_sph = 0;

    bool active = false, exhausted = false;          // exhausted is warp-uniform
    unsigned long long in_next = 0, in_end = 0;      // warp-uniform: the warp's current input chunk
    unsigned long long hit_pos = 0, hit_end = 0, miss_pos = 0, miss_end = 0;   // ... and its output chunks
    uint32_t slot = 0;
    float3 ro = f3(0.f, 0.f, 0.f), rd = ro;
    RaySlab slab;
    slab.inv_xy = slab.noi_xy = slab.ainv_xy = slab.pi_xy = 0ull; slab.inv_z = slab.noi_z = slab.ainv_z = slab.pi_z = 0.f;
    const unsigned magic = node_magic();
    HitRec best; best.dst = RT_FLT_MAX; best.u = best.v = 0.f; best.id = RT_MISS;
    float cull = RT_FLT_MAX;                         // best.dst with slack (RT_CULL_SLACK, rt_kernels.h)
    int node = WV_EMPTY, sp = 1;
    uint32_t leaf_first = 0, leaf_count = 0;

    for (;;) {
        // ---- lanes without a ray take the next queued ones (the warp draws them in chunks) ----
        unsigned want = exhausted ? 0u : __ballot_sync(0xffffffffu, !active);
        while (want) {
            if (in_next == in_end) {
                unsigned long long g = 0;
                if (lane == 0) g = atomicAdd(wv.counts + WC_FETCH, (unsigned long long)WV_CHUNK);
                g = __shfl_sync(0xffffffffu, g, 0);
                if (g >= n_queued) { exhausted = true; break; }
                in_next = g; in_end = (g + WV_CHUNK < n_queued) ? g + WV_CHUNK : n_queued;
            }
            const uint32_t rank = __popc(want & ((1u << lane) - 1u));
            const unsigned long long avail = in_end - in_next;
            if (((want >> lane) & 1u) && rank < avail) {
                slot = qcur[in_next + rank];
                float4 a, b;
                ld256(wv.ray + 2 * (size_t)slot, a, b);
                ro = xyz(a); rd = xyz(b);
                best.dst = RT_FLT_MAX; best.u = 0.0f; best.v = 0.0f; best.id = RT_MISS; cull = RT_FLT_MAX;
                sp = 1; node = WV_EMPTY; leaf_count = 0; leaf_first = 0;
                if (BRUTE) leaf_count = sc.n_prims;
                else if (sc.root >= 0) {
                    node = sc.root;
                    if (sc.packed) slab_setup_packed(slab, ro, rd, sc.max_abs, sc.pad_scale, sc.grid_o, sc.grid_c);
                    else slab_setup(slab, ro, rd, sc.max_abs, sc.pad_scale);
                }
                active = true;
                n_rays++;
            }
            in_next += min((unsigned long long)__popc(want), avail);
            want = __ballot_sync(0xffffffffu, !active);
        }
        const int n_act = __popc(__ballot_sync(0xffffffffu, active));
        if (n_act == 0) break;
        const int f_min = exhausted ? 1 : min(refill_min, n_act);
        // ---- traverse until enough lanes have finished to make a refill worth its cost ----
        for (;;) {
            const int cnt0 = __popc(__ballot_sync(0xffffffffu, active && node >= 0 && node != WV_EMPTY));
            const int thr = max(1, min(n_stay, (cnt0 + 1) >> 1));
            while (__popc(__ballot_sync(0xffffffffu, active && node >= 0 && node != WV_EMPTY)) >= thr) {
                if (active && node >= 0 && node != WV_EMPTY) {
                    if (DETAIL) n_nodes++;
                    float ltn, ltf, rtn, rtf; int cl, cr;
                    if (sc.packed) node_test_packed(sc.nodes + 2 * (size_t)node, slab, cull, magic, ltn, ltf, rtn, rtf, cl, cr);
                    else node_test(sc.nodes + 4 * (size_t)node, slab, cull, ltn, ltf, rtn, rtf, cl, cr);
                    const bool hl = ltn <= ltf, hr = rtn <= rtf;
                    if (hl && hr) {
                        const bool left_first = ltn <= rtn;
                        stack[sp++] = left_first ? cr : cl;
                        node = left_first ? cl : cr;
                    } else if (hl) node = cl;
                    else if (hr) node = cr;
                    else node = stack[--sp];
                    if (node < 0 && leaf_count == 0u) {
                        const uint32_t leaf = (uint32_t)(~node);
                        leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                        node = stack[--sp];
                    }
                }
            }
            while (__ballot_sync(0xffffffffu, active && leaf_count != 0u)) {
                if (active && leaf_count != 0u) {
                    const float4* pr = sc.prims + 4 * (size_t)leaf_first;
                    float4 q0, q1;
                    ldg256(pr, q0, q1);
                    const uint32_t id = __float_as_uint(q0.w);
                    if (id < sc.n_spheres) {
                        if (DETAIL) n_sph++;
                        test_sphere(ro, rd, q0, q1, id, best);
                    } else {
                        if (DETAIL) n_tri++;
                        float det;
                        if (triangle_faces_ray(rd, xyz(q0), det)) {
                            float4 q2, q3;
                            ldg256(pr + 2, q2, q3);
                            test_triangle(ro, rd, xyz(q0), det, xyz(q1), f3(q1.w, q2.x, q2.y), f3(q2.z, q2.w, q3.x), id, best);
                        }
                    }
                    cull = best.dst * RT_CULL_SLACK;
                    leaf_first++;
                    if (--leaf_count == 0u && node < 0) {
                        const uint32_t leaf = (uint32_t)(~node);
                        leaf_first = leaf >> 3; leaf_count = (leaf & 7u) + 1u;
                        node = stack[--sp];
                    }
                }
            }
            if (__popc(__ballot_sync(0xffffffffu, active && node == WV_EMPTY)) >= f_min) break;
        }
        // ---- finished rays: store the closest hit, queue the slot for shading ----
        const bool fin = active && node == WV_EMPTY;
        if (fin) {
            wv.hit[slot] = make_float4(best.dst, best.u, best.v, __uint_as_float(best.id));
            active = false;
        }
        const bool is_hit = fin && best.id != RT_MISS;
        chunk_push(is_hit, slot, wv.q_hit, wv.counts + WC_HIT, hit_pos, hit_end);
        chunk_push(fin && !is_hit, slot | (WV_KIND_MISS << 28), wv.q_miss, wv.counts + WC_MISS, miss_pos, miss_end);
    }
    chunk_close(wv.q_hit, hit_pos, hit_end);
    chunk_close(wv.q_miss, miss_pos, miss_end);

    unsigned long long r = n_rays, nn = n_nodes, nt = n_tri, ns = n_sph;
    for (int o = 16; o > 0; o >>= 1) {
        r += __shfl_xor_sync(0xffffffffu, r, o);
        if (DETAIL) { nn += __shfl_xor_sync(0xffffffffu, nn, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); ns += __shfl_xor_sync(0xffffffffu, ns, o); }
    }
    if (lane == 0) {
        atomicAdd(wv.stats + 0, r);
        if (DETAIL) { atomicAdd(wv.stats + 2, nn); atomicAdd(wv.stats + 3, nt); atomicAdd(wv.stats + 4, ns); }
    }
}

// ---- hit shading: wgsl:236-273 for every slot in the hit queue ---------------------------------
template <bool DETAIL, bool TEX>
__global__ void __launch_bounds__(256)
wave_hit(SceneDev sc, WaveDev wv, LaunchParams lp, uint32_t* __restrict__ primary_ids)
{
    __shared__ BlockScratch sm;
    const unsigned long long n = wv.counts[WC_HIT];
    uint32_t n_hits = 0;
    for (unsigned long long tile = (unsigned long long)blockIdx.x * 256ull; tile < n; tile += (unsigned long long)gridDim.x * 256ull) {
        const unsigned long long i = tile + threadIdx.x;
        const uint32_t slot = (i < n) ? wv.q_hit[i] : WV_NONE;
        const bool valid = slot != WV_NONE;
        bool done = false;
        if (valid) {
            float4 ra, rb, inc, col;
            ld256(wv.ray + 2 * (size_t)slot, ra, rb);
            ld256(wv.rad + 2 * (size_t)slot, inc, col);
            float4 misc = wv.acc[2 * (size_t)slot + 1];
            const float4 h4 = wv.hit[slot];
            HitRec h; h.dst = h4.x; h.u = h4.y; h.v = h4.z; h.id = __float_as_uint(h4.w);
            float3 ro = xyz(ra), rd = xyz(rb);
            int32_t seg = __float_as_int(ra.w);
            uint32_t rng = __float_as_uint(misc.x);
            if (primary_ids != nullptr && __float_as_uint(misc.z) == 0u && __float_as_int(misc.w) == 0 && seg == 0)
                primary_ids[__float_as_uint(misc.y)] = h.id;
            shade_hit<TEX>(sc, h, ro, rd, rng, inc, col);
            seg++;
            if (DETAIL) n_hits++;
            misc.x = __uint_as_float(rng);
            stg256(wv.rad + 2 * (size_t)slot, inc, col);
            wv.acc[2 * (size_t)slot + 1] = misc;
            done = seg > lp.bounces;
            if (!done) stg256(wv.ray + 2 * (size_t)slot, make_float4(ro.x, ro.y, ro.z, __int_as_float(seg)), make_float4(rd.x, rd.y, rd.z, 0.f));
        }
        const unsigned long long at_next = block_reserve(valid && !done, wv.counts + WC_NEXT, sm);
        if (valid && !done) wv.q_next[at_next] = slot;
        const unsigned long long at_miss = block_reserve(valid && done, wv.counts + WC_MISS, sm);
        if (valid && done) wv.q_miss[at_miss] = slot | (WV_KIND_DONE << 28);
    }
    if (DETAIL) {
        unsigned long long c = n_hits;
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((threadIdx.x & 31u) == 0 && c) atomicAdd(wv.stats + 5, c);
    }
}

// ---- path end, frag store, next item, camera ray: wgsl:274-279, 306-327 ------------------------
__global__ void __launch_bounds__(256)
wave_miss(CameraDev cam, WaveDev wv, LaunchParams lp, float4* __restrict__ staging, uint32_t* __restrict__ primary_ids)
{
    __shared__ BlockScratch sm;
    const unsigned long long n = wv.counts[WC_MISS];
    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull;
    const size_t plane = (size_t)lp.width * lp.height;
    const float sw = (float)lp.width, sh = (float)lp.height, rppf = (float)lp.rpp;
    uint32_t n_samples = 0;
    for (unsigned long long tile = (unsigned long long)blockIdx.x * 256ull; tile < n; tile += (unsigned long long)gridDim.x * 256ull) {
        const unsigned long long i = tile + threadIdx.x;
        const uint32_t entry = (i < n) ? wv.q_miss[i] : WV_NONE;
        const uint32_t slot = entry & WV_SLOT_MASK, kind = entry >> 28;
        float4 inc = make_float4(0.f, 0.f, 0.f, 0.f), col = inc, tot = inc, misc = inc;
        bool need_ray = entry != WV_NONE;                 // this thread still owes its slot a ray (or the end of all work)
        bool has_item = need_ray && kind != WV_KIND_FRESH;
        if (has_item) {
            ld256(wv.rad + 2 * (size_t)slot, inc, col);
            ld256(wv.acc + 2 * (size_t)slot, tot, misc);
        }
        uint32_t rng = __float_as_uint(misc.x), pix = __float_as_uint(misc.y), fi = __float_as_uint(misc.z);
        int32_t j = __float_as_int(misc.w);
        if (has_item && kind == WV_KIND_MISS) {
            float4 ra, rb;
            ld256(wv.ray + 2 * (size_t)slot, ra, rb);
            if (primary_ids != nullptr && fi == 0 && j == 0 && __float_as_int(ra.w) == 0) primary_ids[pix] = RT_MISS;
            if (lp.toggle != 0) inc = inc + environment_light(xyz(rb)) * col;              // wgsl:274-279
        }
        if (has_item) { tot = tot + inc; j++; }                                            // wgsl:323
        bool alive = false;
        float3 ro = f3(0.f, 0.f, 0.f), rd = ro;
        for (;;) {                                                                         // block-uniform trip count
            if (need_ray && has_item && j > lp.rpp) {
                staging[(size_t)fi * plane + pix] = frag_of_total(lp, tot);                // wgsl:326
                has_item = false;
            }
            // next work items: same numbering as the megakernel (tile-major, dispatch-minor, 32 pixels at a time)
            const bool want = need_ray && !has_item;
            const unsigned long long g = block_reserve(want, wv.counts + WC_ITEM, sm);
            if (want) {
                if (g >= n_items) need_ray = false;                                        // no work left: the slot dies
                else {
                    uint32_t px, py;
                    if (decode_item<false>(lp, g, px, py, fi)) {                                  // else: outside the image, ask again
                        pix = py * lp.width + px;
                        rng = pixel_seed(px, py, lp.width, lp.height, lp.first_frame + (int32_t)(fi * lp.frame_world), lp.seed_mode);
                        j = 0; has_item = true;
                        tot = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (lp.rpp < 0) {                                                  // wgsl:312 loop never runs
                            staging[(size_t)fi * plane + pix] = make_float4(0.0f / rppf, 0.0f / rppf, 0.0f / rppf, 0.0f / rppf);
                            has_item = false;
                        }
                    }
                }
            }
            if (need_ray && has_item) {
                // camera ray, wgsl:313-321
                const uint32_t py = lp.div_width.div(pix), px = pix - py * lp.width;
                float aax = rand01(rng), aay = rand01(rng);
                float posx = ((float)px + aax) / sw, posy = ((float)py + aay) / sh;
                float rdx, rdy;
                for (;;) {                                   // rand_in_unit_disk, wgsl:216-227
                    float r1 = rand01(rng) * 2.0f, r2 = rand01(rng) * 2.0f;
                    rdx = r1 - 1.0f; rdy = r2 - 1.0f;
                    if (dot3(f3(rdx, rdy, 0.0f), f3(rdx, rdy, 0.0f)) <= 1.00000011920928955078125f) break;   // == length(p) <= 1.0, see rt_trace.cu
                }
                rdx = rdx * cam.lens_radius; rdy = rdy * cam.lens_radius;
                float3 offset = fma3(cam.v, rdy, cam.u * rdx);
                ro = cam.origin + offset;
                rd = fma3(cam.vertical, posy, fma3(cam.horizontal, posx, cam.llc)) - ro;
                inc = make_float4(0.f, 0.f, 0.f, 0.f);
                col = make_float4(1.f, 1.f, 1.f, 1.f);
                n_samples++;
                if (lp.bounces < 0) { tot = tot + inc; j++; }                              // wgsl:233 loop never runs
                else { need_ray = false; alive = true; }
            }
            if (!__syncthreads_or(need_ray)) break;
        }
        if (alive) {
            misc = make_float4(__uint_as_float(rng), __uint_as_float(pix), __uint_as_float(fi), __int_as_float(j));
            stg256(wv.ray + 2 * (size_t)slot, make_float4(ro.x, ro.y, ro.z, __int_as_float(0)), make_float4(rd.x, rd.y, rd.z, 0.f));
            stg256(wv.rad + 2 * (size_t)slot, inc, col);
            stg256(wv.acc + 2 * (size_t)slot, tot, misc);
        }
        const unsigned long long at = block_reserve(alive, wv.counts + WC_NEXT, sm);
        if (alive) wv.q_next[at] = slot;
    }
    unsigned long long s = n_samples;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31u) == 0 && s) atomicAdd(wv.stats + 1, s);
}

// round bookkeeping: next queue becomes current (the host swaps the two pointers to match)
__global__ void wave_advance(WaveDev wv)
{
    wv.counts[WC_CUR] = wv.counts[WC_NEXT];
    wv.counts[WC_NEXT] = 0; wv.counts[WC_HIT] = 0; wv.counts[WC_MISS] = 0; wv.counts[WC_FETCH] = 0;
}

__global__ void wave_seed(WaveDev wv, uint32_t n_slots)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_slots) wv.q_miss[i] = i | (WV_KIND_FRESH << 28);
    if (i == 0) {
        wv.counts[WC_CUR] = 0; wv.counts[WC_NEXT] = 0; wv.counts[WC_HIT] = 0; wv.counts[WC_MISS] = n_slots;
        wv.counts[WC_FETCH] = 0; wv.counts[WC_ITEM] = 0;
    }
}

__global__ void wave_flush_stats(WaveDev wv, unsigned long long* counters, bool detail)
{
    atomicAdd(counters + 0, wv.stats[0]); atomicAdd(counters + 1, wv.stats[1]);
    if (detail) { atomicAdd(counters + 2, wv.stats[2]); atomicAdd(counters + 3, wv.stats[3]); atomicAdd(counters + 4, wv.stats[4]); atomicAdd(counters + 5, wv.stats[5]); }
    for (int k = 0; k < 6; k++) wv.stats[k] = 0;
}

void WavePool::release()
{
    void* ps[] = {ray, hit, rad, acc, q[0], q[1], q_hit, q_miss, counts, stats};
    for (void* p : ps) if (p) cudaFree(p);
    ray = hit = rad = acc = nullptr; q[0] = q[1] = q_hit = q_miss = nullptr; counts = stats = nullptr; n_slots = 0;
    if (h_live) { cudaFreeHost(h_live); h_live = nullptr; }
}

cudaError_t WavePool::ensure(uint32_t slots)
{
    if (n_slots == slots) return cudaSuccess;
    release();
    cudaError_t e;
#define WV_TRY(x) do { if ((e = (x)) != cudaSuccess) { release(); return e; } } while (0)
    WV_TRY(cudaMalloc((void**)&ray, 2 * (size_t)slots * sizeof(float4)));
    WV_TRY(cudaMalloc((void**)&hit, (size_t)slots * sizeof(float4)));
    WV_TRY(cudaMalloc((void**)&rad, 2 * (size_t)slots * sizeof(float4)));
    WV_TRY(cudaMalloc((void**)&acc, 2 * (size_t)slots * sizeof(float4)));
    // q_hit / q_miss also hold the WV_NONE padding of the trace warps' last chunks (at most 16 blocks x 4 warps per SM)
    const size_t padded = (size_t)slots + 2 * (size_t)WV_CHUNK * 256 * 16 * 4;
    for (int k = 0; k < 2; k++) WV_TRY(cudaMalloc((void**)&q[k], (size_t)slots * sizeof(uint32_t)));
    WV_TRY(cudaMalloc((void**)&q_hit, padded * sizeof(uint32_t)));
    WV_TRY(cudaMalloc((void**)&q_miss, padded * sizeof(uint32_t)));
    WV_TRY(cudaMalloc((void**)&counts, WC_COUNT * sizeof(unsigned long long)));
    WV_TRY(cudaMalloc((void**)&stats, 8 * sizeof(unsigned long long)));
    WV_TRY(cudaMemset(stats, 0, 8 * sizeof(unsigned long long)));
    WV_TRY(cudaMallocHost((void**)&h_live, sizeof(unsigned long long)));
#undef WV_TRY
    n_slots = slots;
    return cudaSuccess;
}

// One batch of fused dispatches through the wavefront kernels (then the same blend kernel).
cudaError_t launch_wave(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, WavePool& pool, float4* staging,
                        uint32_t* primary_ids, unsigned long long* counters, bool brute, bool detail, int n_sms, cudaStream_t stream)
{
    if (lp.n_warp_tiles == 0 || lp.n_frames == 0) return cudaSuccess;
    const bool tex = sc.c3 != nullptr && sc.n_tex != 0u;
    const uint64_t n_items = (uint64_t)lp.n_warp_tiles * lp.n_frames * 32ull;
    const uint32_t live_slots = (uint32_t)(n_items < pool.n_slots ? n_items : pool.n_slots);
    WaveDev wv;
    wv.ray = pool.ray; wv.hit = pool.hit; wv.rad = pool.rad; wv.acc = pool.acc;
    wv.q_hit = pool.q_hit; wv.q_miss = pool.q_miss; wv.counts = pool.counts; wv.stats = pool.stats;
    int cur = 0;
    wv.q_cur = pool.q[cur]; wv.q_next = pool.q[cur ^ 1];
    wave_seed<<<(live_slots + 255) / 256, 256, 0, stream>>>(wv, live_slots);

    int tb = 0;
    if (brute) { if (detail) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tb, wave_trace<true, true>, RT_BLOCK_THREADS, 0); else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tb, wave_trace<true, false>, RT_BLOCK_THREADS, 0); }
    else { if (detail) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tb, wave_trace<false, true>, RT_BLOCK_THREADS, 0); else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tb, wave_trace<false, false>, RT_BLOCK_THREADS, 0); }
    if (tb < 1) tb = 1;
    const unsigned trace_blocks = (unsigned)(n_sms * tb), shade_blocks = (unsigned)(n_sms * 8);
    const int n_stay = lp.sched_n_stay, refill_min = lp.sched_l_stay < 1 ? 1 : lp.sched_l_stay;

    cudaError_t e;
    for (uint64_t round = 0;; round++) {
        // miss / fresh entries -> path ends, new items, camera rays (fills q_next); first round only seeds
        wave_miss<<<shade_blocks, 256, 0, stream>>>(cam, wv, lp, staging, primary_ids);
        wave_advance<<<1, 1, 0, stream>>>(wv);
        cur ^= 1; wv.q_cur = pool.q[cur]; wv.q_next = pool.q[cur ^ 1];
        if ((round & 7u) == 7u) {                     // poll the live count now and then
            if ((e = cudaMemcpyAsync(pool.h_live, pool.counts + WC_CUR, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
            if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
            if (*pool.h_live == 0) break;
        }
        if (brute) { if (detail) wave_trace<true, true><<<trace_blocks, RT_BLOCK_THREADS, 0, stream>>>(sc, wv, n_stay, refill_min); else wave_trace<true, false><<<trace_blocks, RT_BLOCK_THREADS, 0, stream>>>(sc, wv, n_stay, refill_min); }
        else { if (detail) wave_trace<false, true><<<trace_blocks, RT_BLOCK_THREADS, 0, stream>>>(sc, wv, n_stay, refill_min); else wave_trace<false, false><<<trace_blocks, RT_BLOCK_THREADS, 0, stream>>>(sc, wv, n_stay, refill_min); }
        if (detail) { if (tex) wave_hit<true, true><<<shade_blocks, 256, 0, stream>>>(sc, wv, lp, primary_ids); else wave_hit<true, false><<<shade_blocks, 256, 0, stream>>>(sc, wv, lp, primary_ids); }
        else { if (tex) wave_hit<false, true><<<shade_blocks, 256, 0, stream>>>(sc, wv, lp, primary_ids); else wave_hit<false, false><<<shade_blocks, 256, 0, stream>>>(sc, wv, lp, primary_ids); }
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    wave_flush_stats<<<1, 1, 0, stream>>>(wv, counters, detail);
    return cudaGetLastError();
}

} // namespace rt
