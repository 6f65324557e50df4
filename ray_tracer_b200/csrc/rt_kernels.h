// rt_kernels.h — host-callable launchers of the CUDA kernels (internal to librt_b200.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rt_device.cuh"
#include "../../include/rt_b200.h"

#define RT_BLOCK_THREADS 128
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 8            // resident blocks per SM the trace kernel is compiled for (64 registers per thread)
#endif
#define RT_STACK 64                 // traversal stack entries per lane (BVH depth limit checked at upload)


// Subtrees are skipped when the ray enters their (padded) box beyond best.dst * RT_CULL_SLACK, not beyond
// best.dst itself.  The reference's computed distances carry rounding error that grows as a ray grazes a
// primitive: a sphere's dst = (-b - sqrt(disc)) / 2a is off by up to 2^-12 relative when disc ~ 0 (the square
// root of the rounding error of b*b - 4ac), a triangle's dst = dot(ao, n) / det by 2^-23 / sin(grazing angle).
// So a primitive's computed distance can be smaller than the distance at which the ray enters its box; in the
// reference's exhaustive scan it still wins, here it would be skipped whenever a slightly farther hit was
// found first -- and whether it was depends on when a parked leaf gets tested, i.e. on the warp schedule.
// Seen on C5 at 4K with the radix-tree builder: 2 pixels of 8.3 M changed from render to render
// (scripts/order_probe.py); with the slack 30 renders and 12 rebuilds agree and the 4K image equals exhaustive
// mode at every pixel (scripts/brute_check.py).  2^-9 covers the sphere bound 8 times over.
#ifndef RT_CULL_SLACK
#define RT_CULL_SLACK 1.001953125f  // 1 + 2^-9 (2^-7 costs 2.7 % of the rays per second on C3, 2^-9 and 2^-11 0.6 %)
#endif

namespace rt {

#define RT_PF_MAX_PRIMS 262144u      // scenes up to this many primitives: push-time node prefetch, 64-byte float nodes; above: packed 32-byte nodes
#ifndef RT_PACKED_NODES
#define RT_PACKED_NODES 1            // 0 (tuning builds): 64-byte float nodes for flat scenes too
#endif
#ifndef RT_QUEUE_CHUNK
#define RT_QUEUE_CHUNK 256ull     // work items a warp takes from the global queue per atomic (while plenty remain): 8 dispatches of one 8x4 tile.
                                  // Fewer tiles in flight per SM = warmer L1: 256 against 128 C3 +1.9 %, C4 +2.6 %, C2 +5.5 %, C5 +0.2 %; 512 C3 -5 %; 32 C3 -9 %
#endif
#ifndef RT_QUEUE_GUIDE
#define RT_QUEUE_GUIDE 20ull      // grab size = remaining items / (warps x this), clamped to [32, RT_QUEUE_CHUNK]; scenes of at least 2,048 primitives
                                  // use twice this (rt_api.cu: LaunchParams.queue_guide)
#endif

// trace (persistent warps over the item queue).  fold = true: items are pixels, the kernel applies the running mean
// to `accum` itself (staging unused, may be null); fold = false: items are (pixel, dispatch) pairs written to
// `staging` (lp.n_frames planes of W*H float4) and blend_kernel folds them into `accum`.
cudaError_t launch_trace(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, float4* staging, float4* accum,
                         uint32_t* primary_ids, unsigned long long* counters, unsigned long long* queue,
                         bool brute, bool detail, bool fold, int n_sms, cudaStream_t stream, bool with_blend = true);

// The staged trace launch + blend with the scratch (queue head, staging planes) in a static device array of
// RT_STATIC_SCRATCH_BYTES (rt_trace.cu): nothing is allocated.  `device` must be current.  cudaErrorInvalidValue when
// lp.n_frames planes of W*H float4 exceed the array.
#ifndef RT_STATIC_SCRATCH_BYTES
#define RT_STATIC_SCRATCH_BYTES (32u << 20)      // == rt_b200.h
#endif
size_t static_scratch_bytes();
cudaError_t launch_trace_static(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, float4* accum,
                                uint32_t* primary_ids, unsigned long long* counters, bool brute, bool detail,
                                int n_sms, int device, cudaStream_t stream);

// Device-resident scene: SoA primitive arrays + LBVH, built on the GPU from the AoS records.
struct SceneBuffers {
    float4* prims = nullptr;
    float4 *c0 = nullptr, *c1 = nullptr, *c2 = nullptr;
    float4 *c3 = nullptr, *c4 = nullptr, *c5 = nullptr;   // texture extension (uvs + edges), only when has_tex
    bool has_tex = false;
    cudaStream_t stream = nullptr;  // the arrays are stream-ordered allocations on this stream
    cudaMemPool_t pool = nullptr;   // ... out of the context's private pool (nullptr: the device's default pool)
    // Build state kept for refit_scene (rt_update_records): the device copies of the caller's records,
    // the id-order hot records and bounds, the Morton-order permutation and the hierarchy's topology.
    struct Keep {
        rt_sphere* sph = nullptr; rt_vertex* vtx = nullptr; uint32_t* idx = nullptr; rt_mesh* mesh = nullptr; uint32_t* tstart = nullptr;
        float4 *hot = nullptr, *bmin = nullptr, *bmax = nullptr, *lmin = nullptr, *lmax = nullptr, *imin = nullptr, *imax = nullptr;
        int* bounds = nullptr; uint32_t *ids2 = nullptr, *new_slot = nullptr;
        int2 *children = nullptr, *ranges = nullptr; int *pin = nullptr, *plf = nullptr, *flags = nullptr;
        uint32_t *flat_start = nullptr, *inst_mesh = nullptr; int* range_of_mesh = nullptr; float4* range_info = nullptr;
        uint32_t nm = 0, leaf_size = 0; int parent_shift = 0;
    } keep;
    float4 *m0 = nullptr, *m1 = nullptr, *m2 = nullptr;
    float4* nodes = nullptr;
    // Instancing: Mesh records that alias one (first, triangles, offset) range share ONE object-space hierarchy; `prims` then
    // holds a proxy per such record, `iprims` the ranges' raw vertex triples in tree order (4 float4 per triangle:
    // (a.xyz, b.x) (b.yz, c.xy) (c.z, triangle index, 0, 0) (unused)), `nodes` the top level followed by the ranges' nodes.
    float4* iprims = nullptr;
    uint32_t n_inst = 0, n_iprims = 0, n_cold = 0, n_top_nodes = 0;
    float bounds_lo[3] = {0, 0, 0}, bounds_hi[3] = {0, 0, 0};       // exact bounds of the top level
    bool packed = false;                                            // `nodes` holds packed 32-byte nodes (flat scenes with finite bounds; node_test_packed); else 64-byte float nodes
    float grid_o[3] = {0, 0, 0}, grid_c[3] = {1, 1, 1};             // their 16-bit grid: origin and cell size per axis
    uint32_t n_prims = 0, n_spheres = 0, n_tris = 0, n_mats = 0, n_nodes = 0, max_depth = 0;
    int32_t root = -1;
    int builder = 0;               // hierarchy actually built: 0 = PLOC, 1 = radix tree
    float max_abs = 0.0f;
    void release();
};

// Repack + LBVH build. All pointers are host pointers to the caller's records (already
// validated). Synchronous with respect to `stream` on return.
cudaError_t build_scene(SceneBuffers& out, const rt_sphere* spheres, uint32_t ns, const rt_vertex* vertices,
                        uint32_t nv, const uint32_t* indices, uint32_t ni, const rt_mesh* meshes, uint32_t nm,
                        uint32_t leaf_size, int builder, int ploc_radius, int instancing /*0 never, 1 when eligible, 2 auto*/, int node_format /*0 by size, 1 float, 2 packed (flat scenes)*/, cudaStream_t stream);

// Same topology, new positions / radii / materials: re-derives the records and bounds from the
// updated spheres and meshes (index ranges must be unchanged) and refits the hierarchy's boxes
// bottom-up; no sort, no hierarchy build. Synchronous with respect to `stream` on return.
cudaError_t refit_scene(SceneBuffers& sc, const rt_sphere* spheres, const rt_mesh* meshes, cudaStream_t stream);

// ---- wavefront execution (rt_wave.cu), RT_OPT_KERNEL = 1 ---------------------------------------
#ifndef RT_WAVE_MIN_BLOCKS
#define RT_WAVE_MIN_BLOCKS 10      // resident blocks per SM of the wavefront trace kernel (48 registers per thread)
#endif
struct WaveDev {                   // path-slot pool in HBM, 32-byte groups
    float4* ray;                   // [2s] = (origin, segment index), [2s+1] = (direction, 0)
    float4* hit;                   // [s]  = (distance, u, v, id)
    float4* rad;                   // [2s] = incoming light, [2s+1] = ray colour
    float4* acc;                   // [2s] = item total, [2s+1] = (rng, pixel, dispatch, path index)
    uint32_t *q_cur, *q_next, *q_hit, *q_miss;
    unsigned long long* counts;    // queue lengths, fetch cursor, item cursor
    unsigned long long* stats;     // rays, samples, node visits, tri tests, sphere tests, hits
};
struct WavePool {
    float4 *ray = nullptr, *hit = nullptr, *rad = nullptr, *acc = nullptr;
    uint32_t *q[2] = {nullptr, nullptr}, *q_hit = nullptr, *q_miss = nullptr;
    unsigned long long *counts = nullptr, *stats = nullptr, *h_live = nullptr;
    uint32_t n_slots = 0;
    cudaError_t ensure(uint32_t slots);
    void release();
};
// Same result as launch_trace (staging planes, counters, primary ids); the caller runs launch_blend next.
cudaError_t launch_wave(const SceneDev& sc, const CameraDev& cam, const LaunchParams& lp, WavePool& pool, float4* staging,
                        uint32_t* primary_ids, unsigned long long* counters, bool brute, bool detail, int n_sms, cudaStream_t stream);
cudaError_t launch_blend(const LaunchParams& lp, const float4* staging, float4* accum, cudaStream_t stream);

// Heaviest tiles first (rt_order.cu): per-tile weights collected by the trace kernel, sorted into the next call's tile order.
struct TileOrder {
    uint32_t *cost = nullptr, *order[2] = {nullptr, nullptr}, *keys = nullptr, *keys_out = nullptr, *vals = nullptr, *stats = nullptr;
    void* tmp = nullptr; size_t tmp_bytes = 0;
    uint32_t n = 0; uint64_t geom = 0; int cur = 0; bool valid = false;
    bool fed = false;                                   // a launch since the last update collected weights
    uint32_t* h_flat = nullptr; cudaEvent_t ev_flat = nullptr; bool flat_pending = false;
    bool disabled = false;                              // the weights turned out flat: raster order, nothing collected (until rearm())
    void poll();
    void rearm() { disabled = false; }
    cudaError_t ensure(uint32_t n_local, uint64_t geometry, cudaStream_t stream);   // (re)allocates for this rank's tile list; a new geometry starts in raster order
    cudaError_t update(cudaStream_t stream);
    void release();
    const uint32_t* current() const { return valid ? order[cur] : nullptr; }
};

cudaError_t launch_tonemap(const float4* accum, uint8_t* rgba8, size_t n_pixels, cudaStream_t stream);

// ---- multi-GPU image combine over peer memory (rt_combine.cu) -----------------------------------
#define RT_MAX_PEERS 16
struct PeerSrc { const float4* accum; uint32_t tile_rank, tile_world; float weight; };
struct PeerSet { PeerSrc p[RT_MAX_PEERS]; uint32_t n; };
// out (RGBA32F, may alias p[0].accum) or, when rgba8 != nullptr, the tone-mapped image (out is then only read
// for pixels no shard owns); runs on `stream` of the device that owns out / rgba8.
cudaError_t launch_combine(const PeerSet& ps, float4* out, uint8_t* rgba8, uint32_t width, uint32_t height, int n_sms, cudaStream_t stream);

// ---- multi-process image combine over NCCL (rt_comm.cu) ------------------------------------------
struct ShardRec { uint32_t tile_rank, tile_world, frame_rank, frame_world, count, pad[3]; };   // count: frames in the rank's running mean
struct Comm;
const char* comm_unique_id(void* id128);                                   // nullptr on success, else the error text
const char* comm_create(Comm** out, int n_ranks, int rank, const void* id128);
void comm_destroy(Comm* c);
int comm_rank(const Comm* c);
int comm_size(const Comm* c);
const char* comm_combine(Comm* c, const ShardRec& mine, int root, const float4* accum, float4* out, uint8_t* out8,
                         uint32_t width, uint32_t height, int n_sms, cudaStream_t stream);

} // namespace rt
