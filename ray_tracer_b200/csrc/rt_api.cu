// rt_api.cu — the C ABI of include/rt_b200.h over the CUDA kernels.
// Every entry point cites (in the header) the wgpu call of the reference it replaces.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <set>
#include <string>
#include <tuple>
#include <vector>
#include "rt_kernels.h"

static_assert(sizeof(rt_params) == 28, "Params is 7 x 4 B (context.rs:11-21)");
static_assert(sizeof(rt_camera_uniform) == 128 && offsetof(rt_camera_uniform, lower_left_corner) == 16 &&
              offsetof(rt_camera_uniform, horizontal) == 32 && offsetof(rt_camera_uniform, vertical) == 48 &&
              offsetof(rt_camera_uniform, w) == 80 && offsetof(rt_camera_uniform, u) == 96 &&
              offsetof(rt_camera_uniform, v) == 112 && offsetof(rt_camera_uniform, lens_radius) == 124,
              "CameraUniform layout (camera.rs:9-29)");
static_assert(sizeof(rt_sphere) == 64 && offsetof(rt_sphere, radius) == 12 && offsetof(rt_sphere, color) == 16 &&
              offsetof(rt_sphere, emission_color) == 32 && offsetof(rt_sphere, emission_strength) == 48 &&
              offsetof(rt_sphere, smoothness) == 52, "Sphere layout (scene.rs:11-21)");
static_assert(offsetof(rt_mesh, diffuse_tex) == 12 && offsetof(rt_mesh, normal_tex) == 72 && offsetof(rt_vertex, tu) == 12 &&
              offsetof(rt_vertex, tv) == 28, "texture extension lives in the padding words");
static_assert(sizeof(rt_mesh) == 80 && offsetof(rt_mesh, pos) == 16 && offsetof(rt_mesh, color) == 32 &&
              offsetof(rt_mesh, emission_color) == 48 && offsetof(rt_mesh, emission_strength) == 64 &&
              offsetof(rt_mesh, specular) == 68, "Mesh layout (scene.rs:23-37)");
static_assert(sizeof(rt_vertex) == 32 && offsetof(rt_vertex, normal) == 16, "Vertex layout (scene.rs:79-86)");

struct rt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t width = 0, height = 0;
    float4* accum = nullptr;
    uint32_t* primary_ids = nullptr;
    unsigned long long* d_counters = nullptr;     // rays, samples, node visits, tri/sphere tests, hit segments
    unsigned long long* d_queue = nullptr;        // work-queue head of the persistent trace kernel
    float4* combined = nullptr;                   // result of rt_combine / rt_combine_peers (allocated at the first combine; accum is never overwritten by a combine)
    uint8_t* d_rgba8 = nullptr;                   // tone-mapped image of rt_read_rgba8 / the fused combines (kept between calls)
    uint32_t mean_count = 0;                      // dispatches in accum's running mean (the combine weight of a sample shard)
    cudaMemPool_t mempool = nullptr;              // private pool of the stream-ordered scene allocations
    rt::Comm* comm = nullptr;                     // NCCL communicator of the multi-process combine (rt_comm_init)
    float4* staging = nullptr;                    // per-dispatch frag planes awaiting the blend
    uint32_t staging_frames = 0;
    // a render call that needs several launches alternates them between two streams (two halves of the staging planes, two
    // queue heads): the next launch's blocks fill the SMs the previous one's drain leaves idle, so short launches cost no tail
    cudaStream_t stream2 = nullptr;
    unsigned long long* d_queue2 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_blend[2] = {nullptr, nullptr};
    int n_sms = 148;
    int64_t opt_staging_bytes = 4352ll << 20;     // staging planes (RT_OPT_STAGING_MIB): 128 dispatches per launch at 1080p, 32 at 4K
    int64_t opt_tile_order = 1;                   // 1: the tiles that were heaviest in the previous call enter the queue first (RT_OPT_TILE_ORDER)
    rt::TileOrder tile_order;
    int64_t opt_node_format = 0;                  // 0: by scene size, 1: float nodes, 2: packed nodes (RT_OPT_NODE_FORMAT)
    int64_t opt_pipeline = 1;                     // 1: the launches of one render call overlap on two streams (RT_OPT_PIPELINE)
    int64_t opt_instancing = 2;                   // Mesh records that alias one index range share one object-space hierarchy (two-level BVH): 0 never, 1 when eligible, 2 auto
    int64_t opt_item_mode = 0;                    // 0: auto, 1: (pixel, dispatch) items + staging + blend kernel, 2: pixel items, running mean applied in the trace kernel
    int64_t opt_sched = 8 | (1 << 8) | (22 << 16);     // node loop runs while >= 8 lanes can step; shade once >= 22 lanes wait (swept on C3/C5)
    rt::SceneBuffers scene;
    std::vector<rt_mesh> mesh_ranges;             // the uploaded Mesh records (rt_update_records checks the index ranges against them)
    struct Tex { cudaArray_t arr; cudaTextureObject_t obj; uint32_t w, h; };
    std::vector<Tex> textures;                    // texture extension: CUDA arrays + texture objects
    rt::TexDev* d_tex = nullptr;                  // device table of (object, width, height)
    bool has_scene = false, has_camera = false, has_params = false;
    bool scene_too_deep = false;                  // BVH deeper than the traversal stack: launches use the exhaustive scan
    rt_params params{};
    rt_camera_uniform camera{};
    uint32_t tile_rank = 0, tile_world = 1, frame_rank = 0, frame_world = 1;
    int64_t opt_traversal = 0, opt_seed_mode = 0, opt_primary_ids = 0, opt_detail = 0, opt_frames_per_launch = 0, opt_leaf_size = 0, opt_builder = 0, opt_ploc_radius = 0, opt_kernel = 0, opt_wave_slots = 1 << 21;
    rt::WavePool wave;                            // path-slot pool of the wavefront kernels (RT_OPT_KERNEL = 1)
    uint64_t dispatches = 0, launches = 0;
    double kernel_ms = 0.0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending, pool;
    std::string err;
};

static thread_local std::string g_create_err;

namespace {

int fail(rt_ctx* c, int code, const std::string& msg) { if (c) c->err = msg; else g_create_err = msg; return code; }
int cuda_fail(rt_ctx* c, cudaError_t e, const char* what)
{
    return fail(c, e == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define CU(c, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail((c), e_, #x); } while (0)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// New image buffers are allocated first and swapped in on success: a failed rt_resize leaves the context as it was.
int alloc_image(rt_ctx* c, uint32_t w, uint32_t h)
{
    const size_t n = (size_t)w * h;
    float4* accum = nullptr; uint32_t* ids = nullptr;
    cudaError_t e = cudaMalloc((void**)&accum, n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc((void**)&ids, n * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemsetAsync(accum, 0, n * sizeof(float4), c->stream);          // wgpu zero-initialises textures
    if (e == cudaSuccess) e = cudaMemsetAsync(ids, 0xFF, n * sizeof(uint32_t), c->stream);
    if (e != cudaSuccess) { if (accum) cudaFree(accum); if (ids) cudaFree(ids); return cuda_fail(c, e, "image allocation"); }
    if (c->accum) cudaFree(c->accum);
    if (c->primary_ids) cudaFree(c->primary_ids);
    if (c->staging) cudaFree(c->staging);
    if (c->combined) cudaFree(c->combined);
    if (c->d_rgba8) cudaFree(c->d_rgba8);
    c->accum = accum; c->primary_ids = ids; c->staging = nullptr; c->staging_frames = 0; c->combined = nullptr; c->d_rgba8 = nullptr;
    c->width = w; c->height = h; c->mean_count = 0;
    return RT_OK;
}

int ensure_combined(rt_ctx* c)
{
    if (c->combined) return RT_OK;
    const size_t n = (size_t)c->width * c->height;
    CU(c, cudaMalloc((void**)&c->combined, n * sizeof(float4)));
    CU(c, cudaMemsetAsync(c->combined, 0, n * sizeof(float4), c->stream));
    return RT_OK;
}

int ensure_rgba8(rt_ctx* c)
{
    if (c->d_rgba8) return RT_OK;
    CU(c, cudaMalloc((void**)&c->d_rgba8, (size_t)c->width * c->height * 4));
    return RT_OK;
}

void drain_events(rt_ctx* c)
{
    for (auto& p : c->pending) {
        float ms = 0.f;
        if (cudaEventSynchronize(p.second) == cudaSuccess && cudaEventElapsedTime(&ms, p.first, p.second) == cudaSuccess) c->kernel_ms += ms;
        c->pool.push_back(p);
    }
    c->pending.clear();
}

rt::CameraDev to_dev(const rt_camera_uniform& u)
{
    rt::CameraDev d;
    d.origin = make_float3(u.origin[0], u.origin[1], u.origin[2]);
    d.llc = make_float3(u.lower_left_corner[0], u.lower_left_corner[1], u.lower_left_corner[2]);
    d.horizontal = make_float3(u.horizontal[0], u.horizontal[1], u.horizontal[2]);
    d.vertical = make_float3(u.vertical[0], u.vertical[1], u.vertical[2]);
    d.u = make_float3(u.u[0], u.u[1], u.u[2]);
    d.v = make_float3(u.v[0], u.v[1], u.v[2]);
    d.lens_radius = u.lens_radius;
    return d;
}

// Pixel items (one lane owns a pixel for every fused dispatch and applies wgsl:59-66 itself: no staging, no blend
// kernel) whenever asked for, and in auto mode for single-dispatch launches, where the two item kinds coincide.
// For fused launches auto mode keeps (pixel, dispatch) items: a 1080p teapot pixel of 128 dispatches is ~18 ms of one
// lane's time and a lane gets only ~3 of them, so pixel items balance badly (measured: C3 3,985 vs 9,132 Mrays/s,
// C4 / C5 / C2 -10 %); the running mean is order-dependent per pixel, so finer items need the staged frags.
bool fold_mode(const rt_ctx* c, uint32_t n_frames) { return c->opt_kernel == 0 && (c->opt_item_mode == 2 || (c->opt_item_mode == 0 && n_frames == 1)); }

// Dispatches fused per launch: 128 by default, bounded by the staging budget (one W*H float4 plane per dispatch);
// forced pixel items need no staging: up to 4096 per launch unless the option says otherwise.
uint32_t frames_per_launch(const rt_ctx* c)
{
    int64_t f = c->opt_frames_per_launch;
    if (c->opt_kernel == 0 && c->opt_item_mode == 2) return (uint32_t)(f < 1 ? 4096 : f);
    if (f < 1) f = 128;
    int64_t by_mem = c->opt_staging_bytes / (int64_t)((size_t)c->width * c->height * sizeof(float4));
    if (f > by_mem) f = by_mem;
    return (uint32_t)(f < 1 ? 1 : f);
}

int ensure_staging(rt_ctx* c, uint32_t frames)
{
    if (c->staging_frames >= frames) return RT_OK;
    if (c->staging) { CU(c, cudaStreamSynchronize(c->stream)); cudaFree(c->staging); c->staging = nullptr; c->staging_frames = 0; }
    CU(c, cudaMalloc((void**)&c->staging, (size_t)frames * c->width * c->height * sizeof(float4)));
    c->staging_frames = frames;
    return RT_OK;
}

// Launch `n_frames` fused dispatches starting at `first_frame` with stride c->frame_world.  slot < 0: a launch of its own on
// the context's stream, timed by an event pair.  slot 0 / 1: one of a pipelined series (rt_render_frames): stream and staging
// half of that slot, the blend ordered after the previous launch's blend on the other stream; `piped_per` planes per half.
int launch(rt_ctx* c, int32_t first_frame, uint32_t n_frames, bool want_ids, int slot = -1, uint32_t piped_per = 0, bool after_blend = false)
{
    rt::SceneDev sd{};
    const rt::SceneBuffers& s = c->scene;
    sd.prims = s.prims; sd.iprims = s.iprims; sd.c0 = s.c0; sd.c1 = s.c1; sd.c2 = s.c2; sd.c3 = s.c3; sd.c4 = s.c4; sd.c5 = s.c5;
    sd.tex = c->d_tex; sd.n_tex = (uint32_t)c->textures.size();
    sd.m0 = s.m0; sd.m1 = s.m1; sd.m2 = s.m2; sd.nodes = s.nodes;
    sd.n_prims = s.n_prims; sd.n_spheres = s.n_spheres; sd.root = s.root;
    sd.max_abs = s.max_abs; sd.pad_scale = 1.52587890625e-05f;   // 2^-16
    sd.packed = s.packed ? 1 : 0; sd.grid_o = make_float3(s.grid_o[0], s.grid_o[1], s.grid_o[2]); sd.grid_c = make_float3(s.grid_c[0], s.grid_c[1], s.grid_c[2]);
    rt::LaunchParams lp{};
    lp.width = c->width; lp.height = c->height;
    lp.bounces = c->params.number_of_bounces; lp.rpp = c->params.rays_per_pixel; lp.toggle = c->params.skybox;
    lp.first_frame = first_frame; lp.n_frames = n_frames;
    lp.blend0 = first_frame >= 0 ? (int32_t)((uint32_t)first_frame / c->frame_world) : first_frame;
    lp.tile_rank = c->tile_rank; lp.tile_world = c->tile_world; lp.frame_world = c->frame_world;
    lp.tiles_x = (c->width + 7) / 8;
    const uint32_t tiles = lp.tiles_x * ((c->height + 7) / 8);
    lp.n_warp_tiles = 2u * (tiles > c->tile_rank ? (tiles - c->tile_rank + c->tile_world - 1) / c->tile_world : 0);
    lp.seed_mode = (int32_t)c->opt_seed_mode;
    // Guide of the item queue (grab = remaining / (warps x guide)): scenes with real geometry end their launches on finer grabs — their
    // items are long enough to hide a grab's round trip and differ enough in cost for the last grabs to matter (C3 x 32: 18.5 instead of
    // 21.7 ms, x 16 -14 %, x 8 -6 %, x 128 +0.2 %; C4 x 8 -2.4 %, x 32 +0.2 %); a handful of primitives (C2: 429, x 32 +3 %, x 16 +8 % with
    // the finer guide) keeps the coarse one.  scripts/queue_sweep.py, profiles/r2_summary.md.
    lp.queue_guide = (uint32_t)(c->scene.n_prims >= 2048u ? 2ull * RT_QUEUE_GUIDE : RT_QUEUE_GUIDE);
    lp.queue_chunk = (uint32_t)RT_QUEUE_CHUNK;
    if ((c->opt_sched >> 24) & 0xff) lp.queue_chunk = 32u * (uint32_t)((c->opt_sched >> 24) & 0xff);      // tuning: RT_OPT_SCHED bits 24-31 = grab size / 32, bits 32-39 = guide
    if ((c->opt_sched >> 32) & 0xff) lp.queue_guide = (uint32_t)((c->opt_sched >> 32) & 0xff);
    lp.div_frames = rt::make_fastdiv(n_frames); lp.div_tiles_x = rt::make_fastdiv(lp.tiles_x); lp.div_width = rt::make_fastdiv(c->width);
    lp.rpp_scale = (lp.rpp > 0 && (lp.rpp & (lp.rpp - 1)) == 0) ? 1.0f / (float)lp.rpp : 0.0f;
    lp.sched_n_stay = (int32_t)(c->opt_sched & 0xff); if (lp.sched_n_stay < 1) lp.sched_n_stay = 1; lp.sched_l_stay = (int32_t)((c->opt_sched >> 8) & 0xff); lp.sched_s_min = (int32_t)((c->opt_sched >> 16) & 0xff);
    if (c->opt_tile_order != 0 && c->opt_kernel == 0) {          // heaviest tiles first (rt_order.cu); a failure to allocate only means raster order
        const uint64_t geometry = ((uint64_t)c->width << 40) ^ ((uint64_t)c->height << 16) ^ ((uint64_t)c->tile_rank << 8) ^ (uint64_t)c->tile_world;
        if (c->stream2 && (c->tile_order.n != lp.n_warp_tiles / 2u || c->tile_order.geom != geometry)) cudaStreamSynchronize(c->stream2);
        c->tile_order.poll();
        if (c->tile_order.ensure(lp.n_warp_tiles / 2u, geometry, c->stream) == cudaSuccess && c->tile_order.n != 0 && !c->tile_order.disabled) {
            // the learned order pays where a launch's fixed cost dominates — the reference's own pattern of one dispatch per frame — and
            // costs a little (collecting the weights, some locality) where it does not.  Same-box, ms per dispatch, on / off: C3 x 1
            // 1.21 / 1.52, x 2 0.88 / 1.05, x 4 0.75 / 0.81, x 16 0.607 / 0.616, x 128 0.543 / 0.542; C4 at 4K x 1 0.84 / 0.90, x 8 0.734 / 0.724;
            // C1 x 8 0.152 / 0.149.  Hence: launches of at most 4 dispatches.
            if (n_frames <= 4u) { lp.tile_order = c->tile_order.current(); lp.tile_cost = c->tile_order.cost; c->tile_order.fed = true; }
        } else cudaGetLastError();
    }

    const bool fold = fold_mode(c, n_frames);
    if (slot >= 0) {
        cudaStream_t st = slot ? c->stream2 : c->stream;
        float4* stg = c->staging + (size_t)slot * piped_per * c->width * c->height;
        cudaError_t e = rt::launch_trace(sd, to_dev(c->camera), lp, stg, c->accum, want_ids ? c->primary_ids : nullptr, c->d_counters,
                                         slot ? c->d_queue2 : c->d_queue, c->opt_traversal == 1 || c->scene_too_deep, c->opt_detail != 0, false, c->n_sms, st, false);
        if (e == cudaSuccess && after_blend) e = cudaStreamWaitEvent(st, c->ev_blend[1 - slot], 0);     // wgsl:59-66 is order-dependent: blends run in launch order
        if (e == cudaSuccess) e = rt::launch_blend(lp, stg, c->accum, st);
        if (e == cudaSuccess) e = cudaEventRecord(c->ev_blend[slot], st);
        if (e != cudaSuccess) return cuda_fail(c, e, "pipelined trace launch");
        c->launches++;
        c->dispatches += n_frames;
        { const int64_t last = (int64_t)lp.blend0 + (int64_t)n_frames - 1; c->mean_count = last >= 0 ? (uint32_t)(last + 1) : 1u; }
        return RT_OK;
    }
    if (!fold) { int rc = ensure_staging(c, n_frames); if (rc != RT_OK) return rc; }
    std::pair<cudaEvent_t, cudaEvent_t> ev;
    if (!c->pool.empty()) { ev = c->pool.back(); c->pool.pop_back(); }
    else { CU(c, cudaEventCreate(&ev.first)); CU(c, cudaEventCreate(&ev.second)); }
    CU(c, cudaEventRecord(ev.first, c->stream));
    cudaError_t e;
    if (c->opt_kernel == 1) {
        if ((e = c->wave.ensure((uint32_t)c->opt_wave_slots)) != cudaSuccess) return cuda_fail(c, e, "wavefront pool");
        e = rt::launch_wave(sd, to_dev(c->camera), lp, c->wave, c->staging, want_ids ? c->primary_ids : nullptr, c->d_counters,
                            c->opt_traversal == 1 || c->scene_too_deep, c->opt_detail != 0, c->n_sms, c->stream);
        if (e == cudaSuccess) e = rt::launch_blend(lp, c->staging, c->accum, c->stream);
    } else {
        e = rt::launch_trace(sd, to_dev(c->camera), lp, c->staging, c->accum, want_ids ? c->primary_ids : nullptr,
                             c->d_counters, c->d_queue, c->opt_traversal == 1 || c->scene_too_deep, c->opt_detail != 0, fold, c->n_sms, c->stream);
    }
    if (e != cudaSuccess) return cuda_fail(c, e, "trace kernel launch");
    CU(c, cudaEventRecord(ev.second, c->stream));
    c->pending.push_back(ev);
    if (c->pending.size() > 4096) drain_events(c);
    c->launches++;
    c->dispatches += n_frames;
    { const int64_t last = (int64_t)lp.blend0 + (int64_t)n_frames - 1; c->mean_count = last >= 0 ? (uint32_t)(last + 1) : 1u; }   // wgsl:59-66: blend index b = mean of b + 1 dispatches
    return RT_OK;
}

// end of a render call: the tile weights its launches collected become the next call's order (on the context's stream, after every launch)
void finish_call(rt_ctx* c)
{
    if (c->opt_tile_order == 0 || c->opt_kernel != 0 || c->tile_order.n < 2) return;
    if (!c->tile_order.fed) return;                          // no launch of this call collected weights: the order stays
    c->tile_order.fed = false;
    if (c->tile_order.update(c->stream) != cudaSuccess) { cudaGetLastError(); c->tile_order.valid = false; }
}

void free_textures(rt_ctx* c)
{
    for (auto& t : c->textures) { if (t.obj) cudaDestroyTextureObject(t.obj); if (t.arr) cudaFreeArray(t.arr); }
    c->textures.clear();
    if (c->d_tex) { cudaFree(c->d_tex); c->d_tex = nullptr; }
}

int check_ready(rt_ctx* c)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!c->has_scene) return fail(c, RT_ERR_NO_SCENE, "no scene uploaded (rt_upload_scene)");
    if (!c->has_camera) return fail(c, RT_ERR_INVALID_ARGUMENT, "no camera set (rt_set_camera)");
    if (!c->has_params) return fail(c, RT_ERR_INVALID_ARGUMENT, "no params set (rt_set_params)");
    return RT_OK;
}

} // namespace

extern "C" {

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

uint32_t rt_version(void) { return (0u << 16) | 1u; }

int rt_create(rt_ctx** out, int device, uint32_t width, uint32_t height)
{
    if (!out) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_create: out is NULL");
    *out = nullptr;
    if (width == 0 || height == 0) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_create: zero-sized image");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceCount");
    if (device < 0 || device >= ndev) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_create: no such CUDA device");
    rt_ctx* c = new (std::nothrow) rt_ctx();
    if (!c) return fail(nullptr, RT_ERR_OUT_OF_MEMORY, "rt_create: host allocation failed");
    c->device = device;
    DeviceGuard g(device);
    int rc = RT_OK;
    do {
        if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cuda_fail(nullptr, e, "cudaStreamCreate"); break; }
        {   // scene uploads allocate stream-ordered from a pool of this context that keeps freed blocks cached for the
            // next upload; the device's default pool, which the host process may use itself, is not touched
            cudaMemPoolProps props; memset(&props, 0, sizeof props);
            props.allocType = cudaMemAllocationTypePinned; props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice; props.location.id = device;
            unsigned long long keep = ~0ull;
            if (cudaMemPoolCreate(&c->mempool, &props) == cudaSuccess) cudaMemPoolSetAttribute(c->mempool, cudaMemPoolAttrReleaseThreshold, &keep);
            else { c->mempool = nullptr; cudaGetLastError(); }
            c->scene.pool = c->mempool;
        }
        if ((e = cudaMalloc((void**)&c->d_counters, 6 * sizeof(unsigned long long))) != cudaSuccess) { rc = cuda_fail(nullptr, e, "cudaMalloc counters"); break; }
        if ((e = cudaMemsetAsync(c->d_counters, 0, 6 * sizeof(unsigned long long), c->stream)) != cudaSuccess) { rc = cuda_fail(nullptr, e, "memset counters"); break; }
        if ((e = cudaMalloc((void**)&c->d_queue, sizeof(unsigned long long))) != cudaSuccess) { rc = cuda_fail(nullptr, e, "cudaMalloc queue"); break; }
        { cudaDeviceProp prop; if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { rc = cuda_fail(nullptr, e, "cudaGetDeviceProperties"); break; } c->n_sms = prop.multiProcessorCount; }
        rc = alloc_image(c, width, height);
        if (rc != RT_OK) { g_create_err = c->err; break; }
        if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) { rc = cuda_fail(nullptr, e, "cudaStreamSynchronize"); break; }
    } while (0);
    if (rc != RT_OK) { rt_destroy(c); return rc; }
    *out = c;
    return RT_OK;
}

void rt_destroy(rt_ctx* c)
{
    if (!c) return;
    DeviceGuard g(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    drain_events(c);
    for (auto& p : c->pool) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    c->scene.release();
    free_textures(c);
    if (c->comm) { rt::comm_destroy(c->comm); c->comm = nullptr; }
    if (c->combined) cudaFree(c->combined);
    if (c->d_rgba8) cudaFree(c->d_rgba8);
    if (c->accum) cudaFree(c->accum);
    if (c->primary_ids) cudaFree(c->primary_ids);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->d_queue) cudaFree(c->d_queue);
    c->wave.release();
    if (c->staging) cudaFree(c->staging);
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    if (c->d_queue2) cudaFree(c->d_queue2);
    c->tile_order.release();
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (cudaEvent_t e : c->ev_blend) if (e) cudaEventDestroy(e);
    if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    if (c->mempool) cudaMemPoolDestroy(c->mempool);
    delete c;
}

int rt_upload_scene(rt_ctx* c, const rt_sphere* spheres, uint32_t ns, const rt_vertex* vertices, uint32_t nv,
                    const uint32_t* indices, uint32_t ni, const rt_mesh* meshes, uint32_t nm)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if ((ns && !spheres) || (nv && !vertices) || (ni && !indices) || (nm && !meshes))
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_upload_scene: NULL array with non-zero count");
    // The shader indexes unchecked (wgsl:161-170); here every address is validated up front.
    uint64_t total_tris = 0;
    std::set<std::tuple<uint32_t, uint32_t, uint32_t>> checked;
    for (uint32_t m = 0; m < nm; m++) {
        const rt_mesh& me = meshes[m];
        if (me.triangles == 0) continue;
        uint64_t end = (uint64_t)me.first + 3ull * me.triangles;
        if (end > ni) {
            char b[160]; snprintf(b, sizeof b, "rt_upload_scene: mesh %u reads indices[%u..%llu) of %u", m, me.first, (unsigned long long)end, ni);
            return fail(c, RT_ERR_BAD_INDEX, b);
        }
        // records that alias one (first, triangles, offset) range (instances) are checked once
        if (!checked.insert(std::make_tuple(me.first, me.triangles, me.offset)).second) { total_tris += me.triangles; continue; }
        for (uint64_t k = me.first; k < end; k++) {
            if ((uint64_t)me.offset + indices[k] >= nv) {
                char b[160]; snprintf(b, sizeof b, "rt_upload_scene: mesh %u index %llu -> vertex %llu of %u", m, (unsigned long long)k, (unsigned long long)me.offset + indices[k], nv);
                return fail(c, RT_ERR_BAD_INDEX, b);
            }
        }
        total_tris += me.triangles;
    }
    if ((uint64_t)ns + total_tris >= (1ull << 27))         // node addresses are 32-bit byte offsets: 2^27 packed nodes of 32 bytes
        return fail(c, RT_ERR_UNSUPPORTED, "rt_upload_scene: more than 2^27 primitives");
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    // 0 = by size (scripts/builder_sweep.py): leaves of 2 and a +-3 window up to 262,144 primitives; above, where the packed nodes make
    // node visits cheap relative to record fetches, leaves of 1 and a +-2 window (C5: 4,702 against 4,412 Mrays/s)
    const bool large = (uint64_t)ns + total_tris > RT_PF_MAX_PRIMS;
    const uint32_t leaf = (uint32_t)(c->opt_leaf_size < 1 ? (large ? 1 : 2) : (c->opt_leaf_size > 8 ? 8 : c->opt_leaf_size));
    const int radius = (int)(c->opt_ploc_radius < 1 ? (large ? 2 : 3) : c->opt_ploc_radius);
    // Mesh records aliasing one (first, triangles, offset) range become instances of one object-space hierarchy, unless the
    // exhaustive scan or the wavefront kernels (which need flattened records) are selected at upload time
    const int instancing = (c->opt_traversal == 0 && c->opt_kernel == 0) ? (int)c->opt_instancing : 0;
    cudaError_t e = rt::build_scene(c->scene, spheres, ns, vertices, nv, indices, ni, meshes, nm, leaf, (int)c->opt_builder, radius, instancing, (int)c->opt_node_format, c->stream);
    if (e != cudaSuccess) { c->has_scene = false; return cuda_fail(c, e, "scene build"); }
    if (c->scene.n_inst != 0 && c->scene.max_depth > RT_STACK - 2) {          // two levels too deep for the stack: flatten instead
        e = rt::build_scene(c->scene, spheres, ns, vertices, nv, indices, ni, meshes, nm, leaf, (int)c->opt_builder, radius, 0, (int)c->opt_node_format, c->stream);
        if (e != cudaSuccess) { c->has_scene = false; return cuda_fail(c, e, "scene build"); }
    }
    if (!c->scene.packed && (uint64_t)c->scene.n_nodes >= (1ull << 26)) {      // ... and 2^26 float nodes of 64 bytes (forced float nodes, bounds that are not finite, instanced scenes)
        c->has_scene = false;
        return fail(c, RT_ERR_UNSUPPORTED, "rt_upload_scene: more than 2^26 hierarchy nodes in the 64-byte format (RT_OPT_NODE_FORMAT = 1, instancing, or bounds that are not finite)");
    }
    // A hierarchy deeper than the traversal stack (both builders; needs thousands of coincident primitives plus far
    // outliers) is not an error: the reference renders any scene, so such a scene is traced with the exhaustive scan.
    c->scene_too_deep = c->scene.max_depth > RT_STACK - 2;
    c->has_scene = true;
    c->tile_order.rearm();
    c->mesh_ranges.assign(meshes, meshes + nm);
    return RT_OK;
}

int rt_update_records(rt_ctx* c, const rt_sphere* spheres, uint32_t ns, const rt_mesh* meshes, uint32_t nm)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!c->has_scene) return fail(c, RT_ERR_NO_SCENE, "rt_update_records: no scene uploaded");
    if ((ns && !spheres) || (nm && !meshes)) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_records: NULL array with non-zero count");
    if (ns != c->scene.n_spheres || nm != c->mesh_ranges.size())
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_records: sphere / mesh counts differ from the uploaded scene (use rt_upload_scene)");
    for (uint32_t m = 0; m < nm; m++) {
        const rt_mesh &a = meshes[m], &b = c->mesh_ranges[m];
        if (a.first != b.first || a.triangles != b.triangles || a.offset != b.offset) {
            char t[160]; snprintf(t, sizeof t, "rt_update_records: mesh %u addresses a different index range than the uploaded one (use rt_upload_scene)", m);
            return fail(c, RT_ERR_INVALID_ARGUMENT, t);
        }
        if (a.triangles && (a.diffuse_tex | a.normal_tex) && !c->scene.has_tex)
            return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_records: texture handles on a scene uploaded without any (use rt_upload_scene)");
    }
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    cudaError_t e = rt::refit_scene(c->scene, spheres, meshes, c->stream);
    if (e != cudaSuccess) { c->has_scene = false; return cuda_fail(c, e, "scene refit"); }
    if (!c->scene.packed && (uint64_t)c->scene.n_nodes >= (1ull << 26)) {      // the moved bounds are not finite any more and the scene is too large for float nodes
        c->has_scene = false;
        return fail(c, RT_ERR_UNSUPPORTED, "rt_update_records: more than 2^26 hierarchy nodes in the 64-byte format (bounds that are not finite)");
    }
    c->mesh_ranges.assign(meshes, meshes + nm);
    return RT_OK;
}

int rt_upload_textures(rt_ctx* c, const rt_texture* tex, uint32_t n)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (n && !tex) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_upload_textures: NULL array with non-zero count");
    for (uint32_t i = 0; i < n; i++)
        if (tex[i].width == 0 || tex[i].height == 0 || tex[i].width > 32768u || tex[i].height > 32768u || !tex[i].rgba8) {
            char b[128]; snprintf(b, sizeof b, "rt_upload_textures: texture %u has no pixels or a side of 0 / more than 32768", i);
            return fail(c, RT_ERR_INVALID_ARGUMENT, b);
        }
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    free_textures(c);
    std::vector<rt::TexDev> table(n);
    for (uint32_t i = 0; i < n; i++) {
        rt_ctx::Tex t{nullptr, 0, tex[i].width, tex[i].height};
        cudaChannelFormatDesc fd = cudaCreateChannelDesc(8, 8, 8, 8, cudaChannelFormatKindUnsigned);
        cudaError_t e = cudaMallocArray(&t.arr, &fd, t.w, t.h);
        if (e == cudaSuccess) e = cudaMemcpy2DToArray(t.arr, 0, 0, tex[i].rgba8, (size_t)t.w * 4, (size_t)t.w * 4, t.h, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) {
            cudaResourceDesc rd; memset(&rd, 0, sizeof rd);
            rd.resType = cudaResourceTypeArray; rd.res.array.array = t.arr;
            cudaTextureDesc td; memset(&td, 0, sizeof td);
            td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;   // wrap is done in the kernel (exact integer arithmetic)
            td.filterMode = cudaFilterModePoint;                            // bilinear weights are applied in the kernel
            td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
            e = cudaCreateTextureObject(&t.obj, &rd, &td, nullptr);
        }
        if (e != cudaSuccess) {
            if (t.arr) cudaFreeArray(t.arr);
            free_textures(c);
            return cuda_fail(c, e, "rt_upload_textures");
        }
        c->textures.push_back(t);
        table[i].obj = t.obj; table[i].width = t.w; table[i].height = t.h;
    }
    if (n) {
        CU(c, cudaMalloc((void**)&c->d_tex, n * sizeof(rt::TexDev)));
        CU(c, cudaMemcpy(c->d_tex, table.data(), n * sizeof(rt::TexDev), cudaMemcpyHostToDevice));
    }
    return RT_OK;
}

int rt_set_camera(rt_ctx* c, const rt_camera_uniform* cam)
{
    if (!c || !cam) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_camera: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (!c->has_camera || memcmp(&c->camera, cam, sizeof *cam) != 0) c->tile_order.rearm();
    c->camera = *cam; c->has_camera = true;
    return RT_OK;
}

int rt_set_params(rt_ctx* c, const rt_params* p)
{
    if (!c || !p) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_params: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (p->width != c->width || p->height != c->height)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_params: width/height differ from the image (call rt_resize first)");
    c->params = *p; c->has_params = true;
    return RT_OK;
}

int rt_dispatch(rt_ctx* c)
{
    int rc = check_ready(c);
    if (rc != RT_OK) return rc;
    DeviceGuard g(c->device);
    // a single dispatch is never frame-sharded: it is exactly the reference's compute pass
    uint32_t fw = c->frame_world; c->frame_world = 1;
    rc = launch(c, c->params.frames, 1, c->opt_primary_ids != 0);
    c->frame_world = fw;
    if (rc == RT_OK) finish_call(c);
    return rc;
}

int rt_render_frames(rt_ctx* c, int32_t first_frame, uint32_t n_frames)
{
    int rc = check_ready(c);
    if (rc != RT_OK) return rc;
    if (n_frames == 0) return RT_OK;
    if (n_frames > 0x3fffffffu || (int64_t)first_frame + n_frames > 0x7fffffffll) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frames: frame range overflows i32");
    DeviceGuard g(c->device);
    int64_t f = first_frame, end = (int64_t)first_frame + n_frames;
    const int64_t world = c->frame_world;
    if (world > 1) {
        if (first_frame < 0) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frames: frame sharding needs first_frame >= 0");
        int64_t r = f % world;
        f += (c->frame_rank + world - r) % world;       // first frame of this rank
    }
    bool want_ids = c->opt_primary_ids != 0;
    int64_t per = frames_per_launch(c);
    const int64_t mine_total = f < end ? (end - f + world - 1) / world : 0;
    int64_t half = c->opt_staging_bytes / 2 / (int64_t)((size_t)c->width * c->height * sizeof(float4));
    if (half > per) half = per;
    // Several launches: pipeline them over two streams, half of the staging planes each — when a half still holds 8 dispatches.
    // Measured (scripts/pipe_probe.py, 1 GiB budget): C3 8 launches of 15 pipelined 9,494 Mrays/s against 4 serial launches of 31
    // 9,130 (one launch of 128: 10,089); C4 at 4K 16 pipelined launches of 4 24.8 Grays/s against 8 serial of 8 25.5: shorter
    // launches trace slower by themselves (more tiles in flight per SM, colder L1), which below 8 outweighs the hidden tails.
    if (mine_total > per && half >= 8 && c->opt_kernel == 0 && c->opt_item_mode != 2 && c->opt_pipeline != 0) {
        if (!c->stream2) {
            CU(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
            CU(c, cudaMalloc((void**)&c->d_queue2, sizeof(unsigned long long)));
            CU(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
            CU(c, cudaEventCreateWithFlags(&c->ev_blend[0], cudaEventDisableTiming));
            CU(c, cudaEventCreateWithFlags(&c->ev_blend[1], cudaEventDisableTiming));
        }
        { int rc2 = ensure_staging(c, (uint32_t)(2 * half)); if (rc2 != RT_OK) return rc2; }
        std::pair<cudaEvent_t, cudaEvent_t> ev;
        if (!c->pool.empty()) { ev = c->pool.back(); c->pool.pop_back(); }
        else { CU(c, cudaEventCreate(&ev.first)); CU(c, cudaEventCreate(&ev.second)); }
        CU(c, cudaEventRecord(ev.first, c->stream));
        CU(c, cudaEventRecord(c->ev_fork, c->stream));
        CU(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
        int slot = 0, k = 0;
        while (f < end) {
            int64_t mine = (end - f + world - 1) / world;
            uint32_t n = (uint32_t)(mine < half ? mine : half);
            rc = launch(c, (int32_t)f, n, want_ids, slot, (uint32_t)half, k > 0);
            if (rc != RT_OK) return rc;
            want_ids = false;
            f += (int64_t)n * world;
            slot ^= 1; k++;
        }
        if (slot == 0) CU(c, cudaStreamWaitEvent(c->stream, c->ev_blend[1], 0));      // the last launch ran on stream2: join
        CU(c, cudaEventRecord(ev.second, c->stream));
        c->pending.push_back(ev);
        if (c->pending.size() > 4096) drain_events(c);
        finish_call(c);
        return RT_OK;
    }
    while (f < end) {
        int64_t mine = (end - f + world - 1) / world;   // frames f, f+world, ... below end
        uint32_t n = (uint32_t)(mine < per ? mine : per);
        rc = launch(c, (int32_t)f, n, want_ids);
        if (rc != RT_OK) return rc;
        want_ids = false;
        f += (int64_t)n * world;
    }
    finish_call(c);
    return RT_OK;
}

// rt_render_frames for callers that may not allocate: one staged launch whose scratch is the static device array of
// rt_trace.cu.  A separate path on purpose (launch() above is left as it is): its own copy of the launch parameters, no
// event pair (cudaEventCreate allocates), no staging / stream / pool set-up.
int rt_render_frames_static(rt_ctx* c, int32_t first_frame, uint32_t n_frames)
{
    int rc = check_ready(c);
    if (rc != RT_OK) return rc;
    if (n_frames == 0) return RT_OK;
    if (n_frames > 0x3fffffffu || (int64_t)first_frame + n_frames > 0x7fffffffll) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frames_static: frame range overflows i32");
    if (c->frame_world != 1) return fail(c, RT_ERR_UNSUPPORTED, "rt_render_frames_static: sample sharding (frame_world > 1) is not supported; tile sharding is");
    if (c->opt_kernel != 0) return fail(c, RT_ERR_UNSUPPORTED, "rt_render_frames_static: only the persistent trace kernel (RT_OPT_KERNEL = 0)");
    const uint64_t plane = (uint64_t)c->width * c->height * sizeof(float4), limit = rt::static_scratch_bytes();
    if ((uint64_t)n_frames * plane > limit) {
        char b[256];
        snprintf(b, sizeof b, "rt_render_frames_static: %u dispatches of %ux%u need %llu bytes of staging, the static scratch holds %llu (at most %llu dispatches per call at this size)",
                 n_frames, c->width, c->height, (unsigned long long)((uint64_t)n_frames * plane), (unsigned long long)limit, (unsigned long long)(limit / plane));
        return fail(c, RT_ERR_INVALID_ARGUMENT, b);
    }
    const bool brute = c->opt_traversal == 1 || c->scene_too_deep;
    if (brute && c->scene.iprims) return fail(c, RT_ERR_UNSUPPORTED, "rt_render_frames_static: the exhaustive scan needs flattened records (upload with RT_OPT_INSTANCING = 0)");
    DeviceGuard g(c->device);
    rt::SceneDev sd{};
    const rt::SceneBuffers& s = c->scene;
    sd.prims = s.prims; sd.iprims = s.iprims; sd.c0 = s.c0; sd.c1 = s.c1; sd.c2 = s.c2; sd.c3 = s.c3; sd.c4 = s.c4; sd.c5 = s.c5;
    sd.tex = c->d_tex; sd.n_tex = (uint32_t)c->textures.size();
    sd.m0 = s.m0; sd.m1 = s.m1; sd.m2 = s.m2; sd.nodes = s.nodes;
    sd.n_prims = s.n_prims; sd.n_spheres = s.n_spheres; sd.root = s.root;
    sd.max_abs = s.max_abs; sd.pad_scale = 1.52587890625e-05f;   // 2^-16
    sd.packed = s.packed ? 1 : 0; sd.grid_o = make_float3(s.grid_o[0], s.grid_o[1], s.grid_o[2]); sd.grid_c = make_float3(s.grid_c[0], s.grid_c[1], s.grid_c[2]);
    rt::LaunchParams lp{};
    lp.width = c->width; lp.height = c->height;
    lp.bounces = c->params.number_of_bounces; lp.rpp = c->params.rays_per_pixel; lp.toggle = c->params.skybox;
    lp.first_frame = first_frame; lp.n_frames = n_frames; lp.blend0 = first_frame;
    lp.tile_rank = c->tile_rank; lp.tile_world = c->tile_world; lp.frame_world = 1;
    lp.tiles_x = (c->width + 7) / 8;
    const uint32_t tiles = lp.tiles_x * ((c->height + 7) / 8);
    lp.n_warp_tiles = 2u * (tiles > c->tile_rank ? (tiles - c->tile_rank + c->tile_world - 1) / c->tile_world : 0);
    lp.seed_mode = (int32_t)c->opt_seed_mode;
    // Guide of the item queue (grab = remaining / (warps x guide)): scenes with real geometry end their launches on finer grabs — their
    // items are long enough to hide a grab's round trip and differ enough in cost for the last grabs to matter (C3 x 32: 18.5 instead of
    // 21.7 ms, x 16 -14 %, x 8 -6 %, x 128 +0.2 %; C4 x 8 -2.4 %, x 32 +0.2 %); a handful of primitives (C2: 429, x 32 +3 %, x 16 +8 % with
    // the finer guide) keeps the coarse one.  scripts/queue_sweep.py, profiles/r2_summary.md.
    lp.queue_guide = (uint32_t)(c->scene.n_prims >= 2048u ? 2ull * RT_QUEUE_GUIDE : RT_QUEUE_GUIDE);
    lp.queue_chunk = (uint32_t)RT_QUEUE_CHUNK;
    if ((c->opt_sched >> 24) & 0xff) lp.queue_chunk = 32u * (uint32_t)((c->opt_sched >> 24) & 0xff);      // tuning: RT_OPT_SCHED bits 24-31 = grab size / 32, bits 32-39 = guide
    if ((c->opt_sched >> 32) & 0xff) lp.queue_guide = (uint32_t)((c->opt_sched >> 32) & 0xff);
    lp.div_frames = rt::make_fastdiv(n_frames); lp.div_tiles_x = rt::make_fastdiv(lp.tiles_x); lp.div_width = rt::make_fastdiv(c->width);
    lp.rpp_scale = (lp.rpp > 0 && (lp.rpp & (lp.rpp - 1)) == 0) ? 1.0f / (float)lp.rpp : 0.0f;
    lp.sched_n_stay = (int32_t)(c->opt_sched & 0xff); if (lp.sched_n_stay < 1) lp.sched_n_stay = 1; lp.sched_l_stay = (int32_t)((c->opt_sched >> 8) & 0xff); lp.sched_s_min = (int32_t)((c->opt_sched >> 16) & 0xff);
    cudaError_t e = rt::launch_trace_static(sd, to_dev(c->camera), lp, c->accum, c->opt_primary_ids != 0 ? c->primary_ids : nullptr, c->d_counters,
                                            brute, c->opt_detail != 0, c->n_sms, c->device, c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "rt_render_frames_static");
    c->launches++;
    c->dispatches += n_frames;
    { const int64_t last = (int64_t)lp.blend0 + (int64_t)n_frames - 1; c->mean_count = last >= 0 ? (uint32_t)(last + 1) : 1u; }
    return RT_OK;
}

int rt_sync(rt_ctx* c)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_resize(rt_ctx* c, uint32_t w, uint32_t h)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (w == 0 || h == 0) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_resize: zero-sized image");   // context.rs:127 ignores it
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    int rc = alloc_image(c, w, h);
    c->has_params = false;
    return rc;
}

int rt_clear(rt_ctx* c)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    DeviceGuard g(c->device);
    CU(c, cudaMemsetAsync(c->accum, 0, (size_t)c->width * c->height * sizeof(float4), c->stream));
    c->mean_count = 0;
    return RT_OK;
}

int rt_set_shard(rt_ctx* c, uint32_t tile_rank, uint32_t tile_world, uint32_t frame_rank, uint32_t frame_world)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (tile_world == 0 || frame_world == 0 || tile_rank >= tile_world || frame_rank >= frame_world)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_shard: rank must be < world, world >= 1");
    c->tile_rank = tile_rank; c->tile_world = tile_world; c->frame_rank = frame_rank; c->frame_world = frame_world;
    return RT_OK;
}

namespace {
// shared body of rt_combine_peers / rt_combine_peers_rgba8
int combine_peers(rt_ctx* const* ctxs, uint32_t n, const float* weights, uint8_t* host_rgba8, size_t n_bytes, const char* who)
{
    if (!ctxs || n == 0 || !ctxs[0]) return RT_ERR_INVALID_ARGUMENT;
    rt_ctx* c = ctxs[0];
    if (n > RT_MAX_PEERS) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": at most 16 contexts");
    const size_t npx = (size_t)c->width * c->height;
    if (host_rgba8 && n_bytes != npx * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": size must be W*H*4");
    rt::PeerSet ps{};
    ps.n = n;
    for (uint32_t k = 0; k < n; k++) {
        rt_ctx* p = ctxs[k];
        if (!p) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": NULL context");
        if (p->width != c->width || p->height != c->height) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": contexts differ in size");
        for (uint32_t j = 0; j < k; j++) if (ctxs[j] == p) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": a context is listed twice");
        ps.p[k].accum = p->accum; ps.p[k].tile_rank = p->tile_rank; ps.p[k].tile_world = p->tile_world;
    }
    for (uint32_t k = 0; k < n; k++) {
        // default weight: this shard's share of the dispatches its pixels have seen (count-weighted mean of the running means);
        // shards with the same tile partition own the same pixels
        uint64_t total = 0;
        for (uint32_t j = 0; j < n; j++)
            if (ctxs[j]->tile_rank == ctxs[k]->tile_rank && ctxs[j]->tile_world == ctxs[k]->tile_world) total += ctxs[j]->mean_count;
        ps.p[k].weight = weights ? weights[k] : (total == 0 || total == ctxs[k]->mean_count ? 1.0f : (float)ctxs[k]->mean_count / (float)total);
    }
    if (npx == 0) return RT_OK;
    DeviceGuard g(c->device);
    { int rc = ensure_combined(c); if (rc != RT_OK) return rc; }
    if (host_rgba8) { int rc = ensure_rgba8(c); if (rc != RT_OK) return rc; }
    // peer access + ordering: the combine waits for every shard's stream, and the shards wait for the combine
    std::vector<cudaEvent_t> events;
    auto drop = [&]() { for (cudaEvent_t e : events) cudaEventDestroy(e); };
    for (uint32_t k = 1; k < n; k++) {
        rt_ctx* p = ctxs[k];
        if (p->device != c->device) {
            int can = 0;
            CU(c, cudaDeviceCanAccessPeer(&can, c->device, p->device));
            if (!can) { drop(); return fail(c, RT_ERR_UNSUPPORTED, std::string(who) + ": device " + std::to_string(c->device) + " has no peer access to device " + std::to_string(p->device)); }
            cudaError_t e = cudaDeviceEnablePeerAccess(p->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { drop(); return cuda_fail(c, e, "cudaDeviceEnablePeerAccess"); }
        }
        cudaEvent_t ev = nullptr;
        cudaError_t e;
        { DeviceGuard gp(p->device); e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming); if (e == cudaSuccess) { events.push_back(ev); e = cudaEventRecord(ev, p->stream); } }
        if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, ev, 0);
        if (e != cudaSuccess) { drop(); return cuda_fail(c, e, "combine: ordering after a shard's stream"); }
    }
    uint8_t* d8 = host_rgba8 ? c->d_rgba8 : nullptr;
    cudaError_t e = rt::launch_combine(ps, c->combined, d8, c->width, c->height, c->n_sms, c->stream);
    cudaEvent_t done = nullptr;
    if (e == cudaSuccess && n > 1) e = cudaEventCreateWithFlags(&done, cudaEventDisableTiming);
    if (done) events.push_back(done);
    if (e == cudaSuccess && done) e = cudaEventRecord(done, c->stream);
    for (uint32_t k = 1; k < n && e == cudaSuccess; k++) { DeviceGuard gp(ctxs[k]->device); e = cudaStreamWaitEvent(ctxs[k]->stream, done, 0); }
    if (e == cudaSuccess && host_rgba8) e = cudaMemcpyAsync(host_rgba8, d8, npx * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && host_rgba8) e = cudaStreamSynchronize(c->stream);
    drop();
    if (e != cudaSuccess) return cuda_fail(c, e, who);
    return RT_OK;
}
} // namespace

int rt_combine_peers(rt_ctx* const* ctxs, uint32_t n, const float* weights)
{
    return combine_peers(ctxs, n, weights, nullptr, 0, "rt_combine_peers");
}

int rt_combine_peers_rgba8(rt_ctx* const* ctxs, uint32_t n, const float* weights, uint8_t* rgba, size_t n_bytes)
{
    if (!rgba) return (ctxs && n && ctxs[0]) ? fail(ctxs[0], RT_ERR_INVALID_ARGUMENT, "rt_combine_peers_rgba8: NULL") : RT_ERR_INVALID_ARGUMENT;
    return combine_peers(ctxs, n, weights, rgba, n_bytes, "rt_combine_peers_rgba8");
}

// ---- multi-process combine over NCCL (rt_comm.cu) ---------------------------------------------------------------
int rt_comm_unique_id(void* id, size_t n_bytes)
{
    if (!id || n_bytes != RT_COMM_ID_BYTES) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_comm_unique_id: id must be RT_COMM_ID_BYTES bytes");
    if (const char* e = rt::comm_unique_id(id)) return fail(nullptr, RT_ERR_UNSUPPORTED, e);
    return RT_OK;
}

int rt_comm_init(rt_ctx* c, int n_ranks, int rank, const void* id, size_t n_bytes)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!id || n_bytes != RT_COMM_ID_BYTES) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_comm_init: id must be RT_COMM_ID_BYTES bytes");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_comm_init: rank must be in [0, n_ranks)");
    if (c->comm) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_comm_init: the context already has a communicator (rt_comm_destroy)");
    DeviceGuard g(c->device);
    if (const char* e = rt::comm_create(&c->comm, n_ranks, rank, id)) { c->comm = nullptr; return fail(c, RT_ERR_UNSUPPORTED, e); }
    return RT_OK;
}

int rt_comm_destroy(rt_ctx* c)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    DeviceGuard g(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) { rt::comm_destroy(c->comm); c->comm = nullptr; }
    return RT_OK;
}

namespace {
int combine_comm(rt_ctx* c, int root, uint8_t* host_rgba8, size_t n_bytes, const char* who)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!c->comm) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": no communicator (rt_comm_init)");
    if (root < 0 || root >= rt::comm_size(c->comm)) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": root out of range");
    const bool is_root = rt::comm_rank(c->comm) == root;
    const size_t npx = (size_t)c->width * c->height;
    if (is_root && host_rgba8 && n_bytes != npx * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, std::string(who) + ": size must be W*H*4");
    DeviceGuard g(c->device);
    const bool fused = host_rgba8 != nullptr || n_bytes == (size_t)-1;
    if (is_root && !fused) { int rc = ensure_combined(c); if (rc != RT_OK) return rc; }
    if (is_root && fused) { int rc = ensure_rgba8(c); if (rc != RT_OK) return rc; }
    rt::ShardRec mine{c->tile_rank, c->tile_world, c->frame_rank, c->frame_world, c->mean_count, {0, 0, 0}};
    if (const char* e = rt::comm_combine(c->comm, mine, root, c->accum, fused ? nullptr : c->combined, (is_root && fused) ? c->d_rgba8 : nullptr,
                                         c->width, c->height, c->n_sms, c->stream))
        return fail(c, RT_ERR_CUDA, e);
    if (is_root && host_rgba8) {
        CU(c, cudaMemcpyAsync(host_rgba8, c->d_rgba8, npx * 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
    }
    return RT_OK;
}
} // namespace

int rt_combine(rt_ctx* c, int root) { return combine_comm(c, root, nullptr, 0, "rt_combine"); }

int rt_combine_rgba8(rt_ctx* c, int root, uint8_t* rgba, size_t n_bytes)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    const bool is_root = c->comm && rt::comm_rank(c->comm) == root;
    if (is_root && !rgba) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_combine_rgba8: NULL at the root");
    return combine_comm(c, root, is_root ? rgba : nullptr, is_root ? n_bytes : (size_t)-1, "rt_combine_rgba8");
}

int rt_read_combined(rt_ctx* c, float* rgba, size_t n_floats)
{
    if (!c || !rgba) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_combined: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (n_floats != (size_t)c->width * c->height * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_combined: size must be W*H*4");
    if (!c->combined) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_combined: no combine has produced an image on this context");
    DeviceGuard g(c->device);
    CU(c, cudaMemcpyAsync(rgba, c->combined, n_floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

void* rt_combined_device_ptr(rt_ctx* c) { return c ? (void*)c->combined : nullptr; }
uint32_t rt_mean_count(const rt_ctx* c) { return c ? c->mean_count : 0u; }

int rt_set_option(rt_ctx* c, int option, int64_t value)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    switch (option) {
    case RT_OPT_TRAVERSAL: if (value != 0 && value != 1) break;
        if (value == 1 && c->has_scene && c->scene.n_inst != 0) return fail(c, RT_ERR_UNSUPPORTED, "rt_set_option: the exhaustive scan needs flattened records; set RT_OPT_TRAVERSAL = 1 (or RT_OPT_INSTANCING = 0) before rt_upload_scene");
        c->opt_traversal = value; return RT_OK;
    case RT_OPT_INSTANCING: if (value < 0 || value > 2) break; c->opt_instancing = value; return RT_OK;
    case RT_OPT_PIPELINE: if (value != 0 && value != 1) break; c->opt_pipeline = value; return RT_OK;
    case RT_OPT_NODE_FORMAT: if (value < 0 || value > 2) break; c->opt_node_format = value; return RT_OK;
    case RT_OPT_TILE_ORDER: if (value != 0 && value != 1) break; c->opt_tile_order = value; return RT_OK;
    case RT_OPT_SEED_MODE: if (value != 0 && value != 1) break; c->opt_seed_mode = value; return RT_OK;
    case RT_OPT_PRIMARY_IDS: c->opt_primary_ids = value != 0; return RT_OK;
    case RT_OPT_DETAIL_COUNTERS: c->opt_detail = value != 0; return RT_OK;
    case RT_OPT_FRAMES_PER_LAUNCH: if (value < 0 || value > (1 << 20)) break; c->opt_frames_per_launch = value; return RT_OK;
    case RT_OPT_ITEM_MODE: if (value < 0 || value > 2) break; c->opt_item_mode = value; return RT_OK;
    case RT_OPT_LEAF_SIZE: if (value < 0 || value > 8) break; c->opt_leaf_size = value; return RT_OK;
    case RT_OPT_SCHED: c->opt_sched = value & 0xffffffffffll; return RT_OK;
    case RT_OPT_BUILDER: if (value != 0 && value != 1) break; c->opt_builder = value; return RT_OK;
    case RT_OPT_PLOC_RADIUS: if (value < 0 || value > 16) break; c->opt_ploc_radius = value; return RT_OK;
    case RT_OPT_KERNEL: if (value != 0 && value != 1) break;
        if (value == 1 && c->has_scene && c->scene.n_inst != 0) return fail(c, RT_ERR_UNSUPPORTED, "rt_set_option: the wavefront kernels need flattened records; set RT_OPT_KERNEL = 1 (or RT_OPT_INSTANCING = 0) before rt_upload_scene");
        c->opt_kernel = value; return RT_OK;
    case RT_OPT_WAVE_SLOTS: if (value < 1024 || value > (1 << 26)) break; c->opt_wave_slots = value; return RT_OK;
    case RT_OPT_STAGING_MIB: if (value < 16 || value > (1 << 17)) break; c->opt_staging_bytes = value << 20; return RT_OK;
    default: return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_option: unknown option");
    }
    return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_option: value out of range");
}

int rt_get_option(const rt_ctx* c, int option, int64_t* value)
{
    if (!c || !value) return RT_ERR_INVALID_ARGUMENT;
    switch (option) {
    case RT_OPT_TRAVERSAL: *value = c->opt_traversal; return RT_OK;
    case RT_OPT_SEED_MODE: *value = c->opt_seed_mode; return RT_OK;
    case RT_OPT_PRIMARY_IDS: *value = c->opt_primary_ids; return RT_OK;
    case RT_OPT_DETAIL_COUNTERS: *value = c->opt_detail; return RT_OK;
    case RT_OPT_FRAMES_PER_LAUNCH: *value = c->opt_frames_per_launch; return RT_OK;
    case RT_OPT_LEAF_SIZE: *value = c->opt_leaf_size; return RT_OK;
    case RT_OPT_SCHED: *value = c->opt_sched; return RT_OK;
    case RT_OPT_BUILDER: *value = c->opt_builder; return RT_OK;
    case RT_OPT_PLOC_RADIUS: *value = c->opt_ploc_radius; return RT_OK;
    case RT_OPT_KERNEL: *value = c->opt_kernel; return RT_OK;
    case RT_OPT_WAVE_SLOTS: *value = c->opt_wave_slots; return RT_OK;
    case RT_OPT_STAGING_MIB: *value = c->opt_staging_bytes >> 20; return RT_OK;
    case RT_OPT_ITEM_MODE: *value = c->opt_item_mode; return RT_OK;
    case RT_OPT_INSTANCING: *value = c->opt_instancing; return RT_OK;
    case RT_OPT_PIPELINE: *value = c->opt_pipeline; return RT_OK;
    case RT_OPT_NODE_FORMAT: *value = c->opt_node_format; return RT_OK;
    case RT_OPT_TILE_ORDER: *value = c->opt_tile_order; return RT_OK;
    case RT_OPT_NODES_PACKED: *value = (c->has_scene && c->scene.packed) ? 1 : 0; return RT_OK;
    default: return RT_ERR_INVALID_ARGUMENT;
    }
}

int rt_read_accum(rt_ctx* c, float* rgba, size_t n_floats)
{
    if (!c || !rgba) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_accum: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (n_floats != (size_t)c->width * c->height * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_accum: size must be W*H*4");
    DeviceGuard g(c->device);
    CU(c, cudaMemcpyAsync(rgba, c->accum, n_floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_write_accum(rt_ctx* c, const float* rgba, size_t n_floats)
{
    if (!c || !rgba) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_write_accum: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (n_floats != (size_t)c->width * c->height * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_write_accum: size must be W*H*4");
    DeviceGuard g(c->device);
    CU(c, cudaMemcpyAsync(c->accum, rgba, n_floats * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_read_rgba8(rt_ctx* c, uint8_t* rgba, size_t n_bytes)
{
    if (!c || !rgba) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_rgba8: NULL") : RT_ERR_INVALID_ARGUMENT;
    size_t n = (size_t)c->width * c->height;
    if (n_bytes != n * 4) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_rgba8: size must be W*H*4");
    DeviceGuard g(c->device);
    { int rc = ensure_rgba8(c); if (rc != RT_OK) return rc; }
    cudaError_t e = rt::launch_tonemap(c->accum, c->d_rgba8, n, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgba, c->d_rgba8, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return cuda_fail(c, e, "tonemap");
    return RT_OK;
}

int rt_read_primary_ids(rt_ctx* c, uint32_t* ids, size_t n)
{
    if (!c || !ids) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_primary_ids: NULL") : RT_ERR_INVALID_ARGUMENT;
    if (n != (size_t)c->width * c->height) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_read_primary_ids: size must be W*H");
    DeviceGuard g(c->device);
    CU(c, cudaMemcpyAsync(ids, c->primary_ids, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_get_counters(rt_ctx* c, rt_counters* out)
{
    if (!c || !out) return RT_ERR_INVALID_ARGUMENT;
    DeviceGuard g(c->device);
    unsigned long long h[6];
    CU(c, cudaMemcpyAsync(h, c->d_counters, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    drain_events(c);
    out->rays = h[0]; out->samples = h[1]; out->node_visits = h[2]; out->tri_tests = h[3]; out->sphere_tests = h[4]; out->hit_segments = h[5];
    out->dispatches = c->dispatches; out->kernel_ms = c->kernel_ms; out->launches = c->launches;
    return RT_OK;
}

int rt_reset_counters(rt_ctx* c)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    DeviceGuard g(c->device);
    CU(c, cudaStreamSynchronize(c->stream));
    drain_events(c);
    CU(c, cudaMemsetAsync(c->d_counters, 0, 6 * sizeof(unsigned long long), c->stream));
    c->dispatches = c->launches = 0; c->kernel_ms = 0.0;
    return RT_OK;
}

void* rt_accum_device_ptr(rt_ctx* c) { return c ? (void*)c->accum : nullptr; }
void* rt_cuda_stream(rt_ctx* c) { return c ? (void*)c->stream : nullptr; }
int rt_device(const rt_ctx* c) { return c ? c->device : -1; }

int rt_bvh_stats(rt_ctx* c, uint32_t* n_prims, uint32_t* n_nodes, uint32_t* max_depth)
{
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!c->has_scene) return fail(c, RT_ERR_NO_SCENE, "no scene uploaded");
    if (n_prims) *n_prims = c->scene.n_prims;
    if (n_nodes) *n_nodes = c->scene.n_nodes;
    if (max_depth) *max_depth = c->scene.max_depth;
    return RT_OK;
}

} // extern "C"
