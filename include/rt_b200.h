/* rt_b200.h — C ABI of the B200-native path tracer (librt_b200.so).
 *
 * This is the drop-in boundary for ONE path of addiswebb/ray_tracer: the compute entry of
 * shaders/ray_tracer.wgsl plus the wgpu dispatch / accumulation plumbing around it in
 * src/core/{ray_tracer,context,texture}.rs.  The reference has no FFI of its own (its only
 * public symbol is `pub async fn run()`, src/lib.rs:10); each entry point below replaces
 * the wgpu call(s) cited next to it, which is what a Rust `extern "C"` block would bind
 * (see INTEGRATION.md for that stub).
 *
 * Plain C: opaque handle, POD records byte-identical to the reference's #[repr(C)]
 * structs, caller-owned buffers, integer status codes.  No exceptions or panics cross
 * this boundary.  A context is bound to one CUDA device and is not thread-safe (the
 * reference drives wgpu from the single winit thread, src/lib.rs:23-69).
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- records (layouts: SURVEY.md §3.4(17)) ------------------------------------------ */

/* src/core/context.rs:11-21 `Params`; shaders/ray_tracer.wgsl:1-8 reads the first 24 B. */
typedef struct rt_params {
    uint32_t width;
    uint32_t height;
    int32_t number_of_bounces;   /* the shader traces number_of_bounces+1 segments (wgsl:233) */
    int32_t rays_per_pixel;      /* the shader takes rays_per_pixel+1 samples, divides by rays_per_pixel (wgsl:312,326) */
    int32_t skybox;              /* WGSL name: toggle (wgsl:274) */
    int32_t frames;              /* host-owned counter: Context::update, context.rs:180-184 */
    int32_t accumulate;          /* host-only */
} rt_params;

/* src/core/camera.rs:9-29 `CameraUniform` (128 B); wgsl:69-80. */
typedef struct rt_camera_uniform {
    float origin[3];            float _padding1;
    float lower_left_corner[3]; float _padding2;
    float horizontal[3];        float _padding3;
    float vertical[3];          float _padding4;
    float near_;                float far_;      float _padding5[2];
    float w[3];                 float _padding6;
    float u[3];                 float _padding7;
    float v[3];                 float lens_radius;
} rt_camera_uniform;

/* src/core/scene.rs:11-21 `Sphere` (64 B); wgsl:9-20. smoothness == -1.0 selects glass. */
typedef struct rt_sphere {
    float position[3];
    float radius;
    float color[4];
    float emission_color[4];
    float emission_strength;
    float smoothness;
    float _padding[2];
} rt_sphere;

/* src/core/scene.rs:23-37 `Mesh` (80 B); wgsl:27-33. `first` indexes indices[], `offset`
 * is added to every index to address vertices[]; several records may alias one range
 * with different `pos` (translation-only instancing). */
typedef struct rt_mesh {
    uint32_t first;
    uint32_t triangles;
    uint32_t offset;
    uint32_t diffuse_tex;       /* Rust `_padding: f32` (always 0.0 upstream). Texture extension: 0 = none, k = texture k-1 of rt_upload_textures */
    float pos[3];
    float _padding2;
    float color[4];
    float emission_color[4];
    float emission_strength;
    float specular;             /* WGSL: material.smoothness */
    uint32_t normal_tex;        /* Rust `_padding3[0]`. Texture extension: tangent-space normal map, same encoding */
    uint32_t _padding3;
} rt_mesh;

/* src/core/scene.rs:79-86 `Vertex` (32 B); wgsl:22-25. */
typedef struct rt_vertex {
    float pos[3];
    float tu;                   /* Rust `_padding1` (0.0 upstream). Texture extension: texture coordinate u */
    float normal[3];
    float tv;                   /* Rust `_padding2`. Texture extension: texture coordinate v */
} rt_vertex;

/* Texture extension (no reference counterpart: upstream has no UVs and loads no images,
 * scene.rs:81-86, texture.rs; SURVEY.md 8(a')).  Tightly packed RGBA8, row 0 first.
 * Sampling: bilinear, repeat addressing, texel centres at (i+0.5)/size, channel = byte/255,
 * no sRGB decode; diffuse: material colour *= texel; normal map: n = 2*rgb-1 in the
 * Gram-Schmidt tangent frame of the UV mapping around the interpolated vertex normal.
 * The arithmetic is specified by oracle/rt_oracle.cpp (sample_texture, apply_textures).
 * Records whose padding words are all zero render exactly as the reference does. */
typedef struct rt_texture {
    uint32_t width;
    uint32_t height;
    const uint8_t* rgba8;
} rt_texture;

/* Device-side counters since the last rt_reset_counters(). A "ray" is one call of
 * calculate_ray_collions (wgsl:234), i.e. one path segment, primary included. */
typedef struct rt_counters {
    uint64_t rays;            /* path segments traced */
    uint64_t samples;         /* camera paths started */
    uint64_t node_visits;     /* BVH nodes fetched, 2 box tests each (only with RT_OPT_DETAIL_COUNTERS) */
    uint64_t tri_tests;       /* ray-triangle tests (only with RT_OPT_DETAIL_COUNTERS) */
    uint64_t sphere_tests;    /* ray-sphere tests (only with RT_OPT_DETAIL_COUNTERS) */
    uint64_t hit_segments;    /* segments that hit something, i.e. were shaded (only with RT_OPT_DETAIL_COUNTERS) */
    uint64_t dispatches;      /* reference-equivalent dispatches executed */
    double kernel_ms;         /* summed CUDA-event time of the trace kernel launches */
    uint64_t launches;        /* trace kernel launches */
} rt_counters;

/* ---- status codes ------------------------------------------------------------------- */
enum {
    RT_OK = 0,
    RT_ERR_INVALID_ARGUMENT = -1,
    RT_ERR_CUDA = -2,
    RT_ERR_NO_SCENE = -3,
    RT_ERR_OUT_OF_MEMORY = -4,
    RT_ERR_BAD_INDEX = -5,       /* an index/first/offset addresses outside the arrays */
    RT_ERR_UNSUPPORTED = -6
};

/* ---- options ------------------------------------------------------------------------ */
enum {
    RT_OPT_TRAVERSAL = 1,        /* 0 = LBVH (default), 1 = exhaustive brute force (validation) */
    RT_OPT_SEED_MODE = 2,        /* the seed's f32->u32 conversion (wgsl:308) of values >= 2^32, i.e. at 3840x2160 every row from 518 on:
                                    0 = strict (default): saturates, the WGSL rule - those pixels share one RNG stream per dispatch;
                                    1 = wide: wraps modulo 2^32 (a labelled deviation; identical to strict below 2^32, e.g. at 1080p) */
    RT_OPT_PRIMARY_IDS = 3,      /* 1 = record the primary-hit AOV on the next render call */
    RT_OPT_DETAIL_COUNTERS = 4,  /* 1 = count node visits / primitive tests (slower kernel) */
    RT_OPT_FRAMES_PER_LAUNCH = 5,/* max reference dispatches fused into one kernel launch (0 = default: 128, or 4096 with RT_OPT_ITEM_MODE = 2; 1 = one launch per dispatch like the reference) */
    RT_OPT_LEAF_SIZE = 6,        /* max primitives per BVH leaf (1..8; 0 = default: 2 up to 262,144 primitives, 1 above); takes effect at the next rt_upload_scene */
    RT_OPT_SCHED = 7,            /* tuning: warp scheduler thresholds in lanes, node_loop_min | (shade_min << 16); bits 24-31: work items per queue grab / 32, bits 32-39: guide of the guided self-scheduling (0 = defaults: 256 items, guide 40, or 20 for scenes of fewer than 2,048 primitives); results do not depend on it */
    RT_OPT_BUILDER = 8,          /* hierarchy over the Morton order: 0 = PLOC (default), 1 = radix tree (Karras 2012); takes effect at the next rt_upload_scene; results do not depend on it */
    RT_OPT_PLOC_RADIUS = 9,      /* PLOC search window, clusters to either side (1..16; 0 = default: 3 up to 262,144 primitives, 2 above) */
    RT_OPT_KERNEL = 10,          /* 0 = persistent megakernel (default), 1 = wavefront kernels over a path-slot pool; results do not depend on it */
    RT_OPT_WAVE_SLOTS = 11,      /* path slots of the wavefront pool (128 B each; default 2^21) */
    RT_OPT_STAGING_MIB = 12,     /* budget of the staging planes (one W*H float4 plane per fused dispatch), MiB (default 4352: 128 dispatches per launch at 1080p, 32 at 4K); bounds the dispatches per launch together with RT_OPT_FRAMES_PER_LAUNCH; allocated at the first fused launch */
    RT_OPT_INSTANCING = 14,      /* Mesh records that alias one (first, triangles, offset) range (at least two records, 32 triangles, no texture handles) can share ONE object-space hierarchy and appear as proxies in a top-level hierarchy (two-level BVH) instead of being flattened; the triangle test then adds the record's pos per instance with the reference's own operations (wgsl:172-174), so results do not depend on it.  1: instance every eligible range (64x less scene memory on C5, rt_update_records of an instance = a top-level refit; measured 16 % slower than flattened records there); 0: flatten every record; 2 (default): instance only when flattening would exceed 8 M triangles.  Takes effect at the next rt_upload_scene; the exhaustive scan (RT_OPT_TRAVERSAL = 1) and the wavefront kernels (RT_OPT_KERNEL = 1) always flatten and must be selected before the upload of such a scene */
    RT_OPT_NODE_FORMAT = 16,     /* hierarchy nodes of flat (non-instanced) scenes: 0 (default) = by size: 64-byte float nodes up to 262,144 primitives, packed 32-byte nodes (box centres on a 16-bit grid, half extents as bf16, one 256-bit load per node visit) above; 1 = float, 2 = packed; takes effect at the next rt_upload_scene; results do not depend on it */
    RT_OPT_NODES_PACKED = 17,    /* read-only (rt_get_option): 1 when the uploaded scene's hierarchy holds packed nodes (see RT_OPT_NODE_FORMAT; instanced scenes and scenes whose bounds are not finite never do) */
    RT_OPT_TILE_ORDER = 18,      /* 1 (default): in launches of at most 4 dispatches (the reference's one dispatch per frame) the 8x8 tiles that were heaviest in the context's previous such launch enter the work queue first, so that the launch ends on cheap work instead of waiting for its last heavy pixels (C3 single dispatch 1.52 -> 1.21 ms; the first call after a resize / re-shard uses raster order); 0: always raster order; results do not depend on it */
    RT_OPT_PIPELINE = 15,        /* 1 (default): when one rt_render_frames call needs several launches (more dispatches than the staging budget holds), they alternate between two streams, half of the staging planes each, so that a launch's blocks fill the SMs the previous one's drain leaves idle (the blends keep dispatch order); 0: one launch after the other; results do not depend on it */
    RT_OPT_ITEM_MODE = 13        /* work items of the trace kernel: 1 = one (pixel, dispatch) pair, frags staged and folded in dispatch order by a blend kernel; 2 = one pixel with all fused dispatches, the progressive store (wgsl:59-66) applied by the owning lane, no staging memory (coarse items: slower on fused launches); 0 = auto (default): 2 for single-dispatch launches, else 1; results do not depend on it */
};

typedef struct rt_ctx rt_ctx;

/* Last error text of `ctx` (or of the failed rt_create when ctx == NULL). Never NULL. */
const char* rt_last_error(const rt_ctx* ctx);

/* Library/ABI version: (major << 16) | minor. */
uint32_t rt_version(void);

/* Replaces Context::new's device + Texture::new(w,h,Rgba32Float) (context.rs:41-72,101;
 * texture.rs:6-31): binds CUDA device `device`, allocates the zero-filled W*H RGBA32F
 * accumulation image. */
int rt_create(rt_ctx** out, int device, uint32_t width, uint32_t height);

/* Rust Drop of Context/RayTracer/Texture. NULL is allowed. */
void rt_destroy(rt_ctx* ctx);

/* Replaces Scene::{sphere,vertex,index,mesh}_buffer (scene.rs:478-507) and the storage
 * bindings 3..6 of RayTracer::new (ray_tracer.rs:51-94,97-130).  Inputs are borrowed for
 * the call only.  Zero counts (with NULL pointers) and 0-triangle meshes are legal.
 * Validates every index, repacks to the SoA device layout and builds the LBVH on the GPU. */
int rt_upload_scene(rt_ctx* ctx,
                    const rt_sphere* spheres, uint32_t n_spheres,
                    const rt_vertex* vertices, uint32_t n_vertices,
                    const uint32_t* indices, uint32_t n_indices,
                    const rt_mesh* meshes, uint32_t n_meshes);

/* Dynamic scenes (upstream rebuilds everything through RayTracer::new on a
 * scene switch, context.rs:259-283).  Same counts and the same first/triangles/offset per mesh
 * as the last rt_upload_scene; sphere positions / radii, mesh `pos` (the translation the shader
 * adds per vertex, wgsl:172-174) and all material fields may change.  Records and bounds are
 * re-derived with the same operations as at upload and the hierarchy's boxes are refitted
 * bottom-up on the kept topology (no sort, no build): results are exactly those of a fresh
 * rt_upload_scene of the updated records; traversal cost grows as primitives move away from
 * where the topology was built. */
int rt_update_records(rt_ctx* ctx, const rt_sphere* spheres, uint32_t n_spheres,
                      const rt_mesh* meshes, uint32_t n_meshes);

/* Texture extension: replaces the context's texture table (n = 0 clears it). Pixels are copied
 * into CUDA arrays bound to texture objects; handles in rt_mesh address this table (handles
 * above n are treated as "none"). Independent of rt_upload_scene order. */
int rt_upload_textures(rt_ctx* ctx, const rt_texture* textures, uint32_t n_textures);

/* queue.write_buffer(camera.buffer, uniform) — context.rs:185. */
int rt_set_camera(rt_ctx* ctx, const rt_camera_uniform* camera);

/* queue.write_buffer(params_buffer, params) — context.rs:186 (also :138,:145).
 * width/height must equal the context's current size (use rt_resize to change it). */
int rt_set_params(rt_ctx* ctx, const rt_params* params);

/* The compute pass dispatch_workgroups(ceil(W/8),ceil(H/8),1) — context.rs:196-208 — with
 * the currently set params (including params.frames). Asynchronous on the context stream. */
int rt_dispatch(rt_ctx* ctx);

/* n_frames consecutive reference dispatches with params.frames = first_frame,
 * first_frame+1, ... (what Context::update + render do per redraw, context.rs:176-208),
 * fused into as few launches as RT_OPT_FRAMES_PER_LAUNCH allows; results are identical to
 * n_frames rt_dispatch calls. With a shard set (rt_set_shard) only this rank's share is
 * rendered. Asynchronous. */
int rt_render_frames(rt_ctx* ctx, int32_t first_frame, uint32_t n_frames);

/* rt_render_frames without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc at this point (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).
 * Inputs are the context's uploaded scene, camera and params, the output its accumulation image, as
 * for rt_render_frames (same image bit for bit, same counters); ONE launch of n_frames fused
 * dispatches whose scratch — the work-queue head and one W*H float4 staging plane per dispatch —
 * lives in a static device array of RT_STATIC_SCRATCH_BYTES that every GPU loading the library
 * holds until the process exits.  The call performs no cudaMalloc / cudaFree / event or stream
 * creation; everything else it touches was allocated by rt_create / rt_upload_*.
 * Limits: n_frames * width * height * 16 <= RT_STATIC_SCRATCH_BYTES (1 dispatch at 1920x1080, 27 at
 * 320x240), else RT_ERR_INVALID_ARGUMENT with the numbers in rt_last_error and nothing launched;
 * RT_ERR_UNSUPPORTED with sample sharding (frame_world > 1) or RT_OPT_KERNEL = 1.  Not included in
 * rt_get_counters' kernel_ms (no timing events).
 * All calls on one device share the array, which keeps the previous call's contents (the call clears
 * the queue head itself, in stream order): calls that overlap on the GPU — two contexts / streams of
 * one device — race on it; serialise them. */
#define RT_STATIC_SCRATCH_BYTES (32u << 20)
int rt_render_frames_static(rt_ctx* ctx, int32_t first_frame, uint32_t n_frames);

/* Wait for all work queued on the context (the implicit wait of queue.submit + present,
 * context.rs:291-292). */
int rt_sync(rt_ctx* ctx);

/* Context::resize (context.rs:126-142): new zero-filled accumulation image. The caller
 * then sets params (frames = -1 upstream) and camera aspect as the reference does. */
int rt_resize(rt_ctx* ctx, uint32_t width, uint32_t height);

/* Zero-fill the accumulation image (a fresh wgpu texture). */
int rt_clear(rt_ctx* ctx);

/* Multi-GPU partition of the image/sample space (no reference counterpart).
 * tile_rank/tile_world: 8x8-pixel tiles are dealt round-robin, tile t -> rank t % world;
 * pixels of other ranks' tiles are left untouched.  frame_rank/frame_world: of the frames
 * passed to rt_render_frames only those with frame % world == rank are rendered, and the
 * running-mean weight of frame f is 1/(f / world + 1), i.e. counts this rank's own frames.
 * world = 1 disables the respective sharding. */
int rt_set_shard(rt_ctx* ctx, uint32_t tile_rank, uint32_t tile_world,
                 uint32_t frame_rank, uint32_t frame_world);

/* ---- image combine of the shards ------------------------------------------------------------------------------
 * Every combine writes a SEPARATE image, the "combined image" of the context that owns the result (rt_read_combined,
 * rt_combined_device_ptr); no accumulation image is modified, so rendering can go on after a combine and be combined
 * again.  For every pixel the shards whose tile partition owns it enter in array / rank order, each with the weight
 * (dispatches in its running mean) / (dispatches in all the owners' running means) — the count-weighted mean of the
 * sample shards' running means, equal to the reference's sequential running mean up to f32 rounding; a pure tile
 * partition reproduces the single-GPU image bit for bit.  Pixels no shard owns take the first context's own value.
 *
 * In ONE process (one rt_ctx per device, one host thread driving them all, as the reference's event loop would):
 * rt_combine_peers runs ONE kernel on ctxs[0]'s device that reads the other images out of their devices' memory
 * through NVLink peer loads (cudaDeviceEnablePeerAccess; contexts on the same device need none).  weights != NULL
 * overrides the count weights.  Ordered after everything queued on every context's stream; later work on any of them
 * waits for the combine.  All contexts must have the same size; 1 <= n <= 16. */
int rt_combine_peers(rt_ctx* const* ctxs, uint32_t n, const float* weights);
/* The same combine fused with the display transform of rt_read_rgba8 (clamp, sRGB OETF, 8 bit) and the copy to the
 * host.  Synchronous. */
int rt_combine_peers_rgba8(rt_ctx* const* ctxs, uint32_t n, const float* weights, uint8_t* rgba, size_t n_bytes);

/* With ONE PROCESS PER GPU (torchrun, MPI, or a host thread per device): the same combine over NCCL (NVLink 5 /
 * NVSwitch), bound at run time (libnccl.so.2).  One rank obtains an id with rt_comm_unique_id and hands it to the
 * others by whatever channel the host has (the reference has none: it is single-device); every rank then calls
 * rt_comm_init (collective) with its context, whose device the communicator is bound to.
 * rt_combine (collective, asynchronous on each context's stream, no host synchronisation): every rank packs the
 * pixels of ITS tiles (W*H / tile_world of them) and sends them to `root` (grouped ncclSend / ncclRecv; the shard
 * table travels by ncclAllGather), where one kernel applies the count weights, de-interleaves the tiles and writes the
 * combined image — or, rt_combine_rgba8, the tone-mapped bytes, copied to `rgba` at the root (ignored elsewhere;
 * synchronous at the root).  All ranks of a communicator must share one tile_world (rt_set_shard).  Results equal
 * rt_combine_peers' on the same shards bit for bit. */
#define RT_COMM_ID_BYTES 128
int rt_comm_unique_id(void* id, size_t n_bytes);
int rt_comm_init(rt_ctx* ctx, int n_ranks, int rank, const void* id, size_t n_bytes);
int rt_comm_destroy(rt_ctx* ctx);
int rt_combine(rt_ctx* ctx, int root);
int rt_combine_rgba8(rt_ctx* ctx, int root, uint8_t* rgba, size_t n_bytes);

/* The combined image of the last rt_combine_peers (on ctxs[0]) / rt_combine (at the root). Synchronous. */
int rt_read_combined(rt_ctx* ctx, float* rgba, size_t n_floats);          /* W*H*4 */
void* rt_combined_device_ptr(rt_ctx* ctx);                                /* float4[W*H], NULL before the first combine */
/* Dispatches in the context's running mean (wgsl:59-66): the weight its image enters a combine with. */
uint32_t rt_mean_count(const rt_ctx* ctx);

int rt_set_option(rt_ctx* ctx, int option, int64_t value);
int rt_get_option(const rt_ctx* ctx, int option, int64_t* value);

/* Read-out (no upstream counterpart: the texture is only ever sampled by render.wgsl).
 * Row 0 is the BOTTOM image row, x grows along camera `u`. Synchronous. */
int rt_read_accum(rt_ctx* ctx, float* rgba, size_t n_floats);          /* W*H*4 */
int rt_write_accum(rt_ctx* ctx, const float* rgba, size_t n_floats);   /* restore a saved image */
/* What shaders/render.wgsl:34-39 + the sRGB swapchain (context.rs:75-80) show: clamp to
 * [0,1], sRGB OETF, 8 bits per channel. */
int rt_read_rgba8(rt_ctx* ctx, uint8_t* rgba, size_t n_bytes);          /* W*H*4 */
/* Primary-hit AOV of the last render call made with RT_OPT_PRIMARY_IDS = 1: global
 * primitive order index (spheres, then triangles mesh by mesh) or 0xFFFFFFFF. */
int rt_read_primary_ids(rt_ctx* ctx, uint32_t* ids, size_t n);           /* W*H */

int rt_get_counters(rt_ctx* ctx, rt_counters* out);
int rt_reset_counters(rt_ctx* ctx);

/* Raw device access for host-side plumbing (NCCL combine, CUDA-event timing on the
 * stream the kernels are launched on). */
void* rt_accum_device_ptr(rt_ctx* ctx);     /* float4[W*H] */
void* rt_cuda_stream(rt_ctx* ctx);          /* cudaStream_t */
int rt_device(const rt_ctx* ctx);

/* FP32 roofline denominator: runs an FFMA-only microbenchmark (independent chains, every
 * SM busy) on `device` and returns the best achieved TFLOP/s (2 flop per FFMA) of `reps`
 * launches, with the duration of that launch. */
int rt_measure_fp32_peak(int device, int reps, double* tflops, double* seconds);

/* BVH introspection for tests: number of nodes, leaves, max depth of the last build. */
int rt_bvh_stats(rt_ctx* ctx, uint32_t* n_prims, uint32_t* n_nodes, uint32_t* max_depth);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
