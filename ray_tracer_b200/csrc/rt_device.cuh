// rt_device.cuh — device-side scene layout and small vector helpers.
//
// HBM layout (all arrays 16-byte aligned, read through the read-only path as 128-bit loads):
//
//  hot, one 64-byte record (4 x float4) per BVH slot k, primitives in tree order:
//    p[4k+0] = (n.x, n.y, n.z, id)       triangle: n = cross(e1, e2)        | sphere: (centre, id)
//    p[4k+1] = (A.x, A.y, A.z, e1.x)     A = a.pos + mesh.pos, e1 = B - A   | (radius, radius^2, 0, 0)
//    p[4k+2] = (e1.y, e1.z, e2.x, e2.y)  e2 = C - A
//    p[4k+3] = (e2.z, 0, 0, 0)
//    The first 16 bytes decide most tests alone: `id` tells the kinds apart (id < S) and n gives the
//    back-face reject of wgsl:131,143 (about half of all triangle tests); the record is one
//    64-byte-aligned pair of sectors, so the survivors' three further loads hit the same L1 line.
//    `id` = global primitive order index of the reference's brute-force loop
//    (wgsl:149-185): spheres 0..S-1 in array order, then triangles mesh by mesh; it is
//    the tie-break key.
//  cold, indexed by id, fetched once per path segment for the winning hit only:
//    c0[id] = (nA.xyz, material index)  | sphere: (centre, material index)
//    c1[id] = (nB.xyz, 0)  c2[id] = (nC.xyz, 0)
//    texture extension (allocated only when a mesh carries a texture handle):
//    c3[id] = (uvA.xy, uvB.xy)  c4[id] = (uvC.xy, e1.xy)  c5[id] = (e1.z, e2.xyz)
//  materials, indexed by material index (sphere i -> i, mesh m -> S + m):
//    m0 = color, m1 = emission_color * emission_strength,
//    m2 = (smoothness, diffuse texture handle, normal-map handle, 0) (handles as int bits, 0 = none)
//  textures: table of (cudaTextureObject_t, width, height); RGBA8 CUDA arrays fetched with
//    point sampling through the texture path, bilinear weights applied in the kernel so the
//    result is reproducible bit for bit (rt_shade.cuh).
//  BVH nodes, 64 B each (both children's boxes live in the parent, one fetch = two tests):
//    boxes as centre c / half extent h (rounded outward), which makes the slab test select-free;
//    the twelve numbers are ordered in the 64-bit pairs the packed FP32 instructions of sm_100
//    (FFMA2 / FADD2: two independent IEEE operations per issue slot) consume, see node_test():
//    n[4*i+0] = (L.c.x, L.c.y | L.c.z, R.c.z)  n[4*i+1] = (R.c.x, R.c.y | L.h.x, L.h.y)
//    n[4*i+2] = (L.h.z, R.h.z | R.h.x, R.h.y)  n[4*i+3] = (left, right, 0, 0) as int bits
//    child >= 0: inner node index; child < 0: leaf, ~child = (first_slot << 3) | (count - 1)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rt_math.cuh"

namespace rt {

struct TexDev { cudaTextureObject_t obj; uint32_t width, height; };

struct SceneDev {
    const float4* prims;
    const float4* iprims;  // instanced ranges: raw vertex triples in tree order (null without instancing)
    const float4* c0; const float4* c1; const float4* c2;
    const float4* c3; const float4* c4; const float4* c5;   // texture extension, may be null
    const struct TexDev* tex; uint32_t n_tex;
    const float4* m0; const float4* m1; const float4* m2;
    const float4* nodes;
    uint32_t n_prims;      // slots
    uint32_t n_spheres;    // ids below this are spheres
    int32_t root;          // root node index, or -1 when the scene is empty
    float max_abs;         // largest |coordinate| of the scene bounds
    float pad_scale;       // box padding = pad_scale * max(max_abs, |ray origin|)
    // packed nodes (flat scenes; 32 B per node, see node_test_packed): the 16-bit grid the box centres are stored on
    int32_t packed;        // 1: `nodes` holds packed nodes (2 float4 each); 0: centre / half-extent floats (4 float4 each)
    float3 grid_o, grid_c; // grid origin and cell size per axis
};

struct CameraDev {         // the fields of CameraUniform the shader reads (wgsl:69-80)
    float3 origin, llc, horizontal, vertical, u, v;
    float lens_radius;
};

// Exact unsigned 32-bit division by a launch constant (Granlund-Montgomery round-up multiplier):
// q = (t + ((n - t) >> s1)) >> s2 with t = umulhi(m, n); 5 instructions instead of the ~25 of a
// hardware-less integer division, and the item decode runs with few lanes active.
struct FastDiv {
    uint32_t d, m, s1, s2;
    __host__ __device__ __forceinline__ uint32_t div(uint32_t n) const
    {
        const uint32_t t = (uint32_t)(((unsigned long long)m * n) >> 32);    // IMAD.HI
        return (t + ((n - t) >> s1)) >> s2;
    }
};
inline FastDiv make_fastdiv(uint32_t d)
{
    FastDiv f; f.d = d ? d : 1u;
    uint32_t l = 0; while ((1ull << l) < f.d) l++;
    f.m = (uint32_t)((((1ull << l) - f.d) << 32) / f.d + 1ull);
    f.s1 = l < 1u ? l : 1u; f.s2 = l ? l - 1u : 0u;
    return f;
}

struct LaunchParams {
    uint32_t width, height;
    int32_t bounces, rpp, toggle;
    int32_t first_frame;       // params.frames of the first fused dispatch
    int32_t blend0;            // running-mean index of the first fused dispatch (this rank's own count): dispatch fi blends with weight 1/(blend0 + fi + 1) when blend0 + fi >= 1
    uint32_t n_frames;         // fused dispatches (frame stride = frame_world)
    uint32_t tile_rank, tile_world;
    uint32_t frame_world;      // stride between frames rendered by this rank
    uint32_t tiles_x, n_warp_tiles;  // 8x8 shard tiles per row; 8x4 warp tiles owned by this rank (2 per shard tile)
    int32_t seed_mode;
    uint32_t queue_chunk;      // items per grab while plenty remain (a multiple of 32)
    uint32_t queue_guide;      // guided self-scheduling of the item queue: grab = remaining / (warps x queue_guide), see rt_kernels.h
    int32_t sched_n_stay, sched_l_stay, sched_s_min;   // warp scheduler thresholds (lanes)
    FastDiv div_frames, div_tiles_x, div_width;         // item decode without integer divisions
    float rpp_scale;                                    // 1/rpp when rpp is a power of two (x/2^k == x*2^-k exactly), else 0
    // heaviest tiles first (rt_api.cu: tile order): the order in which this rank's 8x8 tiles enter the queue (local tile index -> local tile
    // index; null: raster order) and where the launch's first dispatch leaves every tile's segment count for the next order (or null)
    const uint32_t* tile_order;
    uint32_t* tile_cost;
};

// 256-bit read-only load (sm_100: LDG.E.256): two float4 from a 32-byte-aligned address in one
// request, i.e. half the L1 requests / data-pipe wavefronts of two 128-bit loads for the same sector.
__device__ __forceinline__ void ldg256(const float4* p, float4& a, float4& b)
{
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

// ---- packed binary32 pairs (sm_100a FFMA2 / FADD2): two independent, correctly rounded operations
// per instruction on an aligned 64-bit register pair; a scalar operand is broadcast by the hardware.
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) { return (f2_t)__float_as_uint(lo) | ((f2_t)__float_as_uint(hi) << 32); }
__device__ __forceinline__ float f2_lo(f2_t v) { return __uint_as_float((unsigned)(v & 0xffffffffull)); }
__device__ __forceinline__ float f2_hi(f2_t v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) { f2_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) { f2_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_sub(f2_t a, f2_t b) { f2_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
// the same 256-bit read-only load as ldg256, delivered as four 64-bit pairs
__device__ __forceinline__ void ldg256_f2(const float4* p, f2_t& a, f2_t& b, f2_t& c, f2_t& d)
{
    asm volatile("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}

// Per-ray constants of the slab test against boxes stored as centre c / half extent h: along an axis the ray is
// inside the padded slab for t in [a - b, a + b], a = c * (1/d) - o/d, b = (h + pad) * |1/d|.  Not parity-relevant
// (conservative culling only; DESIGN.md "Conservative culling").
struct RaySlab {
    f2_t inv_xy, noi_xy, ainv_xy, pi_xy;      // (1/d, -o/d, |1/d|, pad*|1/d|) for x and y
    float inv_z, noi_z, ainv_z, pi_z;         // ... and z (broadcast into both halves of a pair by the instruction)
};
__device__ __forceinline__ void slab_setup(RaySlab& s, float3 ro, float3 rd, float max_abs, float pad_scale)
{
    const float ix = __fdividef(1.0f, rd.x), iy = __fdividef(1.0f, rd.y), iz = __fdividef(1.0f, rd.z);
    const float pad = pad_scale * fmaxf(fmaxf(max_abs, fabsf(ro.x)), fmaxf(fabsf(ro.y), fabsf(ro.z)));
    const float ax = fabsf(ix), ay = fabsf(iy), az = fabsf(iz);
    s.inv_xy = f2_pack(ix, iy); s.inv_z = iz;
    s.noi_xy = f2_pack(-(ro.x * ix), -(ro.y * iy)); s.noi_z = -(ro.z * iz);
    s.ainv_xy = f2_pack(ax, ay); s.ainv_z = az;
    s.pi_xy = f2_pack(pad * ax, pad * ay); s.pi_z = pad * az;
}
// One node visit = both children's slab tests: 2 x LDG.256, 6 FFMA2 + 6 FADD2 (instead of 12 FFMA + 12 FADD),
// 4 two-input + 4 three-input min/max.  Entry / exit distances clamped to [0, cull]; a child is hit when tn <= tf.
__device__ __forceinline__ void node_test(const float4* np, const RaySlab& s, float cull,
                                          float& ltn, float& ltf, float& rtn, float& rtf, int& cl, int& cr)
{
    f2_t a0, a1, a2, a3, b0, b1, b2, b3;
    ldg256_f2(np, a0, a1, a2, a3); ldg256_f2(np + 2, b0, b1, b2, b3);
    cl = (int)(unsigned)(b2 & 0xffffffffull); cr = (int)(unsigned)(b2 >> 32);
#ifdef RT_SCALAR_SLABS                               // A/B build: the same test with scalar FFMA / FADD
    {
        const float ix = f2_lo(s.inv_xy), iy = f2_hi(s.inv_xy), nx = f2_lo(s.noi_xy), ny = f2_hi(s.noi_xy);
        const float ax = f2_lo(s.ainv_xy), ay = f2_hi(s.ainv_xy), px = f2_lo(s.pi_xy), py = f2_hi(s.pi_xy);
        const float lax = __fmaf_rn(f2_lo(a0), ix, nx), lay = __fmaf_rn(f2_hi(a0), iy, ny), laz = __fmaf_rn(f2_lo(a1), s.inv_z, s.noi_z);
        const float rax = __fmaf_rn(f2_lo(a2), ix, nx), ray_ = __fmaf_rn(f2_hi(a2), iy, ny), raz = __fmaf_rn(f2_hi(a1), s.inv_z, s.noi_z);
        const float lbx = __fmaf_rn(f2_lo(a3), ax, px), lby = __fmaf_rn(f2_hi(a3), ay, py), lbz = __fmaf_rn(f2_lo(b0), s.ainv_z, s.pi_z);
        const float rbx = __fmaf_rn(f2_lo(b1), ax, px), rby = __fmaf_rn(f2_hi(b1), ay, py), rbz = __fmaf_rn(f2_hi(b0), s.ainv_z, s.pi_z);
        ltn = fmaxf(fmaxf(lax - lbx, lay - lby), fmaxf(laz - lbz, 0.0f));
        ltf = fminf(fminf(lax + lbx, lay + lby), fminf(laz + lbz, cull));
        rtn = fmaxf(fmaxf(rax - rbx, ray_ - rby), fmaxf(raz - rbz, 0.0f));
        rtf = fminf(fminf(rax + rbx, ray_ + rby), fminf(raz + rbz, cull));
        return;
    }
#endif
    const f2_t inv_zz = f2_pack(s.inv_z, s.inv_z), noi_zz = f2_pack(s.noi_z, s.noi_z);
    const f2_t ainv_zz = f2_pack(s.ainv_z, s.ainv_z), pi_zz = f2_pack(s.pi_z, s.pi_z);
    const f2_t P0 = f2_fma(a0, s.inv_xy, s.noi_xy);      // (L.a.x, L.a.y)
    const f2_t P1 = f2_fma(a1, inv_zz, noi_zz);          // (L.a.z, R.a.z)
    const f2_t P2 = f2_fma(a2, s.inv_xy, s.noi_xy);      // (R.a.x, R.a.y)
    const f2_t Q0 = f2_fma(a3, s.ainv_xy, s.pi_xy);      // (L.b.x, L.b.y)
    const f2_t Q1 = f2_fma(b0, ainv_zz, pi_zz);          // (L.b.z, R.b.z)
    const f2_t Q2 = f2_fma(b1, s.ainv_xy, s.pi_xy);      // (R.b.x, R.b.y)
    const f2_t N0 = f2_sub(P0, Q0), F0 = f2_add(P0, Q0), N1 = f2_sub(P1, Q1), F1 = f2_add(P1, Q1), N2 = f2_sub(P2, Q2), F2 = f2_add(P2, Q2);
    ltn = fmaxf(fmaxf(f2_lo(N0), f2_hi(N0)), fmaxf(f2_lo(N1), 0.0f));
    ltf = fminf(fminf(f2_lo(F0), f2_hi(F0)), fminf(f2_lo(F1), cull));
    rtn = fmaxf(fmaxf(f2_lo(N2), f2_hi(N2)), fmaxf(f2_hi(N1), 0.0f));
    rtf = fminf(fminf(f2_lo(F2), f2_hi(F2)), fminf(f2_hi(F1), cull));
}

// ---- packed nodes: 32 bytes, ONE 256-bit load per node visit ------------------------------------------------------
// ncu on C3 (profiles/r2_summary.md): the L1 data pipe is the busiest unit of the kernel (79 % of its wavefront peak) and two
// thirds of its wavefronts are the node loads — every divergent lane costs a wavefront per LDG.256, and a 64-byte node
// is two of them.  A packed node is six words + two links: word (child, axis) = (half extent as bf16, rounded up) << 16 |
// (box centre on a 16-bit grid over the scene bounds).  The traversal uses the WORD ITSELF as the half extent (as a
// float it is the bf16 value plus at most 2^-7 of it: a slightly larger box, still conservative) and turns the centre
// into a float with one logic operation per pair: 0x4B000000 | q is the float 2^23 + q, and the grid's scale and offset
// (and the -2^23) are folded into the per-ray slab constants: a = (2^23 + q) * S + B with S = cell / d,
// B = (grid origin - o) / d - 2^23 * S.  So a node visit is 1 LDG.256 + 3 64-bit LOP + the same 6 FFMA2 + 6 FADD2.
// Error budget (all in units of one grid cell = scene extent / 65535): centre rounding 0.5, rounding of S times 2^23 0.5,
// rounding of B 0.5; the emitted half extent is taken about the grid centre and enlarged by 1.25 cells + bf16 round-up
// (rt_bvh.cu pack_axis).  Not parity-relevant: conservative culling only, the exact tests decide.
__device__ __forceinline__ void slab_setup_packed(RaySlab& s, float3 ro, float3 rd, float max_abs, float pad_scale, float3 go, float3 gc)
{
    const float ix = __fdividef(1.0f, rd.x), iy = __fdividef(1.0f, rd.y), iz = __fdividef(1.0f, rd.z);
    const float pad = pad_scale * fmaxf(fmaxf(max_abs, fabsf(ro.x)), fmaxf(fabsf(ro.y), fabsf(ro.z)));
    const float ax = fabsf(ix), ay = fabsf(iy), az = fabsf(iz);
    const float sx = gc.x * ix, sy = gc.y * iy, sz = gc.z * iz;
    s.inv_xy = f2_pack(sx, sy); s.inv_z = sz;                                                   // S
    s.noi_xy = f2_pack(__fmaf_rn(-8388608.0f, sx, (go.x - ro.x) * ix), __fmaf_rn(-8388608.0f, sy, (go.y - ro.y) * iy));
    s.noi_z = __fmaf_rn(-8388608.0f, sz, (go.z - ro.z) * iz);                                   // B
    s.ainv_xy = f2_pack(ax, ay); s.ainv_z = az;
    s.pi_xy = f2_pack(pad * ax, pad * ay); s.pi_z = pad * az;
}
// 2^23 + q per half: (word & 0xffff) | 0x4B000000.  `magic` = 0x4B000000 must sit in a register for this to be ONE logic
// operation per word (an instruction takes a single immediate): node_magic() hands it out through an opaque asm so that
// the compiler keeps it instead of rebuilding it at every use.
__device__ __forceinline__ unsigned node_magic() { unsigned m; asm volatile("mov.u32 %0, 0x4b000000;" : "=r"(m)); return m; }
__device__ __forceinline__ f2_t f2_magic(f2_t w, unsigned magic)
{
    unsigned lo, hi;
    asm("lop3.b32 %0, %1, 0xffff, %2, 0xea;" : "=r"(lo) : "r"((unsigned)(w & 0xffffffffull)), "r"(magic));
    asm("lop3.b32 %0, %1, 0xffff, %2, 0xea;" : "=r"(hi) : "r"((unsigned)(w >> 32)), "r"(magic));
    return (f2_t)lo | ((f2_t)hi << 32);
}
// np: the node's two float4 = words (L.x, L.y | L.z, R.z) (R.x, R.y | left, right)
__device__ __forceinline__ void node_test_packed(const float4* np, const RaySlab& s, float cull, unsigned magic,
                                                 float& ltn, float& ltf, float& rtn, float& rtf, int& cl, int& cr)
{
    f2_t h0, h1, h2, lk;
    ldg256_f2(np, h0, h1, h2, lk);
    cl = (int)(unsigned)(lk & 0xffffffffull); cr = (int)(unsigned)(lk >> 32);
    const f2_t c0 = f2_magic(h0, magic), c1 = f2_magic(h1, magic), c2 = f2_magic(h2, magic);
    const f2_t inv_zz = f2_pack(s.inv_z, s.inv_z), noi_zz = f2_pack(s.noi_z, s.noi_z);
    const f2_t ainv_zz = f2_pack(s.ainv_z, s.ainv_z), pi_zz = f2_pack(s.pi_z, s.pi_z);
    const f2_t P0 = f2_fma(c0, s.inv_xy, s.noi_xy);      // (L.a.x, L.a.y)
    const f2_t P1 = f2_fma(c1, inv_zz, noi_zz);          // (L.a.z, R.a.z)
    const f2_t P2 = f2_fma(c2, s.inv_xy, s.noi_xy);      // (R.a.x, R.a.y)
    const f2_t Q0 = f2_fma(h0, s.ainv_xy, s.pi_xy);      // (L.b.x, L.b.y): the words themselves as half extents
    const f2_t Q1 = f2_fma(h1, ainv_zz, pi_zz);          // (L.b.z, R.b.z)
    const f2_t Q2 = f2_fma(h2, s.ainv_xy, s.pi_xy);      // (R.b.x, R.b.y)
    const f2_t N0 = f2_sub(P0, Q0), F0 = f2_add(P0, Q0), N1 = f2_sub(P1, Q1), F1 = f2_add(P1, Q1), N2 = f2_sub(P2, Q2), F2 = f2_add(P2, Q2);
    ltn = fmaxf(fmaxf(f2_lo(N0), f2_hi(N0)), fmaxf(f2_lo(N1), 0.0f));
    ltf = fminf(fminf(f2_lo(F0), f2_hi(F0)), fminf(f2_lo(F1), cull));
    rtn = fmaxf(fmaxf(f2_lo(N2), f2_hi(N2)), fmaxf(f2_hi(N1), 0.0f));
    rtf = fminf(fminf(f2_lo(F2), f2_hi(F2)), fminf(f2_hi(F1), cull));
}

// ---- vector helpers (the fused forms are the arithmetic contract's) -----
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float4 operator*(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

__device__ __forceinline__ float dot3(float3 a, float3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ float3 cross3(float3 a, float3 b)
{
    return f3(__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ float3 fma3(float3 a, float s, float3 c)
{
    return f3(__fmaf_rn(a.x, s, c.x), __fmaf_rn(a.y, s, c.y), __fmaf_rn(a.z, s, c.z));
}
__device__ __forceinline__ float length3(float3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ float3 normalize3(float3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return a * inv; }
__device__ __forceinline__ float3 reflect3(float3 i, float3 n) { float t = 2.0f * dot3(n, i); return fma3(n, -t, i); }
__device__ __forceinline__ float mix1(float a, float b, float t) { return __fmaf_rn(b, t, a * (1.0f - t)); }
__device__ __forceinline__ float sign1(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }
__device__ __forceinline__ float smoothstep1(float e0, float e1, float x)
{
    float t = fminf(fmaxf((x - e0) / (e1 - e0), 0.0f), 1.0f);
    return (t * t) * (3.0f - 2.0f * t);
}
__device__ __forceinline__ float3 xyz(float4 v) { return f3(v.x, v.y, v.z); }

// ---- RNG: PCG-RXS-M-XS-32 hash stream (wgsl:205-210), rand in [0,1] (wgsl:187-189) -----
__device__ __forceinline__ uint32_t next_random_number(uint32_t& s)
{
    s = s * 747796405u + 2891336453u;
    uint32_t r = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
    return (r >> 22u) ^ r;
}
__device__ __forceinline__ float rand01(uint32_t& s)
{
    return __uint2float_rn(next_random_number(s)) * 2.3283064365386963e-10f; // exact scaling by 2^-32
}

} // namespace rt
