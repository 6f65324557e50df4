// rt_bvh.cu — scene repack (AoS records -> SoA 128-bit slots) and GPU LBVH build.
//
// No reference counterpart: upstream scans every sphere and triangle per segment
// (shaders/ray_tracer.wgsl:149-185).  The build is result-invisible by construction: the
// values the exact tests consume (A = a.pos + mesh.pos, e1 = B - A, e2 = C - A, r*r) are
// produced here with the same single f32 operations the reference performs per test
// (wgsl:172-174, 127-128, 111), and the hierarchy only decides which tests are skipped.
//
// Pipeline (all on the GPU): primitive setup + bounds -> scene bounds (atomic min/max) ->
// 63-bit Morton keys of box centres -> radix sort (CUB) -> hierarchy over the Morton order ->
// node emission with both child boxes in the parent and subtrees of <= leaf_size primitives
// collapsed into one leaf.  Two hierarchy builders over the same sorted keys:
//   * PLOC (default; Meister & Bittner 2018): clusters start as the primitives in Morton
//     order; every round each cluster looks `radius` neighbours to either side for the one
//     whose union with it has the smallest surface area, mutual nearest neighbours merge,
//     a prefix sum compacts the survivors in order.  Costs ~40-50 rounds of four small
//     kernels and buys ~18 % fewer node visits per ray than the radix tree on the teapot
//     (scripts/bvh_proto.cpp is the CPU study behind the choice).
//   * radix tree (Karras 2012) + bottom-up refit with arrival flags: one pass, the fallback
//     when a PLOC tree comes out deeper than the traversal stack.
// Either way the primitives are then laid out so that every subtree owns a contiguous slot
// range (for PLOC: slot = number of primitives to the left in the final tree).
#include <cub/cub.cuh>
#include <vector>
#include "rt_kernels.h"

namespace rt {

namespace {

struct RawScene {
    const rt_sphere* spheres; const rt_vertex* vertices; const uint32_t* indices; const rt_mesh* meshes;
    const uint32_t* tri_start;      // [nm+1] prefix sum of mesh.triangles: ids of the reference's scan order
    uint32_t ns, nm, nt;            // nt = triangles that get their own (flattened) record
    // Instancing (two-level hierarchy): meshes whose (first, triangles, offset) range is shared by several Mesh records are not
    // flattened; each such record becomes ONE proxy primitive of the top level that points at the range's object-space hierarchy.
    const uint32_t* flat_start;     // [nm+1] prefix sum of mesh.triangles over the flattened meshes only (== tri_start without instancing)
    const uint32_t* inst_mesh;      // [n_inst] mesh index of every proxy
    const int* range_of_mesh;       // [nm] -1: flattened, else index into range_info
    const float4* range_info;       // 3 per range: object-space bounds lo, hi, (root node, first cold index, 0, 0) as int bits
    uint32_t n_inst;
};
// Top-level primitive index t: [0, ns) spheres, [ns, ns + nt) flattened triangles, then the proxies.  hot / cold / bounds arrays
// are indexed by t; the `id` stored in a hot record is the GLOBAL primitive order index (the tie-break key and the AOV).

__device__ __forceinline__ int fenc(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float fdec(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// one thread per top-level primitive: hot/cold records in index order, exact bounds, scene bounds
__global__ void setup_prims(RawScene rs, float4* hot /*4 per index*/, float4* c0, float4* c1, float4* c2, float4* c3, float4* c4, float4* c5,
                            float4* bmin, float4* bmax, int* scene_bounds /*[6] encoded*/)
{
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;             // top-level index t
    uint32_t n = rs.ns + rs.nt + rs.n_inst;
    float3 lo = f3(INFINITY, INFINITY, INFINITY), hi = f3(-INFINITY, -INFINITY, -INFINITY);
    if (id < n) {
        if (id < rs.ns) {
            const rt_sphere& s = rs.spheres[id];
            float r = s.radius;
            float3 c = f3(s.position[0], s.position[1], s.position[2]);
            hot[4 * (size_t)id + 0] = make_float4(c.x, c.y, c.z, __uint_as_float(id));
            hot[4 * (size_t)id + 1] = make_float4(r, r * r, 0.f, 0.f);
            hot[4 * (size_t)id + 2] = hot[4 * (size_t)id + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
            c0[id] = make_float4(c.x, c.y, c.z, __uint_as_float(id));
            c1[id] = c2[id] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c3) c3[id] = c4[id] = c5[id] = make_float4(0.f, 0.f, 0.f, 0.f);
            float ar = fabsf(r);
            lo = f3(c.x - ar, c.y - ar, c.z - ar); hi = f3(c.x + ar, c.y + ar, c.z + ar);
        } else if (id < rs.ns + rs.nt) {
            uint32_t t = id - rs.ns;
            uint32_t a = 0, b = rs.nm;               // flattened mesh m with flat_start[m] <= t < flat_start[m+1] (the last such: empty ranges are skipped)
            while (b - a > 1) { uint32_t mid = (a + b) >> 1; if (rs.flat_start[mid] <= t) a = mid; else b = mid; }
            const rt_mesh& m = rs.meshes[a];
            uint32_t i = t - rs.flat_start[a];
            const uint32_t gid = rs.ns + rs.tri_start[a] + i;         // position in the reference's scan order (wgsl:149-185)
            const rt_vertex& va = rs.vertices[m.offset + rs.indices[m.first + i * 3u]];
            const rt_vertex& vb = rs.vertices[m.offset + rs.indices[m.first + i * 3u + 1u]];
            const rt_vertex& vc = rs.vertices[m.offset + rs.indices[m.first + i * 3u + 2u]];
            float3 mp = f3(m.pos[0], m.pos[1], m.pos[2]);
            float3 A = f3(va.pos[0], va.pos[1], va.pos[2]) + mp;       // wgsl:172-174
            float3 B = f3(vb.pos[0], vb.pos[1], vb.pos[2]) + mp;
            float3 C = f3(vc.pos[0], vc.pos[1], vc.pos[2]) + mp;
            float3 e1 = B - A, e2 = C - A;                              // wgsl:127-128
            float3 nrm = cross3(e1, e2);                                // wgsl:129
            hot[4 * (size_t)id + 0] = make_float4(nrm.x, nrm.y, nrm.z, __uint_as_float(gid));
            hot[4 * (size_t)id + 1] = make_float4(A.x, A.y, A.z, e1.x);
            hot[4 * (size_t)id + 2] = make_float4(e1.y, e1.z, e2.x, e2.y);
            hot[4 * (size_t)id + 3] = make_float4(e2.z, __uint_as_float(id), 0.f, 0.f);     // .y: index of the cold records
            c0[id] = make_float4(va.normal[0], va.normal[1], va.normal[2], __uint_as_float(rs.ns + a));
            c1[id] = make_float4(vb.normal[0], vb.normal[1], vb.normal[2], 0.f);
            c2[id] = make_float4(vc.normal[0], vc.normal[1], vc.normal[2], 0.f);
            if (c3) {                                                   // texture extension: uvs + edges for the tangent frame
                c3[id] = make_float4(va.tu, va.tv, vb.tu, vb.tv);
                c4[id] = make_float4(vc.tu, vc.tv, e1.x, e1.y);
                c5[id] = make_float4(e1.z, e2.x, e2.y, e2.z);
            }
            lo = f3(fminf(A.x, fminf(B.x, C.x)), fminf(A.y, fminf(B.y, C.y)), fminf(A.z, fminf(B.z, C.z)));
            hi = f3(fmaxf(A.x, fmaxf(B.x, C.x)), fmaxf(A.y, fmaxf(B.y, C.y)), fmaxf(A.z, fmaxf(B.z, C.z)));
        } else {
            // proxy of one instance: (pos, 0x80000000 | k), (root node of the range's hierarchy, first global id, first cold index, material)
            const uint32_t k = id - rs.ns - rs.nt;
            const uint32_t mi = rs.inst_mesh[k];
            const rt_mesh& m = rs.meshes[mi];
            const int g = rs.range_of_mesh[mi];
            const float4 rlo = rs.range_info[3 * g + 0], rhi = rs.range_info[3 * g + 1], ri = rs.range_info[3 * g + 2];
            hot[4 * (size_t)id + 0] = make_float4(m.pos[0], m.pos[1], m.pos[2], __uint_as_float(0x80000000u | k));
            hot[4 * (size_t)id + 1] = make_float4(ri.x, __uint_as_float(rs.ns + rs.tri_start[mi]), ri.y, __uint_as_float(rs.ns + mi));
            hot[4 * (size_t)id + 2] = hot[4 * (size_t)id + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
            c0[id] = c1[id] = c2[id] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c3) c3[id] = c4[id] = c5[id] = make_float4(0.f, 0.f, 0.f, 0.f);
            // every vertex the shader forms is fl(a + pos) with a inside [rlo, rhi]: rounding is monotonic, so the directed sums bound it
            lo = f3(__fadd_rd(rlo.x, m.pos[0]), __fadd_rd(rlo.y, m.pos[1]), __fadd_rd(rlo.z, m.pos[2]));
            hi = f3(__fadd_ru(rhi.x, m.pos[0]), __fadd_ru(rhi.y, m.pos[1]), __fadd_ru(rhi.z, m.pos[2]));
        }
        bmin[id] = make_float4(lo.x, lo.y, lo.z, 0.f);
        bmax[id] = make_float4(hi.x, hi.y, hi.z, 0.f);
    }
    // warp-reduce then one atomic per warp; NaN bounds are ignored by fminf/fmaxf
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, o)); hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, o));
        hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomicMin(scene_bounds + 0, fenc(lo.x)); atomicMin(scene_bounds + 1, fenc(lo.y)); atomicMin(scene_bounds + 2, fenc(lo.z));
        atomicMax(scene_bounds + 3, fenc(hi.x)); atomicMax(scene_bounds + 4, fenc(hi.y)); atomicMax(scene_bounds + 5, fenc(hi.z));
    }
}

__global__ void setup_materials(RawScene rs, float4* m0, float4* m1, float4* m2)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rs.ns + rs.nm) return;
    const float* col; const float* emi; float str, smooth; uint32_t dtex = 0, ntex = 0;
    if (i < rs.ns) { const rt_sphere& s = rs.spheres[i]; col = s.color; emi = s.emission_color; str = s.emission_strength; smooth = s.smoothness; }
    else { const rt_mesh& m = rs.meshes[i - rs.ns]; col = m.color; emi = m.emission_color; str = m.emission_strength; smooth = m.specular; dtex = m.diffuse_tex; ntex = m.normal_tex; }
    m0[i] = make_float4(col[0], col[1], col[2], col[3]);
    m1[i] = make_float4(emi[0] * str, emi[1] * str, emi[2] * str, emi[3] * str);   // wgsl:271
    m2[i] = make_float4(smooth, __uint_as_float(dtex), __uint_as_float(ntex), 0.f);
}

__device__ __forceinline__ uint64_t spread21(uint32_t v)
{
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void morton_keys(uint32_t n, const float4* bmin, const float4* bmax, const int* scene_bounds,
                            uint64_t* keys, uint32_t* ids)
{
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    float3 slo = f3(fdec(scene_bounds[0]), fdec(scene_bounds[1]), fdec(scene_bounds[2]));
    float3 shi = f3(fdec(scene_bounds[3]), fdec(scene_bounds[4]), fdec(scene_bounds[5]));
    float4 lo = bmin[id], hi = bmax[id];
    float cx = 0.5f * lo.x + 0.5f * hi.x, cy = 0.5f * lo.y + 0.5f * hi.y, cz = 0.5f * lo.z + 0.5f * hi.z;
    float ex = fmaxf(shi.x - slo.x, 1e-30f), ey = fmaxf(shi.y - slo.y, 1e-30f), ez = fmaxf(shi.z - slo.z, 1e-30f);
    float e = fmaxf(ex, fmaxf(ey, ez));              // cubic cells: same resolution on every axis
    float qx = fminf(fmaxf((cx - slo.x) / e, 0.f), 1.f), qy = fminf(fmaxf((cy - slo.y) / e, 0.f), 1.f), qz = fminf(fmaxf((cz - slo.z) / e, 0.f), 1.f);
    if (!(qx == qx)) qx = 0.f; if (!(qy == qy)) qy = 0.f; if (!(qz == qz)) qz = 0.f;
    uint32_t ix = (uint32_t)fminf(qx * 2097152.0f, 2097151.0f), iy = (uint32_t)fminf(qy * 2097152.0f, 2097151.0f), iz = (uint32_t)fminf(qz * 2097152.0f, 2097151.0f);
    keys[id] = (spread21(ix) << 2) | (spread21(iy) << 1) | spread21(iz);
    ids[id] = id;
}

// primitive boxes in Morton order (what both hierarchy builders consume)
__global__ void gather_boxes(uint32_t n, const uint32_t* sorted_ids, const float4* bmin, const float4* bmax, float4* lmin, float4* lmax)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t id = sorted_ids[k];
    lmin[k] = bmin[id]; lmax[k] = bmax[id];
}

// hot records into their final slots (Morton slot k -> new_slot[k])
__global__ void gather_slots(uint32_t n, const uint32_t* sorted_ids, const uint32_t* new_slot, const float4* hot, float4* prims)
{
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;     // one thread per float4 of a record
    uint32_t k = t >> 2, q = t & 3u;
    if (k >= n) return;
    uint32_t id = sorted_ids[k], s = new_slot[k];
    prims[4 * (size_t)s + q] = hot[4 * (size_t)id + q];
}

__device__ __forceinline__ int delta(const uint64_t* keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// Karras 2012: inner node i covers sorted range [first,last]; children encoded as
// index (inner) or ~slot (leaf); parents recorded for the refit.
__global__ void build_radix_tree(int n, const uint64_t* keys, int2* children, int2* ranges, int* parent_inner, int* parent_leaf)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    int gamma = i + s * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    int left = (first == gamma) ? ~gamma : gamma;
    int right = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    ranges[i] = make_int2(first, last);
    if (left >= 0) parent_inner[left] = i; else parent_leaf[~left] = i;
    if (right >= 0) parent_inner[right] = i; else parent_leaf[~right] = i;
    if (i == 0) parent_inner[0] = -1;
}

// bottom-up boxes of inner nodes; the second thread to arrive at a node owns it
__global__ void refit(int n, const int2* children, const int* parent_inner, const int* parent_leaf,
                      const float4* lmin, const float4* lmax, float4* imin, float4* imax, int* flags, int* max_depth)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = parent_leaf[k];
    int depth = 1;
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(flags + node, 1) == 0) { depth = 0; break; }
        int2 ch = children[node];
        float4 alo = ch.x >= 0 ? ((volatile float4*)imin)[ch.x] : lmin[~ch.x], ahi = ch.x >= 0 ? ((volatile float4*)imax)[ch.x] : lmax[~ch.x];
        float4 blo = ch.y >= 0 ? ((volatile float4*)imin)[ch.y] : lmin[~ch.y], bhi = ch.y >= 0 ? ((volatile float4*)imax)[ch.y] : lmax[~ch.y];
        imin[node] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
        imax[node] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
        node = parent_inner[node];
        depth++;
    }
    (void)depth;
    // depth of this leaf (independent climb; cheap, log n)
    int dd = 0;
    for (int p = parent_leaf[k]; p >= 0; p = parent_inner[p]) dd++;
    atomicMax(max_depth, dd);
}

// ---- PLOC ---------------------------------------------------------------------------------
// cluster i of a round: cid[i] (inner node index, or ~morton_slot for a primitive), box clo/chi
// (clo.w carries the primitive count as int bits).
__global__ void ploc_init(int n, const float4* lmin, const float4* lmax, int* cid, float4* clo, float4* chi)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    float4 lo = lmin[k], hi = lmax[k];
    cid[k] = ~k;
    clo[k] = make_float4(lo.x, lo.y, lo.z, __int_as_float(1));
    chi[k] = hi;
}

__device__ __forceinline__ float union_area(float4 alo, float4 ahi, float4 blo, float4 bhi)
{
    float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y), dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return dx * dy + dy * dz + dz * dx;
}

#define PLOC_THREADS 256
#define PLOC_MAX_RADIUS 16
#define PLOC_ROUNDS_PER_CHECK 8
// nearest neighbour inside the window; ties: the parity partner i^1 first (so that runs of
// identical boxes pair up), then the closer index, then the smaller one.  NaN areas never win.
// The round loop runs from device-side state st = {clusters left, nodes created, force flag}: the host
// launches rounds in batches for an upper bound of the cluster count and reads the state back once
// per batch, so an upload costs a handful of host round trips instead of one per round; rounds
// launched after the last merge return at once.
__global__ void ploc_nn(const int* __restrict__ st, int radius, const float4* __restrict__ clo, const float4* __restrict__ chi, int* __restrict__ nn)
{
    __shared__ float4 slo[PLOC_THREADS + 2 * PLOC_MAX_RADIUS], shi[PLOC_THREADS + 2 * PLOC_MAX_RADIUS];
    const int m = st[0], force_pairs = st[2];
    if (m <= st[4] || (int)(blockIdx.x * PLOC_THREADS) >= m) return;  // block-uniform; st[4]: cluster count at which ploc_tail takes over (1 = never)
    const int base = blockIdx.x * PLOC_THREADS - radius;
    for (int t = threadIdx.x; t < PLOC_THREADS + 2 * radius; t += PLOC_THREADS) {
        int g = base + t;
        if (g >= 0 && g < m) { slo[t] = clo[g]; shi[t] = chi[g]; }
    }
    __syncthreads();
    const int i = blockIdx.x * PLOC_THREADS + threadIdx.x;
    if (i >= m) return;
    int partner = i ^ 1;
    if (partner >= m) partner = i - 1;                 // m >= 2 here
    if (force_pairs) { nn[i] = ((i ^ 1) < m) ? (i ^ 1) : -1; return; }
    const float4 alo = slo[threadIdx.x + radius], ahi = shi[threadIdx.x + radius];
    float best = INFINITY; int bj = partner; int brank = 0x7fffffff;
    const int j0 = max(0, i - radius), j1 = min(m - 1, i + radius);
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        const int t = j - base;
        const float a = union_area(alo, ahi, slo[t], shi[t]);
        const int rank = (j == (i ^ 1)) ? 0 : 2 * abs(j - i) + (j > i ? 1 : 0);
        if (a < best || (a == best && rank < brank)) { best = a; bj = j; brank = rank; }
    }
    nn[i] = bj;
}

// keep flag (high word) and merge flag (low word) of every cluster, for one 64-bit prefix sum
__global__ void ploc_flags(const int* __restrict__ st, const int* __restrict__ nn, unsigned long long* __restrict__ flags)
{
    const int m = st[0];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= st[4] || i >= m) return;
    int j = nn[i];
    bool mutual = j >= 0 && nn[j] == i;
    unsigned long long keep = (!mutual || i < j) ? 1ull : 0ull, merge = (mutual && i < j) ? 1ull : 0ull;
    flags[i] = (keep << 32) | merge;
}

// parents are stored as (node << 1) | is_right_child, -1 for the root
__global__ void ploc_scatter(const int* __restrict__ st, const int* __restrict__ nn, const unsigned long long* __restrict__ pos,
                             const int* __restrict__ cid, const float4* __restrict__ clo, const float4* __restrict__ chi,
                             int* __restrict__ cid2, float4* __restrict__ clo2, float4* __restrict__ chi2,
                             int2* __restrict__ children, float4* __restrict__ imin, float4* __restrict__ imax,
                             int* __restrict__ par_inner, int* __restrict__ par_leaf, unsigned int* __restrict__ totals)
{
    const int m = st[0], node_base = st[1];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= st[4] || i >= m) return;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    const unsigned long long p = pos[i];
    const unsigned int out = (unsigned int)(p >> 32), mrg = (unsigned int)(p & 0xffffffffu);
    if (!mutual) {
        cid2[out] = cid[i]; clo2[out] = clo[i]; chi2[out] = chi[i];
    } else if (i < j) {
        const int node = node_base + (int)mrg;
        const int a = cid[i], b = cid[j];
        const float4 alo = clo[i], ahi = chi[i], blo = clo[j], bhi = chi[j];
        const int cnt = __float_as_int(alo.w) + __float_as_int(blo.w);
        const float4 lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(cnt));
        const float4 hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
        children[node] = make_int2(a, b);
        imin[node] = lo; imax[node] = hi;
        if (a >= 0) par_inner[a] = node << 1; else par_leaf[~a] = node << 1;
        if (b >= 0) par_inner[b] = (node << 1) | 1; else par_leaf[~b] = (node << 1) | 1;
        cid2[out] = node; clo2[out] = lo; chi2[out] = hi;
    }
    if (i == m - 1) {
        const bool keep = !mutual || i < j, merge = mutual && i < j;
        totals[0] = out + (keep ? 1u : 0u);
        totals[1] = mrg + (merge ? 1u : 0u);
    }
}

// end of a round: survivors become the cluster list, new nodes are counted, a round without a mutual
// pair (only possible with tied areas) forces neighbour pairs next time
__global__ void ploc_advance(int* st, const unsigned int* __restrict__ totals)
{
    if (st[0] <= st[4]) return;
    st[1] += (int)totals[1];
    st[2] = totals[1] == 0u;
    st[0] = (int)totals[0];
    st[3] += 1;
}

// The last rounds in ONE block: once the cluster list fits PLOC_TAIL threads it lives in shared memory and the rounds
// (same neighbour search, same mutual-pair rule, same numbering of survivors and new nodes as ploc_nn / ploc_flags /
// scan / ploc_scatter / ploc_advance, so the tree is the same) run back to back with block barriers between them instead
// of five launches each: the teapot's ~40 rounds are ~6 launched rounds + this kernel.
#define PLOC_TAIL 1024
__global__ void __launch_bounds__(PLOC_TAIL) ploc_tail(int* st, int radius, const int* __restrict__ cid, const float4* __restrict__ clo, const float4* __restrict__ chi,
                                                       int2* __restrict__ children, float4* __restrict__ imin, float4* __restrict__ imax,
                                                       int* __restrict__ par_inner, int* __restrict__ par_leaf)
{
    __shared__ float4 slo[PLOC_TAIL], shi[PLOC_TAIL];
    __shared__ int scid[PLOC_TAIL], snn[PLOC_TAIL];
    __shared__ unsigned int swarp[PLOC_TAIL / 32];
    __shared__ int s_m, s_base, s_force, s_rounds;
    const int i = threadIdx.x, lane = i & 31, warp = i >> 5;
    if (i == 0) { s_m = st[0]; s_base = st[1]; s_force = st[2]; s_rounds = 0; }
    __syncthreads();
    if (s_m > PLOC_TAIL) return;
    if (i < s_m) { scid[i] = cid[i]; slo[i] = clo[i]; shi[i] = chi[i]; }
    __syncthreads();
    while (s_m > 1) {
        const int m = s_m, node_base = s_base, force_pairs = s_force;
        // nearest neighbour (ploc_nn)
        if (i < m) {
            int bj;
            if (force_pairs) bj = ((i ^ 1) < m) ? (i ^ 1) : -1;
            else {
                int partner = i ^ 1;
                if (partner >= m) partner = i - 1;
                const float4 alo = slo[i], ahi = shi[i];
                float best = INFINITY; bj = partner; int brank = 0x7fffffff;
                const int j0 = max(0, i - radius), j1 = min(m - 1, i + radius);
                for (int j = j0; j <= j1; j++) {
                    if (j == i) continue;
                    const float a = union_area(alo, ahi, slo[j], shi[j]);
                    const int rank = (j == (i ^ 1)) ? 0 : 2 * abs(j - i) + (j > i ? 1 : 0);
                    if (a < best || (a == best && rank < brank)) { best = a; bj = j; brank = rank; }
                }
            }
            snn[i] = bj;
        }
        __syncthreads();
        // keep / merge flags and their exclusive prefix sums (ploc_flags + scan): keep in the high half, merge in the low half
        int j = -1; bool mutual = false;
        unsigned int flag = 0u;
        if (i < m) {
            j = snn[i];
            mutual = j >= 0 && snn[j] == i;
            flag = ((!mutual || i < j) ? 0x10000u : 0u) | ((mutual && i < j) ? 1u : 0u);
        }
        unsigned int incl = flag;
        for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) swarp[warp] = incl;
        // what this thread will write, read before anything is overwritten
        int a = 0, b = 0; float4 alo = make_float4(0.f, 0.f, 0.f, 0.f), ahi = alo, blo = alo, bhi = alo;
        if (i < m) { a = scid[i]; alo = slo[i]; ahi = shi[i]; if (mutual && i < j) { b = scid[j]; blo = slo[j]; bhi = shi[j]; } }
        __syncthreads();
        if (warp == 0) {
            unsigned int w = swarp[lane], winc = w;
            for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += t; }
            swarp[lane] = winc - w;                       // exclusive offset of every warp
            if (lane == 31) { s_m = (int)(winc >> 16); s_base = node_base + (int)(winc & 0xffffu); s_force = (winc & 0xffffu) == 0u; s_rounds++; }
        }
        __syncthreads();
        if (i < m) {
            const unsigned int excl = swarp[warp] + incl - flag;
            const unsigned int out = excl >> 16, mrg = excl & 0xffffu;
            if (!mutual) {
                scid[out] = a; slo[out] = alo; shi[out] = ahi;
            } else if (i < j) {                            // ploc_scatter
                const int node = node_base + (int)mrg;
                const int cnt = __float_as_int(alo.w) + __float_as_int(blo.w);
                const float4 lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(cnt));
                const float4 hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
                children[node] = make_int2(a, b);
                imin[node] = lo; imax[node] = hi;
                if (a >= 0) par_inner[a] = node << 1; else par_leaf[~a] = node << 1;
                if (b >= 0) par_inner[b] = (node << 1) | 1; else par_leaf[~b] = (node << 1) | 1;
                scid[out] = node; slo[out] = lo; shi[out] = hi;
            }
        }
        __syncthreads();
    }
    if (i == 0) { st[0] = s_m; st[1] = s_base; st[2] = s_force; st[3] += s_rounds; }
}

// Final primitive layout of a PLOC tree: a node's first slot = number of primitives to its
// left = sum of the left siblings' counts over the ancestors entered through their right link.
__global__ void ploc_layout(int n, int root, const int2* __restrict__ children, const float4* __restrict__ imin,
                            const int* __restrict__ par_inner, const int* __restrict__ par_leaf,
                            uint32_t* __restrict__ new_slot, int2* __restrict__ ranges, int* __restrict__ max_depth)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * n - 1) return;
    const bool is_leaf = t < n;
    const int self = is_leaf ? t : t - n;             // morton slot | inner node
    int p = is_leaf ? par_leaf[self] : (self == root ? -1 : par_inner[self]);
    int off = 0, depth = 0;
    while (p >= 0) {
        const int node = p >> 1;
        if (p & 1) { int l = children[node].x; off += (l < 0) ? 1 : __float_as_int(imin[l].w); }
        depth++;
        p = (node == root) ? -1 : par_inner[node];
    }
    if (is_leaf) { new_slot[self] = (uint32_t)off; atomicMax(max_depth, depth); }
    else ranges[self] = make_int2(off, off + __float_as_int(imin[self].w) - 1);
}

// Bottom-up boxes over a kept topology (either builder): parents are stored shifted left by
// `shift` (PLOC keeps the is-right-child bit below the index); the second thread to arrive at
// a node owns it; the climb ends at `root`, whose parent word is never read.
__global__ void refit_boxes(int n, int root, int shift, const int2* children, const int* pin, const int* plf,
                            const float4* lmin, const float4* lmax, float4* imin, float4* imax, int* flags)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = plf[k] >> shift;
    for (;;) {
        __threadfence();
        if (atomicAdd(flags + node, 1) == 0) break;
        int2 ch = children[node];
        float4 alo = ch.x >= 0 ? ((volatile float4*)imin)[ch.x] : lmin[~ch.x], ahi = ch.x >= 0 ? ((volatile float4*)imax)[ch.x] : lmax[~ch.x];
        float4 blo = ch.y >= 0 ? ((volatile float4*)imin)[ch.y] : lmin[~ch.y], bhi = ch.y >= 0 ? ((volatile float4*)imax)[ch.y] : lmax[~ch.y];
        imin[node] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
        imax[node] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
        if (node == root) break;
        node = pin[node] >> shift;
    }
}

__global__ void identity_slots(uint32_t n, uint32_t* new_slot)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) new_slot[k] = k;
}

// Boxes are emitted as centre / half extent (the traversal's slab test is then select-free).
// The half extent is rounded outward so that [c - h, c + h] contains [lo, hi].
__device__ __forceinline__ float3 box_centre(float4 lo, float4 hi) { return f3(0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z); }
__device__ __forceinline__ float half_up(float lo, float hi, float c)
{
    float h = fmaxf(hi - c, c - lo);
    return __fmaf_ru(h, 1.1920929e-07f, h);          // one ulp-scale step outward covers the rounding of the two differences
}
__device__ __forceinline__ float3 box_half(float4 lo, float4 hi, float3 c) { return f3(half_up(lo.x, hi.x, c.x), half_up(lo.y, hi.y, c.y), half_up(lo.z, hi.z, c.z)); }

// child link of an emitted node: inner index, or leaf ~((first_slot << 3) | (count - 1))
__device__ __forceinline__ int encode_child(int c, const int2* ranges, const uint32_t* new_slot, int leaf_size)
{
    if (c < 0) return ~((int)(new_slot[~c] << 3) | 0);
    int2 r = ranges[c];
    int cnt = r.y - r.x + 1;
    if (cnt <= leaf_size) return ~((r.x << 3) | (cnt - 1));
    return c;
}

// `children` leaf links are ~morton_slot (lmin/lmax are in Morton order); `ranges` are final slots
__global__ void emit_nodes(int n_inner, const int2* children, const int2* ranges, const uint32_t* new_slot, const float4* lmin, const float4* lmax,
                           const float4* imin, const float4* imax, int leaf_size, float4* nodes)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_inner) return;
    int2 ch = children[i];
    float4 alo = ch.x >= 0 ? imin[ch.x] : lmin[~ch.x], ahi = ch.x >= 0 ? imax[ch.x] : lmax[~ch.x];
    float4 blo = ch.y >= 0 ? imin[ch.y] : lmin[~ch.y], bhi = ch.y >= 0 ? imax[ch.y] : lmax[~ch.y];
    int el = encode_child(ch.x, ranges, new_slot, leaf_size), er = encode_child(ch.y, ranges, new_slot, leaf_size);
    const float3 ac = box_centre(alo, ahi), ah = box_half(alo, ahi, ac), bc = box_centre(blo, bhi), bh = box_half(blo, bhi, bc);
    // pair order of node_test() (rt_device.cuh): (L.c.x,L.c.y | L.c.z,R.c.z) (R.c.x,R.c.y | L.h.x,L.h.y) (L.h.z,R.h.z | R.h.x,R.h.y)
    nodes[4 * (size_t)i + 0] = make_float4(ac.x, ac.y, ac.z, bc.z);
    nodes[4 * (size_t)i + 1] = make_float4(bc.x, bc.y, ah.x, ah.y);
    nodes[4 * (size_t)i + 2] = make_float4(ah.z, bh.z, bh.x, bh.y);
    nodes[4 * (size_t)i + 3] = make_float4(__int_as_float(el), __int_as_float(er), 0.f, 0.f);
}

// ---- packed nodes (node_test_packed, rt_device.cuh) -------------------------------------------------------------------
// grid = bounds[6..11] as float bits: origin xyz, cell xyz (cell = extent / 65535, at least 1e-30)
__global__ void make_grid(int* bounds)
{
    const int k = threadIdx.x;
    if (k >= 3) return;
    const float lo = fdec(bounds[k]), hi = fdec(bounds[k + 3]);
    bounds[6 + k] = __float_as_int(lo);
    bounds[9 + k] = __float_as_int(fmaxf((hi - lo) * (1.0f / 65535.0f), 1e-30f));
}
// One (child, axis) word: the box centre on the grid in the low half, the half extent ABOUT THAT GRID POINT, enlarged by
// 1.25 cells (the traversal's reconstruction error, see node_test_packed) and rounded up to bf16, in the high half.
// A half extent that is not finite becomes +inf: the axis then never culls.
__device__ __forceinline__ uint32_t pack_axis(float lo, float hi, float o, float cell)
{
    const float g = (0.5f * lo + 0.5f * hi - o) / cell;
    const float q = fminf(fmaxf(rintf(g), 0.0f), 65535.0f);        // a NaN centre lands on 0; its half extent below is NaN -> inf
    const float cq = __fmaf_rn(q, cell, o);
    float h = fmaxf(hi - cq, cq - lo) + 1.25f * cell;
    h = __fmaf_ru(h, 4.76837158e-07f, h);                          // 2^-21: the roundings of cq and of the two differences
    uint32_t hb = __float_as_uint(h);
    if (!(h <= 3.0e38f) || !(hi - cq == hi - cq) || !(cq - lo == cq - lo)) hb = 0x7f800000u;
    else hb = (hb + 0xffffu) & 0xffff0000u;
    return hb | (uint32_t)q;
}
__global__ void emit_nodes_packed(int n_inner, const int2* children, const int2* ranges, const uint32_t* new_slot, const float4* lmin, const float4* lmax,
                                  const float4* imin, const float4* imax, int leaf_size, const int* bounds, float4* nodes)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_inner) return;
    const float3 o = f3(__int_as_float(bounds[6]), __int_as_float(bounds[7]), __int_as_float(bounds[8]));
    const float3 c = f3(__int_as_float(bounds[9]), __int_as_float(bounds[10]), __int_as_float(bounds[11]));
    int2 ch = children[i];
    float4 alo = ch.x >= 0 ? imin[ch.x] : lmin[~ch.x], ahi = ch.x >= 0 ? imax[ch.x] : lmax[~ch.x];
    float4 blo = ch.y >= 0 ? imin[ch.y] : lmin[~ch.y], bhi = ch.y >= 0 ? imax[ch.y] : lmax[~ch.y];
    int el = encode_child(ch.x, ranges, new_slot, leaf_size), er = encode_child(ch.y, ranges, new_slot, leaf_size);
    // word order of node_test_packed(): (L.x, L.y | L.z, R.z) (R.x, R.y | left, right)
    nodes[2 * (size_t)i + 0] = make_float4(__uint_as_float(pack_axis(alo.x, ahi.x, o.x, c.x)), __uint_as_float(pack_axis(alo.y, ahi.y, o.y, c.y)),
                                           __uint_as_float(pack_axis(alo.z, ahi.z, o.z, c.z)), __uint_as_float(pack_axis(blo.z, bhi.z, o.z, c.z)));
    nodes[2 * (size_t)i + 1] = make_float4(__uint_as_float(pack_axis(blo.x, bhi.x, o.x, c.x)), __uint_as_float(pack_axis(blo.y, bhi.y, o.y, c.y)),
                                           __int_as_float(el), __int_as_float(er));
}
__global__ void emit_single_packed(const float4* lmin, const float4* lmax, const int* bounds, float4* nodes)
{
    const float3 o = f3(__int_as_float(bounds[6]), __int_as_float(bounds[7]), __int_as_float(bounds[8]));
    const float3 c = f3(__int_as_float(bounds[9]), __int_as_float(bounds[10]), __int_as_float(bounds[11]));
    float4 lo = lmin[0], hi = lmax[0];
    const uint32_t never = 0xff800000u;                            // right child: half extent -inf on every axis (never hit)
    nodes[0] = make_float4(__uint_as_float(pack_axis(lo.x, hi.x, o.x, c.x)), __uint_as_float(pack_axis(lo.y, hi.y, o.y, c.y)),
                           __uint_as_float(pack_axis(lo.z, hi.z, o.z, c.z)), __uint_as_float(never));
    nodes[1] = make_float4(__uint_as_float(never), __uint_as_float(never), __int_as_float(~0), __int_as_float(~0));
}

__global__ void emit_single(const float4* lmin, const float4* lmax, float4* nodes)
{
    float4 lo = lmin[0], hi = lmax[0];
    const float3 c = box_centre(lo, hi), h = box_half(lo, hi, c);
    nodes[0] = make_float4(c.x, c.y, c.z, 0.f);
    nodes[1] = make_float4(0.f, 0.f, h.x, h.y);
    nodes[2] = make_float4(h.z, -INFINITY, -INFINITY, -INFINITY);     // right child: empty box (never hit)
    nodes[3] = make_float4(__int_as_float(~0), __int_as_float(~0), 0.f, 0.f);
}

// Stream-ordered allocation from the device's default pool (rt_create raises its release
// threshold, so repeated uploads reuse the same blocks instead of paying cudaMalloc/cudaFree).
static thread_local cudaStream_t g_alloc_stream = nullptr;
static thread_local cudaMemPool_t g_alloc_pool = nullptr;     // the context's own pool (the device's default pool is left alone)
cudaError_t pool_alloc(void** p, size_t bytes, cudaStream_t s)
{
    return g_alloc_pool ? cudaMallocFromPoolAsync(p, bytes ? bytes : 1, g_alloc_pool, s) : cudaMallocAsync(p, bytes ? bytes : 1, s);
}
template <typename T>
cudaError_t dalloc(T** p, size_t n) { return pool_alloc((void**)p, (n ? n : 1) * sizeof(T), g_alloc_stream); }
cudaError_t cudaMallocFromPoolAsyncOrDefault(void** p, size_t bytes, cudaMemPool_t pool, cudaStream_t s)
{
    return pool ? cudaMallocFromPoolAsync(p, bytes, pool, s) : cudaMallocAsync(p, bytes, s);
}

} // namespace

void SceneBuffers::release()
{
    float4** ps[] = {&prims, &c0, &c1, &c2, &c3, &c4, &c5, &m0, &m1, &m2, &nodes, &iprims};
    for (auto p : ps) { if (*p) cudaFreeAsync(*p, stream); *p = nullptr; }
    {
        void* kp[] = {keep.sph, keep.vtx, keep.idx, keep.mesh, keep.tstart, keep.hot, keep.bmin, keep.bmax, keep.lmin, keep.lmax, keep.imin, keep.imax,
                      keep.bounds, keep.ids2, keep.new_slot, keep.children, keep.ranges, keep.pin, keep.plf, keep.flags,
                      keep.flat_start, keep.inst_mesh, keep.range_of_mesh, keep.range_info};
        for (void* p : kp) if (p) cudaFreeAsync(p, stream);
        keep = Keep();
    }
    n_prims = n_spheres = n_tris = n_mats = n_nodes = max_depth = 0; root = -1; max_abs = 0.f; has_tex = false;
    n_inst = n_iprims = n_cold = n_top_nodes = 0;
    packed = false;
}

#define RT_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

// What instancing adds to a top-level build (nullptr / empty: every mesh is flattened)
struct InstPlan {
    std::vector<int> range_of_mesh;                 // per mesh: -1 flattened, else range index
    std::vector<uint32_t> inst_mesh;                // proxies in mesh order
    std::vector<float4> range_info;                 // 3 per range (setup_prims)
    uint32_t blas_nodes = 0, blas_tris = 0;         // totals appended behind the top level
};

// grid of the packed nodes out of the copied-back bounds block (make_grid); false when it is not finite
static bool read_grid(const int* hb, float* go, float* gc)
{
    bool ok = true;
    for (int k = 0; k < 3; k++) {
        memcpy(&go[k], &hb[6 + k], 4); memcpy(&gc[k], &hb[9 + k], 4);
        ok = ok && go[k] == go[k] && fabsf(go[k]) <= 3.0e38f && gc[k] == gc[k] && gc[k] <= 3.0e38f;
    }
    return ok;
}

static cudaError_t build_scene_impl(SceneBuffers& out, const rt_sphere* spheres, uint32_t ns, const rt_vertex* vertices,
                        uint32_t nv, const uint32_t* indices, uint32_t ni, const rt_mesh* meshes, uint32_t nm,
                        uint32_t leaf_size, int builder, int ploc_radius, const InstPlan* plan, int want_packed /*-1: by size*/, cudaStream_t stream)
{
    out.release();
    out.stream = stream;
    g_alloc_stream = stream; g_alloc_pool = out.pool;
    cudaError_t err = cudaSuccess;
    std::vector<uint32_t> tri_start(nm + 1, 0), flat_start(nm + 1, 0);
    std::vector<int> range_of_mesh(nm, -1);
    if (plan) range_of_mesh = plan->range_of_mesh;
    for (uint32_t m = 0; m < nm; m++) {
        tri_start[m + 1] = tri_start[m] + meshes[m].triangles;
        flat_start[m + 1] = flat_start[m] + (range_of_mesh[m] < 0 ? meshes[m].triangles : 0u);
    }
    const uint32_t n_inst = plan ? (uint32_t)plan->inst_mesh.size() : 0u;
    const uint32_t nt = flat_start[nm], n = ns + nt + n_inst, nmat = ns + nm;
    const uint32_t n_cold = n + (plan ? plan->blas_tris : 0u);
    bool packed = want_packed < 0 ? n > RT_PF_MAX_PRIMS : want_packed != 0;   // by size: small scenes keep float nodes (rt_trace.cu: the prefetching kernel variants)
    uint32_t* d_flat_start = nullptr; uint32_t* d_inst_mesh = nullptr; int* d_range_of_mesh = nullptr; float4* d_range_info = nullptr;
    const int T = 256;
    if (ploc_radius < 1) ploc_radius = 1;
    if (ploc_radius > PLOC_MAX_RADIUS) ploc_radius = PLOC_MAX_RADIUS;

    rt_sphere* d_sph = nullptr; rt_vertex* d_vtx = nullptr; uint32_t* d_idx = nullptr; rt_mesh* d_mesh = nullptr;
    uint32_t* d_tstart = nullptr;
    float4 *hot = nullptr, *bmin = nullptr, *bmax = nullptr, *lmin = nullptr, *lmax = nullptr, *imin = nullptr, *imax = nullptr;
    int* d_bounds = nullptr; uint64_t *keys = nullptr, *keys2 = nullptr; uint32_t *ids = nullptr, *ids2 = nullptr, *new_slot = nullptr;
    int2 *children = nullptr, *ranges = nullptr; int *pin = nullptr, *plf = nullptr, *flags = nullptr, *d_depth = nullptr;
    int *cidA = nullptr, *cidB = nullptr, *nn = nullptr; float4 *cloA = nullptr, *chiA = nullptr, *cloB = nullptr, *chiB = nullptr;
    unsigned long long *pflags = nullptr, *ppos = nullptr; unsigned int* d_tot = nullptr; int* d_state = nullptr;
    void* cub_tmp = nullptr; size_t cub_bytes = 0;
    void* scan_tmp = nullptr; size_t scan_bytes = 0;
    int hb[12]; int h_init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    int h_depth = 0;

    RT_TRY(dalloc(&d_sph, ns)); RT_TRY(dalloc(&d_vtx, nv)); RT_TRY(dalloc(&d_idx, ni)); RT_TRY(dalloc(&d_mesh, nm));
    RT_TRY(dalloc(&d_tstart, nm + 1));
    RT_TRY(dalloc(&d_flat_start, nm + 1)); RT_TRY(dalloc(&d_inst_mesh, n_inst)); RT_TRY(dalloc(&d_range_of_mesh, nm));
    RT_TRY(dalloc(&d_range_info, plan ? plan->range_info.size() : 0));
    RT_TRY(cudaMemcpyAsync(d_flat_start, flat_start.data(), (nm + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    if (nm) RT_TRY(cudaMemcpyAsync(d_range_of_mesh, range_of_mesh.data(), nm * sizeof(int), cudaMemcpyHostToDevice, stream));
    if (n_inst) {
        RT_TRY(cudaMemcpyAsync(d_inst_mesh, plan->inst_mesh.data(), n_inst * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
        RT_TRY(cudaMemcpyAsync(d_range_info, plan->range_info.data(), plan->range_info.size() * sizeof(float4), cudaMemcpyHostToDevice, stream));
    }
    if (ns) RT_TRY(cudaMemcpyAsync(d_sph, spheres, ns * sizeof(rt_sphere), cudaMemcpyHostToDevice, stream));
    if (nv) RT_TRY(cudaMemcpyAsync(d_vtx, vertices, nv * sizeof(rt_vertex), cudaMemcpyHostToDevice, stream));
    if (ni) RT_TRY(cudaMemcpyAsync(d_idx, indices, ni * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    if (nm) RT_TRY(cudaMemcpyAsync(d_mesh, meshes, nm * sizeof(rt_mesh), cudaMemcpyHostToDevice, stream));
    RT_TRY(cudaMemcpyAsync(d_tstart, tri_start.data(), (nm + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));

    RT_TRY(dalloc(&out.prims, 4 * (size_t)n));
    RT_TRY(dalloc(&out.c0, n_cold)); RT_TRY(dalloc(&out.c1, n_cold)); RT_TRY(dalloc(&out.c2, n_cold));
    RT_TRY(dalloc(&out.m0, nmat)); RT_TRY(dalloc(&out.m1, nmat)); RT_TRY(dalloc(&out.m2, nmat));
    out.n_prims = n; out.n_spheres = ns; out.n_tris = nt; out.n_mats = nmat; out.n_inst = n_inst; out.n_cold = n_cold;
    for (uint32_t m = 0; m < nm; m++) if (meshes[m].triangles && (meshes[m].diffuse_tex | meshes[m].normal_tex)) out.has_tex = true;
    if (out.has_tex) { RT_TRY(dalloc(&out.c3, n_cold)); RT_TRY(dalloc(&out.c4, n_cold)); RT_TRY(dalloc(&out.c5, n_cold)); }
    {
        RawScene rs{d_sph, d_vtx, d_idx, d_mesh, d_tstart, ns, nm, nt, d_flat_start, d_inst_mesh, d_range_of_mesh, d_range_info, n_inst};
        if (nmat) setup_materials<<<(nmat + T - 1) / T, T, 0, stream>>>(rs, out.m0, out.m1, out.m2);
        if (n == 0) { RT_TRY(cudaStreamSynchronize(stream)); goto done; }

        RT_TRY(dalloc(&hot, 4 * (size_t)n)); RT_TRY(dalloc(&bmin, n)); RT_TRY(dalloc(&bmax, n));
        RT_TRY(dalloc(&lmin, n)); RT_TRY(dalloc(&lmax, n)); RT_TRY(dalloc(&imin, n)); RT_TRY(dalloc(&imax, n));
        RT_TRY(dalloc(&d_bounds, 12)); RT_TRY(dalloc(&keys, n)); RT_TRY(dalloc(&keys2, n)); RT_TRY(dalloc(&ids, n)); RT_TRY(dalloc(&ids2, n));
        RT_TRY(dalloc(&children, n)); RT_TRY(dalloc(&ranges, n)); RT_TRY(dalloc(&pin, n)); RT_TRY(dalloc(&plf, n)); RT_TRY(dalloc(&flags, n));
        RT_TRY(dalloc(&d_depth, 1)); RT_TRY(dalloc(&new_slot, n));
        RT_TRY(cudaMemcpyAsync(d_bounds, h_init, sizeof(h_init), cudaMemcpyHostToDevice, stream));

        const int nb = (n + T - 1) / T;
        setup_prims<<<nb, T, 0, stream>>>(rs, hot, out.c0, out.c1, out.c2, out.c3, out.c4, out.c5, bmin, bmax, d_bounds);
        make_grid<<<1, 32, 0, stream>>>(d_bounds);
        morton_keys<<<nb, T, 0, stream>>>(n, bmin, bmax, d_bounds, keys, ids);
        RT_TRY(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys, keys2, ids, ids2, (int)n, 0, 63, stream));
        RT_TRY(pool_alloc(&cub_tmp, cub_bytes, stream));
        RT_TRY(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys2, ids, ids2, (int)n, 0, 63, stream));
        gather_boxes<<<nb, T, 0, stream>>>(n, ids2, bmin, bmax, lmin, lmax);
        if (n == 1) {
            RT_TRY(dalloc(&out.nodes, 4 * (size_t)(1 + (plan ? plan->blas_nodes : 0u))));
            if (packed) emit_single_packed<<<1, 1, 0, stream>>>(lmin, lmax, d_bounds, out.nodes);
            else emit_single<<<1, 1, 0, stream>>>(lmin, lmax, out.nodes);
            identity_slots<<<1, T, 0, stream>>>(n, new_slot);
            out.n_nodes = 1; out.root = 0; h_depth = 1;
        } else {
            RT_TRY(dalloc(&out.nodes, 4 * (size_t)(n - 1 + (plan ? plan->blas_nodes : 0u))));
            bool use_ploc = builder == 0;
            int root = 0;
            if (use_ploc) {
                RT_TRY(dalloc(&cidA, n)); RT_TRY(dalloc(&cidB, n)); RT_TRY(dalloc(&nn, n));
                RT_TRY(dalloc(&cloA, n)); RT_TRY(dalloc(&chiA, n)); RT_TRY(dalloc(&cloB, n)); RT_TRY(dalloc(&chiB, n));
                RT_TRY(dalloc(&pflags, n)); RT_TRY(dalloc(&ppos, n)); RT_TRY(dalloc(&d_tot, 2));
                RT_TRY(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, pflags, ppos, (int)n, stream));
                RT_TRY(pool_alloc(&scan_tmp, scan_bytes, stream));
                ploc_init<<<nb, T, 0, stream>>>((int)n, lmin, lmax, cidA, cloA, chiA);
                RT_TRY(dalloc(&d_state, 5));
                const int tail = ploc_radius <= PLOC_MAX_RADIUS ? PLOC_TAIL : 1;
                int h_state[5] = {(int)n, 0, 0, 0, tail};        // clusters left, nodes created, force flag, rounds done, cluster count at which ploc_tail takes over
                RT_TRY(cudaMemcpyAsync(d_state, h_state, sizeof h_state, cudaMemcpyHostToDevice, stream));
                int* const cid0[2] = {cidA, cidB}; float4* const clo0[2] = {cloA, cloB}; float4* const chi0[2] = {chiA, chiB};
                for (int checks = 0; h_state[0] > tail; checks++) {
                    if (checks > 4096) { RT_TRY(cudaErrorUnknown); }   // cannot happen: every round merges at least one pair
                    const int m_ub = h_state[0];                 // the cluster count only shrinks: grids and the scan are sized for the last known one
                    const int mb = (m_ub + PLOC_THREADS - 1) / PLOC_THREADS;
                    for (int r = 0; r < PLOC_ROUNDS_PER_CHECK; r++) {
                        ploc_nn<<<mb, PLOC_THREADS, 0, stream>>>(d_state, ploc_radius, cloA, chiA, nn);
                        ploc_flags<<<mb, PLOC_THREADS, 0, stream>>>(d_state, nn, pflags);
                        RT_TRY(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, pflags, ppos, m_ub, stream));
                        ploc_scatter<<<mb, PLOC_THREADS, 0, stream>>>(d_state, nn, ppos, cidA, cloA, chiA, cidB, cloB, chiB,
                                                                      children, imin, imax, pin, plf, d_tot);
                        ploc_advance<<<1, 1, 0, stream>>>(d_state, d_tot);
                        std::swap(cidA, cidB); std::swap(cloA, cloB); std::swap(chiA, chiB);
                    }
                    RT_TRY(cudaMemcpyAsync(h_state, d_state, sizeof h_state, cudaMemcpyDeviceToHost, stream));
                    RT_TRY(cudaStreamSynchronize(stream));
                }
                if (tail > 1) {                                  // rounds launched after the list fell to `tail` returned at once: the list is where round h_state[3] left it
                    const int cur = h_state[3] & 1;
                    ploc_tail<<<1, PLOC_TAIL, 0, stream>>>(d_state, ploc_radius, cid0[cur], clo0[cur], chi0[cur], children, imin, imax, pin, plf);
                }
                root = (int)n - 2;                               // the last node created
                RT_TRY(cudaMemsetAsync(d_depth, 0, sizeof(int), stream));
                ploc_layout<<<(2 * n - 1 + T - 1) / T, T, 0, stream>>>((int)n, root, children, imin, pin, plf, new_slot, ranges, d_depth);
                RT_TRY(cudaMemcpyAsync(&h_depth, d_depth, sizeof(int), cudaMemcpyDeviceToHost, stream));
                RT_TRY(cudaStreamSynchronize(stream));
                if (h_depth > RT_STACK - 2) use_ploc = false;    // too deep for the traversal stack: take the radix tree instead
            }
            if (!use_ploc) {
                root = 0;
                RT_TRY(cudaMemsetAsync(flags, 0, n * sizeof(int), stream));
                RT_TRY(cudaMemsetAsync(d_depth, 0, sizeof(int), stream));
                build_radix_tree<<<nb, T, 0, stream>>>((int)n, keys2, children, ranges, pin, plf);
                refit<<<nb, T, 0, stream>>>((int)n, children, pin, plf, lmin, lmax, imin, imax, flags, d_depth);
                identity_slots<<<nb, T, 0, stream>>>(n, new_slot);
                RT_TRY(cudaMemcpyAsync(&h_depth, d_depth, sizeof(int), cudaMemcpyDeviceToHost, stream));
            }
            if (packed) emit_nodes_packed<<<nb, T, 0, stream>>>((int)n - 1, children, ranges, new_slot, lmin, lmax, imin, imax, (int)leaf_size, d_bounds, out.nodes);
            else emit_nodes<<<nb, T, 0, stream>>>((int)n - 1, children, ranges, new_slot, lmin, lmax, imin, imax, (int)leaf_size, out.nodes);
            out.n_nodes = n - 1; out.root = root; out.builder = use_ploc ? 0 : 1;
        }
        gather_slots<<<(4 * n + T - 1) / T, T, 0, stream>>>(n, ids2, new_slot, hot, out.prims);
        RT_TRY(cudaMemcpyAsync(hb, d_bounds, sizeof(hb), cudaMemcpyDeviceToHost, stream));
        RT_TRY(cudaStreamSynchronize(stream));
        RT_TRY(cudaGetLastError());
        if (packed && !read_grid(hb, out.grid_o, out.grid_c)) {      // bounds that are not finite have no grid: emit float nodes instead
            packed = false;
            if (n == 1) emit_single<<<1, 1, 0, stream>>>(lmin, lmax, out.nodes);
            else emit_nodes<<<nb, T, 0, stream>>>((int)n - 1, children, ranges, new_slot, lmin, lmax, imin, imax, (int)leaf_size, out.nodes);
            RT_TRY(cudaStreamSynchronize(stream));
            RT_TRY(cudaGetLastError());
        }
        out.packed = packed;
        out.max_depth = (uint32_t)h_depth;
        float m = 0.f;
        for (int k = 0; k < 6; k++) {
            int e = hb[k]; int b = e >= 0 ? e : e ^ 0x7fffffff; float f; memcpy(&f, &b, 4);
            if (f == f && fabsf(f) <= 3.0e38f && fabsf(f) > m) m = fabsf(f);
        }
        out.max_abs = m;
        for (int q = 0; q < 3; q++) {
            int e0 = hb[q], e1 = hb[q + 3]; int b0 = e0 >= 0 ? e0 : e0 ^ 0x7fffffff, b1 = e1 >= 0 ? e1 : e1 ^ 0x7fffffff;
            memcpy(&out.bounds_lo[q], &b0, 4); memcpy(&out.bounds_hi[q], &b1, 4);
        }
        out.n_top_nodes = out.n_nodes;
        // keep what refit_scene needs (ownership moves to `out`; the rest is freed below)
        SceneBuffers::Keep& k = out.keep;
        k.flat_start = d_flat_start; k.inst_mesh = d_inst_mesh; k.range_of_mesh = d_range_of_mesh; k.range_info = d_range_info;
        d_flat_start = nullptr; d_inst_mesh = nullptr; d_range_of_mesh = nullptr; d_range_info = nullptr;
        k.sph = d_sph; k.vtx = d_vtx; k.idx = d_idx; k.mesh = d_mesh; k.tstart = d_tstart; k.hot = hot; k.bmin = bmin; k.bmax = bmax;
        k.lmin = lmin; k.lmax = lmax; k.imin = imin; k.imax = imax; k.bounds = d_bounds; k.ids2 = ids2; k.new_slot = new_slot;
        k.children = children; k.ranges = ranges; k.pin = pin; k.plf = plf; k.flags = flags;
        k.nm = nm; k.leaf_size = leaf_size; k.parent_shift = out.builder == 0 ? 1 : 0;
        d_sph = nullptr; d_vtx = nullptr; d_idx = nullptr; d_mesh = nullptr; d_tstart = nullptr; hot = bmin = bmax = lmin = lmax = imin = imax = nullptr;
        d_bounds = nullptr; ids2 = new_slot = nullptr; children = ranges = nullptr; pin = plf = flags = nullptr;
    }
done:
    {
        void* tmp[] = {d_flat_start, d_inst_mesh, d_range_of_mesh, d_range_info, d_sph, d_vtx, d_idx, d_mesh, d_tstart, hot, bmin, bmax, lmin, lmax, imin, imax, d_bounds,
                       keys, keys2, ids, ids2, children, ranges, pin, plf, flags, d_depth, cub_tmp, new_slot,
                       cidA, cidB, nn, cloA, chiA, cloB, chiB, pflags, ppos, d_tot, d_state, scan_tmp};
        for (void* p : tmp) if (p) cudaFreeAsync(p, stream);
    }
    if (err != cudaSuccess) out.release();
    return err;
}

// ---- instancing: import of a range's object-space hierarchy behind the top level ----------------------------------
// nodes: inner links move by node_base, leaf links by slot_base (leaf = ~((first_slot << 3) | (count - 1)))
__global__ void import_nodes(uint32_t n_nodes, const float4* __restrict__ src, float4* __restrict__ dst, int node_base, uint32_t slot_base)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    dst[4 * (size_t)i + 0] = src[4 * (size_t)i + 0]; dst[4 * (size_t)i + 1] = src[4 * (size_t)i + 1]; dst[4 * (size_t)i + 2] = src[4 * (size_t)i + 2];
    const float4 l = src[4 * (size_t)i + 3];
    int c[2] = {__float_as_int(l.x), __float_as_int(l.y)};
    for (int q = 0; q < 2; q++) {
        if (c[q] >= 0) c[q] += node_base;
        else { const uint32_t leaf = (uint32_t)(~c[q]); c[q] = ~(int)((((leaf >> 3) + slot_base) << 3) | (leaf & 7u)); }
    }
    dst[4 * (size_t)i + 3] = make_float4(__int_as_float(c[0]), __int_as_float(c[1]), 0.f, 0.f);
}
// raw vertex triples in the range's tree order (the exact test re-derives A = a + pos, e1, e2, n per instance with the
// reference's own operations, wgsl:172-174,127-129) and the vertex normals by range-local triangle index
__global__ void import_tris(uint32_t n_tris, const float4* __restrict__ blas_prims, const rt_vertex* __restrict__ vertices, const uint32_t* __restrict__ indices,
                            uint32_t first, uint32_t offset, float4* __restrict__ iprims, const float4* __restrict__ b0, const float4* __restrict__ b1,
                            const float4* __restrict__ b2, float4* __restrict__ c0, float4* __restrict__ c1, float4* __restrict__ c2)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_tris) return;
    const uint32_t tri = __float_as_uint(blas_prims[4 * (size_t)s].w);        // the sub-scene has no spheres and one mesh: id = triangle index
    const rt_vertex& va = vertices[offset + indices[first + tri * 3u]];
    const rt_vertex& vb = vertices[offset + indices[first + tri * 3u + 1u]];
    const rt_vertex& vc = vertices[offset + indices[first + tri * 3u + 2u]];
    iprims[4 * (size_t)s + 0] = make_float4(va.pos[0], va.pos[1], va.pos[2], vb.pos[0]);
    iprims[4 * (size_t)s + 1] = make_float4(vb.pos[1], vb.pos[2], vc.pos[0], vc.pos[1]);
    iprims[4 * (size_t)s + 2] = make_float4(vc.pos[2], __uint_as_float(tri), 0.f, 0.f);
    iprims[4 * (size_t)s + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
    c0[s] = make_float4(b0[s].x, b0[s].y, b0[s].z, 0.f); c1[s] = b1[s]; c2[s] = b2[s];     // cold records by triangle index (s runs over them too)
}

// Meshes sharing one (first, triangles, offset) range become instances of ONE object-space hierarchy when that pays: at least two
// records, a range worth a hierarchy of its own, no texture handles (the texture extension's tangent frame is per flattened record).
#define RT_INST_MIN_TRIS 32u
#define RT_INST_AUTO_TRIS 8000000ull               // ~1.4 GB of flattened records + nodes
cudaError_t build_scene(SceneBuffers& out, const rt_sphere* spheres, uint32_t ns, const rt_vertex* vertices,
                        uint32_t nv, const uint32_t* indices, uint32_t ni, const rt_mesh* meshes, uint32_t nm,
                        uint32_t leaf_size, int builder, int ploc_radius, int instancing, int node_format, cudaStream_t stream)
{
    struct Range { uint32_t first, triangles, offset, members; };
    std::vector<Range> ranges;
    InstPlan plan;
    plan.range_of_mesh.assign(nm, -1);
    if (instancing) {
        std::vector<int> key_of(nm, -1);
        for (uint32_t m = 0; m < nm; m++) {
            const rt_mesh& me = meshes[m];
            if (me.triangles < RT_INST_MIN_TRIS || (me.diffuse_tex | me.normal_tex)) continue;
            for (size_t g = 0; g < ranges.size(); g++)
                if (ranges[g].first == me.first && ranges[g].triangles == me.triangles && ranges[g].offset == me.offset) { key_of[m] = (int)g; break; }
            if (key_of[m] < 0) { key_of[m] = (int)ranges.size(); ranges.push_back(Range{me.first, me.triangles, me.offset, 0}); }
            ranges[key_of[m]].members++;
        }
        std::vector<int> remap(ranges.size(), -1);
        std::vector<Range> kept;
        for (size_t g = 0; g < ranges.size(); g++) if (ranges[g].members >= 2) { remap[g] = (int)kept.size(); kept.push_back(ranges[g]); }
        ranges.swap(kept);
        // auto mode: flattened records trace faster while they fit comfortably (C5, 1.0 M instanced triangles: 4,069 Mrays/s flattened
        // against 3,403 instanced: the per-instance test re-derives A, e1, e2 and the normal and loses the 16-byte back-face reject);
        // instancing is for scenes whose flattened form would not: more than RT_INST_AUTO_TRIS instanced triangles
        if (instancing == 2) {
            uint64_t flat = 0;
            for (const Range& r : ranges) flat += (uint64_t)r.members * r.triangles;
            if (flat <= RT_INST_AUTO_TRIS) ranges.clear();
        }
        if (!ranges.empty())
            for (uint32_t m = 0; m < nm; m++)
                if (key_of[m] >= 0 && remap[key_of[m]] >= 0) { plan.range_of_mesh[m] = remap[key_of[m]]; plan.inst_mesh.push_back(m); }
    }
    if (ranges.empty()) return build_scene_impl(out, spheres, ns, vertices, nv, indices, ni, meshes, nm, leaf_size, builder, ploc_radius, nullptr, RT_PACKED_NODES ? (node_format == 0 ? -1 : node_format == 2) : 0, stream);

    // 1. one object-space hierarchy per range (a sub-scene of one Mesh record with pos = 0: the tested builder as it is)
    cudaError_t err = cudaSuccess;
    std::vector<SceneBuffers> blas(ranges.size());
    uint32_t depth_max = 0;
    uint32_t n_top = ns + (uint32_t)plan.inst_mesh.size();
    for (uint32_t m = 0; m < nm; m++) if (plan.range_of_mesh[m] < 0) n_top += meshes[m].triangles;
    const uint32_t n_top_nodes = n_top > 1 ? n_top - 1 : 1;
    for (size_t g = 0; g < ranges.size() && err == cudaSuccess; g++) {
        rt_mesh sub; memset(&sub, 0, sizeof sub);
        sub.first = ranges[g].first; sub.triangles = ranges[g].triangles; sub.offset = ranges[g].offset;
        blas[g].pool = out.pool;
        err = build_scene_impl(blas[g], nullptr, 0, vertices, nv, indices, ni, &sub, 1, leaf_size, builder, ploc_radius, nullptr, 0, stream);
        if (err != cudaSuccess) break;
        const SceneBuffers& b = blas[g];
        plan.range_info.push_back(make_float4(b.bounds_lo[0], b.bounds_lo[1], b.bounds_lo[2], 0.f));
        plan.range_info.push_back(make_float4(b.bounds_hi[0], b.bounds_hi[1], b.bounds_hi[2], 0.f));
        const int root = (int)(n_top_nodes + plan.blas_nodes) + b.root;
        const uint32_t cold_base = n_top + plan.blas_tris;
        int ri[2] = {root, (int)cold_base}; float4 info; memcpy(&info.x, &ri[0], 4); memcpy(&info.y, &ri[1], 4); info.z = info.w = 0.f;
        plan.range_info.push_back(info);
        plan.blas_nodes += b.n_nodes; plan.blas_tris += b.n_prims;
        if (b.max_depth > depth_max) depth_max = b.max_depth;
    }
    // 2. the top level over spheres, flattened triangles and one proxy per instance
    if (err == cudaSuccess) err = build_scene_impl(out, spheres, ns, vertices, nv, indices, ni, meshes, nm, leaf_size, builder, ploc_radius, &plan, 0, stream);
    // 3. the ranges' nodes / raw triangles / normals behind it
    if (err == cudaSuccess) err = cudaMallocFromPoolAsyncOrDefault((void**)&out.iprims, 4 * (size_t)(plan.blas_tris ? plan.blas_tris : 1) * sizeof(float4), out.pool, stream);
    if (err == cudaSuccess) {
        const int T = 256;
        uint32_t node_off = 0, tri_off = 0;
        for (size_t g = 0; g < ranges.size(); g++) {
            const SceneBuffers& b = blas[g];
            import_nodes<<<(b.n_nodes + T - 1) / T, T, 0, stream>>>(b.n_nodes, b.nodes, out.nodes + 4 * (size_t)(out.n_top_nodes + node_off),
                                                                    (int)(out.n_top_nodes + node_off), tri_off);
            import_tris<<<(b.n_prims + T - 1) / T, T, 0, stream>>>(b.n_prims, b.prims, b.keep.vtx, b.keep.idx, ranges[g].first, ranges[g].offset,
                                                                   out.iprims + 4 * (size_t)tri_off, b.c0, b.c1, b.c2,
                                                                   out.c0 + out.n_prims + tri_off, out.c1 + out.n_prims + tri_off, out.c2 + out.n_prims + tri_off);
            node_off += b.n_nodes; tri_off += b.n_prims;
        }
        out.n_nodes = out.n_top_nodes + plan.blas_nodes; out.n_iprims = plan.blas_tris;
        out.max_depth += depth_max + 3;                 // the deepest stack: top level, rest of a leaf + continuation + marker, then the range's own depth
        err = cudaStreamSynchronize(stream);
        if (err == cudaSuccess) err = cudaGetLastError();
    }
    for (SceneBuffers& b : blas) { b.stream = stream; b.release(); }
    if (err != cudaSuccess) out.release();
    return err;
}

cudaError_t refit_scene(SceneBuffers& sc, const rt_sphere* spheres, const rt_mesh* meshes, cudaStream_t stream)
{
    SceneBuffers::Keep& k = sc.keep;
    const uint32_t n = sc.n_prims, ns = sc.n_spheres, nm = k.nm, nmat = sc.n_mats;
    const int T = 256;
    cudaError_t err = cudaSuccess;
    if (n == 0) return cudaSuccess;                    // nothing is ever hit: no record is read
    int hb[12]; int h_init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    if (ns) RT_TRY(cudaMemcpyAsync(k.sph, spheres, ns * sizeof(rt_sphere), cudaMemcpyHostToDevice, stream));
    if (nm) RT_TRY(cudaMemcpyAsync(k.mesh, meshes, nm * sizeof(rt_mesh), cudaMemcpyHostToDevice, stream));
    {
        RawScene rs{k.sph, k.vtx, k.idx, k.mesh, k.tstart, ns, nm, sc.n_tris, k.flat_start, k.inst_mesh, k.range_of_mesh, k.range_info, sc.n_inst};
        if (nmat) setup_materials<<<(nmat + T - 1) / T, T, 0, stream>>>(rs, sc.m0, sc.m1, sc.m2);
        if (n == 0) { RT_TRY(cudaStreamSynchronize(stream)); goto done; }
        const int nb = (n + T - 1) / T;
        RT_TRY(cudaMemcpyAsync(k.bounds, h_init, sizeof(h_init), cudaMemcpyHostToDevice, stream));
        setup_prims<<<nb, T, 0, stream>>>(rs, k.hot, sc.c0, sc.c1, sc.c2, sc.c3, sc.c4, sc.c5, k.bmin, k.bmax, k.bounds);
        make_grid<<<1, 32, 0, stream>>>(k.bounds);
        gather_boxes<<<nb, T, 0, stream>>>(n, k.ids2, k.bmin, k.bmax, k.lmin, k.lmax);
        if (n == 1) {
            if (sc.packed) emit_single_packed<<<1, 1, 0, stream>>>(k.lmin, k.lmax, k.bounds, sc.nodes);
            else emit_single<<<1, 1, 0, stream>>>(k.lmin, k.lmax, sc.nodes);
        } else {
            RT_TRY(cudaMemsetAsync(k.flags, 0, n * sizeof(int), stream));
            refit_boxes<<<nb, T, 0, stream>>>((int)n, sc.root, k.parent_shift, k.children, k.pin, k.plf, k.lmin, k.lmax, k.imin, k.imax, k.flags);
            if (sc.packed) emit_nodes_packed<<<nb, T, 0, stream>>>((int)n - 1, k.children, k.ranges, k.new_slot, k.lmin, k.lmax, k.imin, k.imax, (int)k.leaf_size, k.bounds, sc.nodes);
            else emit_nodes<<<nb, T, 0, stream>>>((int)n - 1, k.children, k.ranges, k.new_slot, k.lmin, k.lmax, k.imin, k.imax, (int)k.leaf_size, sc.nodes);
        }
        gather_slots<<<(4 * n + T - 1) / T, T, 0, stream>>>(n, k.ids2, k.new_slot, k.hot, sc.prims);
        RT_TRY(cudaMemcpyAsync(hb, k.bounds, sizeof(hb), cudaMemcpyDeviceToHost, stream));
        RT_TRY(cudaStreamSynchronize(stream));
        RT_TRY(cudaGetLastError());
        if (sc.packed && !read_grid(hb, sc.grid_o, sc.grid_c)) {     // the moved records' bounds are not finite: float nodes from now on (the buffer is sized for them)
            sc.packed = false;
            if (n == 1) emit_single<<<1, 1, 0, stream>>>(k.lmin, k.lmax, sc.nodes);
            else emit_nodes<<<nb, T, 0, stream>>>((int)n - 1, k.children, k.ranges, k.new_slot, k.lmin, k.lmax, k.imin, k.imax, (int)k.leaf_size, sc.nodes);
            RT_TRY(cudaStreamSynchronize(stream));
            RT_TRY(cudaGetLastError());
        }
        float m = 0.f;
        for (int q = 0; q < 6; q++) {
            int e = hb[q]; int b = e >= 0 ? e : e ^ 0x7fffffff; float f; memcpy(&f, &b, 4);
            if (f == f && fabsf(f) <= 3.0e38f && fabsf(f) > m) m = fabsf(f);
        }
        sc.max_abs = m;
        for (int q = 0; q < 3; q++) {
            int e0 = hb[q], e1 = hb[q + 3]; int b0 = e0 >= 0 ? e0 : e0 ^ 0x7fffffff, b1 = e1 >= 0 ? e1 : e1 ^ 0x7fffffff;
            memcpy(&sc.bounds_lo[q], &b0, 4); memcpy(&sc.bounds_hi[q], &b1, 4);
        }
    }
done:
    return err;
}

} // namespace rt
