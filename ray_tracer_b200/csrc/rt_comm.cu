// rt_comm.cu — the multi-PROCESS image combine of the multi-GPU partition: one process per
// GPU (the launch model of torchrun / MPI, or a Rust host with one thread per device), the shards' images
// gathered over NCCL (NVLink 5 / NVSwitch) at the rank that reads the image, and ONE consumer kernel there that
// applies the count weights of the sample shards, de-interleaves the tile shards and — optionally fused — the
// display transform.  The reference is single-device and has no counterpart; the partition is rt_set_shard's.
//
// What moves: every rank packs the pixels of ITS tiles (W*H / tile_world of them, tile-major, 16 B each) and
// sends that compact buffer to the root (grouped ncclSend / ncclRecv); a tile partition therefore moves 16 B per
// pixel in total, a sample partition 16 B per pixel per sample group — nothing is zero-padded and there is no
// per-step clear.  The shard table (tile rank / world, frame rank / world, frames in the running mean) travels by
// ncclAllGather into device memory and is read by the consumer kernel, so the whole combine is stream-ordered:
// no host synchronisation, it overlaps whatever the host does next.
//
// Arithmetic = rt_combine_peers' (rt_combine.cu): owners of a pixel in rank order, each value times its weight
// (count / total count of the pixel's owners), separately rounded multiply and add; a pure tile partition
// reproduces the single-GPU image bit for bit.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): a single-GPU host needs no NCCL installed, and inside a
// process that already loaded one (torch) the same library instance is used.
#include <dlfcn.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include "rt_kernels.h"

namespace rt {

// ---- the NCCL entry points used, with the prototypes of nccl.h 2.x -----------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclFloat32 = 7 };
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    std::string err;
};
static NcclApi g_nccl;

const char* nccl_load()
{
    NcclApi& n = g_nccl;
    if (n.handle) return nullptr;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) { n.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (n.handle) break; }
    if (!n.handle) { n.err = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : ""); return n.err.c_str(); }
    struct { const char* name; void** fn; } syms[] = {
        {"ncclGetUniqueId", (void**)&n.GetUniqueId}, {"ncclCommInitRank", (void**)&n.CommInitRank}, {"ncclCommDestroy", (void**)&n.CommDestroy},
        {"ncclGroupStart", (void**)&n.GroupStart}, {"ncclGroupEnd", (void**)&n.GroupEnd}, {"ncclSend", (void**)&n.Send}, {"ncclRecv", (void**)&n.Recv},
        {"ncclAllGather", (void**)&n.AllGather}, {"ncclGetErrorString", (void**)&n.GetErrorString}, {"ncclGetVersion", (void**)&n.GetVersion}};
    for (auto& s : syms) {
        *s.fn = dlsym(n.handle, s.name);
        if (!*s.fn) { n.err = std::string("NCCL symbol missing: ") + s.name; dlclose(n.handle); n.handle = nullptr; return n.err.c_str(); }
    }
    return nullptr;
}

// ---- kernels ----------------------------------------------------------------------------------------------------
// this rank's pixels, tile-major (local tile lt = t / tile_world, 64 pixels each, row-major inside the tile); pixels of
// an edge tile that lie outside the image are written as zeros
__global__ void __launch_bounds__(256)
pack_tiles_kernel(const float4* __restrict__ accum, float4* __restrict__ out, uint32_t width, uint32_t height, uint32_t tiles_x,
                  uint32_t tile_rank, uint32_t tile_world, size_t n_local)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_local; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t lt = (uint32_t)(i >> 6), in_tile = (uint32_t)(i & 63u);
        const uint32_t t = lt * tile_world + tile_rank;
        const uint32_t px = (t % tiles_x) * 8u + (in_tile & 7u), py = (t / tiles_x) * 8u + (in_tile >> 3);
        out[i] = (px < width && py < height) ? accum[(size_t)py * width + px] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

__device__ __forceinline__ uint32_t srgb8_c(float c)          // same arithmetic as tonemap_kernel (rt_tonemap.cu)
{
    c = fminf(fmaxf(c, 0.0f), 1.0f);
    float s = (c <= 0.0031308f) ? 12.92f * c : __fmaf_rn(1.055f, spec_powf(c, 1.0f / 2.4f), -0.055f);
    return (uint32_t)__float2int_rn(fminf(fmaxf(s, 0.0f), 1.0f) * 255.0f);
}

// The consumer at the root: for every pixel, the owners in rank order, weight = count / (sum of the owners' counts).
// `recv` holds one compact buffer per rank at recv_stride pixels apart; the root's own pixels are read from its image.
template <bool TONEMAP>
__global__ void __launch_bounds__(256)
gather_combine_kernel(const ShardRec* __restrict__ shards, uint32_t n_ranks, uint32_t root, const float4* __restrict__ recv, size_t recv_stride,
                      const float4* __restrict__ own, float4* __restrict__ out, uchar4* __restrict__ out8, uint32_t width, uint32_t height, uint32_t tiles_x)
{
    const size_t n = (size_t)width * height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t py = (uint32_t)(i / width), px = (uint32_t)(i - (size_t)py * width);
        const uint32_t t = (py >> 3) * tiles_x + (px >> 3), in_tile = ((py & 7u) << 3) | (px & 7u);
        uint32_t total = 0;
        for (uint32_t k = 0; k < n_ranks; k++) {
            const ShardRec s = shards[k];
            if (t % s.tile_world == s.tile_rank) total += s.count;
        }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        bool any = false;
        for (uint32_t k = 0; k < n_ranks; k++) {
            const ShardRec s = shards[k];
            if (t % s.tile_world != s.tile_rank || s.count == 0u) continue;
            const float4 v = (k == root) ? own[i] : recv[(size_t)k * recv_stride + (size_t)(t / s.tile_world) * 64u + in_tile];
            const float wgt = (s.count == total) ? 1.0f : (float)s.count / (float)total;
            const float4 w = make_float4(v.x * wgt, v.y * wgt, v.z * wgt, v.w * wgt);
            acc = any ? make_float4(acc.x + w.x, acc.y + w.y, acc.z + w.z, acc.w + w.w) : w;
            any = true;
        }
        if (TONEMAP) {
            const float a = fminf(fmaxf(acc.w, 0.0f), 1.0f);
            out8[i] = make_uchar4((unsigned char)srgb8_c(acc.x), (unsigned char)srgb8_c(acc.y), (unsigned char)srgb8_c(acc.z),
                                  (unsigned char)__float2int_rn(a * 255.0f));
        } else {
            out[i] = acc;
        }
    }
}

// ---- host side ----------------------------------------------------------------------------------------------------
struct Comm {
    ncclComm_t comm = nullptr;
    int n_ranks = 0, rank = -1;
    ShardRec* d_shards = nullptr;        // n_ranks records (all-gathered before every combine)
    ShardRec* d_mine = nullptr;
    ShardRec h_mine{};
    float4* send = nullptr; size_t send_px = 0;
    float4* recv = nullptr; size_t recv_px = 0;      // root only: n_ranks compact buffers
};

static std::string nccl_msg(const char* what, ncclResult_t r)
{
    return std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error") + " (" + std::to_string(r) + ")";
}

const char* comm_unique_id(void* id128)
{
    if (const char* e = nccl_load()) return e;
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != 0) { g_nccl.err = nccl_msg("ncclGetUniqueId", r); return g_nccl.err.c_str(); }
    memcpy(id128, &id, sizeof id);
    return nullptr;
}

const char* comm_create(Comm** out, int n_ranks, int rank, const void* id128)
{
    if (const char* e = nccl_load()) return e;
    Comm* c = new Comm();
    ncclUniqueId id; memcpy(&id, id128, sizeof id);
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, n_ranks, id, rank);
    if (r != 0) { delete c; g_nccl.err = nccl_msg("ncclCommInitRank", r); return g_nccl.err.c_str(); }
    c->n_ranks = n_ranks; c->rank = rank;
    cudaError_t e = cudaMalloc((void**)&c->d_shards, (size_t)n_ranks * sizeof(ShardRec));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_mine, sizeof(ShardRec));
    if (e != cudaSuccess) { comm_destroy(c); g_nccl.err = std::string("rt_comm_init: ") + cudaGetErrorString(e); return g_nccl.err.c_str(); }
    *out = c;
    return nullptr;
}

void comm_destroy(Comm* c)
{
    if (!c) return;
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->d_shards) cudaFree(c->d_shards);
    if (c->d_mine) cudaFree(c->d_mine);
    if (c->send) cudaFree(c->send);
    if (c->recv) cudaFree(c->recv);
    delete c;
}

int comm_rank(const Comm* c) { return c ? c->rank : -1; }
int comm_size(const Comm* c) { return c ? c->n_ranks : 0; }

// Everything is enqueued on `stream`.  out (W*H float4) and / or out8 (W*H uchar4) are written at the root only.
const char* comm_combine(Comm* c, const ShardRec& mine, int root, const float4* accum, float4* out, uint8_t* out8,
                         uint32_t width, uint32_t height, int n_sms, cudaStream_t stream)
{
    NcclApi& N = g_nccl;
    if (!c || !c->comm) return "rt_combine: no communicator (rt_comm_init)";
    if (root < 0 || root >= c->n_ranks) return "rt_combine: root out of range";
    const uint32_t tiles_x = (width + 7u) / 8u, tiles = tiles_x * ((height + 7u) / 8u);
    const size_t stride_px = (size_t)tiles * 64u;                                     // a rank owns at most every tile
    const size_t my_tiles = tiles > mine.tile_rank ? (tiles - mine.tile_rank + mine.tile_world - 1) / mine.tile_world : 0;
    const size_t my_px = my_tiles * 64u;
    // One job, one tile rule: every rank of the communicator uses the same tile_world (rt_set_shard), so a message is
    // the largest share under it (tile rank 0's) and the root can post its receives without knowing who owns what.
    const size_t msg_px = (size_t)((tiles + mine.tile_world - 1) / mine.tile_world) * 64u;
    cudaError_t e = cudaSuccess;
    auto cuda_err = [&](const char* what) { N.err = std::string(what) + ": " + cudaGetErrorString(e); return N.err.c_str(); };
    const bool is_root = c->rank == root;
    if (!is_root && c->send_px < msg_px) {
        if (c->send) { cudaStreamSynchronize(stream); cudaFree(c->send); c->send = nullptr; c->send_px = 0; }
        if ((e = cudaMalloc((void**)&c->send, (msg_px ? msg_px : 1) * sizeof(float4))) != cudaSuccess) return cuda_err("rt_combine: send buffer");
        c->send_px = msg_px;
    }
    if (is_root && c->recv_px < stride_px * (size_t)c->n_ranks) {
        if (c->recv) { cudaStreamSynchronize(stream); cudaFree(c->recv); c->recv = nullptr; c->recv_px = 0; }
        if ((e = cudaMalloc((void**)&c->recv, stride_px * (size_t)c->n_ranks * sizeof(float4))) != cudaSuccess) return cuda_err("rt_combine: receive buffer");
        c->recv_px = stride_px * (size_t)c->n_ranks;
    }
    // 1. the shard table
    c->h_mine = mine;                                    // pageable source: the copy is staged before cudaMemcpyAsync returns
    if ((e = cudaMemcpyAsync(c->d_mine, &c->h_mine, sizeof mine, cudaMemcpyHostToDevice, stream)) != cudaSuccess) return cuda_err("rt_combine: shard record");
    ncclResult_t r = N.AllGather(c->d_mine, c->d_shards, sizeof(ShardRec), ncclInt8, c->comm, stream);
    if (r != 0) { N.err = nccl_msg("ncclAllGather", r); return N.err.c_str(); }
    // 2. every other rank's tiles to the root
    const size_t wave = (size_t)(n_sms > 0 ? n_sms : 148) * 8;
    if (!is_root && my_px) {
        size_t blocks = (my_px + 255) / 256; if (blocks > wave) blocks = wave;
        pack_tiles_kernel<<<(unsigned)blocks, 256, 0, stream>>>(accum, c->send, width, height, tiles_x, mine.tile_rank, mine.tile_world, my_px);
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_err("pack_tiles_kernel");
    }
    if ((r = N.GroupStart()) != 0) { N.err = nccl_msg("ncclGroupStart", r); return N.err.c_str(); }
    if (is_root) {
        for (int k = 0; k < c->n_ranks && r == 0; k++)
            if (k != root) r = N.Recv(c->recv + (size_t)k * stride_px, msg_px * 4u, ncclFloat32, k, c->comm, stream);
    } else {
        r = N.Send(c->send, msg_px * 4u, ncclFloat32, root, c->comm, stream);
    }
    ncclResult_t r2 = N.GroupEnd();
    if (r != 0 || r2 != 0) { N.err = nccl_msg("ncclSend/ncclRecv", r != 0 ? r : r2); return N.err.c_str(); }
    // 3. the consumer
    if (is_root) {
        const size_t n = (size_t)width * height;
        size_t blocks = (n + 255) / 256; if (blocks > wave) blocks = wave;
        if (out8) gather_combine_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(c->d_shards, (uint32_t)c->n_ranks, (uint32_t)root, c->recv, stride_px, accum, out, (uchar4*)out8, width, height, tiles_x);
        else gather_combine_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(c->d_shards, (uint32_t)c->n_ranks, (uint32_t)root, c->recv, stride_px, accum, out, nullptr, width, height, tiles_x);
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_err("gather_combine_kernel");
    }
    return nullptr;
}

} // namespace rt
