// rt_combine.cu — the image combine of the multi-GPU partition as ONE kernel over
// NVLink peer memory: the device that owns the result reads every shard's accumulation image
// straight out of the peers' HBM (P2P loads through NVSwitch, no staging copies, no host round
// trip), applies the shard weights, and writes either the combined RGBA32F image or — fused — the
// tone-mapped RGBA8 one (what render.wgsl:34-39 + the sRGB swapchain show).  The reference is
// single-device and has no counterpart; the partition itself is rt_set_shard's.
//
//   tile shards   : tile t (8x8 pixels) belongs to the shard with t % tile_world == tile_rank; the
//                   kernel reads only the owners of a pixel, so a pure tile partition moves
//                   16 B per pixel over NVLink in total and its result is the owner's value
//                   bit for bit;
//   sample shards : every owner's running mean enters with its weight (default 1 / frame_world),
//                   summed in shard order with separately rounded multiply and add.
// NVLink-bound: (owners per pixel) x 16 B in, 16 B (or 4 B) out per pixel; grid-stride over
// 128-bit accesses, one CTA wave sized to the SM count.
#include "rt_kernels.h"

namespace rt {

// same arithmetic as tonemap_kernel (rt_tonemap.cu)
__device__ __forceinline__ uint32_t srgb8_of(float c)
{
    c = fminf(fmaxf(c, 0.0f), 1.0f);
    float s = (c <= 0.0031308f) ? 12.92f * c : __fmaf_rn(1.055f, spec_powf(c, 1.0f / 2.4f), -0.055f);
    return (uint32_t)__float2int_rn(fminf(fmaxf(s, 0.0f), 1.0f) * 255.0f);
}

template <bool TONEMAP>
__global__ void __launch_bounds__(256)
combine_kernel(PeerSet ps, float4* __restrict__ out, uchar4* __restrict__ out8, uint32_t width, uint32_t height, uint32_t tiles_x)
{
    const size_t n = (size_t)width * height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t py = (uint32_t)(i / width), px = (uint32_t)(i - (size_t)py * width);
        const uint32_t t = (py >> 3) * tiles_x + (px >> 3);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        bool any = false;
        #pragma unroll 4
        for (uint32_t k = 0; k < ps.n; k++) {
            const PeerSrc s = ps.p[k];
            if (s.tile_world != 1u && t % s.tile_world != s.tile_rank) continue;
            const float4 v = s.accum[i];                                   // peer load (NVLink) unless the shard lives on this device
            const float4 w = make_float4(v.x * s.weight, v.y * s.weight, v.z * s.weight, v.w * s.weight);
            acc = any ? make_float4(acc.x + w.x, acc.y + w.y, acc.z + w.z, acc.w + w.w) : w;
            any = true;
        }
        if (!any) acc = ps.p[0].accum[i];                                  // no shard owns this pixel: the first context's own value
        if (TONEMAP) {
            const float a = fminf(fmaxf(acc.w, 0.0f), 1.0f);
            out8[i] = make_uchar4((unsigned char)srgb8_of(acc.x), (unsigned char)srgb8_of(acc.y), (unsigned char)srgb8_of(acc.z),
                                  (unsigned char)__float2int_rn(a * 255.0f));
        } else {
            out[i] = acc;
        }
    }
}

cudaError_t launch_combine(const PeerSet& ps, float4* out, uint8_t* rgba8, uint32_t width, uint32_t height, int n_sms, cudaStream_t stream)
{
    const size_t n = (size_t)width * height;
    if (n == 0 || ps.n == 0) return cudaSuccess;
    const uint32_t tiles_x = (width + 7u) / 8u;
    size_t blocks = (n + 255) / 256;
    const size_t wave = (size_t)(n_sms > 0 ? n_sms : 148) * 8;            // 8 CTAs of 256 threads per SM
    if (blocks > wave) blocks = wave;
    if (rgba8) combine_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(ps, out, (uchar4*)rgba8, width, height, tiles_x);
    else combine_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(ps, out, nullptr, width, height, tiles_x);
    return cudaGetLastError();
}

} // namespace rt
