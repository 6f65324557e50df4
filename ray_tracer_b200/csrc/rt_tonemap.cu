// rt_tonemap.cu — what the reference's display path does to the accumulation texture:
// shaders/render.wgsl:34-39 returns the texel unchanged and the sRGB swapchain format
// (src/core/context.rs:75-80) clamps to [0,1] and applies the sRGB OETF in hardware.
// One thread per pixel, 16 B in / 4 B out: HBM-bound, coalesced 128-bit loads.
#include "rt_kernels.h"

namespace rt {

__device__ __forceinline__ uint32_t srgb8(float c)
{
    c = fminf(fmaxf(c, 0.0f), 1.0f);                 // NaN -> 0
    float s = (c <= 0.0031308f) ? 12.92f * c : __fmaf_rn(1.055f, spec_powf(c, 1.0f / 2.4f), -0.055f);
    return (uint32_t)__float2int_rn(fminf(fmaxf(s, 0.0f), 1.0f) * 255.0f);
}

__global__ void tonemap_kernel(const float4* __restrict__ accum, uchar4* __restrict__ out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 c = accum[i];
    float a = fminf(fmaxf(c.w, 0.0f), 1.0f);          // alpha is linear
    out[i] = make_uchar4((unsigned char)srgb8(c.x), (unsigned char)srgb8(c.y), (unsigned char)srgb8(c.z),
                         (unsigned char)__float2int_rn(a * 255.0f));
}

cudaError_t launch_tonemap(const float4* accum, uint8_t* rgba8, size_t n_pixels, cudaStream_t stream)
{
    if (n_pixels == 0) return cudaSuccess;
    const int T = 256;
    tonemap_kernel<<<(unsigned)((n_pixels + T - 1) / T), T, 0, stream>>>(accum, (uchar4*)rgba8, n_pixels);
    return cudaGetLastError();
}

} // namespace rt
