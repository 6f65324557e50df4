// rt_order.cu — heaviest tiles first: the order in which a rank's 8x8 tiles enter the trace kernel's item queue.
//
// A launch ends when its last heavy items do: one teapot pixel per lane, on warps that by then run alone (0.3-0.9 ms, the
// fixed cost of every launch; profiles/r2_summary.md, tail probe).  If the heavy tiles are handed out FIRST, the queue ends
// on cheap background instead and that latency overlaps with the rest of the launch.  Which tiles are heavy is learned from
// the previous launch of the same context: during a launch's first dispatch every finished camera path adds its segment
// count to its tile (rt_trace.cu), and after the call the tiles are sorted by that weight for the next one — progressive
// rendering shows the same view many times over (the reference accumulates while the camera rests, context.rs:176-208).
// The first call after a resize / re-shard runs in raster order; a stale order (the camera moved) is still a valid order.
// Results do not depend on it: the order is a permutation of the rank's own tile list.  No reference counterpart.
#include <cub/cub.cuh>
#include "rt_kernels.h"

namespace rt {

namespace {
// heaviest tile and the sum of all weights: stats[0] (max), stats[4..5] (64-bit sum)
__global__ void order_max(uint32_t n, const uint32_t* __restrict__ cost, uint32_t* __restrict__ out)
{
    uint32_t m = 0; unsigned long long sum = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) { m = max(m, cost[i]); sum += cost[i]; }
    for (int o = 16; o > 0; o >>= 1) { m = max(m, __shfl_xor_sync(0xffffffffu, m, o)); sum += __shfl_xor_sync(0xffffffffu, sum, o); }
    if ((threadIdx.x & 31u) == 0u && m) { atomicMax(out, m); atomicAdd((unsigned long long*)(out + 4), sum); }
}
#ifndef RT_ORDER_FLAT_X4
#define RT_ORDER_FLAT_X4 8u           // flat: heaviest tile below this / 4 times the mean tile
#endif
#ifndef RT_ORDER_LEVELS
#define RT_ORDER_LEVELS 32u
#endif
// level width = max weight / levels, as a shift (stats[1]), and whether the weights are flat (stats[2]: heaviest tile below
// twice the mean): then there is nothing to reorder
__global__ void order_shift(uint32_t* m, uint32_t n)
{
    uint32_t s = 0;
    while ((m[0] >> s) > RT_ORDER_LEVELS - 1u) s++;
    const unsigned long long sum = *(unsigned long long*)(m + 4);
    m[2] = ((unsigned long long)m[0] * n * 4ull < (unsigned long long)RT_ORDER_FLAT_X4 * sum) ? 1u : 0u;
    m[1] = s; m[0] = 0u; *(unsigned long long*)(m + 4) = 0ull;
}
// weight -> sort key: RT_ORDER_LEVELS levels of descending weight; inside a level the stable sort keeps raster order
// (neighbouring tiles share hierarchy nodes, which the L1 of the SMs working on them likes); the weights are cleared
__global__ void order_keys_dyn(uint32_t n, uint32_t* __restrict__ cost, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, const uint32_t* __restrict__ m)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t level = min(cost[i] >> m[1], RT_ORDER_LEVELS - 1u);
    keys[i] = RT_ORDER_LEVELS - 1u - level; vals[i] = i;
    cost[i] = 0u;
}
} // namespace

void TileOrder::release()
{
    void* p[] = {cost, order[0], order[1], keys, keys_out, vals, stats, tmp};
    for (void* q : p) if (q) cudaFree(q);
    if (h_flat) cudaFreeHost(h_flat);
    if (ev_flat) cudaEventDestroy(ev_flat);
    h_flat = nullptr; ev_flat = nullptr; flat_pending = false; disabled = false;
    cost = order[0] = order[1] = keys = keys_out = vals = stats = nullptr; tmp = nullptr; tmp_bytes = 0; n = 0; valid = false; cur = 0;
}

cudaError_t TileOrder::ensure(uint32_t n_local, uint64_t geometry, cudaStream_t stream)
{
    if (n_local == n && geometry == geom) return cudaSuccess;
    cudaError_t e = cudaStreamSynchronize(stream);               // launches in flight read the old tables
    if (e != cudaSuccess) return e;
    if (n_local != n) {
        release();
        if (n_local == 0) { geom = geometry; return cudaSuccess; }
        uint32_t** arrs[] = {&cost, &order[0], &order[1], &keys, &keys_out, &vals};
        for (uint32_t** a : arrs) if ((e = cudaMalloc((void**)a, (size_t)n_local * sizeof(uint32_t))) != cudaSuccess) { release(); return e; }
        if ((e = cudaMalloc((void**)&stats, 8 * sizeof(uint32_t))) != cudaSuccess) { release(); return e; }
        if ((e = cudaHostAlloc((void**)&h_flat, sizeof(uint32_t), cudaHostAllocDefault)) != cudaSuccess) { release(); return e; }
        if ((e = cudaEventCreateWithFlags(&ev_flat, cudaEventDisableTiming)) != cudaSuccess) { release(); return e; }
        *h_flat = 0u;
        if ((e = cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys_out, vals, order[0], (int)n_local, 0, 5, stream)) != cudaSuccess) { release(); return e; }
        if ((e = cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1)) != cudaSuccess) { release(); return e; }
        n = n_local;
    }
    geom = geometry; valid = false; fed = false; disabled = false; flat_pending = false;
    if ((e = cudaMemsetAsync(cost, 0, (size_t)n * sizeof(uint32_t), stream)) != cudaSuccess) return e;
    return cudaMemsetAsync(stats, 0, 8 * sizeof(uint32_t), stream);
}

// A scene whose tiles all weigh about the same (C1's spheres filling the frame) gains nothing and pays for collecting the weights
// (-1.3 .. -3.5 %): once the device has said so, the context stops collecting until its tile list, scene or camera changes.
void TileOrder::poll()
{
    if (!flat_pending || cudaEventQuery(ev_flat) != cudaSuccess) { cudaGetLastError(); return; }
    flat_pending = false;
    if (*h_flat) { disabled = true; valid = false; }
}

// after a render call: the weights its launches collected become the next call's order
cudaError_t TileOrder::update(cudaStream_t stream)
{
    if (n < 2) return cudaSuccess;
    const uint32_t T = 256, nb = (n + T - 1) / T;
    order_max<<<nb < 64 ? nb : 64, T, 0, stream>>>(n, cost, stats);
    order_shift<<<1, 1, 0, stream>>>(stats, n);
    order_keys_dyn<<<nb, T, 0, stream>>>(n, cost, keys, vals, stats);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys_out, vals, order[cur ^ 1], (int)n, 0, 5, stream);
    if (e != cudaSuccess) return e;
    cur ^= 1; valid = true;
    // the flatness verdict travels to the host without a wait: poll() looks at it when the copy has landed
    if (cudaMemcpyAsync(h_flat, stats + 2, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream) == cudaSuccess && cudaEventRecord(ev_flat, stream) == cudaSuccess) flat_pending = true;
    return cudaGetLastError();
}

} // namespace rt
