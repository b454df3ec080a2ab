// binning.cu -- K3 duplicateWithKeys and K5 identifyTileRanges.
// key = (tile_id << 32) | float_bits(depth); value = Gaussian id; emission order inside one
// Gaussian is y-major over its tile rectangle.  The instance count lives on the device
// (status[NUM_RENDERED]); every kernel clamps to `capacity`.
#include "common.cuh"

namespace lsv2 {

__global__ void __launch_bounds__(256)
k_duplicate(int P, int gx, int gy, const int32_t* __restrict__ radii, const float2* __restrict__ xy,
            const float* __restrict__ depth, const uint32_t* __restrict__ offsets, uint64_t capacity,
            uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ status)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    uint64_t off = (i == 0) ? 0u : offsets[i - 1];
    const uint32_t end = offsets[i];
    if ((uint64_t)end > capacity) status[LSV2_STATUS_OVERFLOW] = 1u;
    const float2 p = xy[i];
    int minx, miny, maxx, maxy;
    tile_rect(p.x, p.y, (float)r, gx, gy, minx, miny, maxx, maxy);
    const uint64_t dbits = (uint64_t)__float_as_uint(depth[i]);
    for (int y = miny; y < maxy; ++y)
        for (int x = minx; x < maxx; ++x) {
            if (off < capacity) {
                keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
                vals[off] = (uint32_t)i;
            }
            ++off;
        }
}

int launch_duplicate(const lsv2_view& v, int P, const lsv2_geom& g, const lsv2_binning& b, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    k_duplicate<<<ceil_div(P, 256), 256, 0, st>>>(P, gx, gy, g.radii, reinterpret_cast<const float2*>(g.xy),
                                                  g.depth, g.offsets, b.capacity, b.keys_alt, b.vals_alt,
                                                  b.status);
    LSV2_LAUNCH_CHECK("k_duplicate");
    return LSV2_OK;
}

__global__ void __launch_bounds__(256)
k_ranges(uint64_t capacity, const uint32_t* __restrict__ status, const uint64_t* __restrict__ keys,
         uint2* __restrict__ ranges)
{
    const uint64_t n = min((uint64_t)status[LSV2_STATUS_NUM_RENDERED], capacity);
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
        ranges[t].x = 0u;
    } else {
        const uint32_t pt = (uint32_t)(keys[i - 1] >> 32);
        if (t != pt) {
            ranges[pt].y = (uint32_t)i;
            ranges[t].x = (uint32_t)i;
        }
    }
    if (i == n - 1) ranges[t].y = (uint32_t)n;
}

int launch_ranges(const lsv2_view& v, const lsv2_binning& b, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    LSV2_CUDA_TRY(cudaMemsetAsync(b.ranges, 0, (size_t)gx * gy * 2 * sizeof(uint32_t), st));
    if (b.capacity == 0) return LSV2_OK;
    k_ranges<<<ceil_div((int64_t)b.capacity, 256), 256, 0, st>>>(b.capacity, b.status, b.keys,
                                                                reinterpret_cast<uint2*>(b.ranges));
    LSV2_LAUNCH_CHECK("k_ranges");
    return LSV2_OK;
}

}  // namespace lsv2
