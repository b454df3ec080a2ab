// binning.cu -- K3 duplicateWithKeys and K5 identifyTileRanges.
// An instance is emitted as ONE packed word (tile_id << 32) | Gaussian id; emission order inside one Gaussian is
// y-major over its tile rectangle and the Gaussians emit in depth order (radix_sort.cu, launch_depth_order), so the
// instance sort only has the tile word left to do.  The reference's pair is key = (tile_id << 32) |
// float_bits(depth), value = Gaussian id: same order, and the key is (word & ~0xffffffff) | bits(depth[id]).  The instance count lives on the device
// (status[NUM_RENDERED]); every kernel clamps to `capacity`.
#include "common.cuh"

namespace lsv2 {

// Slot i of the emission order is Gaussian order[i] (its depth rank; order == nullptr: i itself) and owns the
// instance range [offsets[i-1], offsets[i]).  A warp emits the ranges of its 32 slots as one contiguous run.
__global__ void __launch_bounds__(256)
k_duplicate(int P, int gx, int gy, const int32_t* __restrict__ radii, const float2* __restrict__ xy,
            const uint64_t* __restrict__ order, const uint32_t* __restrict__ offsets, uint64_t capacity,
            uint64_t* __restrict__ packed, uint32_t* __restrict__ status)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t gid = 0, off = 0, cnt = 0;
    int minx = 0, miny = 0, w = 1;
    if (i < P) {
        gid = order ? (uint32_t)order[i] : (uint32_t)i;
        const int r = radii[gid];
        if (r > 0) {
            off = (i == 0) ? 0u : offsets[i - 1];
            const uint32_t end = offsets[i];
            if ((uint64_t)end > capacity) status[LSV2_STATUS_OVERFLOW] = 1u;
            const float2 p = xy[gid];
            int maxx, maxy;
            tile_rect(p.x, p.y, (float)r, gx, gy, minx, miny, maxx, maxy);
            w = max(maxx - minx, 1);
            cnt = end - off;                      // == (maxx - minx) * (maxy - miny) = tiles_touched
        }
    }
    // The 32 ranges of a warp are adjacent (culled slots are empty), so the warp writes ONE run of `total` words
    // starting at its first slot's offset; word k belongs to the slot j with excl[j] <= k < incl[j] (binary search
    // over the warp's inclusive counts).
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const uint32_t excl = incl - cnt;
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    // offset of word 0 of the run: the start of the first non-empty slot
    const int first = __ffs(__ballot_sync(0xffffffffu, cnt > 0)) - 1;
    const uint64_t run = (uint64_t)__shfl_sync(0xffffffffu, off, first);
    for (uint32_t k0 = 0; k0 < total; k0 += 32) {
        const uint32_t k = min(k0 + lane, total - 1);
        int j = 0;
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const uint32_t v = __shfl_sync(0xffffffffu, incl, j + step - 1);
            if (v <= k) j += step;
        }
        const uint32_t e = __shfl_sync(0xffffffffu, excl, j);
        const uint32_t g = __shfl_sync(0xffffffffu, gid, j);
        const int x0 = __shfl_sync(0xffffffffu, minx, j);
        const int y0 = __shfl_sync(0xffffffffu, miny, j);
        const int ww = __shfl_sync(0xffffffffu, w, j);
        const uint64_t dst = run + k0 + lane;
        if (k0 + lane < total && dst < capacity) {
            const int kl = (int)(k - e);
            const int y = y0 + kl / ww, x = x0 + kl % ww;                  // y-major over the rectangle
            packed[dst] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | g;
        }
    }
}

int launch_duplicate(const lsv2_view& v, int P, const lsv2_geom& g, const lsv2_binning& b, const uint64_t* order,
                     const uint32_t* offsets, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    k_duplicate<<<ceil_div(P, 256), 256, 0, st>>>(P, gx, gy, g.radii, reinterpret_cast<const float2*>(g.xy),
                                                  order, offsets, b.capacity, b.keys_alt, b.status);
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
