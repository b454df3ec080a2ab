// scan.cu -- K2: single-pass inclusive prefix sum of tiles_touched (chained scan with
// decoupled look-back).  Replaces cub::DeviceScan::InclusiveSum + the blocking D2H copy of
// the last element in the reference's rasterizer: the total is left on the
// device in status[LSV2_STATUS_NUM_RENDERED].
#include "common.cuh"

namespace lsv2 {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_IPT = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;

constexpr uint64_t FLAG_AGG = 1ull << 62;
constexpr uint64_t FLAG_INC = 2ull << 62;
constexpr uint64_t FLAG_MASK = 3ull << 62;

size_t scan_scratch_bytes(int P)
{
    const size_t tiles = (size_t)ceil_div(P > 0 ? P : 1, SCAN_TILE);
    return 16 + tiles * sizeof(uint64_t);
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan(int P, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint32_t* __restrict__ total_out,
       uint32_t* counter, volatile uint64_t* status)
{
    __shared__ uint32_t s_tile, s_prefix;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)tid * SCAN_IPT;

    uint32_t v[SCAN_IPT];
    if (base + SCAN_IPT <= P) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int k = 0; k < SCAN_IPT / 4; ++k) {
            const uint4 q = src[k];
            v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_IPT; ++k) v[k] = (base + k < P) ? in[base + k] : 0u;
    }
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; ++k) { sum += v[k]; v[k] = sum; }

    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const uint32_t t = s_warp[w];
        if (w < warp) warp_off += t;
        block_total += t;
    }

    if (tid == 0) {
        uint32_t prefix = 0;
        if (tile == 0) {
            status[0] = FLAG_INC | block_total;
        } else {
            status[tile] = FLAG_AGG | block_total;
            for (int64_t j = (int64_t)tile - 1; j >= 0; --j) {
                uint64_t s;
                do { s = status[j]; } while ((s & FLAG_MASK) == 0);
                prefix += (uint32_t)s;
                if ((s & FLAG_MASK) == FLAG_INC) break;
            }
            status[tile] = FLAG_INC | (uint64_t)(uint32_t)(prefix + block_total);
        }
        s_prefix = prefix;
        if ((int64_t)(tile + 1) * SCAN_TILE >= P) *total_out = prefix + block_total;
    }
    __syncthreads();
    const uint32_t off = s_prefix + warp_off + (incl - sum);
    if (base + SCAN_IPT <= P) {
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int k = 0; k < SCAN_IPT / 4; ++k)
            dst[k] = make_uint4(v[4 * k] + off, v[4 * k + 1] + off, v[4 * k + 2] + off, v[4 * k + 3] + off);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_IPT; ++k)
            if (base + k < P) out[base + k] = v[k] + off;
    }
}

int launch_scan(int P, const uint32_t* in, uint32_t* out_inclusive, uint32_t* total_out, void* scratch,
                size_t scratch_bytes, cudaStream_t st)
{
    if (P <= 0) {
        LSV2_CUDA_TRY(cudaMemsetAsync(total_out, 0, sizeof(uint32_t), st));
        return LSV2_OK;
    }
    const size_t need = scan_scratch_bytes(P);
    if (scratch_bytes < need) return LSV2_ENOSPC;
    LSV2_CUDA_TRY(cudaMemsetAsync(scratch, 0, need, st));
    uint32_t* counter = reinterpret_cast<uint32_t*>(scratch);
    uint64_t* status = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(scratch) + 16);
    k_scan<<<ceil_div(P, SCAN_TILE), SCAN_THREADS, 0, st>>>(P, in, out_inclusive, total_out, counter, status);
    LSV2_LAUNCH_CHECK("k_scan");
    return LSV2_OK;
}

}  // namespace lsv2
