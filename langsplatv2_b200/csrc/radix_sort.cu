// radix_sort.cu -- K4: hand-written onesweep LSD radix sort of (u64 key, u32 value) pairs.
// Replaces cub::DeviceRadixSort::SortPairs in the reference's rasterizer.
// Stable, least-significant digit first, <= 8-bit digits.  One histogram kernel counts every
// pass's digits in a single read of the keys; each pass is then ONE kernel that ranks its
// tile in shared memory (warp match-any multi-split), resolves its global digit offsets with
// decoupled look-back over the preceding tiles, and scatters keys and values through shared
// memory so global writes come out in runs.
//
// Only the key bits that can differ are sorted on: the depth word is a positive float
// (sign bit clear) and the tile word needs ceil(log2(tiles)) bits -- the result is identical
// to a full 64-bit stable sort.
#include "common.cuh"

namespace lsv2 {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_IPT = 16;
constexpr int RS_TILE = RS_THREADS * RS_IPT;   // 4096 pairs per CTA
constexpr int RS_RADIX = 256;
constexpr int RS_MAX_PASSES = 8;

constexpr uint32_t LB_AGG = 1u << 30;
constexpr uint32_t LB_INC = 2u << 30;
constexpr uint32_t LB_FLAG = 3u << 30;
constexpr uint32_t LB_VAL = ~LB_FLAG;

struct SortPlan {
    int npass;
    int shift[RS_MAX_PASSES];
    int bits[RS_MAX_PASSES];
};

static SortPlan make_plan(int depth_bits, int tile_bits)
{
    SortPlan p;
    p.npass = 0;
    auto add_field = [&](int base, int nbits) {
        if (nbits <= 0) return;
        const int n = (nbits + 7) / 8;
        int done = 0;
        for (int i = 0; i < n; ++i) {
            const int b = (nbits - done + (n - i) - 1) / (n - i);   // spread evenly, larger first
            p.shift[p.npass] = base + done;
            p.bits[p.npass] = b;
            ++p.npass;
            done += b;
        }
    };
    add_field(0, depth_bits);
    add_field(32, tile_bits);
    return p;
}

static size_t rs_tiles(uint64_t capacity) { return (size_t)((capacity + RS_TILE - 1) / RS_TILE); }

// scratch layout: [hist: RS_MAX_PASSES*256 u32][counters: RS_MAX_PASSES u32 (+pad)][lookback]
static size_t rs_header_bytes() { return (size_t)RS_MAX_PASSES * RS_RADIX * 4 + 64; }

size_t sort_scratch_bytes(uint64_t capacity)
{
    return rs_header_bytes() + (size_t)RS_MAX_PASSES * rs_tiles(capacity) * RS_RADIX * sizeof(uint32_t);
}

// `count` points at the device word that holds the number of pairs (status[NUM_RENDERED] for the
// instance list, a scratch word for the per-Gaussian depth sort)
__device__ __forceinline__ uint64_t clamp_count(const uint32_t* count, uint64_t capacity)
{
    return min((uint64_t)count[0], capacity);
}

// ---- all-pass digit histogram ------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS)
k_rs_histogram(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ count, uint64_t capacity,
               SortPlan plan, uint32_t* __restrict__ g_hist)
{
    __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
    const uint64_t n = clamp_count(count, capacity);
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += RS_THREADS) s_hist[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
    for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
        const uint64_t k = keys[i];
#pragma unroll
        for (int p = 0; p < RS_MAX_PASSES; ++p) {
            if (p < plan.npass) {
                const uint32_t d = (uint32_t)(k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u);
                atomicAdd(&s_hist[p * RS_RADIX + d], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += RS_THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&g_hist[i], c);
    }
}

// ---- one onesweep pass -------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS)
k_rs_onesweep(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
              uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
              const uint32_t* __restrict__ n_ptr, uint64_t capacity, int shift, int bits,
              const uint32_t* __restrict__ g_hist /*[256] of this pass*/, uint32_t* counter,
              volatile uint32_t* lookback /*[tiles][256] of this pass*/)
{
    __shared__ uint64_t s_keys[RS_TILE];
    __shared__ uint32_t s_hist[RS_WARPS][RS_RADIX];
    __shared__ uint32_t s_start[RS_RADIX];   // first sorted position of each digit inside the tile
    __shared__ uint32_t s_gbase[RS_RADIX];   // global destination of sorted position 0 of the digit run
    __shared__ uint32_t s_scan[RS_WARPS];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t n = clamp_count(n_ptr, capacity);
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) s_hist[w][tid] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t tile_base = (uint64_t)tile * RS_TILE;
    if (tile_base >= n) return;
    const uint32_t tile_n = (uint32_t)min((uint64_t)RS_TILE, n - tile_base);
    const uint32_t mask = (1u << bits) - 1u;

    // global exclusive offset of this thread's digit (exclusive scan of the pass histogram)
    uint32_t g_excl;
    {
        const uint32_t c = g_hist[tid];
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) if (w < warp) woff += s_scan[w];
        g_excl = woff + incl - c;
        __syncthreads();
    }

    // load: warp-striped, item j of lane l = warp chunk + j*32 + l
    uint64_t key[RS_IPT];
    uint32_t val[RS_IPT];
    const uint32_t wbase = warp * (32 * RS_IPT) + lane;
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const uint32_t li = wbase + j * 32;
        if (li < tile_n) {
            key[j] = keys_in[tile_base + li];
            val[j] = vals_in[tile_base + li];
        } else {
            key[j] = ~0ull;
            val[j] = 0u;
        }
    }

    // rank inside the warp chunk (stable: row j before row j+1, lane order inside a row)
    uint32_t rank[RS_IPT];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (lane == leader) pre = s_hist[warp][d];
        pre = __shfl_sync(0xffffffffu, pre, leader);
        if (lane == leader) s_hist[warp][d] = pre + __popc(peers);
        rank[j] = pre + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // thread d: per-warp offsets inside digit d, tile count of digit d
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
        const uint32_t t = s_hist[w][tid];
        s_hist[w][tid] = count;
        count += t;
    }
    // exclusive scan of the tile's digit counts -> start of each digit run in sorted order
    {
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) if (w < warp) woff += s_scan[w];
        s_start[tid] = woff + incl - count;
    }

    // decoupled look-back for digit `tid`.  Padding items of the last tile carry the all-ones
    // key; they inflate its count of the top digit, which no later tile ever reads.
    {
        uint32_t excl = 0;
        volatile uint32_t* mine = lookback + (size_t)tile * RS_RADIX + tid;
        if (tile == 0) {
            *mine = LB_INC | count;
        } else {
            *mine = LB_AGG | count;
            for (int64_t t = (int64_t)tile - 1; t >= 0; --t) {
                uint32_t s;
                do { s = lookback[(size_t)t * RS_RADIX + tid]; } while ((s & LB_FLAG) == 0);
                excl += s & LB_VAL;
                if ((s & LB_FLAG) == LB_INC) break;
            }
            *mine = LB_INC | ((excl + count) & LB_VAL);
        }
        s_gbase[tid] = g_excl + excl - s_start[tid];
    }
    __syncthreads();

    // scatter keys into tile-sorted order
    uint32_t pos[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        pos[j] = s_start[d] + s_hist[warp][d] + rank[j];
        s_keys[pos[j]] = key[j];
    }
    __syncthreads();
    uint32_t dst[RS_IPT];
#pragma unroll
    for (int k = 0; k < RS_IPT; ++k) {
        const uint32_t i = tid + k * RS_THREADS;
        if (i < tile_n) {
            const uint64_t kk = s_keys[i];
            const uint32_t d = (uint32_t)(kk >> shift) & mask;
            dst[k] = s_gbase[d] + i;
            keys_out[dst[k]] = kk;
        }
    }
    __syncthreads();
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) s_vals[pos[j]] = val[j];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_IPT; ++k) {
        const uint32_t i = tid + k * RS_THREADS;
        if (i < tile_n) vals_out[dst[k]] = s_vals[i];
    }
}

// Sorts `count[0]` (<= capacity) pairs.  The input lives in (k_a, v_a) when the plan has an even number of
// passes and in (k_b, v_b) when odd, so the result always ends in (k_a, v_a).
static int sort_pairs(uint64_t* k_a, uint32_t* v_a, uint64_t* k_b, uint32_t* v_b, const uint32_t* count,
                      uint64_t capacity, const SortPlan& plan, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    if (capacity == 0 || plan.npass == 0) return LSV2_OK;
    if (capacity >= (1ull << 30)) return LSV2_EUNSUPPORTED;
    const size_t tiles = rs_tiles(capacity);
    const size_t need = rs_header_bytes() + (size_t)plan.npass * tiles * RS_RADIX * sizeof(uint32_t);
    if (scratch_bytes < need) return LSV2_ENOSPC;
    LSV2_CUDA_TRY(cudaMemsetAsync(scratch, 0, need, st));
    uint32_t* g_hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* counters = g_hist + RS_MAX_PASSES * RS_RADIX;
    uint32_t* lookback = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(scratch) + rs_header_bytes());

    const bool odd = plan.npass & 1;
    uint64_t* k_src = odd ? k_b : k_a;
    uint32_t* v_src = odd ? v_b : v_a;
    uint64_t* k_dst = odd ? k_a : k_b;
    uint32_t* v_dst = odd ? v_a : v_b;

    const int hist_grid = (int)min((size_t)NUM_SMS * 8, tiles * (RS_TILE / RS_THREADS));
    k_rs_histogram<<<hist_grid, RS_THREADS, 0, st>>>(k_src, count, capacity, plan, g_hist);
    LSV2_LAUNCH_CHECK("k_rs_histogram");
    for (int p = 0; p < plan.npass; ++p) {
        k_rs_onesweep<<<(unsigned)tiles, RS_THREADS, 0, st>>>(
            k_src, v_src, k_dst, v_dst, count, capacity, plan.shift[p], plan.bits[p],
            g_hist + p * RS_RADIX, counters + p, lookback + (size_t)p * tiles * RS_RADIX);
        LSV2_LAUNCH_CHECK("k_rs_onesweep");
        uint64_t* tk = k_src; k_src = k_dst; k_dst = tk;
        uint32_t* tv = v_src; v_src = v_dst; v_dst = tv;
    }
    return LSV2_OK;
}

// The unsorted pairs were emitted into whichever buffer makes the last pass land in (keys, point_list).
int launch_sort(const lsv2_binning& b, int depth_bits, int tile_bits, void* scratch, size_t scratch_bytes,
                cudaStream_t st)
{
    return sort_pairs(b.keys, b.point_list, b.keys_alt, b.vals_alt, b.status + LSV2_STATUS_NUM_RENDERED, b.capacity,
                      make_plan(depth_bits, tile_bits), scratch, scratch_bytes, st);
}

// number of passes for a (depth_bits, tile_bits) key -- lsv2_bin uses it to pick the buffer
// the unsorted pairs are emitted into.
int sort_num_passes(int depth_bits, int tile_bits) { return make_plan(depth_bits, tile_bits).npass; }

// ---- depth order of the Gaussians ----------------------------------------------------------
// A stable sort of the instance list on (tile | depth) equals: stable-sort the P Gaussians on depth, emit their
// instances in that order, stable-sort the instances on the tile word alone.  The first sort moves P pairs instead
// of R (R/P ~ 12 at the metric's size), which takes the 31 depth bits (4 of the 6 passes) off the instance list.
// Equal (tile, depth) keys keep Gaussian-index order in both formulations.
__global__ void __launch_bounds__(256)
k_depth_keys(int P, const int32_t* __restrict__ radii, const float* __restrict__ depth, uint64_t* __restrict__ gk,
             uint32_t* __restrict__ gv, uint32_t* __restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) count[0] = (uint32_t)P;
    if (i >= P) return;
    gk[i] = radii[i] > 0 ? (uint64_t)__float_as_uint(depth[i]) : 0ull;    // culled Gaussians emit nothing: any slot
    gv[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256)
k_gather_touched(int P, const uint32_t* __restrict__ order, const int32_t* __restrict__ radii,
                 const uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t g = order[i];
    out[i] = radii[g] > 0 ? tiles_touched[g] : 0u;
}

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t depth_order_scratch_bytes(int P)
{
    if (P <= 0) return 0;
    const size_t p = (size_t)P;
    return al256(sort_scratch_bytes(p)) + 2 * al256(p * 8) + 4 * al256(p * 4) + al256(scan_scratch_bytes(P)) + 256;
}

int launch_depth_order(int P, const lsv2_geom& g, int depth_bits, void* scratch, size_t scratch_bytes,
                       const uint32_t** order, const uint32_t** offsets, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (scratch_bytes < depth_order_scratch_bytes(P)) return LSV2_ENOSPC;
    const size_t p = (size_t)P;
    char* c = reinterpret_cast<char*>(scratch);
    void* sort_scr = c;                                   c += al256(sort_scratch_bytes(p));
    uint64_t* gk_a = reinterpret_cast<uint64_t*>(c);      c += al256(p * 8);
    uint64_t* gk_b = reinterpret_cast<uint64_t*>(c);      c += al256(p * 8);
    uint32_t* gv_a = reinterpret_cast<uint32_t*>(c);      c += al256(p * 4);
    uint32_t* gv_b = reinterpret_cast<uint32_t*>(c);      c += al256(p * 4);
    uint32_t* touched = reinterpret_cast<uint32_t*>(c);   c += al256(p * 4);
    uint32_t* offs = reinterpret_cast<uint32_t*>(c);      c += al256(p * 4);
    void* scan_scr = c;                                   c += al256(scan_scratch_bytes(P));
    uint32_t* count = reinterpret_cast<uint32_t*>(c);     // [0] = P, [1] = total (unused)

    const SortPlan plan = make_plan(depth_bits, 0);
    const bool odd = plan.npass & 1;
    k_depth_keys<<<ceil_div(P, 256), 256, 0, st>>>(P, g.radii, g.depth, odd ? gk_b : gk_a, odd ? gv_b : gv_a, count);
    LSV2_LAUNCH_CHECK("k_depth_keys");
    int rc = sort_pairs(gk_a, gv_a, gk_b, gv_b, count, p, plan, sort_scr, al256(sort_scratch_bytes(p)), st);
    if (rc) return rc;
    k_gather_touched<<<ceil_div(P, 256), 256, 0, st>>>(P, gv_a, g.radii, g.tiles_touched, touched);
    LSV2_LAUNCH_CHECK("k_gather_touched");
    rc = launch_scan(P, touched, offs, count + 1, scan_scr, al256(scan_scratch_bytes(P)), st);
    if (rc) return rc;
    *order = gv_a;
    *offsets = offs;
    return LSV2_OK;
}

}  // namespace lsv2
