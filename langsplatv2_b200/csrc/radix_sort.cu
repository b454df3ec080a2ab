// radix_sort.cu -- K4: hand-written onesweep LSD radix sort of packed u64 words (sort field | u32 value).
// Replaces cub::DeviceRadixSort::SortPairs in the reference's rasterizer.
// Stable, least-significant digit first, <= 8-bit digits.  One histogram kernel counts every
// pass's digits in a single read of the keys; each pass is then ONE kernel that ranks its
// tile in shared memory (warp match-any multi-split), resolves its global digit offsets with
// decoupled look-back over the preceding tiles, and scatters keys and values through shared
// memory so global writes come out in runs.
//
// The sorted word carries its payload in the low 32 bits (the Gaussian id) and the field being sorted in the
// high 32 (depth bits for the per-Gaussian sort, tile id for the instance sort), so a pass moves 8 bytes per
// element.  The last pass of the instance sort also writes point_list = id; the order is the one a full 64-bit
// stable sort of the reference's (tile << 32 | depth bits) key produces, and that key is recovered from a sorted
// word as (word & ~0xffffffff) | bits(depth[id]).
// Only the bits that can differ are sorted on: the depth word is a positive float (sign bit clear) and the tile
// word needs ceil(log2(tiles)) bits.
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

// passes over the `nbits` low bits of the high word, spread evenly over ceil(nbits / 8) digits, larger first
static SortPlan make_plan(int nbits)
{
    SortPlan p;
    p.npass = 0;
    if (nbits <= 0) return p;
    const int n = (nbits + 7) / 8;
    int done = 0;
    for (int i = 0; i < n; ++i) {
        const int b = (nbits - done + (n - i) - 1) / (n - i);
        p.shift[p.npass] = 32 + done;
        p.bits[p.npass] = b;
        ++p.npass;
        done += b;
    }
    return p;
}

static size_t rs_tiles(uint64_t capacity) { return (size_t)((capacity + RS_TILE - 1) / RS_TILE); }

constexpr int RS_TILE_HIST = 8192;             // image tiles the histogram kernel can count in shared memory

// scratch layout: [hist: RS_MAX_PASSES*256 u32][counters: RS_MAX_PASSES u32 (+pad)][tile hist][lookback]
static size_t rs_header_bytes() { return (size_t)RS_MAX_PASSES * RS_RADIX * 4 + 64 + (size_t)RS_TILE_HIST * 4; }

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

// ---- all-pass digit histogram (+ the per-image-tile counts the tile ranges come from) --------
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
k_rs_histogram(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ count, uint64_t capacity,
               SortPlan plan, uint32_t* __restrict__ g_hist, uint32_t* __restrict__ g_tiles, int n_tiles)
{
    __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
    __shared__ uint32_t s_tiles[RS_TILE_HIST];
    const uint64_t n = clamp_count(count, capacity);
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += THREADS) s_hist[i] = 0;
    if (g_tiles)
        for (int i = threadIdx.x; i < n_tiles; i += THREADS) s_tiles[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * THREADS;
    uint64_t i = (uint64_t)blockIdx.x * THREADS + threadIdx.x;
    if (g_tiles) {                  // the digit counts follow from the tile counts (k_ranges_from_counts)
        for (; i + 3 * stride < n; i += 4 * stride) {          // four loads in flight per thread
            const uint32_t t0 = (uint32_t)(keys[i] >> 32), t1 = (uint32_t)(keys[i + stride] >> 32);
            const uint32_t t2 = (uint32_t)(keys[i + 2 * stride] >> 32), t3 = (uint32_t)(keys[i + 3 * stride] >> 32);
            if (t0 < (uint32_t)n_tiles) atomicAdd(&s_tiles[t0], 1u);
            if (t1 < (uint32_t)n_tiles) atomicAdd(&s_tiles[t1], 1u);
            if (t2 < (uint32_t)n_tiles) atomicAdd(&s_tiles[t2], 1u);
            if (t3 < (uint32_t)n_tiles) atomicAdd(&s_tiles[t3], 1u);
        }
        for (; i < n; i += stride) {
            const uint32_t t = (uint32_t)(keys[i] >> 32);
            if (t < (uint32_t)n_tiles) atomicAdd(&s_tiles[t], 1u);
        }
    }
    for (; i < n; i += stride) {
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
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&g_hist[i], c);
    }
    if (g_tiles)
        for (int i = threadIdx.x; i < n_tiles; i += THREADS) {
            const uint32_t c = s_tiles[i];
            if (c) atomicAdd(&g_tiles[i], c);
        }
}

// K5 from the tile counts: ranges[t] = [sum of counts below t, + count[t]) for tiles that own instances, (0, 0)
// otherwise -- what identifyTileRanges leaves behind on the sorted list.  The digit histograms of the sort passes
// are sums of the same counts.  One CTA, n_tiles <= RS_TILE_HIST.
__global__ void __launch_bounds__(1024)
k_ranges_from_counts(const uint32_t* __restrict__ g_tiles, int n_tiles, uint2* __restrict__ ranges, SortPlan plan,
                     uint32_t* __restrict__ g_hist)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += 1024) s_hist[i] = 0;
    __syncthreads();
    constexpr int PER = RS_TILE_HIST / 1024;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t c[PER], sum = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int t = tid * PER + k;
        c[k] = t < n_tiles ? g_tiles[t] : 0u;
        sum += c[k];
        if (c[k])
            for (int p = 0; p < plan.npass; ++p)
                atomicAdd(&s_hist[p * RS_RADIX + (((uint32_t)t >> (plan.shift[p] - 32)) & ((1u << plan.bits[p]) - 1u))],
                          c[k]);
    }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        s_warp[lane] = wi - w;
    }
    __syncthreads();
    uint32_t run = s_warp[warp] + incl - sum;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int t = tid * PER + k;
        if (t < n_tiles) ranges[t] = c[k] ? make_uint2(run, run + c[k]) : make_uint2(0u, 0u);
        run += c[k];
    }
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += 1024) g_hist[i] = s_hist[i];
}

// ---- one onesweep pass -------------------------------------------------------------------
// FINAL: the last pass of the instance sort also writes vals_out = low word (point_list).
template <bool FINAL>
__global__ void __launch_bounds__(RS_THREADS)
k_rs_onesweep(const uint64_t* __restrict__ keys_in, uint64_t* __restrict__ keys_out,
              uint32_t* __restrict__ vals_out, const uint32_t* __restrict__ n_ptr, uint64_t capacity, int shift, int bits,
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

    // load: warp-striped, item j of lane l = warp chunk + j*32 + l
    uint64_t key[RS_IPT];
    const uint32_t wbase = warp * (32 * RS_IPT) + lane;
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const uint32_t li = wbase + j * 32;
        key[j] = li < tile_n ? keys_in[tile_base + li] : ~0ull;
    }

    // early per-warp digit counts: the tile's aggregate is published before the ranking, so the look-back of
    // the following tiles never waits on this tile's ranking
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) atomicAdd(&s_hist[warp][(uint32_t)(key[j] >> shift) & mask], 1u);
    __syncthreads();

    // thread d: per-warp exclusive offsets inside digit d, tile count of digit d.  Padding items of the last tile
    // carry the all-ones key; they inflate its count of the top digit, which no later tile ever reads.
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
        const uint32_t t = s_hist[w][tid];
        s_hist[w][tid] = count;
        count += t;
    }
    volatile uint32_t* mine = lookback + (size_t)tile * RS_RADIX + tid;
    *mine = (tile == 0 ? LB_INC : LB_AGG) | count;

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
    __syncthreads();

    // rank inside the warp chunk (stable: row j before row j+1, lane order inside a row) and scatter into
    // tile-sorted order.  The peers of a lane (same digit) come from one ballot per digit bit.
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const uint32_t d = (uint32_t)(key[j] >> shift) & mask;
        uint32_t peers = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            if (b < bits) {
                const bool bit = (d >> b) & 1u;
                const uint32_t v = __ballot_sync(0xffffffffu, bit);
                peers &= bit ? v : ~v;
            }
        }
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (lane == leader) pre = s_hist[warp][d];
        pre = __shfl_sync(0xffffffffu, pre, leader);
        if (lane == leader) s_hist[warp][d] = pre + __popc(peers);
        s_keys[s_start[d] + pre + __popc(peers & lt_mask)] = key[j];
        __syncwarp();
    }

    // decoupled look-back for digit `tid`
    {
        uint32_t excl = 0;
        if (tile != 0) {
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
#pragma unroll
    for (int k = 0; k < RS_IPT; ++k) {
        const uint32_t i = tid + k * RS_THREADS;
        if (i < tile_n) {
            const uint64_t kk = s_keys[i];
            const uint32_t d = (uint32_t)(kk >> shift) & mask;
            const uint32_t dst = s_gbase[d] + i;
            if (FINAL) {
                const uint32_t id = (uint32_t)kk;
                vals_out[dst] = id;
                keys_out[dst] = kk;
            } else {
                keys_out[dst] = kk;
            }
        }
    }
}

// Sorts `count[0]` (<= capacity) packed words.  The input lives in k_a when the plan has an even number of passes
// and in k_b when odd, so the result always ends in k_a.  With vals_out the last pass also writes the low words there.
static int sort_packed(uint64_t* k_a, uint64_t* k_b, uint32_t* vals_out, const uint32_t* count,
                       uint64_t capacity, const SortPlan& plan, uint2* ranges, int n_tiles, void* scratch,
                       size_t scratch_bytes, cudaStream_t st)
{
    if (capacity == 0 || plan.npass == 0) return LSV2_OK;
    if (capacity >= (1ull << 30)) return LSV2_EUNSUPPORTED;
    const size_t tiles = rs_tiles(capacity);
    const size_t need = rs_header_bytes() + (size_t)plan.npass * tiles * RS_RADIX * sizeof(uint32_t);
    if (scratch_bytes < need) return LSV2_ENOSPC;
    LSV2_CUDA_TRY(cudaMemsetAsync(scratch, 0, need, st));
    uint32_t* g_hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* counters = g_hist + RS_MAX_PASSES * RS_RADIX;
    uint32_t* g_tiles = counters + 16;
    uint32_t* lookback = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(scratch) + rs_header_bytes());

    const bool odd = plan.npass & 1;
    uint64_t* k_src = odd ? k_b : k_a;
    uint64_t* k_dst = odd ? k_a : k_b;

    if (ranges) {
        // every CTA flushes up to n_tiles counts with global atomics: few, large CTAs
        const int hist_grid = (int)min((size_t)NUM_SMS, tiles);
        k_rs_histogram<1024><<<hist_grid, 1024, 0, st>>>(k_src, count, capacity, plan, g_hist, g_tiles, n_tiles);
    } else {
        const int hist_grid = (int)min((size_t)NUM_SMS * 8, tiles * (RS_TILE / RS_THREADS));
        k_rs_histogram<RS_THREADS><<<hist_grid, RS_THREADS, 0, st>>>(k_src, count, capacity, plan, g_hist, nullptr, 0);
    }
    LSV2_LAUNCH_CHECK("k_rs_histogram");
    if (ranges) {
        k_ranges_from_counts<<<1, 1024, 0, st>>>(g_tiles, n_tiles, ranges, plan, g_hist);
        LSV2_LAUNCH_CHECK("k_ranges_from_counts");
    }
    for (int p = 0; p < plan.npass; ++p) {
        uint32_t* ctr = counters + p;
        uint32_t* lb = lookback + (size_t)p * tiles * RS_RADIX;
        if (vals_out && p == plan.npass - 1)
            k_rs_onesweep<true><<<(unsigned)tiles, RS_THREADS, 0, st>>>(
                k_src, k_dst, vals_out, count, capacity, plan.shift[p], plan.bits[p], g_hist + p * RS_RADIX,
                ctr, lb);
        else
            k_rs_onesweep<false><<<(unsigned)tiles, RS_THREADS, 0, st>>>(
                k_src, k_dst, nullptr, count, capacity, plan.shift[p], plan.bits[p], g_hist + p * RS_RADIX,
                ctr, lb);
        LSV2_LAUNCH_CHECK("k_rs_onesweep");
        uint64_t* tk = k_src; k_src = k_dst; k_dst = tk;
    }
    return LSV2_OK;
}

// Instance sort on the tile word.  The packed instances were emitted into whichever of (keys, keys_alt) makes
// the last pass land in keys; that pass also writes point_list.  With
// n_tiles <= RS_TILE_HIST the tile ranges come out of the histogram pass (*ranges_done = true).
int launch_sort(const lsv2_binning& b, int tile_bits, int n_tiles, bool* ranges_done, void* scratch,
                size_t scratch_bytes, cudaStream_t st)
{
    const bool from_counts = n_tiles <= RS_TILE_HIST;
    *ranges_done = from_counts && b.capacity > 0;
    return sort_packed(b.keys, b.keys_alt, b.point_list, b.status + LSV2_STATUS_NUM_RENDERED, b.capacity,
                       make_plan(tile_bits), from_counts ? reinterpret_cast<uint2*>(b.ranges) : nullptr, n_tiles,
                       scratch, scratch_bytes, st);
}

// Stable sort of n packed words on the `nbits` low bits of their high word (the Morton order of knn.cu); input in k_b
// when sort_num_passes(nbits) is odd, result in k_a.
int sort_words(uint64_t* k_a, uint64_t* k_b, const uint32_t* count, uint64_t n, int nbits, void* scratch,
               size_t scratch_bytes, cudaStream_t st)
{
    return sort_packed(k_a, k_b, nullptr, count, n, make_plan(nbits), nullptr, 0, scratch, scratch_bytes, st);
}

// number of passes over a `nbits` field -- lsv2_bin uses it to pick the buffer the instances are emitted into.
int sort_num_passes(int nbits) { return make_plan(nbits).npass; }

// ---- depth order of the Gaussians ----------------------------------------------------------
// A stable sort of the instance list on (tile | depth) equals: stable-sort the P Gaussians on depth, emit their
// instances in that order, stable-sort the instances on the tile word alone.  The first sort moves P words instead
// of R (R/P ~ 12 at the metric's size), which takes the 31 depth bits (4 of the 6 passes) off the instance list.
// Equal (tile, depth) keys keep Gaussian-index order in both formulations.
__global__ void __launch_bounds__(256)
k_depth_keys(int P, const int32_t* __restrict__ radii, const float* __restrict__ depth, uint64_t* __restrict__ gk,
             uint32_t* __restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) count[0] = (uint32_t)P;
    if (i >= P) return;
    // culled Gaussians emit nothing, so any slot will do for them
    gk[i] = ((uint64_t)(radii[i] > 0 ? __float_as_uint(depth[i]) : 0u) << 32) | (uint32_t)i;
}

__global__ void __launch_bounds__(256)
k_gather_touched(int P, const uint64_t* __restrict__ order, const int32_t* __restrict__ radii,
                 const uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t g = (uint32_t)order[i];
    out[i] = radii[g] > 0 ? tiles_touched[g] : 0u;
}

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t depth_order_scratch_bytes(int P)
{
    if (P <= 0) return 0;
    const size_t p = (size_t)P;
    return al256(sort_scratch_bytes(p)) + 2 * al256(p * 8) + 2 * al256(p * 4) + al256(scan_scratch_bytes(P)) + 256;
}

// -> order[i] (low word) = the Gaussian of depth rank i; offsets[i] = inclusive scan of tiles_touched in that order
int launch_depth_order(int P, const lsv2_geom& g, int depth_bits, void* scratch, size_t scratch_bytes,
                       const uint64_t** order, const uint32_t** offsets, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (scratch_bytes < depth_order_scratch_bytes(P)) return LSV2_ENOSPC;
    const size_t p = (size_t)P;
    char* c = reinterpret_cast<char*>(scratch);
    void* sort_scr = c;                                   c += al256(sort_scratch_bytes(p));
    uint64_t* gk_a = reinterpret_cast<uint64_t*>(c);      c += al256(p * 8);
    uint64_t* gk_b = reinterpret_cast<uint64_t*>(c);      c += al256(p * 8);
    uint32_t* touched = reinterpret_cast<uint32_t*>(c);   c += al256(p * 4);
    uint32_t* offs = reinterpret_cast<uint32_t*>(c);      c += al256(p * 4);
    void* scan_scr = c;                                   c += al256(scan_scratch_bytes(P));
    uint32_t* count = reinterpret_cast<uint32_t*>(c);     // [0] = P, [1] = total (unused)

    const SortPlan plan = make_plan(depth_bits);
    k_depth_keys<<<ceil_div(P, 256), 256, 0, st>>>(P, g.radii, g.depth, (plan.npass & 1) ? gk_b : gk_a, count);
    LSV2_LAUNCH_CHECK("k_depth_keys");
    int rc = sort_packed(gk_a, gk_b, nullptr, count, p, plan, nullptr, 0, sort_scr,
                         al256(sort_scratch_bytes(p)), st);
    if (rc) return rc;
    k_gather_touched<<<ceil_div(P, 256), 256, 0, st>>>(P, gk_a, g.radii, g.tiles_touched, touched);
    LSV2_LAUNCH_CHECK("k_gather_touched");
    rc = launch_scan(P, touched, offs, count + 1, scan_scr, al256(scan_scratch_bytes(P)), st);
    if (rc) return rc;
    *order = gk_a;
    *offsets = offs;
    return LSV2_OK;
}

}  // namespace lsv2
