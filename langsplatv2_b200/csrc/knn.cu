// Mean squared distance of every point to its three nearest neighbours -- `simple_knn._C.distCUDA2`
// (/root/reference/submodules/simple-knn/simple_knn.cu:148-220, spatial.cu:15-25; called once per scene by
// scene/gaussian_model.py:194 to initialise the scales).  SURVEY.md 8f-4.
//
// What the reference computes, and what is reproduced bit for bit: for point i the three smallest values of
//     d(i, j) = fma(dz, dz, fma(dx, dx, dy * dy)),   d* = p_j - p_i in fp32,   j != i
// (that is what nvcc makes of its `d.x*d.x + d.y*d.y + d.z*d.z` without fast-math), summed in ascending order and
// divided by 3.0f; fewer than three other points leave FLT_MAX in the sum (FLT_MAX / 3 with two others, +inf below).  How the reference
// finds them -- Morton order, boxes of 1024 points, every thread testing every box and gathering `points[indices[i]]`
// one by one -- only prunes; any exact search gives the same three values.
//
// This version: no host round trip (the reference reads the bounding box back twice), the Morton sort is the repo's
// onesweep radix sort over packed (code << 32 | id) words, the points are gathered ONCE into Morton order as float4
// (x, y, z, id) with the bounds of every box of 256, and the search runs one CTA per box of queries:
//   1. the CTA's own box from shared memory gives every query a first triple and the CTA a rejection radius;
//   2. the 256 threads test 256 candidate boxes at a time against the CTA's box (box-to-box gap <= radius) and
//      compact the survivors;
//   3. each survivor is staged with one coalesced 4 KB load and scanned from shared memory (broadcast reads) by the
//      queries whose own point-to-box distance still beats their third-best.
// All pruning tests use the same monotone fp32 expression as d(i, j), so a box is never skipped that holds a closer point.
#include "common.cuh"
#include <cfloat>

namespace lsv2 {

constexpr int KN_BOX = 256;
constexpr int KN_MORTON_BITS = 30;

// floats ordered as signed ints; 0.0f <-> 0, so a zeroed word is the reference's reduction seed {0, 0, 0}
// (simple_knn.cu:190: the bounding box always contains the origin)
__device__ __forceinline__ int kn_ord(float f) { const int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7FFFFFFF; }
__device__ __forceinline__ float kn_unord(int s) { return __int_as_float(s >= 0 ? s : s ^ 0x7FFFFFFF); }

__device__ __forceinline__ float kn_dist2(float dx, float dy, float dz)
{
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

__global__ void __launch_bounds__(256)
k_knn_bounds(int P, const float* __restrict__ pts, int* __restrict__ bounds)
{
    float mn[3] = {0.f, 0.f, 0.f}, mx[3] = {0.f, 0.f, 0.f};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = pts[3 * (size_t)i + a];
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int lo = __reduce_min_sync(0xffffffffu, kn_ord(mn[a]));
        const int hi = __reduce_max_sync(0xffffffffu, kn_ord(mx[a]));
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds + a, lo);
            atomicMax(bounds + 3 + a, hi);
        }
    }
}

__device__ __forceinline__ uint32_t kn_spread10(uint32_t x)
{
    x &= 0x3FFu;
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256)
k_knn_morton(int P, const float* __restrict__ pts, const int* __restrict__ bounds, uint64_t* __restrict__ words,
             uint32_t* __restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) count[0] = (uint32_t)P;
    if (i >= P) return;
    uint32_t code = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = kn_unord(bounds[a]), hi = kn_unord(bounds[3 + a]);
        const float ext = hi - lo;
        float u = ext > 0.0f ? (pts[3 * (size_t)i + a] - lo) / ext * 1023.0f : 0.0f;
        u = fminf(fmaxf(u, 0.0f), 1023.0f);                    // NaN -> 0
        code |= kn_spread10((uint32_t)u) << a;
    }
    words[i] = ((uint64_t)code << 32) | (uint32_t)i;
}

// Morton order -> float4 (x, y, z, id) + per-box bounds: boxes[2b] = min, boxes[2b + 1] = max
__global__ void __launch_bounds__(KN_BOX)
k_knn_boxes(int P, const float* __restrict__ pts, const uint64_t* __restrict__ order, float4* __restrict__ sorted,
            float4* __restrict__ boxes)
{
    __shared__ float s_red[6][KN_BOX / 32];
    const int t = threadIdx.x, i = blockIdx.x * KN_BOX + t;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t id = (uint32_t)order[i];
        const float x = pts[3 * (size_t)id], y = pts[3 * (size_t)id + 1], z = pts[3 * (size_t)id + 2];
        sorted[i] = make_float4(x, y, z, __uint_as_float(id));
        mn[0] = mx[0] = x; mn[1] = mx[1] = y; mn[2] = mx[2] = z;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
        if ((t & 31) == 0) { s_red[a][t >> 5] = mn[a]; s_red[3 + a][t >> 5] = mx[a]; }
    }
    __syncthreads();
    if (t == 0) {
        float r[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            r[a] = s_red[a][0];
            for (int w = 1; w < KN_BOX / 32; ++w) r[a] = a < 3 ? fminf(r[a], s_red[a][w]) : fmaxf(r[a], s_red[a][w]);
        }
        boxes[2 * blockIdx.x] = make_float4(r[0], r[1], r[2], 0.f);
        boxes[2 * blockIdx.x + 1] = make_float4(r[3], r[4], r[5], 0.f);
    }
}

struct KnBest { float a, b, c; };                // ascending

__device__ __forceinline__ void kn_update(KnBest& k, const float4& q, const float4& p)
{
    const float d = kn_dist2(p.x - q.x, p.y - q.y, p.z - q.z);
    if (d < k.c) {                               // strict, as the reference's `knn[j] > dist`
        if (d < k.b) {
            k.c = k.b;
            if (d < k.a) { k.b = k.a; k.a = d; } else k.b = d;
        } else {
            k.c = d;
        }
    }
}

// lower bound (in the arithmetic of kn_dist2) of d(q, p) over the points p of a box
__device__ __forceinline__ float kn_point_box(const float4& q, const float4& lo, const float4& hi)
{
    const float dx = fmaxf(0.0f, fmaxf(lo.x - q.x, q.x - hi.x));
    const float dy = fmaxf(0.0f, fmaxf(lo.y - q.y, q.y - hi.y));
    const float dz = fmaxf(0.0f, fmaxf(lo.z - q.z, q.z - hi.z));
    return kn_dist2(dx, dy, dz);
}

// ... over the points p of box B and the queries q of box A
__device__ __forceinline__ float kn_box_box(const float4& alo, const float4& ahi, const float4& blo, const float4& bhi)
{
    const float dx = fmaxf(0.0f, fmaxf(blo.x - ahi.x, alo.x - bhi.x));
    const float dy = fmaxf(0.0f, fmaxf(blo.y - ahi.y, alo.y - bhi.y));
    const float dz = fmaxf(0.0f, fmaxf(blo.z - ahi.z, alo.z - bhi.z));
    return kn_dist2(dx, dy, dz);
}

__global__ void __launch_bounds__(KN_BOX)
k_knn_search(int P, int nboxes, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
             float* __restrict__ out)
{
    __shared__ float4 s_pts[KN_BOX];
    __shared__ uint32_t s_list[KN_BOX];
    __shared__ uint32_t s_n;
    __shared__ uint32_t s_rad[KN_BOX / 32];
    const int t = threadIdx.x, b = blockIdx.x;
    const int i = b * KN_BOX + t;
    const bool live = i < P;
    const int own = min(KN_BOX, P - b * KN_BOX);
    if (live) s_pts[t] = sorted[i];
    __syncthreads();
    const float4 q = live ? s_pts[t] : make_float4(0.f, 0.f, 0.f, 0.f);
    KnBest k{FLT_MAX, FLT_MAX, FLT_MAX};
    if (live) {
        for (int j = 0; j < own; ++j)
            if (j != t) kn_update(k, q, s_pts[j]);
    }
    // the CTA's rejection radius: the largest third-best (non-negative floats order as their bit patterns)
    const uint32_t wr = __reduce_max_sync(0xffffffffu, live ? __float_as_uint(k.c) : 0u);
    if ((t & 31) == 0) s_rad[t >> 5] = wr;
    if (t == 0) s_n = 0;
    __syncthreads();
    uint32_t rb = s_rad[0];
#pragma unroll
    for (int w = 1; w < KN_BOX / 32; ++w) rb = max(rb, s_rad[w]);
    const float radius = __uint_as_float(rb);
    const float4 alo = boxes[2 * b], ahi = boxes[2 * b + 1];

    for (int base = 0; base < nboxes; base += KN_BOX) {
        const int c = base + t;
        if (c < nboxes && c != b && kn_box_box(alo, ahi, boxes[2 * c], boxes[2 * c + 1]) <= radius)
            s_list[atomicAdd(&s_n, 1u)] = (uint32_t)c;        // any order: the three smallest values do not depend on it
        __syncthreads();
        const int n = (int)s_n;
        for (int e = 0; e < n; ++e) {
            const int cb = (int)s_list[e];
            const bool mine = live && kn_point_box(q, boxes[2 * cb], boxes[2 * cb + 1]) <= k.c;
            if (!__syncthreads_or(mine)) continue;             // also: everybody is done with the previous s_pts
            const int cnt = min(KN_BOX, P - cb * KN_BOX);
            if (t < cnt) s_pts[t] = sorted[(size_t)cb * KN_BOX + t];
            __syncthreads();
            if (mine) {
                for (int j = 0; j < cnt; ++j) kn_update(k, q, s_pts[j]);
            }
        }
        __syncthreads();                                       // s_list / s_n / s_pts are rewritten next round
        if (t == 0) s_n = 0;
        __syncthreads();
    }
    if (live) out[__float_as_uint(q.w)] = __fdiv_rn(__fadd_rn(__fadd_rn(k.a, k.b), k.c), 3.0f);
}

static size_t kn_al(size_t x) { return (x + 255) & ~(size_t)255; }

size_t knn_scratch_bytes(int P)
{
    if (P <= 0) return 0;
    const size_t p = (size_t)P, nb = (p + KN_BOX - 1) / KN_BOX;
    return kn_al(sort_scratch_bytes(p)) + 2 * kn_al(p * 8) + kn_al(p * 16) + kn_al(nb * 32) + 256;
}

int launch_dist2_knn3(int P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                      cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (!points || !mean_dist2 || !scratch) return LSV2_EINVAL;
    if (scratch_bytes < knn_scratch_bytes(P)) return LSV2_ENOSPC;
    const size_t p = (size_t)P;
    const int nboxes = (P + KN_BOX - 1) / KN_BOX;
    char* c = reinterpret_cast<char*>(scratch);
    void* sort_scr = c;                                   c += kn_al(sort_scratch_bytes(p));
    uint64_t* w_a = reinterpret_cast<uint64_t*>(c);       c += kn_al(p * 8);
    uint64_t* w_b = reinterpret_cast<uint64_t*>(c);       c += kn_al(p * 8);
    float4* sorted = reinterpret_cast<float4*>(c);        c += kn_al(p * 16);
    float4* boxes = reinterpret_cast<float4*>(c);         c += kn_al((size_t)nboxes * 32);
    int* bounds = reinterpret_cast<int*>(c);              // [0..5] bounds, [8] count
    uint32_t* count = reinterpret_cast<uint32_t*>(bounds + 8);

    LSV2_CUDA_TRY(cudaMemsetAsync(bounds, 0, 64, st));
    k_knn_bounds<<<min(ceil_div(P, 256), NUM_SMS * 8), 256, 0, st>>>(P, points, bounds);
    LSV2_LAUNCH_CHECK("k_knn_bounds");
    const bool odd = sort_num_passes(KN_MORTON_BITS) & 1;     // the sort wants its input where the result ends in w_a
    k_knn_morton<<<ceil_div(P, 256), 256, 0, st>>>(P, points, bounds, odd ? w_b : w_a, count);
    LSV2_LAUNCH_CHECK("k_knn_morton");
    const int rc = sort_words(w_a, w_b, count, p, KN_MORTON_BITS, sort_scr, kn_al(sort_scratch_bytes(p)), st);
    if (rc) return rc;
    k_knn_boxes<<<nboxes, KN_BOX, 0, st>>>(P, points, w_a, sorted, boxes);
    LSV2_LAUNCH_CHECK("k_knn_boxes");
    k_knn_search<<<nboxes, KN_BOX, 0, st>>>(P, nboxes, sorted, boxes, mean_dist2);
    LSV2_LAUNCH_CHECK("k_knn_search");
    return LSV2_OK;
}

}  // namespace lsv2
