// common.cuh -- shared constants and device helpers of liblsv2 (sm_100a only).
//
// NUMERIC POLICY: the library is compiled with --fmad=false; every fp32 multiply/add is
// rounded separately unless fmaf()/__fmaf_rn() is written out.  This is the same policy as
// the CPU checker (gcc -ffp-contract=off), which is what lets the integer outputs (radii,
// tiles_touched, sort keys, ranges, n_contrib) be compared bit for bit.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lsv2.h"

#if !defined(__CUDA_ARCH__) || (__CUDA_ARCH__ >= 1000)
#else
#error "liblsv2 is written for sm_100a (B200) only"
#endif

namespace lsv2 {

// Constants of the 3DGS tile rasterizer that the LangSplatV2 fork chain shares.  None of
// them can be read from /root/reference (the submodule is absent), so they live in one place.
constexpr int TILE = LSV2_TILE;
constexpr int TILE_PIX = TILE * TILE;
constexpr float NEAR_Z = 0.2f;
constexpr float W_GUARD = 0.0000001f;
constexpr float FRUSTUM_CLAMP = 1.3f;
constexpr float LOWPASS = 0.3f;
constexpr float EIG_FLOOR = 0.1f;
constexpr float RADIUS_SIGMA = 3.0f;
constexpr float ALPHA_MAX = 0.99f;
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float T_MIN = 0.0001f;

constexpr int NUM_SMS = 148;

// Thread-local text of the last CUDA failure (host side).
void set_cuda_error(cudaError_t e, const char* what);

#define LSV2_CUDA_TRY(expr)                                   \
    do {                                                      \
        cudaError_t _e = (expr);                              \
        if (_e != cudaSuccess) {                              \
            ::lsv2::set_cuda_error(_e, #expr);                \
            return LSV2_ECUDA;                                \
        }                                                     \
    } while (0)

#define LSV2_LAUNCH_CHECK(name)                               \
    do {                                                      \
        cudaError_t _e = cudaGetLastError();                  \
        if (_e != cudaSuccess) {                              \
            ::lsv2::set_cuda_error(_e, name);                 \
            return LSV2_ECUDA;                                \
        }                                                     \
    } while (0)

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// exp(x) for x <= 0 as a fixed-order polynomial: the same sequence of binary32 operations on
// every machine, so the alpha thresholds of the blend (and with them n_contrib) reproduce.
// <= 2 ulp from the correctly rounded value on [-87, 0].
__device__ __forceinline__ float repro_expf(float x)
{
    if (!(x >= -87.0f)) return 0.0f;
    if (x > 0.0f) x = 0.0f;
    const float t = __fmul_rn(x, 1.44269504088896341f);
    const float n = __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723212e-6f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    const float r2 = __fmul_rn(r, r);
    float y = __fmaf_rn(p, r2, r);
    y = __fadd_rn(y, 1.0f);
    return __int_as_float(__float_as_int(y) + (((int)n) << 23));
}

// Same operations as repro_expf without the range checks, for callers that already know -87 <= x <= 0
// (or discard the result otherwise): branch-free, so several evaluations interleave.
__device__ __forceinline__ float repro_expf_inrange(float x)
{
    const float t = __fmul_rn(x, 1.44269504088896341f);
    const float n = __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723212e-6f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    const float r2 = __fmul_rn(r, r);
    float y = __fmaf_rn(p, r2, r);
    y = __fadd_rn(y, 1.0f);
    return __int_as_float(__float_as_int(y) + (((int)n) << 23));
}

// Four independent evaluations of repro_expf_inrange, written step by step across the four so that the dependent
// chains interleave (same operations per element, hence bit-identical results).  The empty asm at the end pins the
// four results as unconditionally computed: without it the compiler sinks each chain under the branch of its only
// use and the four exponentials run back to back, each a ~20-instruction dependent chain.
__device__ __forceinline__ void repro_expf_inrange4(const float (&x)[4], float (&y)[4])
{
    float t[4], n[4], r[4], p[4], r2[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) t[u] = __fmul_rn(x[u], 1.44269504088896341f);
#pragma unroll
    for (int u = 0; u < 4; ++u) n[u] = __fsub_rn(__fadd_rn(t[u], 12582912.0f), 12582912.0f);
#pragma unroll
    for (int u = 0; u < 4; ++u) r[u] = __fmaf_rn(n[u], -0.693145751953125f, x[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) r[u] = __fmaf_rn(n[u], -1.42860682030941723212e-6f, r[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = __fmaf_rn(1.9875691500e-4f, r[u], 1.3981999507e-3f);
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = __fmaf_rn(p[u], r[u], 8.3334519073e-3f);
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = __fmaf_rn(p[u], r[u], 4.1665795894e-2f);
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = __fmaf_rn(p[u], r[u], 1.6666665459e-1f);
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = __fmaf_rn(p[u], r[u], 5.0000001201e-1f);
#pragma unroll
    for (int u = 0; u < 4; ++u) r2[u] = __fmul_rn(r[u], r[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) y[u] = __fadd_rn(__fmaf_rn(p[u], r2[u], r[u]), 1.0f);
#pragma unroll
    for (int u = 0; u < 4; ++u) y[u] = __int_as_float(__float_as_int(y[u]) + (((int)n[u]) << 23));
    asm volatile("" : "+f"(y[0]), "+f"(y[1]), "+f"(y[2]), "+f"(y[3]));
}

// Power cut-off of a splat: alpha = op*exp(power) < 1/255 for every power below it.  The margin
// (0.01 in the exponent, ~1 % in alpha) dwarfs the 2-ulp error of repro_expf, so skipping the
// exponential (or a whole pixel strip) below the cut-off never changes a decision.  A cut-off
// >= 0 means the splat can never contribute.
__device__ __forceinline__ float power_cutoff(float op)
{
    if (!(op > 0.0f)) return 1.0f;
    return -__logf(255.0f * op) - 0.01f;
}

// Half-extents (in pixels, padded: x 1.0001 + 0.02) of the axis-aligned box around the ellipse power >= thr of a
// splat with conic (a, b, c) given as ax = -a/2, bb = -b, az = -c/2 (the packed blend record): no pixel outside
// that box can reach alpha >= 1/255.  dy < 0: the splat can never contribute; 1e30: no bound on that axis.
__device__ __forceinline__ void cutoff_extents(float ax, float bb, float az, float thr, float& dy, float& dx)
{
    dy = -1.0f;
    dx = 1e30f;
    if (!(thr < 0.0f)) return;
    const float qa = -2.0f * ax, qc = -2.0f * az;
    const float det = qa * qc - bb * bb;
    dy = 1e30f;
    if (det > 0.0f && qa > 0.0f) dy = sqrtf((-2.0f * thr) * (qa / det)) * 1.0001f + 0.02f;
    if (!(dy == dy)) dy = 1e30f;
    if (det > 0.0f && qc > 0.0f) dx = sqrtf((-2.0f * thr) * (qc / det)) * 1.0001f + 0.02f;
    if (!(dx == dx)) dx = 1e30f;
}

// Tile rectangle covered by a splat of integer radius `rad` centred at (px, py).
__device__ __forceinline__ void tile_rect(float px, float py, float rad, int gx, int gy,
                                          int& minx, int& miny, int& maxx, int& maxy)
{
    minx = min(gx, max(0, (int)((px - rad) / (float)TILE)));
    miny = min(gy, max(0, (int)((py - rad) / (float)TILE)));
    maxx = min(gx, max(0, (int)((px + rad + (float)(TILE - 1)) / (float)TILE)));
    maxy = min(gy, max(0, (int)((py + rad + (float)(TILE - 1)) / (float)TILE)));
}

// Streaming (evict-first, no L1 allocate) global stores for write-once outputs.
__device__ __forceinline__ void st_stream_f32(float* p, float v)
{
    asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_stream_f32x4(float* p, float a, float b, float c, float d)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

}  // namespace lsv2

// ---- internal launchers (one per .cu) --------------------------------------------------
namespace lsv2 {
int launch_preprocess(const lsv2_view& v, const lsv2_scene& s, const lsv2_geom& g, cudaStream_t st);
int launch_scan(int P, const uint32_t* in, uint32_t* out_inclusive, uint32_t* total_out,
                void* scratch, size_t scratch_bytes, cudaStream_t st);
size_t scan_scratch_bytes(int P);
int launch_duplicate(const lsv2_view& v, int P, const lsv2_geom& g, const lsv2_binning& b, const uint64_t* order,
                     const uint32_t* offsets, cudaStream_t st);
size_t depth_order_scratch_bytes(int P);
int launch_depth_order(int P, const lsv2_geom& g, int depth_bits, void* scratch, size_t scratch_bytes,
                       const uint64_t** order, const uint32_t** offsets, cudaStream_t st);
int launch_ranges(const lsv2_view& v, const lsv2_binning& b, cudaStream_t st);
size_t sort_scratch_bytes(uint64_t capacity);
// stable sort of the packed instances (tile << 32 | id) on the tile word into keys; the last pass also writes
// point_list.  *ranges_done: the tile ranges were derived from the tile counts.
int launch_sort(const lsv2_binning& b, int tile_bits, int n_tiles, bool* ranges_done, void* scratch,
                size_t scratch_bytes, cudaStream_t st);
bool launch_gemm_nt_tf32x3(const float* X, const float* Y, float* out, int D, int Lc, int64_t HW, cudaStream_t st,
                           int* rc);
bool launch_gemm_h_tf32x3(const float* C, const float* gt, float* h, int D, int Lc, int64_t HW, cudaStream_t st,
                          int* rc);
int launch_compact_coeff(int P, const lsv2_geom& g, const lsv2_features& f, cudaStream_t st);
int launch_render_forward(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b,
                          const lsv2_features& f, const lsv2_image& img, cudaStream_t st);
int launch_render_backward(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b,
                           const lsv2_features& f, const lsv2_image& img, const float* dL_dwmap,
                           const float* dL_dcolor, int dense_grad, float* dL_dcoeff, float* dL_dcolors,
                           cudaStream_t st, int row_stride = 0);
int launch_topk_soft_code_bwd_rows(int64_t N, const float* logits, const float* dense, const uint32_t* sp_count,
                                   const uint16_t* sp_index, const float* d_rows, int row_stride, float* d_logits,
                                   cudaStream_t st);
int launch_mark_visible(int P, const float* means3d, const float* viewmatrix, uint8_t* present, cudaStream_t st);
int sort_num_passes(int nbits);
int sort_words(uint64_t* k_a, uint64_t* k_b, const uint32_t* count, uint64_t n, int nbits, void* scratch,
               size_t scratch_bytes, cudaStream_t st);
size_t knn_scratch_bytes(int P);
int launch_dist2_knn3(int P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                      cudaStream_t st);
int launch_topk_soft_code_sparse(int64_t N, int k, const float* logits, float* dense, uint32_t* sp_count,
                                 uint16_t* sp_index, float* sp_weight, uint32_t* packed, cudaStream_t st);
int launch_decode_tc_static(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                            int normalize, int out_dtype, void* out, cudaStream_t st);
int launch_render_backward_rgb(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_image& img,
                               const float* dL_dcolor, float* acc, cudaStream_t st);
int launch_preprocess_backward(const lsv2_view& v, const lsv2_scene& s, const lsv2_geom& g, const float* acc,
                               const lsv2_geom_grads& out, cudaStream_t st);
}  // namespace lsv2
