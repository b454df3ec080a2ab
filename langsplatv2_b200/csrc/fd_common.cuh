// fd_common.cuh -- pieces shared by the fused blend + decode kernels (render_decode.cu, render_decode2.cu)
#pragma once
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace lsv2 {

using namespace tc;

constexpr int FD_BATCH = 64;
constexpr int FD_STAGES = 2;
constexpr int FD_BLEND_WARPS = 8;
constexpr int FD_PROD_THREADS = 64;
constexpr int FD_THREADS = 32 * FD_BLEND_WARPS + FD_PROD_THREADS + 32;      // 352
constexpr int FD_MAX_L = 192;
constexpr int FD_LC = 64;
constexpr int FD_D = 512;
constexpr int FD_CHUNK_N = 128;                          // features per decode MMA
constexpr int FD_CHUNKS = FD_D / FD_CHUNK_N;             // 4
constexpr int FD_TILE16K = 128 * 128;                    // [128 rows][128 B] operand tile
constexpr int FD_STAGE_BYTES = 2 * FD_TILE16K + FD_MAX_L * 128;             // A0 | A1 | S  = 56 KB
constexpr int FD_KMAX = LSV2_SPARSE_CAP;
constexpr uint32_t FD_ROW_DENSE = 0xFFFFFFFFu;

struct __align__(16) FdHeader {
    float4 g0[FD_BATCH];                 // x, y, -a/2, -b
    float4 g1[FD_BATCH];                 // -c/2, power cut-off, opacity, r
    float2 g2[FD_BATCH];                 // g, b
    uint32_t mask[FD_BLEND_WARPS][2];    // per strip: which of the 64 instances can reach it
    int32_t count;                       // >= 0 batch ; -1 end of tile ; -2 no more tiles
    int32_t tile;
    int32_t batches;                     // terminator: batches this tile had
    int32_t pad;
};

constexpr size_t FD_SMEM_RING = 0;
constexpr size_t FD_SMEM_B = FD_SMEM_RING + (size_t)FD_STAGES * FD_STAGE_BYTES;         // 112 KB
constexpr int FD_G_BYTES = FD_LC * 128;                                                 // Gram tile [64 n][64 k] fp16
constexpr size_t FD_SMEM_G = FD_SMEM_B + (size_t)FD_CHUNKS * FD_TILE16K;                // +64 KB
constexpr size_t FD_SMEM_HDR = FD_SMEM_G + (size_t)(FD_MAX_L / FD_LC) * FD_G_BYTES;     // +24 KB
constexpr size_t FD_SMEM_BAR = FD_SMEM_HDR + FD_STAGES * sizeof(FdHeader);
constexpr size_t FD_SMEM_TOTAL = FD_SMEM_BAR + 32 * sizeof(uint64_t) + 64 + 1024;       // + alignment slack

struct FdArgs {
    int W, H, gx, n_tiles;
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;
    const float* bg;
    int mode, L, K, levels;
    const float* coeff_dense;
    const float* wq;
    const float* iq;
    const uint32_t* sp_count;
    const uint16_t* sp_index;
    const float* sp_weight;
    float* color;
    float* wmap;                 // optional fp32 [L][H][W]
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* status;
    const uint8_t* bimg;         // fp16 codebook image [levels][512 rows][64 k], 128B-swizzled
    const uint8_t* gimg;         // fp16 Gram images  s_l * C_l C_l^T  [levels][64 n][64 k], 128B-swizzled
    const float* gscale_inv;     // [levels] 1 / s_l
    void* out;                   // [levels][512][H][W]
    int normalize;
    int dbg;                     // timing experiments only (LSV2_FD_DEBUG): 1 = no feature stores, 2 = no blend math
    uint32_t* tile_counter;
};

__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi)
{
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

// Instruction descriptor, kind::f16 with F16 operands: D = f32 (bit 4), A/B format 0 = F16, both K-major.
__host__ __device__ constexpr uint32_t umma_idesc_f16_f32(int M, int N)
{
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void named_bar_sync(int id, int count)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// see render_forward.cu: rows of the tile the alpha cut-off ellipse can reach -> mask over the 8 two-row strips.
// x_rel: the splat's x relative to the tile's first column; when the ellipse's bounding box misses the tile's 16
// columns altogether no pixel of the tile can reach alpha >= 1/255 and the mask is empty (conservative: the same
// 1.0001 / +0.02 margins as the row test, so no blend decision changes).
__device__ __forceinline__ uint32_t fd_strip_mask(float y_rel, float ax, float b, float az, float thr,
                                                  float x_rel = 7.5f)
{
    if (!(thr < 0.0f)) return 0u;
    const float qa = -2.0f * ax, qc = -2.0f * az;
    const float det = qa * qc - b * b;
    float dy = 1e30f;
    if (det > 0.0f && qa > 0.0f) dy = sqrtf((-2.0f * thr) * (qa / det)) * 1.0001f + 0.02f;
    if (!(dy == dy)) dy = 1e30f;
    if (det > 0.0f && qc > 0.0f) {
        const float dx = sqrtf((-2.0f * thr) * (qc / det)) * 1.0001f + 0.02f;
        if (dx == dx && (x_rel + dx < 0.0f || x_rel - dx > 15.0f)) return 0u;
    }
    const float lo_f = ceilf(y_rel - dy), hi_f = floorf(y_rel + dy);
    if (!(hi_f >= 0.0f) || !(lo_f <= 15.0f)) {
        if (lo_f != lo_f || hi_f != hi_f) return 0xFFu;
        return 0u;
    }
    const int lo = lo_f <= 0.0f ? 0 : (int)lo_f;
    const int hi = hi_f >= 15.0f ? 15 : (int)hi_f;
    if (hi < lo) return 0u;
    const int wlo = lo >> 1, whi = hi >> 1;
    return ((1u << (whi + 1)) - 1u) & ~((1u << wlo) - 1u);
}

// barrier with an OR reduction over the participating threads (uniform result)
__device__ __forceinline__ bool named_bar_or(int id, int count, bool pred)
{
    uint32_t r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.u32 q, %3, 0;\n\t"
        "bar.red.or.pred p, %1, %2, q;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(r)
        : "r"(id), "r"(count), "r"((uint32_t)pred)
        : "memory");
    return r != 0;
}

// mbarrier wait for the helper warps (producer, MMA issuer): backs off between polls so the spin does not
// take issue slots from the pixel warps
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) __nanosleep(40);
}

}  // namespace lsv2
