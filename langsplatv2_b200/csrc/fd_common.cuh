// fd_common.cuh -- layout constants and small helpers of the fused blend + decode kernel (render_decode.cu)
#pragma once
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace lsv2 {

using namespace tc;

// ---- geometry of one CTA ------------------------------------------------------------------------------------
// Work unit: a 32 x 8 pixel band = the upper (or lower) halves of two horizontally adjacent 16x16 tiles.  Each half
// (128 pixels = one MMA M block) blends its own tile's depth-sorted list; the decode then sees the band as 8 rows of
// 32 pixels, so every store instruction of the epilogue writes one full 128-byte line.
constexpr int FD_BATCH = 16;                             // depth-sorted instances per batch (one MMA k-slice)
constexpr int FD_STAGES = 2;
constexpr int FD_BLEND_WARPS = 8;                        // warps 0-7: one thread per pixel; warp = 8x4 block of a half band
constexpr int FD_EPI_WARPS = 4;                          // warps 8-11: F slots (TMEM) -> scaled 128-byte-line stores
constexpr int FD_WARP_PROD = 12;                         // warps 12, 13: one per tile list -- gather instances, build S
constexpr int FD_PROD_THREADS = 64;
constexpr int FD_WARP_MB = 14;                           // issues the blend MMAs (and owns the TMEM allocation)
constexpr int FD_WARP_MD = 15;                           // issues the Gram / decode MMAs, fetches codebook chunks
constexpr int FD_THREADS = 16 * 32;                      // 512
constexpr int FD_THREADS_NODECODE = 15 * 32;             // weight-map-only variant: no MD warp (epilogue warps idle)
constexpr int FD_MAX_L = 192;
constexpr int FD_LC = 64;
constexpr int FD_D = 512;
constexpr int FD_KMAX = LSV2_SPARSE_CAP;
constexpr uint32_t FD_ROW_DENSE = 0xFFFFFFFFu;
constexpr int FD_UNIT_W = 32, FD_UNIT_H = 8;
// what follows the blend: nothing (the weight map is the product), the codebook decode, or the relevancy query
constexpr int FD_PATH_WMAP = 0, FD_PATH_DECODE = 1, FD_PATH_QUERY = 2;
constexpr int FD_CREC_WORDS = LSV2_PACKED_WORDS;         // 16 (hi | lo << 16) words, 16 channel bytes, count, pad

// blend operands: 64-byte rows [hi fp16 x16 | lo fp16 x16], 64-byte swizzle, K-major
constexpr int FD_AHALF_BYTES = 128 * 64;                 // alpha*T of one 128-pixel half                8 KB
constexpr int FD_S_BYTES = FD_MAX_L * 64;                // coefficients [channel][instance] of one list 12 KB
constexpr int FD_STAGE_BYTES = 2 * FD_AHALF_BYTES + 2 * FD_S_BYTES;                    // A0 | A1 | S0 | S1   40 KB
// decode operands
constexpr int FD_WDEC_LEVEL = 256 * 128;                 // fp16 weights of one level, [256 px][64 codes] 32 KB
constexpr int FD_CB_BYTES = 128 * 128;                   // codebook chunk [128 features][64 codes]     16 KB
constexpr int FD_CB_SLOTS = 2;
constexpr int FD_CHUNKS = FD_D / 128;                    // 4 chunks of 128 features per level
#ifndef FD_NFEAT_OPT
#define FD_NFEAT_OPT 64
#endif
constexpr int FD_NFEAT = FD_NFEAT_OPT;                   // features per decode MMA / F slot
constexpr int FD_FGROUPS = 128 / FD_NFEAT;               // feature groups per chunk
constexpr int FD_FSLOTS = 128 / FD_NFEAT;                // F slots in the 128 spare TMEM columns
constexpr int FD_TMEM_F = 2 * FD_MAX_L;                  // first F column (W halves occupy [0, 384))

struct __align__(16) FdHeader {
    float4 g0[2 * FD_BATCH];             // x, y, -a/2, -b             (list 0: [0,16), list 1: [16,32))
    float4 g1[2 * FD_BATCH];             // -c/2, power cut-off, opacity, r
    float2 g2[2 * FD_BATCH];             // g, b
    uint32_t mask[FD_BLEND_WARPS];       // per blend warp (half*4 + block): which of its list's 16 instances can reach its block
    int32_t count;                       // >= 0 batch ; -1 end of unit ; -2 no more units
    int32_t unit;
    int32_t batches;                     // terminator: batches this unit had
    int32_t pad;
};

struct FdTileInfo {
    int32_t unit;                        // < 0: no more units
    int32_t nb;
};

constexpr size_t FD_SMEM_RING = 0;
constexpr size_t FD_SMEM_HDR = FD_SMEM_RING + (size_t)FD_STAGES * FD_STAGE_BYTES;                   // 80 KB
constexpr size_t FD_SMEM_BAR = FD_SMEM_HDR + FD_STAGES * sizeof(FdHeader);
constexpr size_t FD_SMEM_INV = FD_SMEM_BAR + 48 * sizeof(uint64_t);                                 // f32 [2][3][256]
constexpr size_t FD_SMEM_MISC_END = (FD_SMEM_INV + 2 * 3 * 256 * sizeof(float) + 1023) & ~(size_t)1023;
constexpr size_t FD_SMEM_NODECODE_TOTAL = FD_SMEM_MISC_END + 1024;                                  // + alignment slack
constexpr size_t FD_SMEM_WDEC = FD_SMEM_MISC_END;
constexpr size_t FD_SMEM_CB = FD_SMEM_WDEC + (size_t)(FD_MAX_L / FD_LC) * FD_WDEC_LEVEL;            // + 96 KB
constexpr size_t FD_SMEM_TOTAL = FD_SMEM_CB + (size_t)FD_CB_SLOTS * FD_CB_BYTES + 1024;             // + 32 KB
static_assert(FD_SMEM_TOTAL <= 227 * 1024, "fused kernel shared memory budget");

struct FdArgs {
    int W, H, gx, n_tiles;       // gx: 16x16 tiles per row
    int ux, n_units;             // 32x8 units per band row, units in the image
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;
    const float* bg;
    int mode, L, K, levels;
    const float* coeff_dense;    // rows whose packed count is FD_ROW_DENSE are read from here
    const uint32_t* crec;        // packed coefficients u32[P][FD_CREC_WORDS] (lsv2_pack_coeff)
    float* color;
    float* wmap;                 // optional fp32 [L][H][W]
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* status;
    const uint8_t* bimg;         // fp16 codebook image [levels][512 rows][64 k], 128B-swizzled (scaled by cscale[l])
    const uint8_t* gimg;         // fp16 Gram images  s_l C_l C_l^T  [levels][hi 64 rows | lo 64 rows][64 k], 128B-swizzled
    const float* gscale_inv;     // [levels] 1 / s_l
    void* out;                   // [levels][512][H][W]  ([levels][H][W][512] when hwc)
    int hwc;
    const uint8_t* qimg;         // query: fp16 images of M_l = C_l E^T  [levels][hi 64 rows (phrase) | lo 64 rows][64 k], 128B-swizzled
    float* qout;                 // query: relevancy [levels][n_pos][H][W]
    int n_pos, n_neg;
    int normalize;
    uint32_t* tile_counter;
    long long* prof;             // -DFD_PROF builds only: [grid][4 roles][8] cycle counters
};

__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi)
{
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

// (a, b) -> fp16 pair hi and the fp16 pair of the remainders: a = hi.x + lo.x to ~2^-22 relative
__device__ __forceinline__ void split_f16x2(float a, float b, uint32_t& hi, uint32_t& lo)
{
    const __half2 h = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(h);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = pack_f16x2(__fsub_rn(a, hf.x), __fsub_rn(b, hf.y));
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

// Same mask from the extents stored in the blend record (cutoff_extents, common.cuh): a handful of instructions per
// instance in the producer instead of two divisions and two square roots.
__device__ __forceinline__ uint32_t fd_strip_mask_ext(float y_rel, float x_rel, float dy, float dx)
{
    if (!(dy >= 0.0f)) return 0u;
    if (x_rel + dx < 0.0f || x_rel - dx > 15.0f) return 0u;
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

// The fused kernel's blend warps each own an 8 x 4 pixel block of a 16 x 8 half band (bit = 2 * (row block) + (column
// block)): a splat is skipped by every warp whose block its alpha >= 1/255 bounding box misses.  Blocks instead of
// 16 x 2 strips: for a splat of radius r the boxes touched shrink from (2r + 16)(2r + 2) to (2r + 8)(2r + 4) pixels --
// 15-20 % fewer (pixel, instance) pairs at the radii of the metric's scene.  Conservative like fd_strip_mask_ext (the
// extents are rounded outwards, an integer row must lie inside the y range, the x test is on the continuous range),
// so no blend decision changes.
__device__ __forceinline__ uint32_t fd_block_mask_ext(float y_rel, float x_rel, float dy, float dx)
{
    if (!(dy >= 0.0f)) return 0u;
    const float lo_f = ceilf(y_rel - dy), hi_f = floorf(y_rel + dy);
    if (lo_f != lo_f || hi_f != hi_f) return 0xFu;
    if (hi_f < lo_f) return 0u;
    const float xlo = x_rel - dx, xhi = x_rel + dx;
    const uint32_t rows = ((lo_f <= 3.0f && hi_f >= 0.0f) ? 1u : 0u) | ((lo_f <= 7.0f && hi_f >= 4.0f) ? 2u : 0u);
    const uint32_t cols = (!(xhi < 0.0f || xlo > 7.0f) ? 1u : 0u) | (!(xhi < 8.0f || xlo > 15.0f) ? 2u : 0u);
    // bit 2 * by + bx
    return ((rows & 1u) ? cols : 0u) | ((rows & 2u) ? (cols << 2) : 0u);
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

// mbarrier wait for the helper warps (producer, MMA issuers, epilogue): backs off between polls so the spin does
// not take issue slots from the pixel warps
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) __nanosleep(32);
}

}  // namespace lsv2
