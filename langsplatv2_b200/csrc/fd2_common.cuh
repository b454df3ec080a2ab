// fd2_common.cuh -- pieces shared by the chunk-streaming fused kernels (render_decode2.cu, render_decode3.cu):
// codebook-image geometry, the tile hand-off record and the producer's gather / pack helpers.
#pragma once
#include "fd_common.cuh"

namespace lsv2 {

constexpr int F2_NCH = 32;                               // features per decode chunk / F slot
constexpr int F2_CB_BYTES = F2_NCH * 128;                // 4 KB
constexpr int F2_CB_SLOTS = 3;
constexpr int F2_G_CHUNKS = 4;                           // G_hi rows 0-31, 32-63, G_lo rows 0-31, 32-63
constexpr int F2_C_CHUNKS = FD_D / F2_NCH;               // 16
constexpr int F2_LVL_CHUNKS = F2_G_CHUNKS + F2_C_CHUNKS; // 20 chunks = 640 rows per level
constexpr int F2_LVL_BYTES = F2_LVL_CHUNKS * F2_CB_BYTES;

// builds the per-level images [G_hi | G_lo | codebook^T] (fp16, 128B-swizzled) and 1/s per level (render_decode2.cu)
int f2_prepare_images(int levels, const float* codebooks, int normalize, uint8_t* cimg, float* gscale_inv,
                      cudaStream_t st);
inline size_t f2_scratch_bytes(int levels) { return 512 + (size_t)(levels > 0 ? levels : 1) * F2_LVL_BYTES + 512; }

struct F2TileInfo {
    int32_t tile;        // < 0: no more tiles
    int32_t nb;          // batches blended (0: empty tile, W is all zero)
};

// One instance as gathered (raw, still in flight) and as scattered (packed fp16 weight | channel << 16).  The raw
// form is only converted after the previous batch has been assembled, so the loads' latency stays off the
// per-batch critical path.
struct F2Raw {
    uint32_t g, cnt;
    float4 r0, r1, r2;
    float wv[FD_KMAX], iv[FD_KMAX];
};
struct F2Packed {
    uint32_t g, cnt;
    float4 r0, r1, r2;
    uint32_t wc[FD_KMAX];
    bool ascending, bad;
};

template <int MODE>
__device__ __forceinline__ void f2_gather(const FdArgs& a, uint32_t g, F2Raw& o)
{
    o.g = g;
    o.cnt = 0;
    if (g == 0xFFFFFFFFu) return;
    o.r0 = __ldg(a.rec + 3 * (size_t)g);
    o.r1 = __ldg(a.rec + 3 * (size_t)g + 1);
    o.r2 = __ldg(a.rec + 3 * (size_t)g + 2);
    if (MODE == LSV2_MODE_QUICK) {
        o.cnt = (uint32_t)a.K;
        const float* wsrc = a.wq + (size_t)g * a.K;
        const float* isrc = a.iq + (size_t)g * a.K;
        if ((a.K & 3) == 0) {
#pragma unroll
            for (int k4 = 0; k4 < FD_KMAX / 4; ++k4)
                if (k4 * 4 < a.K) {
                    const float4 w4 = __ldg(reinterpret_cast<const float4*>(wsrc) + k4);
                    const float4 i4 = __ldg(reinterpret_cast<const float4*>(isrc) + k4);
                    o.wv[4 * k4] = w4.x; o.wv[4 * k4 + 1] = w4.y; o.wv[4 * k4 + 2] = w4.z; o.wv[4 * k4 + 3] = w4.w;
                    o.iv[4 * k4] = i4.x; o.iv[4 * k4 + 1] = i4.y; o.iv[4 * k4 + 2] = i4.z; o.iv[4 * k4 + 3] = i4.w;
                }
        } else {
#pragma unroll
            for (int k = 0; k < FD_KMAX; ++k)
                if (k < a.K) { o.wv[k] = __ldg(wsrc + k); o.iv[k] = __ldg(isrc + k); }
        }
    } else {
        o.cnt = a.sp_count[g];
        if (o.cnt != FD_ROW_DENSE) {
#pragma unroll
            for (int k = 0; k < FD_KMAX; ++k)
                if ((uint32_t)k < o.cnt) {
                    o.wv[k] = a.sp_weight[(size_t)g * FD_KMAX + k];
                    o.iv[k] = (float)a.sp_index[(size_t)g * FD_KMAX + k];
                }
        }
    }
}

__device__ __forceinline__ void f2_pack(const F2Raw& r, int L, F2Packed& o)
{
    o.g = r.g; o.cnt = r.cnt; o.r0 = r.r0; o.r1 = r.r1; o.r2 = r.r2;
    o.ascending = true;
    o.bad = false;
    if (r.g == 0xFFFFFFFFu || r.cnt == FD_ROW_DENSE) return;
    float prev = -1.0f;
#pragma unroll
    for (int k = 0; k < FD_KMAX; ++k)
        if ((uint32_t)k < r.cnt) {
            const int c = (int)r.iv[k];
            if (c < 0 || c >= L) { o.bad = true; o.wc[k] = 0xFFFF0000u; }
            else o.wc[k] = (uint32_t)__half_as_ushort(__float2half_rn(r.wv[k])) | ((uint32_t)c << 16);
            o.ascending &= r.iv[k] > prev;
            prev = r.iv[k];
        }
}

}  // namespace lsv2
