// gemm_tf32x3.cu -- out[k][d] += sum_p Y[k][p] * X[d][p]   (k < 64, d < D, p < HW)
//
// The codebook gradient of the decode (scene/gaussian_model.py:538, dC = W dF^T) and the W W^T product of the
// feature-map-free loss (loss.cu) are this one "NT" product with the PIXEL axis as the contraction: both operands
// are K-major as they lie in HBM ([rows][HW], pixels contiguous), which is the layout tcgen05.mma wants.
//
// Precision: the reference runs these products as fp32 GEMMs.  Each fp32 operand is split in place into
// hi = x with the low 13 mantissa bits cleared (exactly a tf32 number) and lo = x - hi (exact), and the product is
// accumulated in fp32 as hi*hi + hi*lo + lo*hi on the tf32 tensor-core path ("3xTF32"); the dropped lo*lo term and
// the truncation of lo are both <= 2^-20 relative per product.
//
// The tensor core's fp32 accumulate does not round to nearest: a long chain in TMEM drifts (measured 7e-5 relative
// over 3864 accumulations).  So TMEM only ever holds GX_CHUNK steps (24 accumulations); four epilogue warps drain
// each chunk into round-to-nearest fp32 register accumulators while the next chunk runs into the other TMEM
// buffer -- measured error then equals the CUDA-core fp32 kernel's.
//
// One CTA = a slab of pixels x MT*128 rows of X.  Warps 0-7 stream 32-pixel steps of X and Y into 128B-swizzled
// tiles with 16-byte async copies, split the chunks they copied themselves (no cross-thread hand-off), and
// arrive on the stage's barrier; the next warp issues the MMAs (2 x MT x [128 x 64] fp32 in TMEM); 4 warps per M tile
// accumulate the chunks and at the end add their registers to `out` (split-K over the slabs: fp32 atomics).
#include "common.cuh"
#include "tc_common.cuh"

namespace lsv2 {
using namespace tc;

constexpr int GX_STAGES = 3;                       // raw (-> hi) tiles in flight; the lo tiles have 2 buffers
constexpr int GX_KSTEP = 32;                       // pixels per stage: 32 fp32 = one 128-byte swizzled row
constexpr int GX_LOADERS = 256;                    // 8 warps copy + split (they are the critical path of a step)
constexpr int GX_LW = GX_LOADERS / 32;             // the MMA warp id
constexpr int gx_threads(int MT) { return GX_LOADERS + 32 + 128 * MT; }   // + MMA warp + 4 epilogue warps per M tile
constexpr int GX_CHUNK = 2;                        // steps accumulated in TMEM before the registers take over
constexpr int GX_Y_BYTES = 64 * 128;               // 8 KB
constexpr int GX_XT_BYTES = 128 * 128;             // 16 KB per 128-row tile

template <int MT> struct GxLayout {
    static constexpr int X_BYTES = MT * GX_XT_BYTES;
    static constexpr int STAGE = X_BYTES + GX_Y_BYTES;                 // X | Y (raw, then hi in place); lo: same shape
    static constexpr int LO_OFF = GX_STAGES * STAGE;
    static constexpr int BAR_OFF = LO_OFF + 2 * STAGE;
    static constexpr int TOTAL = BAR_OFF + 1024 /*align*/ + 128 /*barriers*/;
    static constexpr int TMEM_COLS = 2 * MT * 64;                      // two chunk buffers: 128 or 256 columns
};

// kind::tf32: D = f32 (bit 4), A/B format 2 = TF32, both K-major
__host__ __device__ constexpr uint32_t umma_idesc_tf32_f32(int M, int N)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int MT>
__global__ void __launch_bounds__(gx_threads(MT), 1)
k_gemm_nt_tf32x3(const float* __restrict__ X, const float* __restrict__ Y, float* __restrict__ out, int D, int64_t HW,
                 int64_t slab)
{
    using L = GxLayout<MT>;
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* full = bars;                          // [GX_STAGES] 128 loader arrivals
    uint64_t* empty = bars + GX_STAGES;             // [GX_STAGES] MMA commit of the step that read the stage
    uint64_t* acc_full = bars + 2 * GX_STAGES;      // [2] MMA commit of a chunk
    uint64_t* acc_free = acc_full + 2;              // [2] 128 * MT epilogue arrivals
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(acc_free + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t p_beg = (int64_t)blockIdx.x * slab;
    const int64_t p_end = min(HW, p_beg + slab);
    const int d0 = blockIdx.y * (MT * 128);
    const int nsteps = (int)((p_end - p_beg + GX_KSTEP - 1) / GX_KSTEP);

    if (tid == 0) {
        for (int s = 0; s < GX_STAGES; ++s) { mbar_init(full + s, GX_LOADERS); mbar_init(empty + s, 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(acc_full + b, 1); mbar_init(acc_free + b, 128 * MT); }
        mbar_fence_init();
    }
    if (warp == GX_LW) tmem_alloc(s_tmem, L::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp < GX_LW) {
        // ---- loaders: chunk c = tid + 128*i of a tile -> row c>>3, 16-byte chunk c&7 -----------------
        constexpr int XCH = MT * 1024 / GX_LOADERS, YCH = 512 / GX_LOADERS;
        auto issue = [&](int s) {
            unsigned char* st = smem + (s % GX_STAGES) * L::STAGE;
            const int64_t p0 = p_beg + (int64_t)s * GX_KSTEP;
#pragma unroll
            for (int i = 0; i < XCH + YCH; ++i) {
                const bool isx = i < XCH;
                const int c = tid + GX_LOADERS * (isx ? i : i - XCH);
                const int row = c >> 3, ch = c & 7;
                const int64_t p = p0 + ch * 4;
                const int grow = isx ? d0 + row : row;
                const bool ok = p < p_end && (!isx || grow < D);
                const float* src = (isx ? X : Y) + (ok ? (size_t)grow * HW + p : 0);
                unsigned char* dst = st + (isx ? 0 : L::X_BYTES) + sw128_offset(row, ch);
                cp_async16(smem_u32(dst), src, ok ? 16u : 0u);
            }
            cp_async_commit();
        };
        if (nsteps > 0) issue(0);
        if (nsteps > 1) issue(1);
        for (int s = 0; s < nsteps; ++s) {
            if (s + 1 < nsteps) cp_async_wait<1>();
            else cp_async_wait<0>();
            // lo buffer s%2 was last read by step s-2
            if (s >= 2) mbar_wait(empty + (s - 2) % GX_STAGES, ((s - 2) / GX_STAGES) & 1);
            unsigned char* st = smem + (s % GX_STAGES) * L::STAGE;
#pragma unroll
            for (int i = 0; i < XCH + YCH; ++i) {
                const bool isx = i < XCH;
                const int c = tid + GX_LOADERS * (isx ? i : i - XCH);
                const uint32_t off = (isx ? 0 : L::X_BYTES) + sw128_offset(c >> 3, c & 7);
                unsigned char* hi = st + off;
                unsigned char* lo = smem + L::LO_OFF + (s & 1) * L::STAGE + off;
                const float4 v = *reinterpret_cast<const float4*>(hi);
                float4 h, l;
                h.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l.x = v.x - h.x;
                h.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l.y = v.y - h.y;
                h.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l.z = v.z - h.z;
                h.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l.w = v.w - h.w;
                // the head is not written back: kind::tf32 truncates its fp32 operands itself (see k_gemm_h_tf32x3)
                *reinterpret_cast<float4*>(lo) = l;
            }
            fence_async_smem();
            mbar_arrive(full + s % GX_STAGES);
            // step s-1's MMAs (running under the split above) done: stage (s+2)%3 == (s-1)%3 takes the next loads
            if (s + 2 < nsteps) {
                if (s >= 1) mbar_wait(empty + (s - 1) % GX_STAGES, ((s - 1) / GX_STAGES) & 1);
                issue(s + 2);
            }
        }
    } else if (warp > GX_LW) {
        // ---- epilogue warps: TMEM lane = X row inside the tile, column = k ---------------------------
        // warp 5 + 4*m + j drains M tile m; a warp may only read the TMEM lane quarter (warp id % 4)
        const int q = warp & 3, m = (warp - GX_LW - 1) >> 2;
        const int nchunks = (nsteps + GX_CHUNK - 1) / GX_CHUNK;
        float acc[64];
#pragma unroll
        for (int k = 0; k < 64; ++k) acc[k] = 0.0f;
        for (int c = 0; c < nchunks; ++c) {
            const int bsel = c & 1;
            mbar_wait(acc_full + bsel, (c >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t r[32];
                tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + bsel * (MT * 64) + m * 64 + half * 32, r);
                tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < 32; ++k) acc[half * 32 + k] += __uint_as_float(r[k]);
            }
            tc_fence_before();
            mbar_arrive(acc_free + bsel);
        }
        const int d = d0 + m * 128 + q * 32 + lane;
        if (d < D) {
#pragma unroll
            for (int k = 0; k < 64; ++k)
                if (acc[k] != 0.0f) atomicAdd(out + (size_t)k * D + d, acc[k]);
        }
    } else if (lane == 0) {
        // ---- MMA issuer ---------------------------------------------------------------------------
        const uint32_t idesc = umma_idesc_tf32_f32(128, 64);
        for (int s = 0; s < nsteps; ++s) {
            const int c = s / GX_CHUNK, bsel = c & 1;
            const bool first = s % GX_CHUNK == 0;
            if (first && c >= 2) {                                       // the registers took chunk c - 2
                mbar_wait(acc_free + bsel, ((c >> 1) - 1) & 1);
                tc_fence_after();
            }
            mbar_wait(full + s % GX_STAGES, (s / GX_STAGES) & 1);
            tc_fence_after();
            const uint32_t st = smem_u32(smem + (s % GX_STAGES) * L::STAGE);
            const uint32_t lo = smem_u32(smem + L::LO_OFF + (s & 1) * L::STAGE);
            const uint32_t y_hi = st + L::X_BYTES, y_lo = lo + L::X_BYTES;
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const uint32_t x_hi = st + m * GX_XT_BYTES, x_lo = lo + m * GX_XT_BYTES;
#pragma unroll
                for (int k8 = 0; k8 < GX_KSTEP / 8; ++k8) {
                    const uint64_t koff = (uint64_t)((k8 * 32) >> 4);
                    const uint64_t a_hi = umma_desc_k_sw128(x_hi) + koff, a_lo = umma_desc_k_sw128(x_lo) + koff;
                    const uint64_t b_hi = umma_desc_k_sw128(y_hi) + koff, b_lo = umma_desc_k_sw128(y_lo) + koff;
                    const uint32_t d_t = tmem + bsel * (MT * 64) + m * 64;
                    umma_tf32(d_t, a_lo, b_hi, idesc, (!first || k8 > 0) ? 1u : 0u);     // small terms first
                    umma_tf32(d_t, a_hi, b_lo, idesc, 1u);
                    umma_tf32(d_t, a_hi, b_hi, idesc, 1u);
                }
            }
            umma_commit(empty + s % GX_STAGES);
            if (s % GX_CHUNK == GX_CHUNK - 1 || s == nsteps - 1) umma_commit(acc_full + bsel);
        }
    }
    __syncthreads();
    if (warp == GX_LW) {
        tc_fence_after();
        tmem_dealloc(tmem, L::TMEM_COLS);
    }
}

// ---- h[k][p] = sum_d C[k][d] * gt[d][p]  (k < 64, d < D, p < HW) ---------------------------------------
// The other 64 x D x HW product of the training step (dL/dW = C dF of scene/gaussian_model.py:538's backward; h = C gt
// of the feature-map-free loss).  The contraction runs over the feature axis, so the big operand gt[d][p] is
// MN-major (pixels contiguous): A = gt tile, M = 128 pixels, MN-major 128B-swizzled panels of 32 pixels x 8 d;
// B = codebook tile [64 codes][32 d], K-major.  Same 3xTF32 split, same 2-step TMEM chunks drained into registers;
// a CTA walks its pixel tiles with the 16 d-steps of a tile as the inner loop and stores h at the end of each tile.
constexpr int GH_A_BYTES = 32 * 512;               // 32 d rows x 128 pixels fp32 = 4 panels of 4 KB
constexpr int GH_B_BYTES = 64 * 128;               // 64 codes x 32 d
constexpr int GH_STAGE = GH_A_BYTES + GH_B_BYTES;  // 24 KB
// deeper ring than the pixel-axis kernels: this one streams 1.56 GB with one CTA per SM, and what it keeps in flight
// (GH_PF stages x 16 KB) bounds its bandwidth: 2 stages in flight measured 2.6 TB/s (profiles/r02r)
constexpr int GH_STAGES = 5;
constexpr int GH_PF = 4;                           // stages requested ahead of the one being split
constexpr int GH_LO_OFF = GH_STAGES * GH_STAGE;
constexpr int GH_BAR_OFF = GH_LO_OFF + 2 * GH_STAGE;
constexpr int GH_TOTAL = GH_BAR_OFF + 1024 + 128;
constexpr int GH_THREADS = GX_LOADERS + 32 + 128;
constexpr int GH_ACH = 1024 / GX_LOADERS, GH_BCH = 512 / GX_LOADERS;   // 16-byte chunks per loader thread and step

// MN-major operand of 32-bit elements: layout type 1 = SWIZZLE_128B_BASE32B (the plain 128B swizzle yields zeros for
// an MN-major tf32 operand).  Atom = 128-byte rows (32 fp32 along MN) x 4 k; inside an atom the 32-byte unit index is
// XORed with the row (byte-address bits [7,9) into [5,7)).  `lbo` = bytes between the 32-element MN blocks, `sbo` =
// bytes between 4-k groups.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128_32b(uint32_t smem_addr, uint32_t lbo, uint32_t sbo)
{
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(lbo >> 4) << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;
    return d;
}
// byte offset of (k row, 16-byte chunk) inside a [rows][128 B] panel with that swizzle
__device__ __forceinline__ uint32_t sw128_32b_offset(uint32_t row, uint32_t chunk)
{
    return row * 128u + ((((chunk >> 1) ^ (row & 3u)) << 5) | ((chunk & 1u) << 4));
}

__global__ void __launch_bounds__(GH_THREADS, 1)
k_gemm_h_tf32x3(const float* __restrict__ C, const float* __restrict__ gt, float* __restrict__ h, int D, int64_t HW,
                int ntiles, uint32_t lbo, uint32_t sbo)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + GH_BAR_OFF);
    uint64_t* full = bars;
    uint64_t* empty = bars + GH_STAGES;
    uint64_t* acc_full = bars + 2 * GH_STAGES;
    uint64_t* acc_free = acc_full + 2;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(acc_free + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int dsteps = D / GX_KSTEP;                                     // 16 for D = 512 (even, see the launcher)
    int my_tiles = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) ++my_tiles;
    const int nsteps = my_tiles * dsteps;

    if (tid == 0) {
        for (int s = 0; s < GH_STAGES; ++s) { mbar_init(full + s, GX_LOADERS); mbar_init(empty + s, 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(acc_full + b, 1); mbar_init(acc_free + b, 128); }
        mbar_fence_init();
    }
    if (warp == GX_LW) tmem_alloc(s_tmem, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp < GX_LW) {
        auto issue = [&](int s) {
            unsigned char* st = smem + (s % GH_STAGES) * GH_STAGE;
            const int64_t p0 = ((int64_t)blockIdx.x + (int64_t)(s / dsteps) * gridDim.x) * 128;
            const int d0 = (s % dsteps) * GX_KSTEP;
#pragma unroll
            for (int i = 0; i < GH_ACH; ++i) {                           // A: row d = c >> 5, 16-byte chunk c & 31
                const int c = tid + GX_LOADERS * i;
                const int row = c >> 5, ch = c & 31;
                const int64_t p = p0 + ch * 4;
                const bool ok = p < HW;
                const float* src = gt + (ok ? (size_t)(d0 + row) * HW + p : 0);
                unsigned char* dst = st + (ch >> 3) * 4096 + sw128_32b_offset(row, ch & 7);
                cp_async16(smem_u32(dst), src, ok ? 16u : 0u);
            }
#pragma unroll
            for (int i = 0; i < GH_BCH; ++i) {                           // B: row k = c >> 3, chunk c & 7
                const int c = tid + GX_LOADERS * i;
                const int row = c >> 3, ch = c & 7;
                cp_async16(smem_u32(st + GH_A_BYTES + sw128_offset(row, ch)), C + (size_t)row * D + d0 + ch * 4, 16u);
            }
            cp_async_commit();
        };
#pragma unroll
        for (int p = 0; p < GH_PF; ++p)
            if (nsteps > p) issue(p);
        for (int s = 0; s < nsteps; ++s) {
            // copies of step s have landed when at most min(GH_PF - 1, steps still to come) groups are outstanding
            const int later = nsteps - 1 - s;
            if (later >= 3) cp_async_wait<3>();
            else if (later == 2) cp_async_wait<2>();
            else if (later == 1) cp_async_wait<1>();
            else cp_async_wait<0>();
            if (s >= 2) mbar_wait(empty + (s - 2) % GH_STAGES, ((s - 2) / GH_STAGES) & 1);
            unsigned char* st = smem + (s % GH_STAGES) * GH_STAGE;
            unsigned char* lo_st = smem + GH_LO_OFF + (s & 1) * GH_STAGE;
#pragma unroll
            for (int i = 0; i < GH_ACH + GH_BCH; ++i) {
                const int c = tid + GX_LOADERS * (i < GH_ACH ? i : i - GH_ACH);
                const uint32_t off = i < GH_ACH ? (uint32_t)((c & 31) >> 3) * 4096u + sw128_32b_offset(c >> 5, c & 7)
                                           : (uint32_t)GH_A_BYTES + sw128_offset(c >> 3, c & 7);
                const float4 v = *reinterpret_cast<const float4*>(st + off);
                float4 hh, ll;
                hh.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); ll.x = v.x - hh.x;
                hh.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); ll.y = v.y - hh.y;
                hh.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); ll.z = v.z - hh.z;
                hh.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); ll.w = v.w - hh.w;
                // `hh` is not written back: kind::tf32 ignores the 13 low mantissa bits of its fp32 operands (truncation),
                // so the raw tile IS the head; only the remainder needs a tile of its own (one STS.128 less per chunk;
                // tests/test_train_gpu.py::test_codebook_gradient_tensor_core_3xtf32_matches_fp64 would see a rounding
                // tensor core as a 100x larger error)
                *reinterpret_cast<float4*>(lo_st + off) = ll;
            }
            fence_async_smem();
            mbar_arrive(full + s % GH_STAGES);
            if (s + GH_PF < nsteps) {
                // the stage of step s + GH_PF was last read by the MMAs of step s + GH_PF - GH_STAGES = s - 1
                if (s >= 1) mbar_wait(empty + (s - 1) % GH_STAGES, ((s - 1) / GH_STAGES) & 1);
                issue(s + GH_PF);
            }
        }
    } else if (warp > GX_LW) {
        const int q = warp & 3;
        const int nchunks = nsteps / GX_CHUNK;                           // dsteps is even: chunks never straddle tiles
        const int chunks_per_tile = dsteps / GX_CHUNK;
        float acc[64];
#pragma unroll
        for (int k = 0; k < 64; ++k) acc[k] = 0.0f;
        for (int c = 0; c < nchunks; ++c) {
            const int bsel = c & 1;
            mbar_wait(acc_full + bsel, (c >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t r[32];
                tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + bsel * 64 + half * 32, r);
                tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < 32; ++k) acc[half * 32 + k] += __uint_as_float(r[k]);
            }
            tc_fence_before();
            mbar_arrive(acc_free + bsel);
            if ((c + 1) % chunks_per_tile == 0) {
                const int64_t p = ((int64_t)blockIdx.x + (int64_t)(c / chunks_per_tile) * gridDim.x) * 128 + q * 32 + lane;
                if (p < HW) {
#pragma unroll
                    for (int k = 0; k < 64; ++k) h[(size_t)k * HW + p] = acc[k];
                }
#pragma unroll
                for (int k = 0; k < 64; ++k) acc[k] = 0.0f;
            }
        }
    } else if (lane == 0) {
        // A: MN-major (bit 15 of the instruction descriptor), B: K-major
        const uint32_t idesc = umma_idesc_tf32_f32(128, 64) | (1u << 15);
        for (int s = 0; s < nsteps; ++s) {
            const int c = s / GX_CHUNK, bsel = c & 1;
            const bool first = s % GX_CHUNK == 0;
            if (first && c >= 2) {
                mbar_wait(acc_free + bsel, ((c >> 1) - 1) & 1);
                tc_fence_after();
            }
            mbar_wait(full + s % GH_STAGES, (s / GH_STAGES) & 1);
            tc_fence_after();
            const uint32_t st = smem_u32(smem + (s % GH_STAGES) * GH_STAGE);
            const uint32_t lo = smem_u32(smem + GH_LO_OFF + (s & 1) * GH_STAGE);
#pragma unroll
            for (int k8 = 0; k8 < GX_KSTEP / 8; ++k8) {
                const uint64_t a_hi = umma_desc_mn_sw128_32b(st + k8 * 1024, lbo, sbo);
                const uint64_t a_lo = umma_desc_mn_sw128_32b(lo + k8 * 1024, lbo, sbo);
                const uint64_t koff = (uint64_t)((k8 * 32) >> 4);
                const uint64_t b_hi = umma_desc_k_sw128(st + GH_A_BYTES) + koff;
                const uint64_t b_lo = umma_desc_k_sw128(lo + GH_A_BYTES) + koff;
                const uint32_t d_t = tmem + bsel * 64;
                umma_tf32(d_t, a_lo, b_hi, idesc, (!first || k8 > 0) ? 1u : 0u);
                umma_tf32(d_t, a_hi, b_lo, idesc, 1u);
                umma_tf32(d_t, a_hi, b_hi, idesc, 1u);
            }
            umma_commit(empty + s % GH_STAGES);
            if (s % GX_CHUNK == GX_CHUNK - 1) umma_commit(acc_full + bsel);
        }
    }
    __syncthreads();
    if (warp == GX_LW) {
        tc_fence_after();
        tmem_dealloc(tmem, 128);
    }
}

bool launch_gemm_h_tf32x3(const float* C, const float* gt, float* h, int D, int Lc, int64_t HW, cudaStream_t st, int* rc)
{
    static const bool off = [] { const char* e = getenv("LSV2_DC_SIMT"); return e && e[0] == '1'; }();
    static const int dbg = [] { const char* e = getenv("LSV2_GH_DESC"); return e ? atoi(e) : 0; }();
    *rc = LSV2_OK;
    if (off || Lc != 64 || D < 64 || D % 64 != 0 || HW < 128 || (HW & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(C) & 15) != 0 || (reinterpret_cast<uintptr_t>(gt) & 15) != 0)
        return false;
    const int ntiles = (int)ceil_div(HW, (int64_t)128);
    const int grid = min(ntiles, NUM_SMS);
    // panels of 32 pixels x 32 d (4 KB) side by side; 4-d groups 512 B apart inside a panel
    uint32_t lbo = 4096, sbo = 512;
    if (dbg == 1) { lbo = 512; sbo = 4096; }
    cudaError_t e = cudaFuncSetAttribute(k_gemm_h_tf32x3, cudaFuncAttributeMaxDynamicSharedMemorySize, GH_TOTAL);
    if (e == cudaSuccess) k_gemm_h_tf32x3<<<grid, GH_THREADS, GH_TOTAL, st>>>(C, gt, h, D, HW, ntiles, lbo, sbo);
    if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) *rc = LSV2_ECUDA;
    return true;
}

// true when the tensor-core path took the product; false -> the caller runs its CUDA-core kernel
bool launch_gemm_nt_tf32x3(const float* X, const float* Y, float* out, int D, int Lc, int64_t HW, cudaStream_t st,
                           int* rc)
{
    static const bool off = [] { const char* e = getenv("LSV2_DC_SIMT"); return e && e[0] == '1'; }();
    *rc = LSV2_OK;
    if (off || Lc != 64 || D <= 0 || HW < GX_KSTEP || (HW & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(X) & 15) != 0 || (reinterpret_cast<uintptr_t>(Y) & 15) != 0)
        return false;
    const int mt = D > 128 ? 2 : 1;
    const int ny = ceil_div(D, mt * 128);
    int nslab = max(1, NUM_SMS / ny);
    int64_t slab = ceil_div(HW, (int64_t)nslab);
    slab = (slab + GX_KSTEP - 1) / GX_KSTEP * GX_KSTEP;
    nslab = (int)ceil_div(HW, slab);
    dim3 grid((unsigned)nslab, (unsigned)ny);
    cudaError_t e;
    if (mt == 2) {
        e = cudaFuncSetAttribute(k_gemm_nt_tf32x3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GxLayout<2>::TOTAL);
        if (e == cudaSuccess) k_gemm_nt_tf32x3<2><<<grid, gx_threads(2), GxLayout<2>::TOTAL, st>>>(X, Y, out, D, HW, slab);
    } else {
        e = cudaFuncSetAttribute(k_gemm_nt_tf32x3<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GxLayout<1>::TOTAL);
        if (e == cudaSuccess) k_gemm_nt_tf32x3<1><<<grid, gx_threads(1), GxLayout<1>::TOTAL, st>>>(X, Y, out, D, HW, slab);
    }
    if (e != cudaSuccess) { *rc = LSV2_ECUDA; return true; }
    if (cudaGetLastError() != cudaSuccess) { *rc = LSV2_ECUDA; }
    return true;
}

}  // namespace lsv2
