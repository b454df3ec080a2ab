// decode_tc.cu -- K7: global codebook decode + L2 normalisation on the 5th-gen tensor cores.
// Replaces the einsum + norm + divide that every caller of the reference runs after the
// rasterizer (/root/reference/eval_lerf.py:214-218, scene/gaussian_model.py:533-549;
// SURVEY.md row R9):  F_l[512, HW] = C_l^T[512 x 64] . W_l[64 x HW],  F_l /= (||F_l||_2 + 1e-10).
//
// Formulation: D[pixel, feature] = A[pixel, k] . B[k, feature] with M = 128 pixels (one TMEM
// lane per pixel), N = 512 features (the whole TMEM: 512 fp32 columns), K = 64 codes.
//   * lane == pixel makes the L2 norm over the 512 features a thread-local sum over TMEM
//     columns, and makes every store of the channel-major [512][HW] output a fully coalesced
//     128-byte line per warp (32 consecutive pixels of one feature).
//   * B (one level's codebook, bf16, 64 KB, K-major, 128B-swizzled, pre-arranged by a tiny
//     prep kernel) is fetched ONCE per CTA with a bulk async copy and stays in shared memory;
//     CTA c serves level c % levels.
//   * warp roles: 4 producer warps read the fp32 weight-map tile (coalesced along pixels),
//     round to bf16 and write the swizzled A tile (4-stage ring, mbarrier hand-off);
//     one thread issues tcgen05.mma (2 halves x 4 K-steps, M128 N256 K16, accumulators in
//     TMEM); 4 epilogue warps read TMEM twice (sum of squares, then scale + store).
// The output write (levels*512*HW*4 B) is the roofline; everything else hides under it.
#include "common.cuh"
#include "tc_common.cuh"

namespace lsv2 {

using namespace tc;

constexpr int DT_LC = 64;          // codes per level  (K)
constexpr int DT_D = 512;          // feature dim      (N)
constexpr int DT_PIX = 128;        // pixels per unit  (M)
constexpr int DT_STAGES = 4;
constexpr int DT_A_BYTES = DT_PIX * 128;            // 16 KB: 128 rows x 64 bf16
constexpr int DT_B_BYTES = DT_D * 128;              // 64 KB: 512 rows x 64 bf16
constexpr int DT_THREADS = 288;                     // 4 epilogue + 4 producer + 1 MMA warp
constexpr size_t DT_SMEM = 1024 + DT_B_BYTES + DT_STAGES * DT_A_BYTES + 256;

size_t decode_tc_scratch_bytes(int levels, int Lc, int D)
{
    return (size_t)(levels > 0 ? levels : 1) * DT_B_BYTES + 256;
}

// codebooks f32 [levels][64][512] -> bf16 B image [levels][512 rows n][64 k], 128B-swizzled
__global__ void k_decode_prep_b(int levels, const float* __restrict__ codebooks, uint8_t* __restrict__ bimg)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;      // (level, n, chunk)
    if (t >= levels * DT_D * 8) return;
    const int chunk = t & 7, n = (t >> 3) % DT_D, l = t / (DT_D * 8);
    const float* src = codebooks + ((size_t)l * DT_LC + chunk * 8) * DT_D + n;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = pack_bf16x2(src[(size_t)(2 * i) * DT_D], src[(size_t)(2 * i + 1) * DT_D]);
    *reinterpret_cast<uint4*>(bimg + (size_t)l * DT_B_BYTES + sw128_offset(n, chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
}

template <bool OUT_BF16>
__global__ void __launch_bounds__(DT_THREADS, 1)
k_decode_tc(int levels, int64_t HW, const float* __restrict__ wmap, const uint8_t* __restrict__ bimg, int normalize,
            void* __restrict__ out_)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* s_b = smem;
    unsigned char* s_a = smem + DT_B_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_a + DT_STAGES * DT_A_BYTES);
    uint64_t* a_full = bars;                  // [stages]  128 producer arrivals
    uint64_t* a_empty = bars + DT_STAGES;     // [stages]  tcgen05.commit
    uint64_t* d_full = bars + 2 * DT_STAGES;  // tcgen05.commit
    uint64_t* d_empty = d_full + 1;           // 128 epilogue arrivals
    uint64_t* b_full = d_full + 2;            // bulk copy tx
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(d_full + 3);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int level = blockIdx.x % levels;
    const int rank = blockIdx.x / levels;
    const int per_level = (gridDim.x - level + levels - 1) / levels;    // CTAs serving this level
    const int64_t n_units = (HW + DT_PIX - 1) / DT_PIX;

    if (tid == 0) {
        for (int s = 0; s < DT_STAGES; ++s) { mbar_init(a_full + s, 128); mbar_init(a_empty + s, 1); }
        mbar_init(d_full, 1);
        mbar_init(d_empty, 128);
        mbar_init(b_full, 1);
        mbar_fence_init();
    }
    if (warp == 8) tmem_alloc(s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp >= 4 && warp < 8) {
        // ===== A producers: thread = pixel row of the tile =====
        const int row = tid - 128;
        const float* W = wmap + (size_t)level * DT_LC * HW;
        int it = 0;
        for (int64_t u = rank; u < n_units; u += per_level, ++it) {
            const int s = it % DT_STAGES, use = it / DT_STAGES;
            if (use > 0) mbar_wait(a_empty + s, (use - 1) & 1);
            const int64_t p = u * DT_PIX + row;
            unsigned char* dst = s_a + s * DT_A_BYTES;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = p < HW ? __ldg(W + (size_t)(c * 8 + i) * HW + p) : 0.0f;
                const uint4 q = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                           pack_bf16x2(v[6], v[7]));
                *reinterpret_cast<uint4*>(dst + sw128_offset(row, c)) = q;
            }
            fence_async_smem();
            mbar_arrive(a_full + s);
        }
    } else if (warp == 8) {
        // ===== B fetch + MMA issue (one thread) =====
        if (lane == 0) {
            mbar_arrive_expect_tx(b_full, DT_B_BYTES);
            for (int q = 0; q < 4; ++q)
                bulk_g2s(s_b + q * (DT_B_BYTES / 4), bimg + (size_t)level * DT_B_BYTES + q * (DT_B_BYTES / 4),
                         DT_B_BYTES / 4, b_full);
            mbar_wait(b_full, 0);
            constexpr uint32_t idesc = umma_idesc_bf16_f32(128, 256);
            const uint32_t b_addr = smem_u32(s_b), a_addr0 = smem_u32(s_a);
            int it = 0;
            for (int64_t u = rank; u < n_units; u += per_level, ++it) {
                const int s = it % DT_STAGES, use = it / DT_STAGES;
                mbar_wait(a_full + s, use & 1);
                if (it > 0) mbar_wait(d_empty, (it - 1) & 1);
                tc_fence_after();
                const uint32_t a_addr = a_addr0 + s * DT_A_BYTES;
#pragma unroll
                for (int nh = 0; nh < 2; ++nh)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_bf16(tmem + nh * 256, umma_desc_k_sw128(a_addr + ks * 32),
                                  umma_desc_k_sw128(b_addr + nh * 256 * 128 + ks * 32), idesc, ks > 0 ? 1u : 0u);
                umma_commit(a_empty + s);
                umma_commit(d_full);
            }
        }
    } else {
        // ===== epilogue: thread = TMEM lane = pixel =====
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        int it = 0;
        for (int64_t u = rank; u < n_units; u += per_level, ++it) {
            mbar_wait(d_full, it & 1);
            tc_fence_after();
            const int64_t p = u * DT_PIX + tid;
            float inv = 1.0f;
            if (normalize) {
                float ss = 0.0f;
#pragma unroll 1
                for (int c = 0; c < DT_D / 32; ++c) {
                    uint32_t r[32];
                    tmem_ld32(lane_base + c * 32, r);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) ss = fmaf(__uint_as_float(r[i]), __uint_as_float(r[i]), ss);
                }
                inv = 1.0f / (sqrtf(ss) + 1e-10f);
            }
#pragma unroll 1
            for (int c = 0; c < DT_D / 32; ++c) {
                uint32_t r[32];
                tmem_ld32(lane_base + c * 32, r);
                tmem_wait_ld();
                if (p < HW) {
                    if (OUT_BF16) {
                        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out_) + ((size_t)level * DT_D + c * 32) * HW + p;
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[(size_t)i * HW] = __float2bfloat16_rn(__uint_as_float(r[i]) * inv);
                    } else {
                        float* o = reinterpret_cast<float*>(out_) + ((size_t)level * DT_D + c * 32) * HW + p;
#pragma unroll
                        for (int i = 0; i < 32; ++i) st_stream_f32(o + (size_t)i * HW, __uint_as_float(r[i]) * inv);
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(d_empty);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, 512);
}

int launch_decode_tc(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks, int normalize,
                     int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    if (levels <= 0 || HW <= 0) return LSV2_OK;
    if (Lc != DT_LC || D != DT_D) return LSV2_EUNSUPPORTED;       // callers fall back to LSV2_DECODE_FP32
    if (!scratch || scratch_bytes < decode_tc_scratch_bytes(levels, Lc, D)) return LSV2_ENOSPC;
    uint8_t* bimg = reinterpret_cast<uint8_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    k_decode_prep_b<<<ceil_div(levels * DT_D * 8, 256), 256, 0, st>>>(levels, codebooks, bimg);
    LSV2_LAUNCH_CHECK("k_decode_prep_b");
    const int64_t n_units = (HW + DT_PIX - 1) / DT_PIX;
    int grid = (int)min((int64_t)NUM_SMS, n_units * levels);
    if (grid < levels) grid = levels;
    if (out_dtype == 0) {
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_decode_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DT_SMEM));
        k_decode_tc<false><<<grid, DT_THREADS, DT_SMEM, st>>>(levels, HW, wmap, bimg, normalize, out);
    } else {
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_decode_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DT_SMEM));
        k_decode_tc<true><<<grid, DT_THREADS, DT_SMEM, st>>>(levels, HW, wmap, bimg, normalize, out);
    }
    LSV2_LAUNCH_CHECK("k_decode_tc");
    return LSV2_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// lsv2_decode_static: the tensor-core decode for callers that may not allocate (CUDA-graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The only scratch the decode needs -- the bf16, 128B-swizzled image of the codebooks, 64 KB per level -- lives in a
// file-scope device array of LSV2_DECODE_STATIC_MAX_LEVELS levels (192 KB, held by every GPU that loads the library
// until the process exits).  k_decode_prep_b rewrites every byte the decode kernel reads on every call, so nothing
// has to start at zero; the array keeps the last call's codebook image in between.  All calls on a device share the
// array: two calls that overlap on the GPU (different streams, no ordering between them) race on it.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) uint8_t g_decode_static_scratch[LSV2_DECODE_STATIC_MAX_LEVELS * DT_B_BYTES + 256];

constexpr int DS_MAX_DEVICES = 64;
static void* g_decode_static_ptr[DS_MAX_DEVICES];      // device address of the array, per device; 0 = not looked up yet

// In host code `g_decode_static_scratch` is a host-side stand-in, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, looked up once per device with that device current.
static int decode_static_address(void** out)
{
    int dev = -1;
    LSV2_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= DS_MAX_DEVICES) return LSV2_EUNSUPPORTED;
    void* p = __atomic_load_n(&g_decode_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        LSV2_CUDA_TRY(cudaGetSymbolAddress(&p, g_decode_static_scratch));
        if (!p) return LSV2_ECUDA;
        __atomic_store_n(&g_decode_static_ptr[dev], p, __ATOMIC_RELEASE);     // idempotent: every thread finds the same address
    }
    *out = p;
    return LSV2_OK;
}

int launch_decode_tc_static(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                            int normalize, int out_dtype, void* out, cudaStream_t st)
{
    if (levels <= 0 || HW <= 0) return LSV2_OK;
    if (Lc != DT_LC || D != DT_D) return LSV2_EUNSUPPORTED;
    if (levels > LSV2_DECODE_STATIC_MAX_LEVELS) return LSV2_ENOSPC;           // beyond the built-in scratch
    void* scratch = nullptr;
    const int rc = decode_static_address(&scratch);
    if (rc != LSV2_OK) return rc;
    return launch_decode_tc(levels, Lc, D, HW, wmap, codebooks, normalize, out_dtype, out, scratch,
                            sizeof(g_decode_static_scratch), st);
}

}  // namespace lsv2
