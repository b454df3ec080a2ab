// tma_store_probe.cu -- stand-alone probe for the building blocks of the fused kernel's store path (not part of
// liblsv2.so).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tma_store_probe tma_store_probe.cu
//   1. 64-byte / 32-byte swizzled staging -> 3-D TMA tensor store into [d][H][W]: layout formulas checked bit for bit
//   2. the store rate of that path from 148 SMs (the ceiling of the fused kernel's epilogue)
//   3. tcgen05.mma with K-major SWIZZLE_64B operands (64-byte rows), including mixed k-slice pairing
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <math.h>

#include "../langsplatv2_b200/csrc/tc_common.cuh"
using namespace lsv2::tc;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeFn get_encode()
{
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled\n"); exit(1); }
    return (EncodeFn)fn;
}

// tensor [D][H][W]; map dims ordered (x, d, y) so a box is [y][d][x] in shared memory
static CUtensorMap make_map(void* base, int D, int H, int W, int esz, int bx, int bd, int by, CUtensorMapSwizzle sw)
{
    CUtensorMap m;
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)D, (cuuint64_t)H};
    cuuint64_t strides[2] = {(cuuint64_t)H * W * esz, (cuuint64_t)W * esz};
    cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)bd, (cuuint32_t)by};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = get_encode()(&m, esz == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, base,
                              dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
    return m;
}

// ---------------------------------------------------------------------------------------------------------------
// Store kernel: warp q of a CTA owns features [32q, 32q+32) of every 128-feature chunk; lane = feature; per step
// it writes the box {x 16 (or 32), d 32, y 2 (or 1)} = 4 KB from its own staging ring.  value = f(d, y, x).
template <int NBUF, int WIDE>       // WIDE: box {32, 32, 1} (128-byte rows, two tiles side by side), SWIZZLE_128B
__global__ void __launch_bounds__(128) k_store(const __grid_constant__ CUtensorMap tmap, int D, int H, int W,
                                               uint32_t* counter, int check)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ int s_tile;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gx = (W + 15) / 16, gy = (H + 15) / 16;
    const int units = WIDE ? ((gx + 1) / 2) * gy : gx * gy;
    unsigned char* stg = smem + warp * NBUF * 4096;
    uint32_t n = 0;
    for (;;) {
        if (threadIdx.x == 0) s_tile = (int)atomicAdd(counter, 1u);
        __syncthreads();
        const int u = s_tile;
        __syncthreads();
        if (u >= units) break;
        int x0, y0;
        if (WIDE) { const int ux = (gx + 1) / 2; x0 = (u % ux) * 32; y0 = (u / ux) * 16; }
        else { x0 = (u % gx) * 16; y0 = (u / gx) * 16; }
        for (int dc = 0; dc < D / 128; ++dc) {
            const int d = dc * 128 + warp * 32 + lane;
            const int steps = WIDE ? 16 : 8;
            for (int j = 0; j < steps; ++j, ++n) {
                unsigned char* buf = stg + (n % NBUF) * 4096;
                if (lane == 0 && n >= NBUF) bulk_wait_read<NBUF - 1>();
                __syncwarp();
                // 32 pixels of this feature
                float v[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int x = WIDE ? x0 + c : x0 + (c & 15);
                    const int y = WIDE ? y0 + j : y0 + 2 * j + (c >> 4);
                    v[c] = check ? (float)(((d * 131 + y) * 977 + x) % 1000003) : (float)(n + c);
                }
                if (WIDE) {
                    // [y 1][d 32][x 32] : row = lane, 128-byte rows, SWIZZLE_128B
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        *reinterpret_cast<float4*>(buf + sw128_offset((uint32_t)lane, (uint32_t)q)) =
                            make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                } else {
                    // [y 2][d 32][x 16] : row = y*32 + lane, 64-byte rows, SWIZZLE_64B
#pragma unroll
                    for (int y = 0; y < 2; ++y)
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            *reinterpret_cast<float4*>(buf + sw64_offset((uint32_t)(y * 32 + lane), (uint32_t)q)) =
                                make_float4(v[16 * y + 4 * q], v[16 * y + 4 * q + 1], v[16 * y + 4 * q + 2], v[16 * y + 4 * q + 3]);
                }
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (WIDE) tma_store_3d(&tmap, buf, x0, dc * 128 + warp * 32, y0 + j);
                    else tma_store_3d(&tmap, buf, x0, dc * 128 + warp * 32, y0 + 2 * j);
                    bulk_commit();
                }
            }
        }
    }
    if (lane == 0) bulk_wait_all();
}

// bf16 variant of the narrow box: [y 2][d 32][x 16] bf16, 32-byte rows, SWIZZLE_32B (check only)
__global__ void __launch_bounds__(128) k_store_bf16(const __grid_constant__ CUtensorMap tmap, int D, int H, int W)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gx = (W + 15) / 16;
    const int u = blockIdx.x;
    const int x0 = (u % gx) * 16, y0 = (u / gx) * 16;
    unsigned char* buf = smem + warp * 2048;
    for (int dc = 0; dc < D / 128; ++dc) {
        const int d = dc * 128 + warp * 32 + lane;
        for (int j = 0; j < 8; ++j) {
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
            for (int y = 0; y < 2; ++y)
                for (int q = 0; q < 2; ++q) {
                    uint32_t w[4];
                    for (int i = 0; i < 4; ++i) {
                        const int xa = x0 + q * 8 + 2 * i, yy = y0 + 2 * j + y;
                        const float a = (float)(((d * 13 + yy) * 7 + xa) % 251), b = (float)(((d * 13 + yy) * 7 + xa + 1) % 251);
                        w[i] = pack_bf16x2(a, b);
                    }
                    *reinterpret_cast<uint4*>(buf + sw32_offset((uint32_t)(y * 32 + lane), (uint32_t)q)) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) { tma_store_3d(&tmap, buf, x0, dc * 128 + warp * 32, y0 + 2 * j); bulk_commit(); }
        }
    }
    if (lane == 0) bulk_wait_all();
}

// plain coalesced stores for comparison: lane = pixel (two 64-byte runs per warp instruction), thread loops features
__global__ void __launch_bounds__(256) k_store_lsu(float* out, int D, int H, int W, uint32_t* counter)
{
    __shared__ int s_tile;
    const int gx = (W + 15) / 16, gy = (H + 15) / 16;
    const int lx = threadIdx.x & 15, ly = threadIdx.x >> 4;
    for (;;) {
        if (threadIdx.x == 0) s_tile = (int)atomicAdd(counter, 1u);
        __syncthreads();
        const int u = s_tile;
        __syncthreads();
        if (u >= gx * gy) break;
        const int x = (u % gx) * 16 + lx, y = (u / gx) * 16 + ly;
        if (x < W && y < H) {
            float* p = out + (size_t)y * W + x;
            const size_t HW = (size_t)H * W;
#pragma unroll 8
            for (int d = 0; d < D; ++d) asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p + d * HW), "f"((float)d) : "memory");
        }
    }
}


// one TMA op per [ROWS rows][FEATS features][16 px] box (FEATS*ROWS*64 bytes), issued by one thread after the four
// warps have filled their 32-feature quarters (named barrier), double-buffered.
template <int FEATS, int ROWS>
__global__ void __launch_bounds__(128) k_store_big(const __grid_constant__ CUtensorMap tmap, int D, int H, int W,
                                                   uint32_t* counter)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ int s_tile;
    constexpr int BOX = FEATS * ROWS * 64;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gx = (W + 15) / 16, gy = (H + 15) / 16;
    uint32_t n = 0;
    for (;;) {
        if (threadIdx.x == 0) s_tile = (int)atomicAdd(counter, 1u);
        __syncthreads();
        const int u = s_tile;
        __syncthreads();
        if (u >= gx * gy) break;
        const int x0 = (u % gx) * 16, y0 = (u / gx) * 16;
        for (int dc = 0; dc < D / FEATS; ++dc) {
            for (int j = 0; j < 16 / ROWS; ++j, ++n) {
                unsigned char* buf = smem + (n & 1) * BOX;
                if (threadIdx.x == 0 && n >= 2) bulk_wait_read<1>();
                __syncthreads();
                // each warp fills FEATS/4 features x ROWS rows : lane -> (feature, row) pairs, 4 STS.128 per row
                for (int t = lane; t < (FEATS / 4) * ROWS; t += 32) {
                    const int f = warp * (FEATS / 4) + t % (FEATS / 4), y = t / (FEATS / 4);
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        *reinterpret_cast<float4*>(buf + sw64_offset((uint32_t)(y * FEATS + f), (uint32_t)q)) =
                            make_float4((float)n, (float)f, (float)y, (float)q);
                }
                fence_async_smem();
                __syncthreads();
                if (threadIdx.x == 0) { tma_store_3d(&tmap, buf, x0, dc * FEATS, y0 + ROWS * j); bulk_commit(); }
            }
        }
    }
    if (threadIdx.x == 0) bulk_wait_all();
}

// linear, fully coalesced 16-byte stores over the whole buffer (what a memset does)
__global__ void __launch_bounds__(256) k_store_linear(float4* out, size_t n4)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride)
        asm volatile("st.global.cs.v4.f32 [%0], {%1, %1, %1, %1};" ::"l"(out + i), "f"(1.0f) : "memory");
}

// per-CTA contiguous chunks: CTA b writes bytes [b*chunk, (b+1)*chunk) front to back (page-local write streams)
__global__ void __launch_bounds__(256) k_store_chunked(float4* out, size_t n4)
{
    const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
    const size_t a = (size_t)blockIdx.x * per, b = a + per < n4 ? a + per : n4;
    for (size_t i = a + threadIdx.x; i < b; i += blockDim.x)
        asm volatile("st.global.cs.v4.f32 [%0], {%1, %1, %1, %1};" ::"l"(out + i), "f"(1.0f) : "memory");
}

// lane = pixel of a 32x8 tile (128-byte runs), thread loops features
__global__ void __launch_bounds__(256) k_store_lsu128(float* out, int D, int H, int W, uint32_t* counter)
{
    __shared__ int s_tile;
    const int gx = (W + 31) / 32, gy = (H + 7) / 8;
    const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
    for (;;) {
        if (threadIdx.x == 0) s_tile = (int)atomicAdd(counter, 1u);
        __syncthreads();
        const int u = s_tile;
        __syncthreads();
        if (u >= gx * gy) break;
        const int x = (u % gx) * 32 + lx, y = (u / gx) * 8 + ly;
        if (x < W && y < H) {
            float* p = out + (size_t)y * W + x;
            const size_t HW = (size_t)H * W;
#pragma unroll 8
            for (int d = 0; d < D; ++d) asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p + d * HW), "f"((float)d) : "memory");
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// MMA probe: D[128 x 64] = sum over chosen k-slice pairs of A[128][k16] . B[64][k16]^T ; A, B: 64-byte rows
// (two 16-element k-slices), SWIZZLE_64B K-major.
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N)
{
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__global__ void __launch_bounds__(128) k_mma64(const __half* A, const __half* B, float* Dout, int pairing)
{
    __shared__ __align__(1024) unsigned char sA[128 * 64];
    __shared__ __align__(1024) unsigned char sB[64 * 64];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < 128 * 32; i += 128) {
        const int r = i / 32, k = i % 32;
        *reinterpret_cast<__half*>(sA + sw64_offset(r, k >> 3) + (k & 7) * 2) = A[i];
    }
    for (int i = t; i < 64 * 32; i += 128) {
        const int r = i / 32, k = i % 32;
        *reinterpret_cast<__half*>(sB + sw64_offset(r, k >> 3) + (k & 7) * 2) = B[i];
    }
    if (t == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&s_tmem, 64);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    if (t == 0) {
        const uint32_t a = smem_u32(sA), b = smem_u32(sB);
        const uint32_t id = idesc_f16(128, 64);
        if (pairing == 0) {          // plain K = 32
            umma_bf16(tmem, umma_desc_k_sw64(a), umma_desc_k_sw64(b), id, 0);
            umma_bf16(tmem, umma_desc_k_sw64(a + 32), umma_desc_k_sw64(b + 32), id, 1);
        } else {                     // A0.B0 + A1.B0 + A0.B1 (the hi/lo scheme of the blend)
            umma_bf16(tmem, umma_desc_k_sw64(a), umma_desc_k_sw64(b), id, 0);
            umma_bf16(tmem, umma_desc_k_sw64(a + 32), umma_desc_k_sw64(b), id, 1);
            umma_bf16(tmem, umma_desc_k_sw64(a), umma_desc_k_sw64(b + 32), id, 1);
        }
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    uint32_t r[32];
    for (int c = 0; c < 2; ++c) {
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, r);
        tmem_wait_ld();
        for (int i = 0; i < 32; ++i) Dout[(size_t)t * 64 + c * 32 + i] = __uint_as_float(r[i]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 64);
}

int main()
{
    // ---- 1. layout check, fp32 narrow + wide, odd-ish size with clipped edge tiles ----
    {
        const int D = 256, H = 40, W = 72;     // H not a multiple of 16: last tile row is clipped by the TMA
        float* out;
        uint32_t* ctr;
        CK(cudaMalloc(&out, (size_t)D * H * W * 4));
        CK(cudaMalloc(&ctr, 4));
        for (int wide = 0; wide < 2; ++wide) {
            CK(cudaMemset(out, 0xFF, (size_t)D * H * W * 4));
            CK(cudaMemset(ctr, 0, 4));
            CUtensorMap m = wide ? make_map(out, D, H, W, 4, 32, 32, 1, CU_TENSOR_MAP_SWIZZLE_128B)
                                 : make_map(out, D, H, W, 4, 16, 32, 2, CU_TENSOR_MAP_SWIZZLE_64B);
            if (wide) { CK(cudaFuncSetAttribute(k_store<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000)); k_store<2, 1><<<8, 128, 4 * 2 * 4096 + 1024>>>(m, D, H, W, ctr, 1); }
            else { CK(cudaFuncSetAttribute(k_store<2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000)); k_store<2, 0><<<8, 128, 4 * 2 * 4096 + 1024>>>(m, D, H, W, ctr, 1); }
            CK(cudaDeviceSynchronize());
            std::vector<float> h((size_t)D * H * W);
            CK(cudaMemcpy(h.data(), out, h.size() * 4, cudaMemcpyDeviceToHost));
            size_t bad = 0;
            for (int d = 0; d < D; ++d) for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
                const float e = (float)(((d * 131 + y) * 977 + x) % 1000003);
                if (h[((size_t)d * H + y) * W + x] != e) { if (bad < 5) printf("  mismatch d=%d y=%d x=%d got %f want %f\n", d, y, x, h[((size_t)d * H + y) * W + x], e); ++bad; }
            }
            printf("LAYOUT fp32 %s: %zu mismatches of %zu\n", wide ? "wide{32,32,1}/SW128" : "narrow{16,32,2}/SW64", bad, h.size());
        }
        // bf16 narrow
        {
            __nv_bfloat16* ob;
            const int Wb = 64;
            CK(cudaMalloc(&ob, (size_t)D * H * Wb * 2));
            CK(cudaMemset(ob, 0xFF, (size_t)D * H * Wb * 2));
            CUtensorMap m = make_map(ob, D, H, Wb, 2, 16, 32, 2, CU_TENSOR_MAP_SWIZZLE_32B);
            const int gx = Wb / 16, gy = (H + 15) / 16;
            k_store_bf16<<<gx * gy, 128, 4 * 2048 + 1024>>>(m, D, H, Wb);
            CK(cudaDeviceSynchronize());
            std::vector<__nv_bfloat16> h((size_t)D * H * Wb);
            CK(cudaMemcpy(h.data(), ob, h.size() * 2, cudaMemcpyDeviceToHost));
            size_t bad = 0;
            for (int d = 0; d < D; ++d) for (int y = 0; y < H; ++y) for (int x = 0; x < Wb; ++x) {
                const float e = (float)(((d * 13 + y) * 7 + x) % 251);
                if (__bfloat162float(h[((size_t)d * H + y) * Wb + x]) != e) ++bad;
            }
            printf("LAYOUT bf16 narrow{16,32,2}/SW32: %zu mismatches of %zu\n", bad, h.size());
            cudaFree(ob);
        }
        cudaFree(out);
        cudaFree(ctr);
    }
    // ---- 3. MMA probe ----
    {
        std::vector<__half> A(128 * 32), B(64 * 32);
        srand(1);
        for (auto& v : A) v = __float2half((rand() % 17 - 8) / 8.0f);
        for (auto& v : B) v = __float2half((rand() % 13 - 6) / 4.0f);
        __half *dA, *dB;
        float* dD;
        CK(cudaMalloc(&dA, A.size() * 2)); CK(cudaMalloc(&dB, B.size() * 2)); CK(cudaMalloc(&dD, 128 * 64 * 4));
        CK(cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
        for (int pairing = 0; pairing < 2; ++pairing) {
            k_mma64<<<1, 128>>>(dA, dB, dD, pairing);
            CK(cudaDeviceSynchronize());
            std::vector<float> Dh(128 * 64);
            CK(cudaMemcpy(Dh.data(), dD, Dh.size() * 4, cudaMemcpyDeviceToHost));
            double maxerr = 0;
            for (int m = 0; m < 128; ++m) for (int n = 0; n < 64; ++n) {
                double e = 0;
                auto a = [&](int k) { return (double)__half2float(A[m * 32 + k]); };
                auto b = [&](int k) { return (double)__half2float(B[n * 32 + k]); };
                if (pairing == 0) { for (int k = 0; k < 32; ++k) e += a(k) * b(k); }
                else { for (int k = 0; k < 16; ++k) e += a(k) * b(k) + a(16 + k) * b(k) + a(k) * b(16 + k); }
                maxerr = fmax(maxerr, fabs(e - Dh[m * 64 + n]));
            }
            printf("MMA SW64 pairing %d: max err %.3g\n", pairing, maxerr);
        }
    }
    // ---- 2. store rate at the metric's output size ----
    {
        const int D = 1536, H = 1080, W = 1920;
        const size_t bytes = (size_t)D * H * W * 4;
        float* out;
        uint32_t* ctr;
        CK(cudaMalloc(&out, bytes));
        CK(cudaMalloc(&ctr, 4));
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        auto run = [&](const char* name, auto launch) {
            float best = 1e9f;
            for (int it = 0; it < 4; ++it) {
                CK(cudaMemset(ctr, 0, 4));
                cudaEventRecord(e0);
                launch();
                cudaEventRecord(e1);
                CK(cudaEventSynchronize(e1));
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (it > 0) best = fminf(best, ms);
            }
            printf("RATE %-44s %.3f ms  %.0f GB/s\n", name, best, bytes / best * 1e-6);
        };
        CUtensorMap mn = make_map(out, D, H, W, 4, 16, 32, 2, CU_TENSOR_MAP_SWIZZLE_64B);
        CUtensorMap mw = make_map(out, D, H, W, 4, 32, 32, 1, CU_TENSOR_MAP_SWIZZLE_128B);
        CK(cudaFuncSetAttribute(k_store<2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
        CK(cudaFuncSetAttribute(k_store<4, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
        CK(cudaFuncSetAttribute(k_store<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
        CK(cudaFuncSetAttribute(k_store<4, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
        run("tma narrow{16,32,2} 2 buf/warp, 1 CTA/SM", [&] { k_store<2, 0><<<148, 128, 4 * 2 * 4096 + 1024>>>(mn, D, H, W, ctr, 0); });
        run("tma narrow{16,32,2} 4 buf/warp, 1 CTA/SM", [&] { k_store<4, 0><<<148, 128, 4 * 4 * 4096 + 1024>>>(mn, D, H, W, ctr, 0); });
        run("tma narrow{16,32,2} 2 buf/warp, 2 CTA/SM", [&] { k_store<2, 0><<<296, 128, 4 * 2 * 4096 + 1024>>>(mn, D, H, W, ctr, 0); });
        run("tma wide{32,32,1}   2 buf/warp, 1 CTA/SM", [&] { k_store<2, 1><<<148, 128, 4 * 2 * 4096 + 1024>>>(mw, D, H, W, ctr, 0); });
        run("tma wide{32,32,1}   4 buf/warp, 1 CTA/SM", [&] { k_store<4, 1><<<148, 128, 4 * 4 * 4096 + 1024>>>(mw, D, H, W, ctr, 0); });
        {
            CUtensorMap m128 = make_map(out, D, H, W, 4, 16, 128, 2, CU_TENSOR_MAP_SWIZZLE_64B);
            CUtensorMap m256 = make_map(out, D, H, W, 4, 16, 256, 2, CU_TENSOR_MAP_SWIZZLE_64B);
            CUtensorMap m128x4 = make_map(out, D, H, W, 4, 16, 128, 4, CU_TENSOR_MAP_SWIZZLE_64B);
            CK(cudaFuncSetAttribute(k_store_big<128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000));
            CK(cudaFuncSetAttribute(k_store_big<256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000));
            CK(cudaFuncSetAttribute(k_store_big<128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000));
            run("tma big {16,128,2} 16 KB/op, 1 CTA/SM", [&] { k_store_big<128, 2><<<148, 128, 2 * 16384 + 1024>>>(m128, D, H, W, ctr); });
            run("tma big {16,128,2} 16 KB/op, 2 CTA/SM", [&] { k_store_big<128, 2><<<296, 128, 2 * 16384 + 1024>>>(m128, D, H, W, ctr); });
            run("tma big {16,256,2} 32 KB/op, 1 CTA/SM", [&] { k_store_big<256, 2><<<148, 128, 2 * 32768 + 1024>>>(m256, D, H, W, ctr); });
            run("tma big {16,128,4} 32 KB/op, 1 CTA/SM", [&] { k_store_big<128, 4><<<148, 128, 2 * 32768 + 1024>>>(m128x4, D, H, W, ctr); });
        }
        run("linear st.v4 grid-stride, 148x8 CTAs", [&] { k_store_linear<<<148 * 8, 256>>>((float4*)out, bytes / 16); });
        run("linear st.v4 grid-stride, 148x2 CTAs", [&] { k_store_linear<<<148 * 2, 256>>>((float4*)out, bytes / 16); });
        run("chunked st.v4 per-CTA contiguous, 148x4 CTAs", [&] { k_store_chunked<<<148 * 4, 256>>>((float4*)out, bytes / 16); });
        run("cudaMemsetAsync", [&] { cudaMemsetAsync(out, 0, bytes); });
        run("lsu st.global.cs 128-B runs (32x8 tiles), 1 CTA/SM", [&] { k_store_lsu128<<<148, 256>>>(out, D, H, W, ctr); });
        run("lsu st.global.cs 128-B runs (32x8 tiles), 2 CTA/SM", [&] { k_store_lsu128<<<296, 256>>>(out, D, H, W, ctr); });
        run("lsu st.global.cs 64-B runs, 256 thr, 1 CTA/SM", [&] { k_store_lsu<<<148, 256>>>(out, D, H, W, ctr); });
        run("lsu st.global.cs 64-B runs, 256 thr, 4 CTA/SM", [&] { k_store_lsu<<<592, 256>>>(out, D, H, W, ctr); });
    }
    return 0;
}
