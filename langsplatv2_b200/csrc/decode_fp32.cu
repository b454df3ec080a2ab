// decode_fp32.cu -- K7 (exact variant): global codebook decode + L2 normalisation on CUDA cores
// with fp32 FMA accumulation.  Replaces the Python helpers every caller of the reference runs
// after the rasterizer (/root/reference/eval_lerf.py:214-218, scene/gaussian_model.py:533-549,
// train.py:38-58):
//     F_l[D, HW] = C_l^T[D x Lc] . W_l[Lc x HW];   F_l /= (||F_l||_2 over D + 1e-10)
// The tensor-core variant lives in decode_tc.cu; this one is the fp32-exact option and the
// backward GEMMs of the decode.
#include "common.cuh"
#include <cuda_fp16.h>

namespace lsv2 {

constexpr int DF_THREADS = 256;
constexpr int DF_PIX = 128;    // pixels per CTA
constexpr int DF_FEAT = 128;   // features per chunk
constexpr int DF_KMAX = 32;    // codebook entries staged per sweep

// grid: (ceil(HW/128), levels).  Thread tile 8 px x 8 features.
__global__ void __launch_bounds__(DF_THREADS)
k_decode_fp32(int Lc, int D, int64_t HW, const float* __restrict__ wmap, const float* __restrict__ codebooks,
              int normalize, float* __restrict__ out)
{
    __shared__ __align__(16) float s_w[DF_KMAX][DF_PIX];
    __shared__ __align__(16) float s_c[DF_KMAX][DF_FEAT];
    __shared__ float s_ss[DF_PIX];

    const int tid = threadIdx.x;
    const int level = blockIdx.y;
    const int64_t p0 = (int64_t)blockIdx.x * DF_PIX;
    const int fg = tid & 15, pg = tid >> 4;           // feature group / pixel group (8 each)
    const float* W = wmap + (size_t)level * Lc * HW;
    const float* C = codebooks + (size_t)level * Lc * D;
    float* O = out + (size_t)level * D * HW;

    if (tid < DF_PIX) s_ss[tid] = 0.0f;
    const int passes = normalize ? 2 : 1;
    float inv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) inv[i] = 1.0f;

    for (int pass = 0; pass < passes; ++pass) {
        const bool write = (pass == passes - 1);
        float ss[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) ss[i] = 0.0f;
        for (int f0 = 0; f0 < D; f0 += DF_FEAT) {
            float acc[8][8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
            for (int k0 = 0; k0 < Lc; k0 += DF_KMAX) {
                const int kn = min(DF_KMAX, Lc - k0);
                __syncthreads();
                for (int e = tid; e < kn * DF_PIX; e += DF_THREADS) {
                    const int k = e / DF_PIX, p = e % DF_PIX;
                    s_w[k][p] = (p0 + p < HW) ? W[(size_t)(k0 + k) * HW + p0 + p] : 0.0f;
                }
                for (int e = tid; e < kn * DF_FEAT; e += DF_THREADS) {
                    const int k = e / DF_FEAT, f = e % DF_FEAT;
                    s_c[k][f] = (f0 + f < D) ? C[(size_t)(k0 + k) * D + f0 + f] : 0.0f;
                }
                __syncthreads();
                for (int k = 0; k < kn; ++k) {
                    const float4 wa = *reinterpret_cast<const float4*>(&s_w[k][pg * 8]);
                    const float4 wb = *reinterpret_cast<const float4*>(&s_w[k][pg * 8 + 4]);
                    const float4 ca = *reinterpret_cast<const float4*>(&s_c[k][fg * 8]);
                    const float4 cb = *reinterpret_cast<const float4*>(&s_c[k][fg * 8 + 4]);
                    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
                    const float cv[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
                    for (int i = 0; i < 8; ++i)
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(wv[i], cv[j], acc[i][j]);
                }
            }
            if (!write) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) ss[i] = fmaf(acc[i][j], acc[i][j], ss[i]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int f = f0 + fg * 8 + j;
                    if (f >= D) continue;
                    float* dst = O + (size_t)f * HW + p0 + pg * 8;
                    if (p0 + pg * 8 + 8 <= HW && (HW & 3) == 0) {
                        st_stream_f32x4(dst, acc[0][j] * inv[0], acc[1][j] * inv[1], acc[2][j] * inv[2], acc[3][j] * inv[3]);
                        st_stream_f32x4(dst + 4, acc[4][j] * inv[4], acc[5][j] * inv[5], acc[6][j] * inv[6], acc[7][j] * inv[7]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (p0 + pg * 8 + i < HW) dst[i] = acc[i][j] * inv[i];
                    }
                }
            }
        }
        if (!write) {
#pragma unroll
            for (int i = 0; i < 8; ++i) atomicAdd(&s_ss[pg * 8 + i], ss[i]);
            __syncthreads();
#pragma unroll
            for (int i = 0; i < 8; ++i) inv[i] = 1.0f / (sqrtf(s_ss[pg * 8 + i]) + 1e-10f);
        }
    }
}

int launch_decode_fp32(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                       int normalize, float* out, cudaStream_t st)
{
    if (levels <= 0 || HW <= 0) return LSV2_OK;
    dim3 grid((unsigned)ceil_div(HW, DF_PIX), (unsigned)levels);
    k_decode_fp32<<<grid, DF_THREADS, 0, st>>>(Lc, D, HW, wmap, codebooks, normalize, out);
    LSV2_LAUNCH_CHECK("k_decode_fp32");
    return LSV2_OK;
}

// ---- decode backward (fp32) -------------------------------------------------------------
// dW[l*Lc+k][p] = sum_d C[l][k][d] dF[l][d][p]
// grid: (ceil(HW/128), levels).  CTA tile 128 pixels x 64 codes, thread tile 4 pixels x 8 codes (a warp shares its
// 8 codes, so the codebook reads broadcast and the dF reads are one conflict-free float4 per thread); dF and the
// codebook are staged 32 features at a time.
__global__ void __launch_bounds__(256)
k_decode_bwd_dW(int Lc, int D, int64_t HW, const float* __restrict__ codebooks, const float* __restrict__ dF,
                float* __restrict__ dW)
{
    __shared__ __align__(16) float s_f[32][128];     // [feature][pixel]
    __shared__ __align__(16) float s_c[32][68];      // [feature][code] (+4 pad)
    const int tid = threadIdx.x;
    const int level = blockIdx.y;
    const int pg = tid & 31, kg = tid >> 5;
    const int64_t pix0 = (int64_t)blockIdx.x * 128;
    const float* C = codebooks + (size_t)level * Lc * D;
    const float* F = dF + (size_t)level * D * HW;
    const bool vec = (HW & 3) == 0 && pix0 + 128 <= HW;
    for (int k0 = 0; k0 < Lc; k0 += 64) {
        float acc[4][8];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
        for (int d0 = 0; d0 < D; d0 += 32) {
            __syncthreads();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = tid + i * 256, d = e >> 5, p4 = e & 31;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (d0 + d < D) {
                    const float* src = F + (size_t)(d0 + d) * HW + pix0 + p4 * 4;
                    if (vec) {
                        v = __ldg(reinterpret_cast<const float4*>(src));
                    } else {
                        const int64_t q = pix0 + p4 * 4;
                        if (q < HW) v.x = src[0];
                        if (q + 1 < HW) v.y = src[1];
                        if (q + 2 < HW) v.z = src[2];
                        if (q + 3 < HW) v.w = src[3];
                    }
                }
                *reinterpret_cast<float4*>(&s_f[d][p4 * 4]) = v;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int e = tid + i * 256, k = e >> 5, d = e & 31;
                s_c[d][k] = (k0 + k < Lc && d0 + d < D) ? __ldg(C + (size_t)(k0 + k) * D + d0 + d) : 0.0f;
            }
            __syncthreads();
#pragma unroll 8
            for (int d = 0; d < 32; ++d) {
                const float4 f4 = *reinterpret_cast<const float4*>(&s_f[d][pg * 4]);
                const float4 ca = *reinterpret_cast<const float4*>(&s_c[d][kg * 8]);
                const float4 cb = *reinterpret_cast<const float4*>(&s_c[d][kg * 8 + 4]);
                const float fv[4] = {f4.x, f4.y, f4.z, f4.w};
                const float cv[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(fv[i], cv[j], acc[i][j]);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int k = k0 + kg * 8 + j;
            if (k >= Lc) continue;
            float* dst = dW + ((size_t)level * Lc + k) * HW + pix0 + pg * 4;
            if (vec) {
                *reinterpret_cast<float4*>(dst) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (pix0 + pg * 4 + i < HW) dst[i] = acc[i][j];
            }
        }
    }
}

// dC[l][k][d] += sum_p W[l*Lc+k][p] dF[l][d][p]
// grid: (pixel slabs, D/64, levels); CTA tile 64 k x 64 d, thread tile 4 x 4; one atomic per
// output element per CTA.
constexpr int DC_SLAB = 4096;
__global__ void __launch_bounds__(256)
k_decode_bwd_dC(int Lc, int D, int64_t HW, const float* __restrict__ wmap, const float* __restrict__ dF,
                float* __restrict__ dC)
{
    __shared__ __align__(16) float s_w[64][33];
    __shared__ __align__(16) float s_f[64][33];
    const int tid = threadIdx.x;
    const int level = blockIdx.z;
    const int d0 = blockIdx.y * 64;
    const int64_t pbeg = (int64_t)blockIdx.x * DC_SLAB;
    const int64_t pend = min(HW, pbeg + DC_SLAB);
    const int kt = tid >> 4, dt = tid & 15;          // 16 x 16 threads, 4 x 4 outputs each
    for (int k0 = 0; k0 < Lc; k0 += 64) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
        for (int64_t pp = pbeg; pp < pend; pp += 32) {
            __syncthreads();
            for (int e = tid; e < 64 * 32; e += 256) {
                const int r = e >> 5, q = e & 31;
                const bool okp = pp + q < pend;
                s_w[r][q] = (okp && k0 + r < Lc) ? wmap[((size_t)level * Lc + k0 + r) * HW + pp + q] : 0.0f;
                s_f[r][q] = (okp && d0 + r < D) ? dF[((size_t)level * D + d0 + r) * HW + pp + q] : 0.0f;
            }
            __syncthreads();
#pragma unroll 8
            for (int q = 0; q < 32; ++q) {
                float wv[4], fv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { wv[i] = s_w[kt * 4 + i][q]; fv[i] = s_f[dt * 4 + i][q]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i], fv[j], acc[i][j]);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = k0 + kt * 4 + i, d = d0 + dt * 4 + j;
                if (k < Lc && d < D && acc[i][j] != 0.0f)
                    atomicAdd(dC + ((size_t)level * Lc + k) * D + d, acc[i][j]);
            }
    }
}

int launch_decode_backward(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                           const float* dF, float* dW, float* dC, cudaStream_t st)
{
    if (levels <= 0 || HW <= 0) return LSV2_OK;
    if (dW) {
        bool all_tc = true;
        for (int l = 0; l < levels && all_tc; ++l) {
            int rc = LSV2_OK;
            all_tc = launch_gemm_h_tf32x3(codebooks + (size_t)l * Lc * D, dF + (size_t)l * D * HW,
                                          dW + (size_t)l * Lc * HW, D, Lc, HW, st, &rc);
            if (rc) return rc;
        }
        if (!all_tc) {
            dim3 grid((unsigned)ceil_div(HW, 128), (unsigned)levels);
            k_decode_bwd_dW<<<grid, 256, 0, st>>>(Lc, D, HW, codebooks, dF, dW);
            LSV2_LAUNCH_CHECK("k_decode_bwd_dW");
        }
    }
    if (dC) {
        // tensor-core path (3xTF32, gemm_tf32x3.cu) level by level; CUDA-core kernel for shapes it does not take
        bool all_tc = true;
        for (int l = 0; l < levels && all_tc; ++l) {
            int rc = LSV2_OK;
            all_tc = launch_gemm_nt_tf32x3(dF + (size_t)l * D * HW, wmap + (size_t)l * Lc * HW,
                                           dC + (size_t)l * Lc * D, D, Lc, HW, st, &rc);
            if (rc) return rc;
            if (!all_tc && l > 0) return LSV2_ECUDA;      // cannot happen: the shape test does not depend on l
        }
        if (!all_tc) {
            dim3 grid((unsigned)ceil_div(HW, DC_SLAB), (unsigned)ceil_div(D, 64), (unsigned)levels);
            k_decode_bwd_dC<<<grid, 256, 0, st>>>(Lc, D, HW, wmap, dF, dC);
            LSV2_LAUNCH_CHECK("k_decode_bwd_dC");
        }
    }
    return LSV2_OK;
}

// ---- softmax -> top-k -> renormalise (utils/vq_utils.py:9-40) -----------------------------
// One warp per row; C <= 256.  Ties: the lower channel wins (torch.topk on CPU and the
// checker agree on that for distinct values; exact ties are measure-zero for real logits).
template <int NI>   // NI = ceil(C / 32): elements per lane
__global__ void __launch_bounds__(256)
k_topk_soft_code(int64_t N, int C, int k, const float* __restrict__ logits, float* __restrict__ dense,
                 float* __restrict__ weights, float* __restrict__ indices)
{
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const float* x = logits + (size_t)row * C;
    float v[NI];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        v[i] = c < C ? x[c] : -INFINITY;
        mx = fmaxf(mx, v[i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    // softmax in channel order to follow the checker's summation order is not possible in
    // parallel; the sum is taken as a tree (differs from a serial sum in the last ulps).
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        v[i] = c < C ? expf(v[i] - mx) : 0.0f;
        s += v[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
#pragma unroll
    for (int i = 0; i < NI; ++i) v[i] = v[i] / s;
    uint32_t sel = 0;   // bit i: element i*32+lane selected
    float ts = 0.0f;
    for (int j = 0; j < k; ++j) {
        float best = -1.0f;
        int bc = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int c = i * 32 + lane;
            if (c < C && !((sel >> i) & 1u) && (v[i] > best)) { best = v[i]; bc = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (ob > best || (ob == best && oc < bc)) { best = ob; bc = oc; }
        }
        if ((bc & 31) == lane && bc < C) sel |= 1u << (bc >> 5);
        ts += best;     // same value on every lane
    }
    // the checker sums the selected probabilities in channel order
    float tsum = 0.0f;
    {
        // only k terms are non-zero: walk the set bits of the per-slot ballots in channel order
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            uint32_t m = __ballot_sync(0xffffffffu, (sel >> i) & 1u);
            while (m) {
                const int l = __ffs(m) - 1;
                tsum += __shfl_sync(0xffffffffu, v[i], l);
                m &= m - 1;
            }
        }
    }
    (void)ts;
    const float den = tsum + 1e-10f;
    int before = 0;   // number of selected channels below this lane's element i
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        const uint32_t bits = __ballot_sync(0xffffffffu, (sel >> i) & 1u);
        const int o = before + __popc(bits & ((1u << lane) - 1u));
        const bool on = (sel >> i) & 1u;
        const float val = on ? v[i] / den : 0.0f;
        if (c < C) {
            if (dense) dense[(size_t)row * C + c] = val;
            if (on && o < k) {
                if (weights) weights[(size_t)row * k + o] = val;
                if (indices) indices[(size_t)row * k + o] = (float)c;
            }
        }
        before += __popc(bits);
    }
}

// C == 64 (the LangSplatV2 codebook size): 8 lanes per row, 8 consecutive channels per lane (two 16-byte loads and
// stores), 3-step reductions, 4 rows per warp -- half the instructions per row of the one-warp-per-row kernel.
__device__ __forceinline__ float group8_sum_in_order(float loc, int lane)
{
    float t = 0.0f;
#pragma unroll
    for (int l = 0; l < 8; ++l) t += __shfl_sync(0xffffffffu, loc, (lane & ~7) + l);
    return t;
}

// Optional extra outputs (the producer emits what the blend reads, directly): the compacted row of
// lsv2_compact_coeff (count, u16 index[16], f32 weight[16]: entries != 0 in ascending channel order) and the packed
// row of lsv2_pack_coeff (16 x (fp16 head | fp16 remainder << 16), 16 channel bytes, count, padding), so that a
// training iteration needs neither of those two kernels after its top-k.
struct TopkSparseOut {
    uint32_t* sp_count;      // [N]
    uint16_t* sp_index;      // [N][16]
    float* sp_weight;        // [N][16]
    uint32_t* packed;        // [N][24]
};

__global__ void __launch_bounds__(256)
k_topk_soft_code_c64(int64_t N, int k, const float* __restrict__ logits, float* __restrict__ dense,
                     float* __restrict__ weights, float* __restrict__ indices, TopkSparseOut so)
{
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int lane = threadIdx.x & 31, sub = threadIdx.x & 7;
    const bool live = row < N;
    const int64_t r = live ? row : N - 1;                      // every lane stays in the shuffles
    const float4* x4 = reinterpret_cast<const float4*>(logits + (size_t)r * 64 + sub * 8);
    const float4 a = x4[0], b = x4[1];
    float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    float mx = v[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, v[i]);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = expf(v[i] - mx); s += v[i]; }
    s = group8_sum_in_order(s, lane);                          // channel order up to the 8-wide partials
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = v[i] / s;
    uint32_t sel = 0;                                          // bit i: channel sub*8 + i selected
    for (int j = 0; j < k; ++j) {
        float best = -1.0f;
        int bc = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (!((sel >> i) & 1u) && v[i] > best) { best = v[i]; bc = sub * 8 + i; }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (ob > best || (ob == best && oc < bc)) { best = ob; bc = oc; }
        }
        if ((bc >> 3) == sub && bc < 64) sel |= 1u << (bc & 7);
    }
    float loc = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) loc += ((sel >> i) & 1u) ? v[i] : 0.0f;
    const float den = group8_sum_in_order(loc, lane) + 1e-10f;
    // rank of this lane's first selected channel among the row's selected ones
    const int cnt = __popc(sel);
    int before = 0;
#pragma unroll
    for (int l = 0; l < 8; ++l) {
        const int c = __shfl_sync(0xffffffffu, cnt, (lane & ~7) + l);
        if (l < sub) before += c;
    }
    float y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) y[i] = ((sel >> i) & 1u) ? v[i] / den : 0.0f;
    // for the sparse rows: which kept entries are not exactly 0, and where this lane's go in the row
    uint32_t nz = 0, incl = 0, total = 0;
    if (so.sp_count) {                                         // uniform over the grid
#pragma unroll
        for (int i = 0; i < 8; ++i) nz |= (((sel >> i) & 1u) && y[i] != 0.0f ? 1u : 0u) << i;
        incl = __popc(nz);
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o, 8);
            if (sub >= o) incl += t;
        }
        total = __shfl_sync(0xffffffffu, incl, 7, 8);
    }
    if (!live) return;
    if (dense) {
        float4* d4 = reinterpret_cast<float4*>(dense + (size_t)row * 64 + sub * 8);
        d4[0] = make_float4(y[0], y[1], y[2], y[3]);
        d4[1] = make_float4(y[4], y[5], y[6], y[7]);
    }
    if (weights || indices) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if ((sel >> i) & 1u) {
                const int o = before + __popc(sel & ((1u << i) - 1u));
                if (o < k) {
                    if (weights) weights[(size_t)row * k + o] = y[i];
                    if (indices) indices[(size_t)row * k + o] = (float)(sub * 8 + i);
                }
            }
        }
    }
    if (so.sp_count) {
        // the rows lsv2_compact_coeff / lsv2_pack_coeff would build from `dense`: entries that are exactly 0 (a kept
        // probability that underflowed) are not part of the row, as there.  All 8 lanes of a live row are here.
        uint32_t* prow = so.packed + (size_t)row * LSV2_PACKED_WORDS;
        // zero the 24 words first (3 per lane), then every lane drops its entries in
        prow[3 * sub + 0] = (3 * sub + 0 == 20) ? min(total, 16u) : 0u;
        prow[3 * sub + 1] = (3 * sub + 1 == 20) ? min(total, 16u) : 0u;
        prow[3 * sub + 2] = (3 * sub + 2 == 20) ? min(total, 16u) : 0u;
        __syncwarp(__activemask());                            // whole rows (8 lanes) leave together above
        uint32_t pos = incl - __popc(nz);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if ((nz >> i) & 1u) {
                if (pos < 16u) {
                    const int ch = sub * 8 + i;
                    so.sp_index[(size_t)row * 16 + pos] = (uint16_t)ch;
                    so.sp_weight[(size_t)row * 16 + pos] = y[i];
                    const __half h = __float2half_rn(y[i]);
                    const __half lo = __float2half_rn(__fsub_rn(y[i], __half2float(h)));
                    prow[pos] = (uint32_t)__half_as_ushort(h) | ((uint32_t)__half_as_ushort(lo) << 16);
                    reinterpret_cast<uint8_t*>(prow + 16)[pos] = (uint8_t)ch;
                }
                ++pos;
            }
        }
        if (sub == 0) so.sp_count[row] = total;          // k <= 16 here, so never a "dense" row
    }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int launch_topk_soft_code_sparse(int64_t N, int k, const float* logits, float* dense, uint32_t* sp_count,
                                 uint16_t* sp_index, float* sp_weight, uint32_t* packed, cudaStream_t st)
{
    if (N <= 0) return LSV2_OK;
    if (k <= 0 || k > 16) return LSV2_EUNSUPPORTED;
    if (!logits || !sp_count || !sp_index || !sp_weight || !packed || !aligned16(logits) || (dense && !aligned16(dense)))
        return LSV2_EINVAL;
    TopkSparseOut so{sp_count, sp_index, sp_weight, packed};
    k_topk_soft_code_c64<<<(unsigned)ceil_div(N * 8, 256), 256, 0, st>>>(N, k, logits, dense, nullptr, nullptr, so);
    LSV2_LAUNCH_CHECK("k_topk_soft_code_c64<sparse>");
    return LSV2_OK;
}

int launch_topk_soft_code(int64_t N, int C, int k, const float* logits, float* dense, float* weights,
                          float* indices, cudaStream_t st)
{
    if (N <= 0) return LSV2_OK;
    if (C <= 0 || C > 256 || k <= 0 || k > C) return LSV2_EUNSUPPORTED;
    if (C == 64 && aligned16(logits) && aligned16(dense)) {
        k_topk_soft_code_c64<<<(unsigned)ceil_div(N * 8, 256), 256, 0, st>>>(N, k, logits, dense, weights, indices,
                                                                             TopkSparseOut{nullptr, nullptr, nullptr, nullptr});
        LSV2_LAUNCH_CHECK("k_topk_soft_code_c64");
        return LSV2_OK;
    }
    const unsigned grid = (unsigned)ceil_div(N * 32, 256);
    if (C <= 32) k_topk_soft_code<1><<<grid, 256, 0, st>>>(N, C, k, logits, dense, weights, indices);
    else if (C <= 64) k_topk_soft_code<2><<<grid, 256, 0, st>>>(N, C, k, logits, dense, weights, indices);
    else if (C <= 128) k_topk_soft_code<4><<<grid, 256, 0, st>>>(N, C, k, logits, dense, weights, indices);
    else k_topk_soft_code<8><<<grid, 256, 0, st>>>(N, C, k, logits, dense, weights, indices);
    LSV2_LAUNCH_CHECK("k_topk_soft_code");
    return LSV2_OK;
}

// Backward of softmax -> top-k -> renormalise (autograd of utils/vq_utils.py:9-24): one warp per row.
//   p = softmax(x);  y_i = m_i p_i / s,  s = sum_{j in top-k} p_j + 1e-10   (m = the forward's top-k mask, read back
//   from the forward's dense output: y_i != 0)
//   dL/dp_i = m_i (g_i / s - G / s^2),  G = sum_j m_j g_j p_j ;   dL/dx_i = p_i (dL/dp_i - sum_j dL/dp_j p_j)
template <int NI>
__global__ void __launch_bounds__(256)
k_topk_soft_code_bwd(int64_t N, int C, const float* __restrict__ logits, const float* __restrict__ dense,
                     const float* __restrict__ d_dense, float* __restrict__ d_logits)
{
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const float* x = logits + (size_t)row * C;
    const float* y = dense + (size_t)row * C;
    const float* g = d_dense + (size_t)row * C;
    float v[NI], gm[NI];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        v[i] = c < C ? x[c] : -INFINITY;
        gm[i] = (c < C && y[c] != 0.0f) ? g[c] : 0.0f;
        mx = fmaxf(mx, v[i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        v[i] = c < C ? expf(v[i] - mx) : 0.0f;
        s += v[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float ts = 0.0f, G = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        v[i] = v[i] / s;
        const bool m = c < C && y[c] != 0.0f;
        ts += m ? v[i] : 0.0f;
        G = fmaf(gm[i], v[i], G);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ts += __shfl_xor_sync(0xffffffffu, ts, o);
        G += __shfl_xor_sync(0xffffffffu, G, o);
    }
    ts += 1e-10f;
    const float inv = 1.0f / ts, sub = G * inv * inv;
    float dp[NI], D = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        const bool m = c < C && y[c] != 0.0f;
        dp[i] = m ? gm[i] * inv - sub : 0.0f;
        D = fmaf(dp[i], v[i], D);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) D += __shfl_xor_sync(0xffffffffu, D, o);
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int c = i * 32 + lane;
        if (c < C) d_logits[(size_t)row * C + c] = v[i] * (dp[i] - D);
    }
}

// C == 64 version of the backward: same 8-lanes-per-row layout as k_topk_soft_code_c64.
// ROWS: the upstream gradient arrives in the compacted row form of lsv2_render_backward_rows -- d_rows[row][e] belongs
// to channel sp_index[row][e], e < sp_count[row] (<= row_stride) -- instead of a dense [N, 64] array: 16 B per row
// instead of 256 B for top-4, which is also all a data-parallel exchange has to move.
struct TopkRowsGrad {
    const uint32_t* sp_count;
    const uint16_t* sp_index;      // [N][16]
    const float* d_rows;           // [N][row_stride]
    int row_stride;
};

template <bool ROWS>
__global__ void __launch_bounds__(256)
k_topk_soft_code_bwd_c64(int64_t N, const float* __restrict__ logits, const float* __restrict__ dense,
                         const float* __restrict__ d_dense, float* __restrict__ d_logits, TopkRowsGrad rg)
{
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    const bool live = row < N;
    const int64_t r = live ? row : N - 1;
    const size_t base = (size_t)r * 64 + sub * 8;
    const float4 xa = *reinterpret_cast<const float4*>(logits + base), xb = *reinterpret_cast<const float4*>(logits + base + 4);
    const float4 ya = *reinterpret_cast<const float4*>(dense + base), yb = *reinterpret_cast<const float4*>(dense + base + 4);
    float v[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    const float y[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
    float g[8];
    if (ROWS) {
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = 0.0f;
        const int kmax = min(rg.row_stride, 16);
        const int cnt = min((int)min(rg.sp_count[r], 16u), kmax);
        // lane `sub` fetches entries sub and sub + 8 of its row, the 8 lanes pass them round
        int ch0 = -1, ch1 = -1;
        float w0 = 0.0f, w1 = 0.0f;
        if (sub < cnt) { ch0 = rg.sp_index[(size_t)r * 16 + sub]; w0 = rg.d_rows[(size_t)r * rg.row_stride + sub]; }
        if (sub + 8 < cnt) { ch1 = rg.sp_index[(size_t)r * 16 + sub + 8]; w1 = rg.d_rows[(size_t)r * rg.row_stride + sub + 8]; }
        for (int e = 0; e < kmax; ++e) {                       // kmax is uniform over the grid
            const int ch = __shfl_sync(0xffffffffu, e < 8 ? ch0 : ch1, e & 7, 8);
            const float w = __shfl_sync(0xffffffffu, e < 8 ? w0 : w1, e & 7, 8);
            const bool mine = ch >= 0 && (ch >> 3) == sub;
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = (mine && (ch & 7) == i) ? w : g[i];
        }
    } else {
        const float4 ga = *reinterpret_cast<const float4*>(d_dense + base), gb = *reinterpret_cast<const float4*>(d_dense + base + 4);
        g[0] = ga.x; g[1] = ga.y; g[2] = ga.z; g[3] = ga.w; g[4] = gb.x; g[5] = gb.y; g[6] = gb.z; g[7] = gb.w;
    }
    float mx = v[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, v[i]);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = expf(v[i] - mx); s += v[i]; }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float ts = 0.0f, G = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        v[i] = v[i] / s;
        const bool m = y[i] != 0.0f;
        ts += m ? v[i] : 0.0f;
        G = fmaf(m ? g[i] : 0.0f, v[i], G);
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        ts += __shfl_xor_sync(0xffffffffu, ts, o);
        G += __shfl_xor_sync(0xffffffffu, G, o);
    }
    ts += 1e-10f;
    const float inv = 1.0f / ts, sub2 = G * inv * inv;
    float dp[8], Dd = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        dp[i] = (y[i] != 0.0f) ? g[i] * inv - sub2 : 0.0f;
        Dd = fmaf(dp[i], v[i], Dd);
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) Dd += __shfl_xor_sync(0xffffffffu, Dd, o);
    if (!live) return;
    float4* o4 = reinterpret_cast<float4*>(d_logits + (size_t)row * 64 + sub * 8);
    o4[0] = make_float4(v[0] * (dp[0] - Dd), v[1] * (dp[1] - Dd), v[2] * (dp[2] - Dd), v[3] * (dp[3] - Dd));
    o4[1] = make_float4(v[4] * (dp[4] - Dd), v[5] * (dp[5] - Dd), v[6] * (dp[6] - Dd), v[7] * (dp[7] - Dd));
}

int launch_topk_soft_code_bwd_rows(int64_t N, const float* logits, const float* dense, const uint32_t* sp_count,
                                   const uint16_t* sp_index, const float* d_rows, int row_stride, float* d_logits,
                                   cudaStream_t st)
{
    if (N == 0) return LSV2_OK;
    if (row_stride <= 0 || !logits || !dense || !sp_count || !sp_index || !d_rows || !d_logits) return LSV2_EINVAL;
    if (!aligned16(logits) || !aligned16(dense) || !aligned16(d_logits)) return LSV2_EINVAL;
    k_topk_soft_code_bwd_c64<true><<<(unsigned)ceil_div(N * 8, 256), 256, 0, st>>>(
        N, logits, dense, nullptr, d_logits, TopkRowsGrad{sp_count, sp_index, d_rows, row_stride});
    LSV2_LAUNCH_CHECK("k_topk_soft_code_bwd_c64<rows>");
    return LSV2_OK;
}

int launch_topk_soft_code_bwd(int64_t N, int C, const float* logits, const float* dense, const float* d_dense,
                              float* d_logits, cudaStream_t st)
{
    if (N == 0) return LSV2_OK;
    if (C <= 0 || C > 256) return LSV2_EUNSUPPORTED;
    if (C == 64 && aligned16(logits) && aligned16(dense) && aligned16(d_dense) && aligned16(d_logits)) {
        k_topk_soft_code_bwd_c64<false><<<(unsigned)ceil_div(N * 8, 256), 256, 0, st>>>(N, logits, dense, d_dense, d_logits,
                                                                                        TopkRowsGrad{nullptr, nullptr, nullptr, 0});
        LSV2_LAUNCH_CHECK("k_topk_soft_code_bwd_c64");
        return LSV2_OK;
    }
    const unsigned grid = (unsigned)ceil_div(N * 32, 256);
    if (C <= 32) k_topk_soft_code_bwd<1><<<grid, 256, 0, st>>>(N, C, logits, dense, d_dense, d_logits);
    else if (C <= 64) k_topk_soft_code_bwd<2><<<grid, 256, 0, st>>>(N, C, logits, dense, d_dense, d_logits);
    else if (C <= 128) k_topk_soft_code_bwd<4><<<grid, 256, 0, st>>>(N, C, logits, dense, d_dense, d_logits);
    else k_topk_soft_code_bwd<8><<<grid, 256, 0, st>>>(N, C, logits, dense, d_dense, d_logits);
    LSV2_LAUNCH_CHECK("k_topk_soft_code_bwd");
    return LSV2_OK;
}

}  // namespace lsv2
