// loss.cu -- the feature-stage loss of train.py as two kernels instead of ~28 elementwise passes over [512, H, W]:
//   language_feature / (norm(dim=0) + 1e-10)            /root/reference/train.py:157-159
//   cos_loss(feature * mask, gt * mask)                 /root/reference/train.py:162-164, utils/loss_utils.py:24-25
//     = 1 - mean_p  <a_p / max(|a_p|, 1e-8),  b_p / max(|b_p|, 1e-8)>       (torch.nn.functional.cosine_similarity)
// Forward reads F and gt once (one thread per pixel, coalesced across pixels for every channel) and leaves, per
// pixel, the two scalars of the gradient  dL/dF_p = alpha_p * gt_p + beta_p * F_p ; backward reads F and gt once
// more and writes dL/dF.  Algorithmic bytes: 2 x D x HW x 4 forward, 3 x D x HW x 4 backward.
#include "common.cuh"

namespace lsv2 {

constexpr int LOSS_THREADS = 256;
constexpr float COS_EPS = 1e-8f;      // torch.nn.functional.cosine_similarity default
constexpr float NORM_EPS = 1e-10f;    // train.py:159

__global__ void __launch_bounds__(LOSS_THREADS)
k_cos_loss_fwd(int D, int64_t HW, const float* __restrict__ F, const float* __restrict__ gt,
               const float* __restrict__ mask, int normalize, float* __restrict__ alpha_beta,
               float* __restrict__ partial)
{
    __shared__ float s_red[LOSS_THREADS / 32];
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    float cosv = 0.0f;
    if (p < HW) {
        float n2 = 0.0f, fg = 0.0f, g2 = 0.0f;
        const float* f = F + p;
        const float* g = gt + p;
#pragma unroll 8
        for (int d = 0; d < D; ++d) {
            const float fv = __ldg(f + (size_t)d * HW), gv = __ldg(g + (size_t)d * HW);
            n2 = fmaf(fv, fv, n2);
            fg = fmaf(fv, gv, fg);
            g2 = fmaf(gv, gv, g2);
        }
        const float m = mask ? mask[p] : 1.0f;
        const float n = sqrtf(n2);
        const float s = normalize ? 1.0f / (n + NORM_EPS) : 1.0f;      // a = m s F
        const float na = fabsf(m) * s * n, nb = fabsf(m) * sqrtf(g2);
        const float ca = fmaxf(na, COS_EPS), cb = fmaxf(nb, COS_EPS);
        const float icab = 1.0f / (ca * cb);
        cosv = m * m * s * fg * icab;
        // dL/dcos = -1/HW is applied by the backward kernel together with the upstream gradient
        const float ka = na > COS_EPS ? cosv / (ca * na) : 0.0f;       // coefficient of a in d cos / d a (clamped norm: none)
        // q = d cos/d a = m g icab - ka m s F ;   q.F = m fg icab - ka m s n2
        const float qF = m * fg * icab - ka * m * s * n2;
        float alpha = m * m * s * icab;
        float beta = -ka * m * m * s * s;
        if (normalize && n > 0.0f) beta -= m * qF * s * s / n;         // through the norm in the denominator
        alpha_beta[p] = alpha;
        alpha_beta[HW + p] = beta;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cosv += __shfl_xor_sync(0xffffffffu, cosv, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = cosv;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int i = 0; i < LOSS_THREADS / 32; ++i) t += s_red[i];
        partial[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(LOSS_THREADS)
k_cos_loss_bwd(int D, int64_t HW, const float* __restrict__ F, const float* __restrict__ gt,
               const float* __restrict__ alpha_beta, const float* __restrict__ grad_out, float* __restrict__ dF)
{
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    if (p >= HW) return;
    const float go = -(grad_out ? grad_out[0] : 1.0f) / (float)HW;
    const float alpha = alpha_beta[p] * go, beta = alpha_beta[HW + p] * go;
    const int d0 = blockIdx.y * (D / gridDim.y), d1 = d0 + D / gridDim.y;
#pragma unroll 8
    for (int d = d0; d < d1; ++d) {
        const size_t o = (size_t)d * HW + p;
        dF[o] = fmaf(alpha, __ldg(gt + o), beta * __ldg(F + o));
    }
}

int launch_cos_loss_forward(int D, int64_t HW, const float* F, const float* gt, const float* mask, int normalize,
                            float* alpha_beta, float* partial, cudaStream_t st)
{
    if (HW == 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    k_cos_loss_fwd<<<(unsigned)blocks, LOSS_THREADS, 0, st>>>(D, HW, F, gt, mask, normalize, alpha_beta, partial);
    LSV2_LAUNCH_CHECK("k_cos_loss_fwd");
    return LSV2_OK;
}

int launch_cos_loss_backward(int D, int64_t HW, const float* F, const float* gt, const float* alpha_beta,
                             const float* grad_out, float* dF, cudaStream_t st)
{
    if (HW == 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    const int split = (D % 4 == 0) ? 4 : 1;        // more CTAs in flight for the streaming pass
    k_cos_loss_bwd<<<dim3((unsigned)blocks, split), LOSS_THREADS, 0, st>>>(D, HW, F, gt, alpha_beta, grad_out, dF);
    LSV2_LAUNCH_CHECK("k_cos_loss_bwd");
    return LSV2_OK;
}

// ---- the same loss without ever forming the feature map ------------------------------------
// With F_p = C^T w_p everything the loss needs is a scalar product in code space:
//   |F_p|^2 = w_p^T G w_p  (G = C C^T, 64 x 64),   F_p . gt_p = w_p . h_p  (h = C gt, one GEMM),   |gt_p|^2 = g2_p
// and the gradient dL/dF_p = alpha_p gt_p + beta_p F_p pulls back to
//   dL/dw_p = alpha_p h_p + beta_p G w_p ,      dL/dC = (alpha w) gt^T + ((beta w) w^T) C .
// k_gram_loss_fwd: one thread per pixel, w_p in registers, G in shared memory; writes u = G w (kept for backward),
// the two scalars and the per-block sums of cos_p.  k_gram_loss_bwd: the three [Lc][HW] products of the pull-back.
constexpr int GL_LC = 64;

__device__ __forceinline__ void cos_scalars(float m, float n2, float fg, float g2, int normalize, float& cosv,
                                            float& alpha, float& beta)
{
    const float n = sqrtf(fmaxf(n2, 0.0f));
    const float s = normalize ? 1.0f / (n + NORM_EPS) : 1.0f;
    const float na = fabsf(m) * s * n, nb = fabsf(m) * sqrtf(g2);
    const float ca = fmaxf(na, COS_EPS), cb = fmaxf(nb, COS_EPS);
    const float icab = 1.0f / (ca * cb);
    cosv = m * m * s * fg * icab;
    const float ka = na > COS_EPS ? cosv / (ca * na) : 0.0f;
    const float qF = m * fg * icab - ka * m * s * n2;
    alpha = m * m * s * icab;
    beta = -ka * m * m * s * s;
    if (normalize && n > 0.0f) beta -= m * qF * s * s / n;
}

// SEG: the ground truth is table[seg]; h is then (C table^T)[k][seg_p] (`h` = that [Lc][nseg] matrix), g2 the
// per-segment |table_s|^2 and the mask is seg_p >= 0.
template <bool SEG>
__global__ void __launch_bounds__(LOSS_THREADS)
k_gram_loss_fwd(int64_t HW, const float* __restrict__ wmap, const float* __restrict__ h, const float* __restrict__ g2,
                const float* __restrict__ mask, const float* __restrict__ G, int normalize, float* __restrict__ u,
                float* __restrict__ alpha_beta, float* __restrict__ partial, const int32_t* __restrict__ seg, int nseg)
{
    __shared__ __align__(16) float s_G[GL_LC][GL_LC];
    __shared__ float s_red[LOSS_THREADS / 32];
    for (int e = threadIdx.x; e < GL_LC * GL_LC; e += LOSS_THREADS) s_G[e / GL_LC][e % GL_LC] = G[e];
    __syncthreads();
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    float cosv = 0.0f;
    if (p < HW) {
        float w[GL_LC];
        float fg = 0.0f;
        int sp = 0;
        if (SEG) { sp = seg[p]; if (sp >= nseg) sp = -1; }
        const int sc = max(sp, 0);
#pragma unroll
        for (int k = 0; k < GL_LC; ++k) {
            w[k] = __ldg(wmap + (size_t)k * HW + p);
            fg = fmaf(w[k], SEG ? __ldg(h + (size_t)k * nseg + sc) : __ldg(h + (size_t)k * HW + p), fg);
        }
        float n2 = 0.0f;
#pragma unroll
        for (int k = 0; k < GL_LC; ++k) {
            float acc = 0.0f;
#pragma unroll
            for (int j4 = 0; j4 < GL_LC / 4; ++j4) {
                const float4 g4 = *reinterpret_cast<const float4*>(&s_G[k][4 * j4]);
                acc = fmaf(g4.x, w[4 * j4], acc);
                acc = fmaf(g4.y, w[4 * j4 + 1], acc);
                acc = fmaf(g4.z, w[4 * j4 + 2], acc);
                acc = fmaf(g4.w, w[4 * j4 + 3], acc);
            }
            u[(size_t)k * HW + p] = acc;
            n2 = fmaf(w[k], acc, n2);
        }
        float alpha, beta;
        if (SEG) cos_scalars(sp >= 0 ? 1.0f : 0.0f, n2, fg, g2[sc], normalize, cosv, alpha, beta);
        else cos_scalars(mask ? mask[p] : 1.0f, n2, fg, g2[p], normalize, cosv, alpha, beta);
        alpha_beta[p] = alpha;
        alpha_beta[HW + p] = beta;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cosv += __shfl_xor_sync(0xffffffffu, cosv, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = cosv;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int i = 0; i < LOSS_THREADS / 32; ++i) t += s_red[i];
        partial[blockIdx.x] = t;
    }
}

template <bool SEG>
__global__ void __launch_bounds__(LOSS_THREADS)
k_gram_loss_bwd(int64_t HW, const float* __restrict__ wmap, const float* __restrict__ h, const float* __restrict__ u,
                const float* __restrict__ alpha_beta, const float* __restrict__ grad_out, float* __restrict__ dw,
                float* __restrict__ aw, float* __restrict__ bw, const int32_t* __restrict__ seg, int nseg)
{
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    if (p >= HW) return;
    int sc = 0;
    if (SEG) { sc = seg[p]; sc = (sc < 0 || sc >= nseg) ? 0 : sc; }
    const float go = -(grad_out ? grad_out[0] : 1.0f) / (float)HW;
    const float alpha = alpha_beta[p] * go, beta = alpha_beta[HW + p] * go;
    const int k0 = blockIdx.y * (GL_LC / gridDim.y), k1 = k0 + GL_LC / gridDim.y;
#pragma unroll 4
    for (int k = k0; k < k1; ++k) {
        const size_t o = (size_t)k * HW + p;
        const float wv = __ldg(wmap + o);
        dw[o] = fmaf(alpha, SEG ? __ldg(h + (size_t)k * nseg + sc) : __ldg(h + o), beta * __ldg(u + o));
        aw[o] = alpha * wv;
        bw[o] = beta * wv;
    }
}

__global__ void __launch_bounds__(LOSS_THREADS)
k_colsumsq(int D, int64_t HW, const float* __restrict__ x, float* __restrict__ out)
{
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    if (p >= HW) return;
    float s = 0.0f;
#pragma unroll 8
    for (int d = 0; d < D; ++d) {
        const float v = __ldg(x + (size_t)d * HW + p);
        s = fmaf(v, v, s);
    }
    out[p] = s;
}

int launch_gram_loss_forward(int Lc, int64_t HW, const float* wmap, const float* h, const float* g2, const float* mask,
                             const float* G, int normalize, float* u, float* alpha_beta, float* partial,
                             const int32_t* seg, int nseg, cudaStream_t st)
{
    if (Lc != GL_LC) return LSV2_EUNSUPPORTED;
    if (HW == 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    if (seg)
        k_gram_loss_fwd<true><<<(unsigned)blocks, LOSS_THREADS, 0, st>>>(HW, wmap, h, g2, mask, G, normalize, u,
                                                                         alpha_beta, partial, seg, nseg);
    else
        k_gram_loss_fwd<false><<<(unsigned)blocks, LOSS_THREADS, 0, st>>>(HW, wmap, h, g2, mask, G, normalize, u,
                                                                          alpha_beta, partial, nullptr, 0);
    LSV2_LAUNCH_CHECK("k_gram_loss_fwd");
    return LSV2_OK;
}

int launch_gram_loss_backward(int Lc, int64_t HW, const float* wmap, const float* h, const float* u,
                              const float* alpha_beta, const float* grad_out, float* dw, float* aw, float* bw,
                              const int32_t* seg, int nseg, cudaStream_t st)
{
    if (Lc != GL_LC) return LSV2_EUNSUPPORTED;
    if (HW == 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    if (seg)
        k_gram_loss_bwd<true><<<dim3((unsigned)blocks, 4), LOSS_THREADS, 0, st>>>(HW, wmap, h, u, alpha_beta, grad_out, dw,
                                                                                  aw, bw, seg, nseg);
    else
        k_gram_loss_bwd<false><<<dim3((unsigned)blocks, 4), LOSS_THREADS, 0, st>>>(HW, wmap, h, u, alpha_beta, grad_out, dw,
                                                                                   aw, bw, nullptr, 0);
    LSV2_LAUNCH_CHECK("k_gram_loss_bwd");
    return LSV2_OK;
}

// out[k][s] += sum over the pixels p with seg[p] == s of x[k][p]  (k < Lc; seg < 0: pixel skipped).
// The ground truth of the feature stage is a per-segment table gathered through a segmentation map
// (scene/cameras.py:59-96: point_feature = feature_map[seg]), so (alpha w) gt^T = (segment sums of alpha w) . table.
// A warp takes 32 consecutive pixels, which belong to one or a few segments: it adds ONE value per (segment, k).
__global__ void __launch_bounds__(LOSS_THREADS)
k_segment_sum(int Lc, int64_t HW, int nseg, const float* __restrict__ x, const int32_t* __restrict__ seg,
              float* __restrict__ out)
{
    const int64_t p = (int64_t)blockIdx.x * LOSS_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    int s = p < HW ? seg[p] : -1;
    if (s >= nseg) s = -1;
    // the (few) distinct segments of this warp: up to SS_GROUPS are reduced with shuffles, lanes of any further
    // segment add on their own
    constexpr int SS_GROUPS = 4;
    const uint32_t peers = __match_any_sync(0xffffffffu, s);
    uint32_t leaders = __ballot_sync(0xffffffffu, s >= 0 && lane == __ffs(peers) - 1);
    if (leaders == 0u) return;
    uint32_t gmask[SS_GROUPS];
    int gseg[SS_GROUPS];
    uint32_t covered = 0;
#pragma unroll
    for (int g = 0; g < SS_GROUPS; ++g) {
        gmask[g] = 0; gseg[g] = 0;
        if (leaders) {
            const int l = __ffs(leaders) - 1;
            leaders &= leaders - 1;
            gmask[g] = __shfl_sync(0xffffffffu, peers, l);
            gseg[g] = __shfl_sync(0xffffffffu, s, l);
            covered |= gmask[g];
        }
    }
    const bool own = s >= 0 && !((covered >> lane) & 1u);
    const int k0 = blockIdx.y * (Lc / gridDim.y), k1 = k0 + Lc / gridDim.y;      // (k1 - k0) % 4 == 0 (launcher)
    const bool hi16 = lane & 16, hi8 = lane & 8;
    for (int k = k0; k < k1; k += 4) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = s >= 0 ? __ldg(x + (size_t)(k + j) * HW + p) : 0.0f;
#pragma unroll
        for (int g = 0; g < SS_GROUPS; ++g) {
            if (gmask[g] == 0u) break;                                            // warp-uniform
            const bool in = (gmask[g] >> lane) & 1u;
            const float a0 = in ? v[0] : 0.0f, a1 = in ? v[1] : 0.0f, a2 = in ? v[2] : 0.0f, a3 = in ? v[3] : 0.0f;
            // transposing butterfly: lane 8q ends with the warp sum of channel k + q
            float b0 = hi16 ? a2 : a0, b1 = hi16 ? a3 : a1;
            b0 += __shfl_xor_sync(0xffffffffu, hi16 ? a0 : a2, 16);
            b1 += __shfl_xor_sync(0xffffffffu, hi16 ? a1 : a3, 16);
            float t = hi8 ? b1 : b0;
            t += __shfl_xor_sync(0xffffffffu, hi8 ? b0 : b1, 8);
            t += __shfl_xor_sync(0xffffffffu, t, 4);
            t += __shfl_xor_sync(0xffffffffu, t, 2);
            t += __shfl_xor_sync(0xffffffffu, t, 1);
            if ((lane & 7) == 0 && t != 0.0f) atomicAdd(out + (size_t)(k + (lane >> 3)) * nseg + gseg[g], t);
        }
        if (own) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (v[j] != 0.0f) atomicAdd(out + (size_t)(k + j) * nseg + s, v[j]);
        }
    }
}

int launch_segment_sum(int Lc, int64_t HW, int nseg, const float* x, const int32_t* seg, float* out, cudaStream_t st)
{
    if (HW == 0 || Lc <= 0 || nseg <= 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    if (Lc % 16 != 0) return LSV2_EUNSUPPORTED;      // 4 row blocks of a multiple of 4 channels
    const int ky = 4;
    k_segment_sum<<<dim3((unsigned)blocks, ky), LOSS_THREADS, 0, st>>>(Lc, HW, nseg, x, seg, out);
    LSV2_LAUNCH_CHECK("k_segment_sum");
    return LSV2_OK;
}

int launch_colsumsq(int D, int64_t HW, const float* x, float* out, cudaStream_t st)
{
    if (HW == 0) return LSV2_OK;
    const int64_t blocks = (HW + LOSS_THREADS - 1) / LOSS_THREADS;
    k_colsumsq<<<(unsigned)blocks, LOSS_THREADS, 0, st>>>(D, HW, x, out);
    LSV2_LAUNCH_CHECK("k_colsumsq");
    return LSV2_OK;
}

}  // namespace lsv2
