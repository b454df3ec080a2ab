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

}  // namespace lsv2
