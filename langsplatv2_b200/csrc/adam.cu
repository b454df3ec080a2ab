// adam.cu -- the optimizer step of the feature stage as one streaming kernel over a flat parameter range.
// Replaces torch.optim.Adam(l, lr=0.0, eps=1e-15) stepped once per iteration
// (/root/reference/scene/gaussian_model.py:255, train.py:261-263) on the [N, 64] logits and the codebooks.
//
// Same update as torch's single-tensor Adam (no weight decay, no amsgrad, maximize = False):
//   m = m + (g - m)(1 - b1) ;  v = b2 v + (1 - b2) g g ;  p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// in torch's operation order (lerp, mul + addcmul, sqrt / div / add, addcdiv).
// 16-byte loads/stores, 7 streams (read p, g, m, v; write p, m, v): HBM-bound, 28 B per element.
// A 4-element group whose g, m and v are all zero is left untouched -- that IS the dense update (m and v stay 0,
// p moves by 0 / eps = 0), so skipping its three stores is exact; logits rows of Gaussians no view has reached
// yet cost 12 B per element instead of 28.
#include "common.cuh"

namespace lsv2 {

struct AdamArgs {
    float* p;
    const float* g;
    float* m;
    float* v;
    int64_t n;
    float lr_over_bc1, b1, b2, one_m_b1, one_m_b2, eps, sqrt_bc2;
    const float* dev_scal;   // optional device pair {lr / (1 - b1^t), sqrt(1 - b2^t)} (CUDA-graph replays)
    float* g_zero;      // optional: the gradient is cleared in the same pass (set_to_none=False semantics)
};

__global__ void __launch_bounds__(256) k_adam(AdamArgs a)
{
    const int64_t n4 = a.n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float one_m_b1 = a.one_m_b1, one_m_b2 = a.one_m_b2;
    if (a.dev_scal) { a.lr_over_bc1 = __ldg(a.dev_scal); a.sqrt_bc2 = __ldg(a.dev_scal + 1); }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 g = __ldcs(reinterpret_cast<const float4*>(a.g) + i);
        float4 m = __ldcs(reinterpret_cast<const float4*>(a.m) + i);
        float4 v = __ldcs(reinterpret_cast<const float4*>(a.v) + i);
        const bool idle = g.x == 0.f && g.y == 0.f && g.z == 0.f && g.w == 0.f && m.x == 0.f && m.y == 0.f && m.z == 0.f &&
                          m.w == 0.f && v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f;
        if (idle) continue;
        float4 p = __ldcs(reinterpret_cast<const float4*>(a.p) + i);
#define LSV2_ADAM1(c)                                                                        \
        m.c = fmaf(g.c - m.c, one_m_b1, m.c);                                                \
        v.c = fmaf(one_m_b2 * g.c, g.c, a.b2 * v.c);                                         \
        p.c = p.c - a.lr_over_bc1 * (m.c / (sqrtf(v.c) / a.sqrt_bc2 + a.eps));
        LSV2_ADAM1(x) LSV2_ADAM1(y) LSV2_ADAM1(z) LSV2_ADAM1(w)
#undef LSV2_ADAM1
        __stcs(reinterpret_cast<float4*>(a.p) + i, p);
        __stcs(reinterpret_cast<float4*>(a.m) + i, m);
        __stcs(reinterpret_cast<float4*>(a.v) + i, v);
        if (a.g_zero) __stcs(reinterpret_cast<float4*>(a.g_zero) + i, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    // tail (n not a multiple of 4)
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < a.n) {
        const float g = a.g[t];
        float m = a.m[t], v = a.v[t];
        if (!(g == 0.f && m == 0.f && v == 0.f)) {
            m = fmaf(g - m, one_m_b1, m);
            v = fmaf(one_m_b2 * g, g, a.b2 * v);
            a.p[t] = a.p[t] - a.lr_over_bc1 * (m / (sqrtf(v) / a.sqrt_bc2 + a.eps));
            a.m[t] = m;
            a.v[t] = v;
            if (a.g_zero) a.g_zero[t] = 0.f;
        }
    }
}

int launch_adam(float* p, const float* g, float* m, float* v, int64_t n, double lr, double b1, double b2, double eps,
                int64_t step, int zero_grad, const float* dev_scalars, cudaStream_t st)
{
    if (n == 0) return LSV2_OK;
    if (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) return LSV2_EINVAL;
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.n = n;
    // bias corrections in double on the host, as torch does
    const double bc1 = 1.0 - pow(b1, (double)step), bc2 = 1.0 - pow(b2, (double)step);
    a.lr_over_bc1 = (float)(lr / bc1);
    a.sqrt_bc2 = (float)sqrt(bc2);
    // torch forms 1 - beta in double from the user's double and only then narrows (1 - 0.999f != float(0.001))
    a.b1 = (float)b1; a.b2 = (float)b2; a.one_m_b1 = (float)(1.0 - b1); a.one_m_b2 = (float)(1.0 - b2); a.eps = (float)eps;
    a.dev_scal = dev_scalars;
    a.g_zero = zero_grad ? const_cast<float*>(g) : nullptr;
    const int64_t n4 = (n + 3) >> 2;
    const int blocks = (int)((n4 + 255) / 256 < (int64_t)NUM_SMS * 8 ? (n4 + 255) / 256 : (int64_t)NUM_SMS * 8);
    k_adam<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(a);
    LSV2_LAUNCH_CHECK("k_adam");
    return LSV2_OK;
}

}  // namespace lsv2
