// render_forward.cu -- K6: per-tile front-to-back alpha blending with sparse-coefficient
// accumulation into an L-channel weight map.  Replaces renderCUDA (forward) of the
// reference's rasterizer (call site /root/reference/gaussian_renderer/__init__.py:108-119,
// SURVEY.md row R8).
//
// One CTA per 16x16 tile, one thread per pixel.  Instances are staged through shared memory
// in batches (geometry + each Gaussian's <=16 (channel, weight) pairs); the per-pixel weight
// vector lives in shared memory as acc[channel][pixel] so the data-dependent channel index
// is a conflict-free access (lane == pixel) instead of a local-memory spill.  A warp whose 32
// pixels are all saturated stops blending; the CTA stops when all 8 warps have.
//
// The arithmetic per (pixel, instance) is the checker's, operation for operation, so
// n_contrib, final_T, colour and the weight map reproduce bit for bit.
#include "common.cuh"

namespace lsv2 {

constexpr int RF_THREADS = TILE_PIX;   // 256
constexpr int RF_BATCH = 128;
constexpr int RF_KMAX = LSV2_SPARSE_CAP;
constexpr uint32_t ROW_DENSE = 0xFFFFFFFFu;
constexpr uint16_t IDX_SKIP = 0xFFFFu;

struct RenderArgs {
    int W, H, gx;
    const uint2* ranges;
    const uint32_t* point_list;
    const float2* xy;
    const float4* conic_opacity;
    const float* rgb;
    const float* bg;
    int mode, L, K;
    const float* coeff_dense;
    const float* wq;
    const float* iq;
    const uint32_t* sp_count;
    const uint16_t* sp_index;
    const float* sp_weight;
    float* color;
    float* wmap;
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* status;
};

struct StagedInstance {
    float2 xy;
    float ax, b, az, op;   // -0.5*conic.x, -conic.y, -0.5*conic.z, opacity
    float r, g, bl;
    float thr;             // power below this can never reach alpha >= 1/255
    uint32_t gid;
    uint32_t cnt;
};

// power threshold: alpha = op*exp(power) < 1/255 for every power < thr.  The margin (0.01 in
// the exponent, ~1 % in alpha) dwarfs the 2-ulp error of repro_expf, so skipping the
// exponential below thr never changes a decision.
__device__ __forceinline__ float power_threshold(float op)
{
    if (!(op > 0.0f)) return 1.0f;                // alpha <= 0 always: skip everything
    return -__logf(255.0f * op) - 0.01f;
}

template <bool FEAT>
__global__ void __launch_bounds__(RF_THREADS) k_render_forward(RenderArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    StagedInstance* s_inst = reinterpret_cast<StagedInstance*>(smem_raw);
    uint16_t* s_sidx = reinterpret_cast<uint16_t*>(s_inst + RF_BATCH);
    float* s_sw = reinterpret_cast<float*>(s_sidx + RF_BATCH * RF_KMAX);
    float* acc = s_sw + RF_BATCH * RF_KMAX;        // [L][256]

    const int tid = threadIdx.x;
    const int tile = blockIdx.x;
    const int tx = tile % a.gx, ty = tile / a.gx;
    const int px = tx * TILE + (tid & (TILE - 1));
    const int py = ty * TILE + (tid >> 4);
    const bool inside = px < a.W && py < a.H;
    const float pxf = (float)px, pyf = (float)py;
    const uint2 range = a.ranges[tile];

    if (FEAT) {
        for (int i = tid; i < a.L * RF_THREADS; i += RF_THREADS) acc[i] = 0.0f;
    }

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    uint32_t contributor = 0, last = 0;
    bool done = !inside;

    for (uint32_t start = range.x; start < range.y; start += RF_BATCH) {
        const int batch_n = (int)min((uint32_t)RF_BATCH, range.y - start);
        __syncthreads();                            // previous batch fully consumed (and acc zeroed)
        if (tid < batch_n) {
            const uint32_t g = a.point_list[start + tid];
            const float4 co = a.conic_opacity[g];
            StagedInstance s;
            s.xy = a.xy[g];
            s.ax = -0.5f * co.x; s.b = -co.y; s.az = -0.5f * co.z; s.op = co.w;
            s.r = a.rgb[3 * (size_t)g]; s.g = a.rgb[3 * (size_t)g + 1]; s.bl = a.rgb[3 * (size_t)g + 2];
            s.thr = power_threshold(co.w);
            s.gid = g;
            s.cnt = 0;
            if (FEAT) {
                if (a.mode == LSV2_MODE_QUICK) s.cnt = (uint32_t)a.K;
                else s.cnt = a.sp_count ? a.sp_count[g] : ROW_DENSE;
            }
            s_inst[tid] = s;
        }
        if (FEAT) {
            // two threads per instance fetch its (channel, weight) pairs
            const int inst = tid >> 1;
            if (inst < batch_n) {
                const uint32_t g = a.point_list[start + inst];
                if (a.mode == LSV2_MODE_QUICK) {
                    for (int k = tid & 1; k < a.K; k += 2) {
                        const float fi = a.iq[(size_t)g * a.K + k];
                        const int c = (int)fi;
                        const bool ok = c >= 0 && c < a.L;
                        if (!ok) a.status[LSV2_STATUS_BAD_INDEX] = 1u;
                        s_sidx[inst * RF_KMAX + k] = ok ? (uint16_t)c : IDX_SKIP;
                        s_sw[inst * RF_KMAX + k] = a.wq[(size_t)g * a.K + k];
                    }
                } else if (a.sp_count) {
                    const uint32_t cnt = a.sp_count[g];
                    if (cnt != ROW_DENSE)
                        for (uint32_t k = tid & 1; k < cnt; k += 2) {
                            s_sidx[inst * RF_KMAX + k] = a.sp_index[(size_t)g * RF_KMAX + k];
                            s_sw[inst * RF_KMAX + k] = a.sp_weight[(size_t)g * RF_KMAX + k];
                        }
                }
            }
        }
        __syncthreads();

        if (!__all_sync(0xffffffffu, done)) {
            for (int j = 0; j < batch_n; ++j) {
                if (done) continue;
                ++contributor;
                const StagedInstance& s = s_inst[j];
                const float dx = s.xy.x - pxf, dy = s.xy.y - pyf;
                const float power = fmaf(s.ax * dx, dx, fmaf(s.az * dy, dy, (s.b * dx) * dy));
                if (power > 0.0f || power < s.thr) continue;
                const float alpha = fminf(ALPHA_MAX, s.op * repro_expf(power));
                if (alpha < ALPHA_MIN) continue;
                const float test_T = T * (1.0f - alpha);
                if (test_T < T_MIN) { done = true; continue; }
                const float w = alpha * T;
                C0 = fmaf(s.r, w, C0);
                C1 = fmaf(s.g, w, C1);
                C2 = fmaf(s.bl, w, C2);
                if (FEAT) {
                    const uint32_t cnt = s.cnt;
                    if (cnt != ROW_DENSE) {
                        for (uint32_t k = 0; k < cnt; ++k) {
                            const uint32_t c = s_sidx[j * RF_KMAX + k];
                            if (c == IDX_SKIP) continue;
                            float* p = acc + c * RF_THREADS + tid;
                            *p = fmaf(s_sw[j * RF_KMAX + k], w, *p);
                        }
                    } else {
                        const float* row = a.coeff_dense + (size_t)s.gid * a.L;
                        for (int c = 0; c < a.L; ++c) {
                            float* p = acc + c * RF_THREADS + tid;
                            *p = fmaf(__ldg(row + c), w, *p);
                        }
                    }
                }
                T = test_T;
                last = contributor;
            }
        }
        if (__syncthreads_count(done) == RF_THREADS) break;
    }

    if (inside) {
        const size_t HW = (size_t)a.H * a.W;
        const size_t pix = (size_t)py * a.W + px;
        a.final_T[pix] = T;
        a.n_contrib[pix] = last;
        a.color[pix] = fmaf(T, a.bg[0], C0);
        a.color[HW + pix] = fmaf(T, a.bg[1], C1);
        a.color[2 * HW + pix] = fmaf(T, a.bg[2], C2);
        if (FEAT) {
            float* out = a.wmap + pix;
            for (int c = 0; c < a.L; ++c) st_stream_f32(out + (size_t)c * HW, acc[c * RF_THREADS + tid]);
        }
    }
}

static size_t rf_smem_bytes(bool feat, int L)
{
    size_t b = sizeof(StagedInstance) * RF_BATCH;
    if (feat) b += (size_t)RF_BATCH * RF_KMAX * (sizeof(uint16_t) + sizeof(float)) + (size_t)L * RF_THREADS * sizeof(float);
    return b;
}

int launch_render_forward(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b,
                          const lsv2_features& f, const lsv2_image& img, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    const bool feat = f.mode != LSV2_MODE_RGB;
    if (feat) {
        if (f.L <= 0 || img.wmap == nullptr) return LSV2_EINVAL;
        if (f.mode == LSV2_MODE_QUICK && (f.K <= 0 || f.K > RF_KMAX || !f.weights_quick || !f.indices_quick))
            return f.K > RF_KMAX ? LSV2_EUNSUPPORTED : LSV2_EINVAL;
        if (f.mode == LSV2_MODE_DENSE && !f.coeff_dense) return LSV2_EINVAL;
        if (f.L >= 0xFFFF) return LSV2_EUNSUPPORTED;
    }
    const size_t smem = rf_smem_bytes(feat, f.L);
    if (smem > 227 * 1024) return LSV2_EUNSUPPORTED;
    RenderArgs a;
    a.W = v.width; a.H = v.height; a.gx = gx;
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.xy = reinterpret_cast<const float2*>(g.xy);
    a.conic_opacity = reinterpret_cast<const float4*>(g.conic_opacity);
    a.rgb = g.rgb; a.bg = v.bg;
    a.mode = f.mode; a.L = f.L; a.K = f.K;
    a.coeff_dense = f.coeff_dense; a.wq = f.weights_quick; a.iq = f.indices_quick;
    a.sp_count = f.sparse_count; a.sp_index = f.sparse_index; a.sp_weight = f.sparse_weight;
    if (!(f.sparse_count && f.sparse_index && f.sparse_weight)) a.sp_count = nullptr;
    a.color = img.color; a.wmap = img.wmap; a.final_T = img.final_T; a.n_contrib = img.n_contrib;
    a.status = b.status;
    if (feat) {
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_forward<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_render_forward<true><<<gx * gy, RF_THREADS, smem, st>>>(a);
    } else {
        k_render_forward<false><<<gx * gy, RF_THREADS, smem, st>>>(a);
    }
    LSV2_LAUNCH_CHECK("k_render_forward");
    return LSV2_OK;
}

// ---- dense row compaction ---------------------------------------------------------------
// One warp per visible Gaussian: (channel, weight) pairs of the non-zero entries of its
// coefficient row, ascending channel order; rows with more than RF_KMAX non-zeros are
// flagged and read densely by the blend.
__global__ void __launch_bounds__(256)
k_compact_coeff(int P, int L, const int32_t* __restrict__ radii, const float* __restrict__ coeff,
                uint32_t* __restrict__ sp_count, uint16_t* __restrict__ sp_index, float* __restrict__ sp_weight)
{
    const int g = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= P) return;
    if (radii[g] <= 0) {
        if (lane == 0) sp_count[g] = 0;
        return;
    }
    const float* row = coeff + (size_t)g * L;
    uint32_t count = 0;
    for (int c0 = 0; c0 < L; c0 += 32) {
        const int c = c0 + lane;
        const float v = c < L ? row[c] : 0.0f;
        const uint32_t nz = __ballot_sync(0xffffffffu, v != 0.0f);
        const uint32_t pos = count + __popc(nz & ((1u << lane) - 1u));
        if (v != 0.0f && pos < RF_KMAX) {
            sp_index[(size_t)g * RF_KMAX + pos] = (uint16_t)c;
            sp_weight[(size_t)g * RF_KMAX + pos] = v;
        }
        count += __popc(nz);
    }
    if (lane == 0) sp_count[g] = count > RF_KMAX ? ROW_DENSE : count;
}

int launch_compact_coeff(int P, const lsv2_geom& g, const lsv2_features& f, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (f.mode != LSV2_MODE_DENSE) return LSV2_OK;
    if (!f.coeff_dense || !f.sparse_count || !f.sparse_index || !f.sparse_weight || f.L <= 0) return LSV2_EINVAL;
    k_compact_coeff<<<ceil_div((int64_t)P * 32, 256), 256, 0, st>>>(P, f.L, g.radii, f.coeff_dense, f.sparse_count,
                                                                   f.sparse_index, f.sparse_weight);
    LSV2_LAUNCH_CHECK("k_compact_coeff");
    return LSV2_OK;
}

}  // namespace lsv2
