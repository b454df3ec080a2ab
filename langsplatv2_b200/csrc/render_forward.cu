// render_forward.cu -- K6: per-tile front-to-back alpha blending with sparse-coefficient
// accumulation into an L-channel weight map.  Replaces renderCUDA (forward) of the
// reference's rasterizer (call site /root/reference/gaussian_renderer/__init__.py:108-119,
// SURVEY.md row R8).
//
// One CTA per 16x16 tile: 8 blend warps (one thread per pixel, a warp = a 16x2 pixel strip)
// fed by 2 producer warps through a 4-stage shared-memory ring with mbarrier hand-off.
//   producer: gathers each instance's packed blend record (written by preprocess) and its
//             <=16 (channel, weight) pairs, and builds, per strip, a bit mask of the instances
//             whose alpha cut-off ellipse can reach the strip's two pixel rows.
//   blend   : walks only the set bits of its strip's mask, four instances at a time (the four
//             quadratic forms are evaluated independently before any branch), and accumulates
//             into acc[channel][pixel] in shared memory, so the data-dependent channel index is
//             a conflict-free access (lane == pixel) instead of a local-memory spill.
// A strip whose 32 pixels are all saturated stops blending; when all 8 have, the producers
// stop gathering (the ring carries a terminator).
//
// The arithmetic per (pixel, instance) is the checker's, operation for operation, and skipped
// instances are exactly those the checker skips, so n_contrib, final_T, colour and the weight
// map reproduce bit for bit.
#include "common.cuh"
#include "tc_common.cuh"

namespace lsv2 {

using tc::mbar_arrive;
using tc::mbar_init;
using tc::mbar_wait;

constexpr int RF_PIX_THREADS = TILE_PIX;                 // 256 blend threads
constexpr int RF_PRODUCER_WARPS = 2;
constexpr int RF_THREADS = RF_PIX_THREADS + 32 * RF_PRODUCER_WARPS;
constexpr int RF_BATCH = 64;
constexpr int RF_STAGES = 4;
constexpr int RF_KMAX = LSV2_SPARSE_CAP;
constexpr int RF_STRIPS = 8;
constexpr uint32_t ROW_DENSE = 0xFFFFFFFFu;
constexpr uint32_t CNT_DENSE = 0xFFu;
constexpr uint8_t IDX_SKIP = 0xFFu;
constexpr int RF_MAX_L = 192;

struct __align__(16) Stage {
    float4 g0[RF_BATCH];                 // x, y, -a/2, -b
    float4 g1[RF_BATCH];                 // -c/2, power cut-off, opacity, r
    float4 g2[RF_BATCH];                 // g, b, bits(count), bits(gaussian id)
    float sw[RF_BATCH][RF_KMAX];
    uint8_t sidx[RF_BATCH][RF_KMAX];
    uint32_t mask[RF_STRIPS][2];
    int32_t count;                       // instances in this stage; < 0 terminates
    int32_t pad[3];
};

struct RenderArgs {
    int W, H, gx;
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;
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

// rows (relative to the tile, 0..15) whose pixel centres the cut-off ellipse of a splat can reach
// -> bit mask over the 8 two-row strips.  Conservative by 0.02 px.
__device__ __forceinline__ uint32_t strip_mask(float y_rel, float ax, float b, float az, float thr)
{
    if (!(thr < 0.0f)) return 0u;                       // never reaches alpha >= 1/255
    const float qa = -2.0f * ax, qc = -2.0f * az;       // conic a, c ; conic b = -b
    const float det = qa * qc - b * b;
    float dy = 1e30f;
    if (det > 0.0f && qa > 0.0f) dy = sqrtf((-2.0f * thr) * (qa / det)) * 1.0001f + 0.02f;
    if (!(dy == dy)) dy = 1e30f;
    const float lo_f = ceilf(y_rel - dy), hi_f = floorf(y_rel + dy);
    if (!(hi_f >= 0.0f) || !(lo_f <= 15.0f)) {
        if (lo_f != lo_f || hi_f != hi_f) return 0xFFu;  // NaN geometry: keep everything
        return 0u;
    }
    const int lo = lo_f <= 0.0f ? 0 : (int)lo_f;
    const int hi = hi_f >= 15.0f ? 15 : (int)hi_f;
    if (hi < lo) return 0u;
    const int wlo = lo >> 1, whi = hi >> 1;
    return ((1u << (whi + 1)) - 1u) & ~((1u << wlo) - 1u);
}

template <int MODE>
__global__ void __launch_bounds__(RF_THREADS) k_render_forward(RenderArgs a)
{
    constexpr bool FEAT = MODE != LSV2_MODE_RGB;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Stage* stages = reinterpret_cast<Stage*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + RF_STAGES);
    uint64_t* empty = full + RF_STAGES;
    uint32_t* s_ndone = reinterpret_cast<uint32_t*>(empty + RF_STAGES);
    float* acc = reinterpret_cast<float*>(s_ndone + 4);            // [L][256]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tile = blockIdx.x;
    const int tx = tile % a.gx, ty = tile / a.gx;
    const uint2 range = a.ranges[tile];

    if (tid == 0) {
        for (int s = 0; s < RF_STAGES; ++s) { mbar_init(full + s, 32); mbar_init(empty + s, RF_STRIPS); }
        s_ndone[0] = 0;        // strips whose 32 pixels are all saturated
        s_ndone[1] = 0xFFFFFFFFu;   // smallest batch index at which a terminator has been queued
        tc::mbar_fence_init();
    }
    __syncthreads();

    if (warp >= RF_STRIPS) {
        // =============================== producers ===============================
        const int pw = warp - RF_STRIPS;
        const float y0 = (float)(ty * TILE);
        const uint32_t n_inst = range.y - range.x;
        const uint32_t n_batches = (n_inst + RF_BATCH - 1) / RF_BATCH;
        uint32_t gid_next[2];
        auto load_ids = [&](uint32_t bi, uint32_t (&g)[2]) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const uint32_t i = bi * RF_BATCH + q * 32 + lane;
                g[q] = i < n_inst ? a.point_list[range.x + i] : 0xFFFFFFFFu;
            }
        };
        if ((uint32_t)pw < n_batches) load_ids(pw, gid_next);
        for (uint32_t bi = pw;; bi += RF_PRODUCER_WARPS) {
            const int s = bi % RF_STAGES;
            const uint32_t use = bi / RF_STAGES;
            // Termination protocol: a producer whose turn meets the stop condition (list exhausted, or
            // all 8 strips saturated) queues a terminator at ITS batch index and records the smallest
            // such index in `term`.  The blend warps stop at the first terminator in batch order, so a
            // producer only ever abandons batches that lie beyond a recorded terminator -- every batch
            // before the first terminator is still produced by its owner.
            volatile uint32_t* term = s_ndone + 1;
            if (use > 0) {
                for (;;) {                       // lane 0 polls, the warp follows (keeps the warp converged)
                    uint32_t st_ = 0;
                    if (lane == 0)
                        st_ = tc::mbar_try_wait(empty + s, (use - 1) & 1) ? 1u : (bi > *term ? 2u : 0u);
                    st_ = __shfl_sync(0xffffffffu, st_, 0);
                    if (st_ == 1u) break;
                    if (st_ == 2u) return;
                }
                __syncwarp();
            }
            Stage& S = stages[s];
            uint32_t flags = 0;
            if (lane == 0) flags = (*(volatile uint32_t*)s_ndone >= RF_STRIPS ? 1u : 0u) | (bi > *term ? 2u : 0u);
            flags = __shfl_sync(0xffffffffu, flags, 0);
            if (flags & 2u) return;
            if (bi >= n_batches || (flags & 1u)) {
                if (lane == 0) { S.count = -1; atomicMin(s_ndone + 1, bi); }
                mbar_arrive(full + s);
                return;
            }
            uint32_t gid[2] = {gid_next[0], gid_next[1]};
            if (bi + RF_PRODUCER_WARPS < n_batches) load_ids(bi + RF_PRODUCER_WARPS, gid_next);
            uint32_t m8[2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int slot = q * 32 + lane;
                const uint32_t g = gid[q];
                m8[q] = 0;
                if (g != 0xFFFFFFFFu) {
                    const float4 r0 = __ldg(a.rec + 3 * (size_t)g);
                    const float4 r1 = __ldg(a.rec + 3 * (size_t)g + 1);
                    const float4 r2 = __ldg(a.rec + 3 * (size_t)g + 2);
                    uint32_t cnt = 0;
                    if (MODE == LSV2_MODE_QUICK) {
                        cnt = (uint32_t)a.K;
                        const float* wsrc = a.wq + (size_t)g * a.K;
                        const float* isrc = a.iq + (size_t)g * a.K;
                        for (int k = 0; k < a.K; ++k) {
                            const int c = (int)__ldg(isrc + k);
                            const bool ok = c >= 0 && c < a.L;
                            if (!ok) a.status[LSV2_STATUS_BAD_INDEX] = 1u;
                            S.sidx[slot][k] = ok ? (uint8_t)c : IDX_SKIP;
                            S.sw[slot][k] = __ldg(wsrc + k);
                        }
                    } else if (MODE == LSV2_MODE_DENSE) {
                        const uint32_t c0 = a.sp_count ? a.sp_count[g] : ROW_DENSE;
                        if (c0 == ROW_DENSE) {
                            cnt = CNT_DENSE;
                        } else {
                            cnt = c0;
                            for (uint32_t k = 0; k < c0; ++k) {
                                S.sidx[slot][k] = (uint8_t)a.sp_index[(size_t)g * RF_KMAX + k];
                                S.sw[slot][k] = a.sp_weight[(size_t)g * RF_KMAX + k];
                            }
                        }
                    }
                    S.g0[slot] = r0;
                    S.g1[slot] = r1;
                    S.g2[slot] = make_float4(r2.x, r2.y, __uint_as_float(cnt), __uint_as_float(g));
                    m8[q] = strip_mask(r0.y - y0, r0.z, r0.w, r1.x, r1.y);
                }
            }
#pragma unroll
            for (int w = 0; w < RF_STRIPS; ++w) {
                const uint32_t b0 = __ballot_sync(0xffffffffu, (m8[0] >> w) & 1u);
                const uint32_t b1 = __ballot_sync(0xffffffffu, (m8[1] >> w) & 1u);
                if (lane == 0) { S.mask[w][0] = b0; S.mask[w][1] = b1; }
            }
            if (lane == 0) S.count = (int32_t)min((uint32_t)RF_BATCH, n_inst - bi * RF_BATCH);
            mbar_arrive(full + s);
        }
        return;
    }

    // ================================= blend =================================
    const int px = tx * TILE + (tid & (TILE - 1));
    const int py = ty * TILE + (tid >> 4);
    const bool inside = px < a.W && py < a.H;
    const float pxf = (float)px, pyf = (float)py;

    if (FEAT) {
        for (int c = 0; c < a.L; ++c) acc[c * RF_PIX_THREADS + tid] = 0.0f;   // own column only
    }

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    uint32_t last = 0;
    bool done = !inside;
    bool warp_done = __all_sync(0xffffffffu, done);
    if (warp_done && lane == 0) atomicAdd(s_ndone, 1u);

    for (uint32_t bi = 0;; ++bi) {
        const int s = bi % RF_STAGES;
        mbar_wait(full + s, (bi / RF_STAGES) & 1);
        const Stage& S = stages[s];
        const int count = S.count;
        if (count < 0) break;
        if (!warp_done) {
            uint64_t m = ((uint64_t)S.mask[warp][1] << 32) | S.mask[warp][0];
            const uint32_t base = bi * RF_BATCH;
            while (m) {
                int j[4];
                int n = 0;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (m) { j[u] = __ffsll((long long)m) - 1; m &= m - 1; n = u + 1; }
                    else j[u] = j[0];
                }
                float pw_[4], thr_[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4 g0 = S.g0[j[u]];
                    const float4 g1 = S.g1[j[u]];
                    const float dx = g0.x - pxf, dy = g0.y - pyf;
                    pw_[u] = fmaf(g0.z * dx, dx, fmaf(g1.x * dy, dy, (g0.w * dx) * dy));
                    thr_[u] = g1.y;
                }
                bool hit[4];
                bool any = false;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    hit[u] = (u < n) && !(pw_[u] > 0.0f || pw_[u] < thr_[u]);
                    any |= hit[u];
                }
                if (!any || done) continue;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!hit[u] || done) continue;
                    const int jj = j[u];
                    const float4 g1 = S.g1[jj];
                    const float alpha = fminf(ALPHA_MAX, g1.z * repro_expf(pw_[u]));
                    if (alpha < ALPHA_MIN) continue;
                    const float test_T = T * (1.0f - alpha);
                    if (test_T < T_MIN) { done = true; continue; }
                    const float w = alpha * T;
                    const float4 g2 = S.g2[jj];
                    C0 = fmaf(g1.w, w, C0);
                    C1 = fmaf(g2.x, w, C1);
                    C2 = fmaf(g2.y, w, C2);
                    if (FEAT) {
                        const uint32_t cnt = __float_as_uint(g2.z);
                        if (cnt != CNT_DENSE) {
                            for (uint32_t k = 0; k < cnt; ++k) {
                                const uint32_t c = S.sidx[jj][k];
                                if (c == IDX_SKIP) continue;
                                float* p = acc + c * RF_PIX_THREADS + tid;
                                *p = fmaf(S.sw[jj][k], w, *p);
                            }
                        } else {
                            const float* row = a.coeff_dense + (size_t)__float_as_uint(g2.w) * a.L;
                            for (int c = 0; c < a.L; ++c) {
                                float* p = acc + c * RF_PIX_THREADS + tid;
                                *p = fmaf(__ldg(row + c), w, *p);
                            }
                        }
                    }
                    T = test_T;
                    last = base + jj + 1;
                }
            }
            if (__all_sync(0xffffffffu, done)) {
                warp_done = true;
                if (lane == 0) atomicAdd(s_ndone, 1u);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + s);
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
            for (int c = 0; c < a.L; ++c) st_stream_f32(out + (size_t)c * HW, acc[c * RF_PIX_THREADS + tid]);
        }
    }
}

static size_t rf_smem_bytes(bool feat, int L)
{
    size_t b = sizeof(Stage) * RF_STAGES + 2 * RF_STAGES * sizeof(uint64_t) + 16;
    if (feat) b += (size_t)L * RF_PIX_THREADS * sizeof(float);
    return b;
}

int launch_render_forward(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b,
                          const lsv2_features& f, const lsv2_image& img, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec) return LSV2_EINVAL;
    const bool feat = f.mode != LSV2_MODE_RGB;
    if (feat) {
        if (f.L <= 0 || img.wmap == nullptr) return LSV2_EINVAL;
        if (f.mode == LSV2_MODE_QUICK && (f.K <= 0 || f.K > RF_KMAX || !f.weights_quick || !f.indices_quick))
            return f.K > RF_KMAX ? LSV2_EUNSUPPORTED : LSV2_EINVAL;
        if (f.mode == LSV2_MODE_DENSE && !f.coeff_dense) return LSV2_EINVAL;
        if (f.L > RF_MAX_L) return LSV2_EUNSUPPORTED;
    }
    const size_t smem = rf_smem_bytes(feat, f.L);
    if (smem > 227 * 1024) return LSV2_EUNSUPPORTED;
    RenderArgs a;
    a.W = v.width; a.H = v.height; a.gx = gx;
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.rec = reinterpret_cast<const float4*>(g.blend_rec);
    a.bg = v.bg;
    a.mode = f.mode; a.L = f.L; a.K = f.K;
    a.coeff_dense = f.coeff_dense; a.wq = f.weights_quick; a.iq = f.indices_quick;
    a.sp_count = f.sparse_count; a.sp_index = f.sparse_index; a.sp_weight = f.sparse_weight;
    if (!(f.sparse_count && f.sparse_index && f.sparse_weight)) a.sp_count = nullptr;
    a.color = img.color; a.wmap = img.wmap; a.final_T = img.final_T; a.n_contrib = img.n_contrib;
    a.status = b.status;
#define RF_LAUNCH(M)                                                                                            \
    do {                                                                                                        \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_forward<M>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                           (int)smem));                                                         \
        k_render_forward<M><<<gx * gy, RF_THREADS, smem, st>>>(a);                                              \
    } while (0)
    if (f.mode == LSV2_MODE_QUICK) RF_LAUNCH(LSV2_MODE_QUICK);
    else if (f.mode == LSV2_MODE_DENSE) RF_LAUNCH(LSV2_MODE_DENSE);
    else RF_LAUNCH(LSV2_MODE_RGB);
#undef RF_LAUNCH
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

// L == 64: 8 lanes per row, 8 consecutive channels per lane (two 16-byte loads), 4 rows per warp
__global__ void __launch_bounds__(256)
k_compact_coeff_c64(int P, const int32_t* __restrict__ radii, const float* __restrict__ coeff,
                    uint32_t* __restrict__ sp_count, uint16_t* __restrict__ sp_index, float* __restrict__ sp_weight)
{
    const int g = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3);
    const int sub = threadIdx.x & 7;
    const bool live = g < P && radii[min(g, P - 1)] > 0;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (live) {
        const float4* r4 = reinterpret_cast<const float4*>(coeff + (size_t)g * 64 + sub * 8);
        const float4 a = r4[0], b = r4[1];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    uint32_t nz = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) nz |= (v[i] != 0.0f ? 1u : 0u) << i;
    const uint32_t cnt = __popc(nz);
    uint32_t incl = cnt;                                      // inclusive prefix inside the 8-lane group
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o, 8);
        if (sub >= o) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 7, 8);
    if (g >= P) return;
    if (!live) {
        if (sub == 0) sp_count[g] = 0;
        return;
    }
    uint32_t pos = incl - cnt;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if ((nz >> i) & 1u) {
            if (pos < RF_KMAX) {
                sp_index[(size_t)g * RF_KMAX + pos] = (uint16_t)(sub * 8 + i);
                sp_weight[(size_t)g * RF_KMAX + pos] = v[i];
            }
            ++pos;
        }
    }
    if (sub == 0) sp_count[g] = total > RF_KMAX ? ROW_DENSE : total;
}

int launch_compact_coeff(int P, const lsv2_geom& g, const lsv2_features& f, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (f.mode != LSV2_MODE_DENSE) return LSV2_OK;
    if (!f.coeff_dense || !f.sparse_count || !f.sparse_index || !f.sparse_weight || f.L <= 0) return LSV2_EINVAL;
    if (f.L == 64 && (reinterpret_cast<uintptr_t>(f.coeff_dense) & 15) == 0) {
        k_compact_coeff_c64<<<ceil_div((int64_t)P * 8, 256), 256, 0, st>>>(P, g.radii, f.coeff_dense, f.sparse_count,
                                                                          f.sparse_index, f.sparse_weight);
        LSV2_LAUNCH_CHECK("k_compact_coeff_c64");
        return LSV2_OK;
    }
    k_compact_coeff<<<ceil_div((int64_t)P * 32, 256), 256, 0, st>>>(P, f.L, g.radii, f.coeff_dense, f.sparse_count,
                                                                   f.sparse_index, f.sparse_weight);
    LSV2_LAUNCH_CHECK("k_compact_coeff");
    return LSV2_OK;
}

}  // namespace lsv2
