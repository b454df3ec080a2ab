// render_backward.cu -- K8: gradient of the blend w.r.t. the per-Gaussian sparse coefficients
// (and colours).  Replaces the language-feature part of renderCUDA (backward) in the
// reference's rasterizer for the frozen-geometry feature stage
// (/root/reference/scene/gaussian_model.py:238-243, train.py:173):
//     dL/dcoeff[g][c] = sum_pix alpha*T * dL/dWmap[c][pix]
//
// One CTA per tile.  The tile of dL/dWmap is staged in shared memory once; the tile's
// instance list is replayed with the forward's exact arithmetic up to each pixel's
// n_contrib (read from the forward's packed blend record, so alpha, T and the cut-off are the
// forward's bit for bit); a warp skips instances whose alpha >= 1/255 extent misses its two rows;
// the per-(instance, channel) sums are reduced with warp shuffles into a shared-memory batch
// buffer and flushed with ONE global atomic per (instance, channel) per tile -- not one per
// pixel.  The replay is latency-bound: shared memory is budgeted for three CTAs per SM
// (1.52 -> 1.21 ms at 1M Gaussians / 1008x756, profiles/r02h_train_launches.txt).
#include "common.cuh"

namespace lsv2 {

constexpr int RB_THREADS = TILE_PIX;
constexpr int RB_WARPS = RB_THREADS / 32;
constexpr int RB_BATCH = 64;
constexpr int RB_KMAX = LSV2_SPARSE_CAP;
constexpr uint32_t RB_ROW_DENSE = 0xFFFFFFFFu;

struct BackArgs {
    int W, H, gx, L;
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;               // packed blend record of the forward (preprocess.cu): same alpha, same cut-off
    const uint32_t* n_contrib;
    const uint32_t* sp_count;
    const uint16_t* sp_index;
    const float* coeff_dense;
    const float* dL_dwmap;
    const float* dL_dcolor;
    float* dL_dcoeff;
    float* dL_dcolors;
    int row_stride;                  // > 0: dL_dcoeff is [P][row_stride] in the order of the compacted row (entry e), not [P][L]
};

struct BackInstance {
    float x, y, ax, b;
    float az, thr, op, ext_y;
    uint32_t gid;
    uint32_t cnt;
    uint32_t pad0, pad1;
};

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <bool DENSE_GRAD, bool COLOR>
__global__ void __launch_bounds__(RB_THREADS, 3) k_render_backward(BackArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BackInstance* s_inst = reinterpret_cast<BackInstance*>(smem_raw);
    uint16_t* s_sidx = reinterpret_cast<uint16_t*>(s_inst + RB_BATCH);
    float* s_grad = reinterpret_cast<float*>(s_sidx + RB_BATCH * RB_KMAX);   // [B][KMAX]
    float* s_gradc = s_grad + RB_BATCH * RB_KMAX;                            // [B][4] (COLOR only)
    float* s_dw = s_gradc + (COLOR ? RB_BATCH * 4 : 0);                      // [L][256]
    __shared__ uint32_t s_max;
    // shared-memory budget: at L = 64 the dL/dWmap tile alone is 64 KB; everything else is kept under 10 KB so that
    // THREE CTAs fit an SM -- the replay is latency-bound and time follows resident warps (profiles/r02h)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tx = tile % a.gx, ty = tile / a.gx;
    const int px = tx * TILE + (tid & (TILE - 1));
    const int py = ty * TILE + (tid >> 4);
    const bool inside = px < a.W && py < a.H;
    const float pxf = (float)px, pyf = (float)py;
    const float wy0 = (float)(ty * TILE + 2 * warp), wy1 = wy0 + 1.0f;       // the warp's two pixel rows
    const uint2 range = a.ranges[tile];
    const size_t HW = (size_t)a.H * a.W;
    const size_t pix = inside ? (size_t)py * a.W + px : 0;

    if (tid == 0) s_max = 0;
    __syncthreads();
    const uint32_t last = inside ? a.n_contrib[pix] : 0u;
    atomicMax(&s_max, last);
    for (int c = 0; c < a.L; ++c) s_dw[c * RB_THREADS + tid] = inside ? a.dL_dwmap[(size_t)c * HW + pix] : 0.0f;
    float dc0 = 0.f, dc1 = 0.f, dc2 = 0.f;
    if (COLOR && inside) {
        dc0 = a.dL_dcolor[pix]; dc1 = a.dL_dcolor[HW + pix]; dc2 = a.dL_dcolor[2 * HW + pix];
    }
    __syncthreads();
    const uint32_t n_replay = min(s_max, range.y - range.x);

    float T = 1.0f;
    for (uint32_t off = 0; off < n_replay; off += RB_BATCH) {
        const int batch_n = (int)min((uint32_t)RB_BATCH, n_replay - off);
        __syncthreads();
        if (tid < batch_n) {
            const uint32_t g = a.point_list[range.x + off + tid];
            const float4 r0 = a.rec[3 * (size_t)g + 0], r1 = a.rec[3 * (size_t)g + 1], r2 = a.rec[3 * (size_t)g + 2];
            BackInstance s;
            s.x = r0.x; s.y = r0.y; s.ax = r0.z; s.b = r0.w;
            s.az = r1.x; s.thr = r1.y; s.op = r1.z; s.ext_y = r2.z;
            s.gid = g;
            s.cnt = a.sp_count ? a.sp_count[g] : RB_ROW_DENSE;
            if (a.row_stride > 0 && s.cnt == RB_ROW_DENSE) s.cnt = 0;   // no row form for > 16 non-zeros (never emitted by the producer)
            s.pad0 = s.pad1 = 0;
            s_inst[tid] = s;
        }
        if (!DENSE_GRAD) {
            for (int e = tid; e < RB_BATCH * RB_KMAX; e += RB_THREADS) s_grad[e] = 0.0f;
            const int inst = tid >> 2;                          // 4 threads per instance
            if (inst < batch_n) {
                const uint32_t g = a.point_list[range.x + off + inst];
                const uint32_t cnt = a.sp_count[g];
                if (cnt != RB_ROW_DENSE)
                    for (uint32_t k = tid & 3; k < cnt; k += 4)
                        s_sidx[inst * RB_KMAX + k] = a.sp_index[(size_t)g * RB_KMAX + k];
            }
        }
        if (COLOR) {
            for (int e = tid; e < RB_BATCH * 4; e += RB_THREADS) s_gradc[e] = 0.0f;
        }
        __syncthreads();

        {
#pragma unroll 1
            for (int j = 0; j < batch_n; ++j) {
                const BackInstance& s = s_inst[j];
                // rows this warp owns against the splat's alpha >= 1/255 extent (rounded outwards by the forward)
                if (s.ext_y < 0.0f || s.y + s.ext_y < wy0 || s.y - s.ext_y > wy1) continue;
                float w = 0.0f;
                if (off + j < last) {
                    const float dx = s.x - pxf, dy = s.y - pyf;
                    const float power = fmaf(s.ax * dx, dx, fmaf(s.az * dy, dy, (s.b * dx) * dy));
                    if (!(power > 0.0f || power < s.thr)) {
                        const float alpha = fminf(ALPHA_MAX, s.op * repro_expf(power));
                        if (!(alpha < ALPHA_MIN)) {
                            w = alpha * T;
                            T = T * (1.0f - alpha);
                        }
                    }
                }
                if (__ballot_sync(0xffffffffu, w != 0.0f) == 0u) continue;
                const uint32_t cnt = s.cnt;
                if (DENSE_GRAD || cnt == RB_ROW_DENSE) {
                    // every channel (strict dense gradient, or a row too dense to compact)
                    for (int c = 0; c < a.L; ++c) {
                        // support-only gradient: channels whose coefficient is exactly 0 get none
                        if (!DENSE_GRAD && __ldg(a.coeff_dense + (size_t)s.gid * a.L + c) == 0.0f) continue;
                        const float v = warp_sum(w * s_dw[c * RB_THREADS + tid]);
                        if (lane == 0 && v != 0.0f) atomicAdd(a.dL_dcoeff + (size_t)s.gid * a.L + c, v);
                    }
                } else {
                    // four channels per pass: a transposing butterfly leaves channel k0 + q summed over the warp in
                    // lane 8q (6 shuffles for 4 sums instead of 20)
                    for (uint32_t k0 = 0; k0 < cnt; k0 += 4) {
                        float a0, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
                        a0 = w * s_dw[s_sidx[j * RB_KMAX + k0] * RB_THREADS + tid];
                        if (k0 + 1 < cnt) a1 = w * s_dw[s_sidx[j * RB_KMAX + k0 + 1] * RB_THREADS + tid];
                        if (k0 + 2 < cnt) a2 = w * s_dw[s_sidx[j * RB_KMAX + k0 + 2] * RB_THREADS + tid];
                        if (k0 + 3 < cnt) a3 = w * s_dw[s_sidx[j * RB_KMAX + k0 + 3] * RB_THREADS + tid];
                        const bool hi16 = lane & 16, hi8 = lane & 8;
                        float b0 = hi16 ? a2 : a0, b1 = hi16 ? a3 : a1;
                        b0 += __shfl_xor_sync(0xffffffffu, hi16 ? a0 : a2, 16);
                        b1 += __shfl_xor_sync(0xffffffffu, hi16 ? a1 : a3, 16);
                        float v = hi8 ? b1 : b0;
                        v += __shfl_xor_sync(0xffffffffu, hi8 ? b0 : b1, 8);
                        v += __shfl_xor_sync(0xffffffffu, v, 4);
                        v += __shfl_xor_sync(0xffffffffu, v, 2);
                        v += __shfl_xor_sync(0xffffffffu, v, 1);
                        const uint32_t kq = k0 + (lane >> 3);
                        if ((lane & 7) == 0 && kq < cnt) atomicAdd(&s_grad[j * RB_KMAX + kq], v);
                    }
                }
                if (COLOR) {
                    const float v0 = warp_sum(w * dc0), v1 = warp_sum(w * dc1), v2 = warp_sum(w * dc2);
                    if (lane == 0) {
                        atomicAdd(&s_gradc[j * 4 + 0], v0);
                        atomicAdd(&s_gradc[j * 4 + 1], v1);
                        atomicAdd(&s_gradc[j * 4 + 2], v2);
                    }
                }
            }
        }
        __syncthreads();
        // flush the batch: one global atomic per (instance, channel) of this tile
        if (!DENSE_GRAD) {
            for (int e = tid; e < batch_n * RB_KMAX; e += RB_THREADS) {
                const int j = e / RB_KMAX, k = e % RB_KMAX;
                const uint32_t cnt = s_inst[j].cnt;
                if (cnt != RB_ROW_DENSE && (uint32_t)k < cnt) {
                    const float v = s_grad[e];
                    if (v != 0.0f) {
                        if (a.row_stride > 0) {
                            if (k < a.row_stride) atomicAdd(a.dL_dcoeff + (size_t)s_inst[j].gid * a.row_stride + k, v);
                        } else {
                            atomicAdd(a.dL_dcoeff + (size_t)s_inst[j].gid * a.L + s_sidx[e], v);
                        }
                    }
                }
            }
        }
        if (COLOR) {
            for (int e = tid; e < batch_n * 4; e += RB_THREADS) {
                const int j = e >> 2, k = e & 3;
                if (k < 3) {
                    const float v = s_gradc[e];
                    if (v != 0.0f) atomicAdd(a.dL_dcolors + (size_t)s_inst[j].gid * 3 + k, v);
                }
            }
        }
    }
}

static size_t rb_smem_bytes(int L, bool color)
{
    return sizeof(BackInstance) * RB_BATCH + (size_t)RB_BATCH * RB_KMAX * (sizeof(uint16_t) + sizeof(float)) +
           (color ? (size_t)RB_BATCH * 4 * sizeof(float) : 0) + (size_t)L * RB_THREADS * sizeof(float);
}

int launch_render_backward(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b,
                           const lsv2_features& f, const lsv2_image& img, const float* dL_dwmap,
                           const float* dL_dcolor, int dense_grad, float* dL_dcoeff, float* dL_dcolors,
                           cudaStream_t st, int row_stride)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (f.mode != LSV2_MODE_DENSE || f.L <= 0 || !f.coeff_dense || !dL_dwmap || !dL_dcoeff || !img.n_contrib ||
        !g.blend_rec)
        return LSV2_EINVAL;
    const bool color = dL_dcolor != nullptr && dL_dcolors != nullptr;
    const bool have_sparse = f.sparse_count && f.sparse_index;
    const bool dense = dense_grad != 0 || !have_sparse;
    if (row_stride < 0 || (row_stride > 0 && dense)) return LSV2_EINVAL;      // the row form IS the support-only gradient
    const size_t smem = rb_smem_bytes(f.L, color);
    if (smem > 227 * 1024) return LSV2_EUNSUPPORTED;
    BackArgs a;
    a.W = v.width; a.H = v.height; a.gx = gx; a.L = f.L;
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.rec = reinterpret_cast<const float4*>(g.blend_rec);
    a.n_contrib = img.n_contrib;
    a.sp_count = have_sparse ? f.sparse_count : nullptr;
    a.sp_index = f.sparse_index;
    a.coeff_dense = f.coeff_dense;
    a.dL_dwmap = dL_dwmap; a.dL_dcolor = dL_dcolor; a.dL_dcoeff = dL_dcoeff; a.dL_dcolors = dL_dcolors;
    a.row_stride = row_stride;
#define RB_LAUNCH(D, C)                                                                                      \
    do {                                                                                                     \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_backward<D, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           (int)smem));                                                      \
        k_render_backward<D, C><<<gx * gy, RB_THREADS, smem, st>>>(a);                                       \
    } while (0)
    if (dense) { if (color) RB_LAUNCH(true, true); else RB_LAUNCH(true, false); }
    else       { if (color) RB_LAUNCH(false, true); else RB_LAUNCH(false, false); }
#undef RB_LAUNCH
    LSV2_LAUNCH_CHECK("k_render_backward");
    return LSV2_OK;
}

}  // namespace lsv2
