// preprocess.cu -- K1: per-Gaussian projection, EWA covariance, conic, radius, tile count,
// SH -> RGB.  Replaces preprocessCUDA (forward) of the reference's rasterizer submodule
// (call site /root/reference/gaussian_renderer/__init__.py:108-119).
//
// One thread per Gaussian, 128 threads per CTA.  Geometry first; the SH block (up to 192 B
// per Gaussian, the bulk of this kernel's bytes) is then staged through shared memory with
// fully coalesced row loads, and only for Gaussians that survived culling.
#include "common.cuh"

namespace lsv2 {

constexpr int PRE_THREADS = 128;
constexpr int SH_MAX_FLOATS = 48;              // degree 3: 16 coefficients x RGB
constexpr int SH_ROW_STRIDE = SH_MAX_FLOATS + 1;  // odd stride -> conflict-free per-thread rows

__constant__ float c_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__constant__ float c_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

struct PreArgs {
    int P, W, H, gx, gy, deg, M;
    float tanfovx, tanfovy, fx, fy, mod;
    const float *vm, *pm, *campos;
    const float *means, *opac, *scales, *rots, *cov3d_pre, *shs, *colors;
    int32_t* radii;
    float2* xy;
    float* depth;
    float4* conic_opacity;
    float* rgb;
    uint8_t* clamped;
    uint32_t* tiles;
    float* cov3d;
    int32_t* rect;
    float4* blend_rec;
};

__global__ void __launch_bounds__(PRE_THREADS) k_preprocess(PreArgs a)
{
    __shared__ float s_m[32];                     // viewmatrix[16] | projmatrix[16]
    __shared__ float s_cam[3];
    __shared__ float s_sh[PRE_THREADS * SH_ROW_STRIDE];
    __shared__ int s_vis[PRE_THREADS];

    const int tid = threadIdx.x;
    if (tid < 16) s_m[tid] = a.vm[tid];
    else if (tid < 32) s_m[tid] = a.pm[tid - 16];
    else if (tid < 35) s_cam[tid - 32] = a.campos[tid - 32];
    __syncthreads();

    const int i = blockIdx.x * PRE_THREADS + tid;
    const float* vm = s_m;
    const float* pm = s_m + 16;

    bool vis = false;
    float px = 0.f, py = 0.f, zv = 0.f, rad = 0.f;
    float con0 = 0.f, con1 = 0.f, con2 = 0.f, op = 0.f;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f;
    float c6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int rminx = 0, rminy = 0, rmaxx = 0, rmaxy = 0;

    if (i < a.P) {
        p0 = a.means[3 * (size_t)i + 0];
        p1 = a.means[3 * (size_t)i + 1];
        p2 = a.means[3 * (size_t)i + 2];
        const float vx = ((vm[0] * p0 + vm[4] * p1) + vm[8] * p2) + vm[12];
        const float vy = ((vm[1] * p0 + vm[5] * p1) + vm[9] * p2) + vm[13];
        const float vz = ((vm[2] * p0 + vm[6] * p1) + vm[10] * p2) + vm[14];
        if (vz > NEAR_Z) {
            const float hx = ((pm[0] * p0 + pm[4] * p1) + pm[8] * p2) + pm[12];
            const float hy = ((pm[1] * p0 + pm[5] * p1) + pm[9] * p2) + pm[13];
            const float hw = ((pm[3] * p0 + pm[7] * p1) + pm[11] * p2) + pm[15];
            const float pw = 1.0f / (hw + W_GUARD);
            const float ppx = hx * pw, ppy = hy * pw;

            if (a.cov3d_pre) {
#pragma unroll
                for (int k = 0; k < 6; ++k) c6[k] = a.cov3d_pre[6 * (size_t)i + k];
            } else {
                const float4 q = reinterpret_cast<const float4*>(a.rots)[i];
                const float r = q.x, x = q.y, y = q.z, z = q.w;
                float R[3][3];
                R[0][0] = 1.f - 2.f * (y * y + z * z);
                R[0][1] = 2.f * (x * y - r * z);
                R[0][2] = 2.f * (x * z + r * y);
                R[1][0] = 2.f * (x * y + r * z);
                R[1][1] = 1.f - 2.f * (x * x + z * z);
                R[1][2] = 2.f * (y * z - r * x);
                R[2][0] = 2.f * (x * z - r * y);
                R[2][1] = 2.f * (y * z + r * x);
                R[2][2] = 1.f - 2.f * (x * x + y * y);
                float s[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) s[k] = a.mod * a.scales[3 * (size_t)i + k];
                float Lm[3][3];
#pragma unroll
                for (int r_ = 0; r_ < 3; ++r_)
#pragma unroll
                    for (int k = 0; k < 3; ++k) Lm[r_][k] = R[r_][k] * s[k];
                int o = 0;
#pragma unroll
                for (int r_ = 0; r_ < 3; ++r_)
#pragma unroll
                    for (int c_ = r_; c_ < 3; ++c_)
                        c6[o++] = (Lm[r_][0] * Lm[c_][0] + Lm[r_][1] * Lm[c_][1]) + Lm[r_][2] * Lm[c_][2];
            }

            // EWA projection
            const float limx = FRUSTUM_CLAMP * a.tanfovx, limy = FRUSTUM_CLAMP * a.tanfovy;
            const float txtz = vx / vz, tytz = vy / vz;
            const float tx = fminf(limx, fmaxf(-limx, txtz)) * vz;
            const float ty = fminf(limy, fmaxf(-limy, tytz)) * vz;
            const float J00 = a.fx / vz;
            const float J02 = -(a.fx * tx) / (vz * vz);
            const float J11 = a.fy / vz;
            const float J12 = -(a.fy * ty) / (vz * vz);
            float M0[3], M1[3];
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                M0[j] = J00 * vm[4 * j + 0] + J02 * vm[4 * j + 2];
                M1[j] = J11 * vm[4 * j + 1] + J12 * vm[4 * j + 2];
            }
            const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
            float A0[3], A1[3];
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                A0[j] = (M0[0] * S[0][j] + M0[1] * S[1][j]) + M0[2] * S[2][j];
                A1[j] = (M1[0] * S[0][j] + M1[1] * S[1][j]) + M1[2] * S[2][j];
            }
            const float cv0 = ((A0[0] * M0[0] + A0[1] * M0[1]) + A0[2] * M0[2]) + LOWPASS;
            const float cv1 = (A0[0] * M1[0] + A0[1] * M1[1]) + A0[2] * M1[2];
            const float cv2 = ((A1[0] * M1[0] + A1[1] * M1[1]) + A1[2] * M1[2]) + LOWPASS;

            const float det = cv0 * cv2 - cv1 * cv1;
            if (det != 0.0f) {
                const float det_inv = 1.0f / det;
                con0 = cv2 * det_inv;
                con1 = -cv1 * det_inv;
                con2 = cv0 * det_inv;
                const float mid = 0.5f * (cv0 + cv2);
                const float sq = sqrtf(fmaxf(EIG_FLOOR, mid * mid - det));
                const float l1 = mid + sq, l2 = mid - sq;
                rad = ceilf(RADIUS_SIGMA * sqrtf(fmaxf(l1, l2)));
                px = ((ppx + 1.0f) * (float)a.W - 1.0f) * 0.5f;
                py = ((ppy + 1.0f) * (float)a.H - 1.0f) * 0.5f;
                tile_rect(px, py, rad, a.gx, a.gy, rminx, rminy, rmaxx, rmaxy);
                if ((rmaxx - rminx) * (rmaxy - rminy) > 0) {
                    vis = true;
                    zv = vz;
                    op = a.opac[i];
                }
            }
        }
    }

    // ---- colour ------------------------------------------------------------------------
    float rgb[3] = {0.f, 0.f, 0.f};
    uint8_t clampd[3] = {0, 0, 0};
    if (a.shs) {
        const int row = a.M * 3;                            // floats per Gaussian in memory
        const int ncoef = (a.deg + 1) * (a.deg + 1);
        const int used = min(row, ncoef * 3);               // floats the active degree reads
        s_vis[tid] = vis ? 1 : 0;
        __syncthreads();
        const int warp = tid >> 5, lane = tid & 31;
        const size_t base = (size_t)blockIdx.x * PRE_THREADS;
        if ((row & 3) == 0 && (reinterpret_cast<uintptr_t>(a.shs) & 15) == 0) {
            // the CTA's rows are one contiguous block: 16-byte loads, independent of each other, skipping culled rows
            const int f4_per_row = row >> 2;
            const int total = min(PRE_THREADS, a.P - (int)base) * f4_per_row;
            const float4* src4 = reinterpret_cast<const float4*>(a.shs + base * row);
#pragma unroll 4
            for (int e = tid; e < total; e += PRE_THREADS) {
                const int r = e / f4_per_row, q = e - r * f4_per_row;
                if (!s_vis[r]) continue;
                const float4 v = __ldg(src4 + e);
                float* d = s_sh + r * SH_ROW_STRIDE + 4 * q;
                d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
            }
        } else {
            for (int r = warp; r < PRE_THREADS; r += PRE_THREADS / 32) {
                if (!s_vis[r]) continue;
                const float* src = a.shs + (base + r) * row;
                for (int k = lane; k < used; k += 32) s_sh[r * SH_ROW_STRIDE + k] = __ldg(src + k);
            }
        }
        __syncthreads();
        if (vis) {
            const float* sh = s_sh + tid * SH_ROW_STRIDE;
            const float d0 = p0 - s_cam[0], d1 = p1 - s_cam[1], d2 = p2 - s_cam[2];
            const float len = sqrtf((d0 * d0 + d1 * d1) + d2 * d2);
            const float x = d0 / len, y = d1 / len, z = d2 / len;
            const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
#define SH(k) sh[(k) * 3 + c]
                float res = SH_C0 * SH(0);
                if (a.deg > 0) {
                    res = ((res - (SH_C1 * y) * SH(1)) + (SH_C1 * z) * SH(2)) - (SH_C1 * x) * SH(3);
                    if (a.deg > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z;
                        const float xy = x * y, yz = y * z, xz = x * z;
                        res = ((((res + (c_SH_C2[0] * xy) * SH(4)) + (c_SH_C2[1] * yz) * SH(5)) +
                                (c_SH_C2[2] * ((2.0f * zz - xx) - yy)) * SH(6)) +
                               (c_SH_C2[3] * xz) * SH(7)) +
                              (c_SH_C2[4] * (xx - yy)) * SH(8);
                        if (a.deg > 2) {
                            res = ((((((res + ((c_SH_C3[0] * y) * (3.0f * xx - yy)) * SH(9)) +
                                       ((c_SH_C3[1] * xy) * z) * SH(10)) +
                                      ((c_SH_C3[2] * y) * ((4.0f * zz - xx) - yy)) * SH(11)) +
                                     ((c_SH_C3[3] * z) * ((2.0f * zz - 3.0f * xx) - 3.0f * yy)) * SH(12)) +
                                    ((c_SH_C3[4] * x) * ((4.0f * zz - xx) - yy)) * SH(13)) +
                                   ((c_SH_C3[5] * z) * (xx - yy)) * SH(14)) +
                                  ((c_SH_C3[6] * x) * (xx - 3.0f * yy)) * SH(15);
                        }
                    }
                }
#undef SH
                res = res + 0.5f;
                clampd[c] = (uint8_t)(res < 0.0f);
                rgb[c] = fmaxf(res, 0.0f);
            }
        }
    } else if (vis) {
#pragma unroll
        for (int c = 0; c < 3; ++c) rgb[c] = a.colors[3 * (size_t)i + c];
    }

    if (i >= a.P) return;
    // culled Gaussians leave an all-zero record (radii == 0 is the visibility filter the
    // callers read: gaussian_renderer/__init__.py:127)
    a.radii[i] = vis ? (int32_t)rad : 0;
    a.xy[i] = vis ? make_float2(px, py) : make_float2(0.f, 0.f);
    a.depth[i] = zv;
    a.conic_opacity[i] = vis ? make_float4(con0, con1, con2, op) : make_float4(0.f, 0.f, 0.f, 0.f);
    a.rgb[3 * (size_t)i + 0] = rgb[0];
    a.rgb[3 * (size_t)i + 1] = rgb[1];
    a.rgb[3 * (size_t)i + 2] = rgb[2];
    if (a.clamped) {
        a.clamped[3 * (size_t)i + 0] = clampd[0];
        a.clamped[3 * (size_t)i + 1] = clampd[1];
        a.clamped[3 * (size_t)i + 2] = clampd[2];
    }
    a.tiles[i] = vis ? (uint32_t)((rmaxx - rminx) * (rmaxy - rminy)) : 0u;
    if (a.cov3d) {
        // the checker zeroes cov3d for Gaussians culled at the near plane only
#pragma unroll
        for (int k = 0; k < 6; ++k) a.cov3d[6 * (size_t)i + k] = c6[k];
    }
    if (a.blend_rec) {
        float4* rec = a.blend_rec + 3 * (size_t)i;
        if (vis) {
            rec[0] = make_float4(px, py, -0.5f * con0, -con1);
            const float thr = power_cutoff(op);
            float ext_y, ext_x;
            cutoff_extents(-0.5f * con0, -con1, -0.5f * con2, thr, ext_y, ext_x);
            rec[1] = make_float4(-0.5f * con2, thr, op, rgb[0]);
            rec[2] = make_float4(rgb[1], rgb[2], ext_y, ext_x);
        } else {
            rec[0] = make_float4(0.f, 0.f, 0.f, 0.f);
            rec[1] = make_float4(0.f, 1.0f, 0.f, 0.f);
            rec[2] = make_float4(0.f, 0.f, -1.f, 0.f);
        }
    }
    if (a.rect) {
        reinterpret_cast<int4*>(a.rect)[i] =
            vis ? make_int4(rminx, rminy, rmaxx, rmaxy) : make_int4(0, 0, 0, 0);
    }
}

int launch_preprocess(const lsv2_view& v, const lsv2_scene& s, const lsv2_geom& g, cudaStream_t st)
{
    if (s.P <= 0) return LSV2_OK;
    if ((s.shs == nullptr) == (s.colors_precomp == nullptr)) return LSV2_EINVAL;
    const bool has_sr = s.scales != nullptr && s.rotations != nullptr;
    if (has_sr == (s.cov3d_precomp != nullptr)) return LSV2_EINVAL;
    if (s.shs && (s.sh_coeffs < 1 || s.sh_coeffs * 3 > SH_MAX_FLOATS)) return LSV2_EUNSUPPORTED;
    if (v.sh_degree < 0 || v.sh_degree > 3) return LSV2_EINVAL;
    if (s.shs && (v.sh_degree + 1) * (v.sh_degree + 1) > s.sh_coeffs) return LSV2_EINVAL;
    PreArgs a;
    a.P = s.P; a.W = v.width; a.H = v.height;
    a.gx = (v.width + TILE - 1) / TILE; a.gy = (v.height + TILE - 1) / TILE;
    a.deg = v.sh_degree; a.M = s.sh_coeffs;
    a.tanfovx = v.tanfovx; a.tanfovy = v.tanfovy;
    a.fx = (float)v.width / (2.0f * v.tanfovx);
    a.fy = (float)v.height / (2.0f * v.tanfovy);
    a.mod = v.scale_modifier;
    a.vm = v.viewmatrix; a.pm = v.projmatrix; a.campos = v.campos;
    a.means = s.means3d; a.opac = s.opacities; a.scales = s.scales; a.rots = s.rotations;
    a.cov3d_pre = s.cov3d_precomp; a.shs = s.shs; a.colors = s.colors_precomp;
    a.radii = g.radii; a.xy = reinterpret_cast<float2*>(g.xy); a.depth = g.depth;
    a.conic_opacity = reinterpret_cast<float4*>(g.conic_opacity); a.rgb = g.rgb;
    a.clamped = g.clamped; a.tiles = g.tiles_touched; a.cov3d = g.cov3d; a.rect = g.rect;
    a.blend_rec = reinterpret_cast<float4*>(g.blend_rec);
    k_preprocess<<<ceil_div(s.P, PRE_THREADS), PRE_THREADS, 0, st>>>(a);
    LSV2_LAUNCH_CHECK("k_preprocess");
    return LSV2_OK;
}

// ---- markVisible ---------------------------------------------------------------------
__global__ void k_mark_visible(int P, const float* __restrict__ means, const float* __restrict__ vm,
                               uint8_t* __restrict__ present)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float p0 = means[3 * (size_t)i], p1 = means[3 * (size_t)i + 1], p2 = means[3 * (size_t)i + 2];
    const float vz = ((vm[2] * p0 + vm[6] * p1) + vm[10] * p2) + vm[14];
    present[i] = vz > NEAR_Z ? 1 : 0;
}

int launch_mark_visible(int P, const float* means3d, const float* viewmatrix, uint8_t* present, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    k_mark_visible<<<ceil_div(P, 256), 256, 0, st>>>(P, means3d, viewmatrix, present);
    LSV2_LAUNCH_CHECK("k_mark_visible");
    return LSV2_OK;
}

}  // namespace lsv2
