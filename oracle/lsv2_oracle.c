/*
 * lsv2_oracle.c -- CPU restatement of the LangSplatV2 rasterization hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (langsplatv2_b200/,
 * diff_gaussian_rasterization/) may import, link or execute this file.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, and there only as the checker / the timed CPU arm.
 *
 * PARITY STATUS
 *   - The rasterizer core (preprocess, binning, blend, blend-backward; SURVEY.md
 *     rows R6, R7, R8, R10) lives in the git submodule
 *       submodules/efficient-langsplat-rasterization @ 1305b7c23ff64092cd6c4b0e4996d6557e56c76a
 *     (python package `diff_gaussian_rasterization`), which is ABSENT from
 *     /root/reference (.SUBMODULES.json:9-15).  The reference holds no test,
 *     golden vector or fixture for this path.  Those rows restate the published
 *     3DGS tile-rasterizer algorithm that the fork chain (Inria
 *     diff-gaussian-rasterization -> LangSplat -> LangSplatV2) shares, anchored
 *     on the reference's call site gaussian_renderer/__init__.py:37-54,108-119.
 *     ==> "parity unpinned" for R6-R8, R10.
 *   - The parts whose reference source IS present are pinned by golden vectors
 *     generated from the reference's own Python (tests/golden/make_golden.py):
 *       sparse coefficients  utils/vq_utils.py:9-40
 *       codebook decode      eval_lerf.py:214-218, scene/gaussian_model.py:533-549
 *       camera matrices      utils/graphics_utils.py:38-71, scene/cameras.py:55-58
 *       SH -> RGB            utils/sh_utils.py:57-112, gaussian_renderer/__init__.py:77-81
 *       cov3D                utils/general_utils.py:78-110, scene/gaussian_model.py:28-32
 *       cos / l1 loss        utils/loss_utils.py:18-25
 *
 * NUMERIC POLICY (shared with the CUDA path so integer outputs can be bit-exact)
 *   Compile with -ffp-contract=off.  Every multiply/add below is a separately
 *   rounded IEEE binary32 op unless fmaf() is written explicitly.  exp() in the
 *   blend is the fixed polynomial lsv2o_expf() (<= 2 ulp of libm on [-87,0]); libm
 *   and the GPU's MUFU-based expf differ in the last bits, which would make
 *   n_contrib (an integer output that depends on alpha thresholds) irreproducible.
 *
 * All unverifiable constants of the rasterizer core are collected here:
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define LSV2O_TILE            16        /* BLOCK_X = BLOCK_Y                       */
#define LSV2O_NEAR_Z          0.2f      /* frustum cull: p_view.z <= 0.2 dropped   */
#define LSV2O_W_GUARD         0.0000001f/* 1/(p_hom.w + 1e-7)                      */
#define LSV2O_FRUSTUM_CLAMP   1.3f      /* x/z, y/z clamp = 1.3 * tanfov           */
#define LSV2O_LOWPASS         0.3f      /* cov2D diag += 0.3 px^2                  */
#define LSV2O_EIG_FLOOR       0.1f      /* sqrt(max(0.1, mid^2-det))               */
#define LSV2O_RADIUS_SIGMA    3.0f      /* radius = ceil(3 sqrt(lambda_max))       */
#define LSV2O_ALPHA_MAX       0.99f
#define LSV2O_ALPHA_MIN       (1.0f / 255.0f)
#define LSV2O_T_MIN           0.0001f

/* ------------------------------------------------------------------------- */
/* exp for x <= 0: fixed-order polynomial, reproducible on CPU and GPU.       */
/* ------------------------------------------------------------------------- */
float lsv2o_expf(float x)
{
    if (!(x >= -87.0f)) return 0.0f;               /* also catches NaN           */
    if (x > 0.0f) x = 0.0f;                        /* callers never pass x > 0   */
    const float t = x * 1.44269504088896341f;
    const float n = (t + 12582912.0f) - 12582912.0f;   /* rint(t), RN-even       */
    float r = fmaf(n, -0.693145751953125f, x);
    r = fmaf(n, -1.42860682030941723212e-6f, r);
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    const float r2 = r * r;
    float y = fmaf(p, r2, r);
    y = y + 1.0f;
    union { float f; int32_t i; } u;
    u.f = y;
    u.i += ((int32_t)n) << 23;                      /* y * 2^n, n in [-126,0]    */
    return u.f;
}

void lsv2o_expf_array(int n, const float* x, float* y)
{
    for (int i = 0; i < n; ++i) y[i] = lsv2o_expf(x[i]);
}

/* ------------------------------------------------------------------------- */
/* R6  preprocess (3DGS preprocessCUDA forward)  [LINEAGE]                    */
/* matrices arrive as the reference passes them: world_view_transform and     */
/* full_proj_transform are the TRANSPOSED (row-vector) matrices               */
/* (scene/cameras.py:55-57), so m[4*c + r] is element (r,c) of the usual      */
/* column-vector matrix.                                                      */
/* ------------------------------------------------------------------------- */
static inline void xform4x3(const float* m, const float* p, float* o)
{
    o[0] = ((m[0] * p[0] + m[4] * p[1]) + m[8] * p[2]) + m[12];
    o[1] = ((m[1] * p[0] + m[5] * p[1]) + m[9] * p[2]) + m[13];
    o[2] = ((m[2] * p[0] + m[6] * p[1]) + m[10] * p[2]) + m[14];
}
static inline void xform4x4(const float* m, const float* p, float* o)
{
    o[0] = ((m[0] * p[0] + m[4] * p[1]) + m[8] * p[2]) + m[12];
    o[1] = ((m[1] * p[0] + m[5] * p[1]) + m[9] * p[2]) + m[13];
    o[2] = ((m[2] * p[0] + m[6] * p[1]) + m[10] * p[2]) + m[14];
    o[3] = ((m[3] * p[0] + m[7] * p[1]) + m[11] * p[2]) + m[15];
}

/* cov3D = R diag(mod*s)^2 R^T, R from quaternion (w,x,y,z) used AS GIVEN (the
 * reference normalises in Python: scene/gaussian_model.py:37 rotation_activation;
 * utils/general_utils.py:78-110 is the same R).  Stored as the 6 upper-triangle
 * entries (xx,xy,xz,yy,yz,zz) like strip_lowerdiag (general_utils.py:62-72). */
static void cov3d_from_scale_rot(const float* s, float mod, const float* q, float* c6)
{
    const float r = q[0], x = q[1], y = q[2], z = q[3];
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
    float L[3][3];
    for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k) L[i][k] = R[i][k] * (mod * s[k]);
    int o = 0;
    for (int i = 0; i < 3; ++i)
        for (int j = i; j < 3; ++j)
            c6[o++] = (L[i][0] * L[j][0] + L[i][1] * L[j][1]) + L[i][2] * L[j][2];
}

/* EWA projection: cov2D = J Rv Sigma Rv^T J^T  (+0.3 on the diagonal). */
static void cov2d_project(const float* t_in, float fx, float fy, float tanfovx, float tanfovy,
                          const float* c6, const float* vm, float* cov3)
{
    float t[3] = { t_in[0], t_in[1], t_in[2] };
    const float limx = LSV2O_FRUSTUM_CLAMP * tanfovx;
    const float limy = LSV2O_FRUSTUM_CLAMP * tanfovy;
    const float txtz = t[0] / t[2];
    const float tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    const float J00 = fx / t[2];
    const float J02 = -(fx * t[0]) / (t[2] * t[2]);
    const float J11 = fy / t[2];
    const float J12 = -(fy * t[1]) / (t[2] * t[2]);
    /* Rv[i][j] = W2C[i][j] = vm[4*j + i] */
    float M0[3], M1[3];
    for (int j = 0; j < 3; ++j) {
        M0[j] = J00 * vm[4 * j + 0] + J02 * vm[4 * j + 2];
        M1[j] = J11 * vm[4 * j + 1] + J12 * vm[4 * j + 2];
    }
    const float S[3][3] = { { c6[0], c6[1], c6[2] }, { c6[1], c6[3], c6[4] }, { c6[2], c6[4], c6[5] } };
    float A0[3], A1[3];
    for (int j = 0; j < 3; ++j) {
        A0[j] = (M0[0] * S[0][j] + M0[1] * S[1][j]) + M0[2] * S[2][j];
        A1[j] = (M1[0] * S[0][j] + M1[1] * S[1][j]) + M1[2] * S[2][j];
    }
    cov3[0] = ((A0[0] * M0[0] + A0[1] * M0[1]) + A0[2] * M0[2]) + LSV2O_LOWPASS;
    cov3[1] = (A0[0] * M1[0] + A0[1] * M1[1]) + A0[2] * M1[2];
    cov3[2] = ((A1[0] * M1[0] + A1[1] * M1[1]) + A1[2] * M1[2]) + LSV2O_LOWPASS;
}

/* SH -> RGB, utils/sh_utils.py:57-112 op order, then +0.5 and clamp >= 0
 * (gaussian_renderer/__init__.py:77-81).  sh: [M][3] per Gaussian. */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f };
static const float SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                -0.5900435899266435f };

static void sh_to_rgb(int deg, const float* sh, const float* pos, const float* campos,
                      float* rgb, uint8_t* clamped)
{
    float d[3] = { pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2] };
    const float len = sqrtf((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]);
    const float x = d[0] / len, y = d[1] / len, z = d[2] / len;
    for (int c = 0; c < 3; ++c) {
#define SH(i) sh[(i) * 3 + c]
        float res = SH_C0 * SH(0);
        if (deg > 0) {
            res = ((res - (SH_C1 * y) * SH(1)) + (SH_C1 * z) * SH(2)) - (SH_C1 * x) * SH(3);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z;
                const float xy = x * y, yz = y * z, xz = x * z;
                res = ((((res + (SH_C2[0] * xy) * SH(4)) + (SH_C2[1] * yz) * SH(5)) +
                        (SH_C2[2] * ((2.0f * zz - xx) - yy)) * SH(6)) +
                       (SH_C2[3] * xz) * SH(7)) +
                      (SH_C2[4] * (xx - yy)) * SH(8);
                if (deg > 2) {
                    res = ((((((res + ((SH_C3[0] * y) * (3.0f * xx - yy)) * SH(9)) +
                               ((SH_C3[1] * xy) * z) * SH(10)) +
                              ((SH_C3[2] * y) * ((4.0f * zz - xx) - yy)) * SH(11)) +
                             ((SH_C3[3] * z) * ((2.0f * zz - 3.0f * xx) - 3.0f * yy)) * SH(12)) +
                            ((SH_C3[4] * x) * ((4.0f * zz - xx) - yy)) * SH(13)) +
                           ((SH_C3[5] * z) * (xx - yy)) * SH(14)) +
                          ((SH_C3[6] * x) * (xx - 3.0f * yy)) * SH(15);
                }
            }
        }
#undef SH
        res = res + 0.5f;
        clamped[c] = (uint8_t)(res < 0.0f);
        rgb[c] = fmaxf(res, 0.0f);
    }
}

/*
 * Outputs (all length P unless noted):
 *   radii i32, xy f32[2], depth f32, cov3d f32[6], conic_opacity f32[4],
 *   rgb f32[3], clamped u8[3], tiles_touched u32, rect i32[4]=(minx,miny,maxx,maxy)
 * shs: [P][M][3] or NULL; colors_precomp [P][3] or NULL; scales [P][3], rotations
 * [P][4] or cov3d_precomp [P][6].
 */
int lsv2o_preprocess(int P, int deg, int M,
                     const float* means3d, const float* scales, float scale_modifier,
                     const float* rotations, const float* opacities, const float* shs,
                     const float* colors_precomp, const float* cov3d_precomp,
                     const float* viewmatrix, const float* projmatrix, const float* campos,
                     int W, int H, float tanfovx, float tanfovy,
                     int32_t* radii, float* xy, float* depth, float* cov3d, float* conic_opacity,
                     float* rgb, uint8_t* clamped, uint32_t* tiles_touched, int32_t* rect)
{
    const float fx = (float)W / (2.0f * tanfovx);
    const float fy = (float)H / (2.0f * tanfovy);
    const int gx = (W + LSV2O_TILE - 1) / LSV2O_TILE, gy = (H + LSV2O_TILE - 1) / LSV2O_TILE;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        xy[2 * i] = xy[2 * i + 1] = 0.f;
        depth[i] = 0.f;
        for (int k = 0; k < 4; ++k) { conic_opacity[4 * i + k] = 0.f; rect[4 * i + k] = 0; }
        for (int k = 0; k < 3; ++k) { rgb[3 * i + k] = 0.f; clamped[3 * i + k] = 0; }
        for (int k = 0; k < 6; ++k) cov3d[6 * i + k] = 0.f;

        const float* p = means3d + 3 * i;
        float pv[3];
        xform4x3(viewmatrix, p, pv);
        if (pv[2] <= LSV2O_NEAR_Z) continue;
        float ph[4];
        xform4x4(projmatrix, p, ph);
        const float pw = 1.0f / (ph[3] + LSV2O_W_GUARD);
        const float pp[2] = { ph[0] * pw, ph[1] * pw };

        float c6[6];
        if (cov3d_precomp) memcpy(c6, cov3d_precomp + 6 * i, sizeof c6);
        else cov3d_from_scale_rot(scales + 3 * i, scale_modifier, rotations + 4 * i, c6);
        memcpy(cov3d + 6 * i, c6, sizeof c6);

        float cov[3];
        cov2d_project(pv, fx, fy, tanfovx, tanfovy, c6, viewmatrix, cov);
        const float det = cov[0] * cov[2] - cov[1] * cov[1];
        if (det == 0.0f) continue;
        const float det_inv = 1.0f / det;
        const float conic[3] = { cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv };
        const float mid = 0.5f * (cov[0] + cov[2]);
        const float sq = sqrtf(fmaxf(LSV2O_EIG_FLOOR, mid * mid - det));
        const float l1 = mid + sq, l2 = mid - sq;
        const float rad = ceilf(LSV2O_RADIUS_SIGMA * sqrtf(fmaxf(l1, l2)));
        const float px = ((pp[0] + 1.0f) * (float)W - 1.0f) * 0.5f;
        const float py = ((pp[1] + 1.0f) * (float)H - 1.0f) * 0.5f;
        int rminx = (int)((px - rad) / (float)LSV2O_TILE);
        int rminy = (int)((py - rad) / (float)LSV2O_TILE);
        int rmaxx = (int)((px + rad + (float)(LSV2O_TILE - 1)) / (float)LSV2O_TILE);
        int rmaxy = (int)((py + rad + (float)(LSV2O_TILE - 1)) / (float)LSV2O_TILE);
        rminx = rminx < 0 ? 0 : (rminx > gx ? gx : rminx);
        rminy = rminy < 0 ? 0 : (rminy > gy ? gy : rminy);
        rmaxx = rmaxx < 0 ? 0 : (rmaxx > gx ? gx : rmaxx);
        rmaxy = rmaxy < 0 ? 0 : (rmaxy > gy ? gy : rmaxy);
        if ((rmaxx - rminx) * (rmaxy - rminy) <= 0) continue;

        if (colors_precomp) {
            for (int k = 0; k < 3; ++k) rgb[3 * i + k] = colors_precomp[3 * i + k];
        } else {
            sh_to_rgb(deg, shs + (size_t)i * M * 3, p, campos, rgb + 3 * i, clamped + 3 * i);
        }
        depth[i] = pv[2];
        radii[i] = (int32_t)rad;
        xy[2 * i] = px;
        xy[2 * i + 1] = py;
        conic_opacity[4 * i + 0] = conic[0];
        conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2];
        conic_opacity[4 * i + 3] = opacities[i];
        tiles_touched[i] = (uint32_t)((rmaxx - rminx) * (rmaxy - rminy));
        rect[4 * i + 0] = rminx; rect[4 * i + 1] = rminy; rect[4 * i + 2] = rmaxx; rect[4 * i + 3] = rmaxy;
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* R7  binning: inclusive scan, duplicateWithKeys, stable sort on             */
/* key = (tile_id << 32) | float_bits(depth), identifyTileRanges  [LINEAGE]   */
/* ------------------------------------------------------------------------- */
int64_t lsv2o_scan(int P, const uint32_t* tiles_touched, uint32_t* offsets /* inclusive */)
{
    uint64_t acc = 0;
    for (int i = 0; i < P; ++i) { acc += tiles_touched[i]; offsets[i] = (uint32_t)acc; }
    return (int64_t)acc;
}

static void radix_sort_pairs(uint64_t n, uint64_t* k0, uint32_t* v0, uint64_t* k1, uint32_t* v1)
{
    /* stable LSD radix, 4 passes of 16 bits; result ends in (k0, v0). */
    uint64_t* cnt = (uint64_t*)malloc(65536 * sizeof(uint64_t));
    for (int pass = 0; pass < 4; ++pass) {
        const int sh = 16 * pass;
        memset(cnt, 0, 65536 * sizeof(uint64_t));
        for (uint64_t i = 0; i < n; ++i) cnt[(k0[i] >> sh) & 0xffff]++;
        uint64_t run = 0;
        for (int d = 0; d < 65536; ++d) { uint64_t c = cnt[d]; cnt[d] = run; run += c; }
        for (uint64_t i = 0; i < n; ++i) {
            const uint64_t dst = cnt[(k0[i] >> sh) & 0xffff]++;
            k1[dst] = k0[i]; v1[dst] = v0[i];
        }
        uint64_t* tk = k0; k0 = k1; k1 = tk;
        uint32_t* tv = v0; v0 = v1; v1 = tv;
    }
    free(cnt);
}

/*
 * keys_unsorted/vals_unsorted: emission order (Gaussian index ascending, tiles
 * y-major then x inside one Gaussian).  keys/vals: sorted.  ranges: [tiles][2].
 * scratch_k/scratch_v: R-sized temporaries.
 */
int lsv2o_bin(int P, int W, int H, const int32_t* radii, const int32_t* rect, const float* depth,
              const uint32_t* offsets, int64_t R,
              uint64_t* keys_unsorted, uint32_t* vals_unsorted,
              uint64_t* keys, uint32_t* vals, uint64_t* scratch_k, uint32_t* scratch_v,
              uint32_t* ranges)
{
    const int gx = (W + LSV2O_TILE - 1) / LSV2O_TILE, gy = (H + LSV2O_TILE - 1) / LSV2O_TILE;
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        uint64_t off = (i == 0) ? 0 : offsets[i - 1];
        union { float f; uint32_t u; } d; d.f = depth[i];
        for (int y = rect[4 * i + 1]; y < rect[4 * i + 3]; ++y)
            for (int x = rect[4 * i + 0]; x < rect[4 * i + 2]; ++x) {
                uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
                key = (key << 32) | d.u;
                keys_unsorted[off] = key;
                vals_unsorted[off] = (uint32_t)i;
                ++off;
            }
    }
    memcpy(keys, keys_unsorted, (size_t)R * sizeof(uint64_t));
    memcpy(vals, vals_unsorted, (size_t)R * sizeof(uint32_t));
    radix_sort_pairs((uint64_t)R, keys, vals, scratch_k, scratch_v);
    memset(ranges, 0, (size_t)gx * gy * 2 * sizeof(uint32_t));
    for (int64_t i = 0; i < R; ++i) {
        const uint32_t t = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * t] = 0;
        else {
            const uint32_t pt = (uint32_t)(keys[i - 1] >> 32);
            if (t != pt) { ranges[2 * pt + 1] = (uint32_t)i; ranges[2 * t] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * t + 1] = (uint32_t)R;
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* R8  render forward: front-to-back alpha blend per pixel with sparse        */
/* coefficient accumulation.  [LINEAGE]                                       */
/*   mode 0: RGB only                                                         */
/*   mode 1: include_feature, dense coeff f32[P][L]                            */
/*   mode 2: quick_render, weights f32[P][K] + indices f32[P][K] (float-typed  */
/*           integers in [0,L), eval_lerf.py:333-348)                         */
/* Weight map has NO background term; rgb gets + T*bg.                         */
/* ------------------------------------------------------------------------- */
static inline float blend_power(const float* xy, const float* co, float px, float py,
                                float* dx_out, float* dy_out)
{
    const float dx = xy[0] - px, dy = xy[1] - py;
    const float ax = -0.5f * co[0], az = -0.5f * co[2], b = -co[1];
    *dx_out = dx; *dy_out = dy;
    return fmaf(ax * dx, dx, fmaf(az * dy, dy, (b * dx) * dy));
}

static int render_forward_impl(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                         const float* xy, const float* conic_opacity, const float* rgb,
                         int mode, int L, int K,
                         const float* coeff_dense, const float* wq, const float* iq,
                         const float* bg,
                         float* out_color /*[3][H][W]*/, float* out_wmap /*[L][H][W]*/,
                         float* final_T /*[H][W]*/, uint32_t* n_contrib /*[H][W]*/,
                         int tile_stride, int tile_phase)
{
    const int gx = (W + LSV2O_TILE - 1) / LSV2O_TILE, gy = (H + LSV2O_TILE - 1) / LSV2O_TILE;
    const size_t HW = (size_t)H * W;
#pragma omp parallel
    {
        float* acc = (float*)malloc(sizeof(float) * (size_t)(L > 0 ? L : 1));
#pragma omp for schedule(dynamic, 1)
        for (int tile = 0; tile < gx * gy; ++tile) {
            if (tile % tile_stride != tile_phase) continue;
            const int tx = tile % gx, ty = tile / gx;
            const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
            for (int ly = 0; ly < LSV2O_TILE; ++ly)
                for (int lx = 0; lx < LSV2O_TILE; ++lx) {
                    const int x = tx * LSV2O_TILE + lx, y = ty * LSV2O_TILE + ly;
                    if (x >= W || y >= H) continue;
                    const size_t pix = (size_t)y * W + x;
                    const float pxf = (float)x, pyf = (float)y;
                    float T = 1.0f, C[3] = { 0.f, 0.f, 0.f };
                    for (int c = 0; c < L; ++c) acc[c] = 0.f;
                    uint32_t contributor = 0, last = 0;
                    for (uint32_t j = r0; j < r1; ++j) {
                        ++contributor;
                        const uint32_t g = point_list[j];
                        const float* co = conic_opacity + 4 * (size_t)g;
                        float dx, dy;
                        const float power = blend_power(xy + 2 * (size_t)g, co, pxf, pyf, &dx, &dy);
                        if (power > 0.0f) continue;
                        const float alpha = fminf(LSV2O_ALPHA_MAX, co[3] * lsv2o_expf(power));
                        if (alpha < LSV2O_ALPHA_MIN) continue;
                        const float test_T = T * (1.0f - alpha);
                        if (test_T < LSV2O_T_MIN) break;
                        const float w = alpha * T;
                        for (int c = 0; c < 3; ++c) C[c] = fmaf(rgb[3 * (size_t)g + c], w, C[c]);
                        if (mode == 1) {
                            const float* cf = coeff_dense + (size_t)g * L;
                            for (int c = 0; c < L; ++c) acc[c] = fmaf(cf[c], w, acc[c]);
                        } else if (mode == 2) {
                            for (int k = 0; k < K; ++k) {
                                const int c = (int)iq[(size_t)g * K + k];
                                if (c >= 0 && c < L) acc[c] = fmaf(wq[(size_t)g * K + k], w, acc[c]);
                            }
                        }
                        T = test_T;
                        last = contributor;
                    }
                    final_T[pix] = T;
                    n_contrib[pix] = last;
                    for (int c = 0; c < 3; ++c) out_color[c * HW + pix] = fmaf(T, bg[c], C[c]);
                    for (int c = 0; c < L; ++c) out_wmap[c * HW + pix] = acc[c];
                }
        }
        free(acc);
    }
    return 0;
}

int lsv2o_render_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                         const float* xy, const float* conic_opacity, const float* rgb,
                         int mode, int L, int K,
                         const float* coeff_dense, const float* wq, const float* iq,
                         const float* bg, float* out_color, float* out_wmap, float* final_T,
                         uint32_t* n_contrib)
{
    return render_forward_impl(W, H, ranges, point_list, xy, conic_opacity, rgb, mode, L, K, coeff_dense,
                               wq, iq, bg, out_color, out_wmap, final_T, n_contrib, 1, 0);
}

/* Same blend, but only the tiles with tile % stride == phase are rendered (the other pixels of
 * the outputs are left untouched).  Used by bench.py to time a bounded sample of a large frame. */
int lsv2o_render_forward_tiles(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                               const float* xy, const float* conic_opacity, const float* rgb,
                               int mode, int L, int K,
                               const float* coeff_dense, const float* wq, const float* iq,
                               const float* bg, float* out_color, float* out_wmap, float* final_T,
                               uint32_t* n_contrib, int tile_stride, int tile_phase)
{
    if (tile_stride < 1) tile_stride = 1;
    return render_forward_impl(W, H, ranges, point_list, xy, conic_opacity, rgb, mode, L, K, coeff_dense,
                               wq, iq, bg, out_color, out_wmap, final_T, n_contrib, tile_stride,
                               tile_phase % tile_stride);
}

/* ------------------------------------------------------------------------- */
/* R10 render backward, coefficient gradient only (geometry frozen,          */
/* scene/gaussian_model.py:238-243):                                          */
/*   dL/dcoeff[g][c] = sum_pix alpha*T * dL/dWmap[c][pix]                     */
/* The reference re-blends back-to-front from final_T / n_contrib [LINEAGE];  */
/* the per-(pixel,Gaussian) weight alpha*T is identical either way, so the     */
/* oracle recomputes it front-to-back with the forward's exact arithmetic and */
/* accumulates in float64 (the "truth" the tolerance is measured against).     */
/* dense=1: every channel of L; dense=0: only channels in the Gaussian's        */
/* support (coeff != 0), others left 0.                                        */
/* Also returns dL/dcolors f64[P][3] when dL_dcolor != NULL.                    */
/* ------------------------------------------------------------------------- */
int lsv2o_render_backward_coeff(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                                const float* xy, const float* conic_opacity,
                                const uint32_t* n_contrib, int L, int dense,
                                const float* coeff_dense /* support mask, may be NULL if dense */,
                                const float* dL_dwmap /*[L][H][W]*/,
                                const float* dL_dcolor /*[3][H][W] or NULL*/,
                                double* dL_dcoeff /*[P][L], zero-initialised by caller*/,
                                double* dL_dcolors /*[P][3] or NULL*/)
{
    const int gx = (W + LSV2O_TILE - 1) / LSV2O_TILE, gy = (H + LSV2O_TILE - 1) / LSV2O_TILE;
    const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; ++tile) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile];
        for (int ly = 0; ly < LSV2O_TILE; ++ly)
            for (int lx = 0; lx < LSV2O_TILE; ++lx) {
                const int x = tx * LSV2O_TILE + lx, y = ty * LSV2O_TILE + ly;
                if (x >= W || y >= H) continue;
                const size_t pix = (size_t)y * W + x;
                const float pxf = (float)x, pyf = (float)y;
                float T = 1.0f;
                const uint32_t last = n_contrib[pix];
                for (uint32_t jj = 0; jj < last; ++jj) {
                    const uint32_t g = point_list[r0 + jj];
                    const float* co = conic_opacity + 4 * (size_t)g;
                    float dx, dy;
                    const float power = blend_power(xy + 2 * (size_t)g, co, pxf, pyf, &dx, &dy);
                    if (power > 0.0f) continue;
                    const float alpha = fminf(LSV2O_ALPHA_MAX, co[3] * lsv2o_expf(power));
                    if (alpha < LSV2O_ALPHA_MIN) continue;
                    const float w = alpha * T;
                    for (int c = 0; c < L; ++c) {
                        if (!dense && coeff_dense[(size_t)g * L + c] == 0.0f) continue;
                        const double v = (double)w * (double)dL_dwmap[c * HW + pix];
#pragma omp atomic
                        dL_dcoeff[(size_t)g * L + c] += v;
                    }
                    if (dL_dcolors && dL_dcolor)
                        for (int c = 0; c < 3; ++c) {
                            const double v = (double)w * (double)dL_dcolor[c * HW + pix];
#pragma omp atomic
                            dL_dcolors[(size_t)g * 3 + c] += v;
                        }
                    T = T * (1.0f - alpha);
                }
            }
    }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* R9  codebook decode + L2 normalise                                         */
/*   eval_lerf.py:214-218: F_l[512,HW] = C_l^T[512x64] W_l[64,HW];             */
/*   F_l /= (||F_l||_2 over 512 + 1e-10)                                       */
/*   scene/gaussian_model.py:533-549 (single-layer / all-layer variants: the   */
/*   caller picks `levels` and whether to normalise, train.py:157-159).        */
/* wmap [levels*Lc][HW], codebooks [levels][Lc][D], out [levels][D][HW].        */
/* Accumulation in float64 (truth); the fp32 result is the rounded value.       */
/* ------------------------------------------------------------------------- */
int lsv2o_decode(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                 int normalize, float* out)
{
#pragma omp parallel
    {
        double* f = (double*)malloc(sizeof(double) * (size_t)D);
#pragma omp for schedule(static) collapse(2)
        for (int l = 0; l < levels; ++l)
            for (int64_t p = 0; p < HW; ++p) {
                for (int d = 0; d < D; ++d) f[d] = 0.0;
                for (int k = 0; k < Lc; ++k) {
                    const double w = wmap[((size_t)l * Lc + k) * HW + p];
                    if (w == 0.0) continue;
                    const float* cb = codebooks + ((size_t)l * Lc + k) * D;
                    for (int d = 0; d < D; ++d) f[d] += w * (double)cb[d];
                }
                double inv = 1.0;
                if (normalize) {
                    double ss = 0.0;
                    for (int d = 0; d < D; ++d) ss += f[d] * f[d];
                    /* reference computes the norm in fp32 then adds 1e-10 */
                    inv = 1.0 / ((double)(float)sqrt(ss) + 1e-10);
                }
                for (int d = 0; d < D; ++d) out[((size_t)l * D + d) * HW + p] = (float)(f[d] * inv);
            }
        free(f);
    }
    return 0;
}

/* decode backward (autograd of gaussian_model.py:538 without normalise):
 *   dW[l*Lc+k][p] = sum_d C[l][k][d] dF[l][d][p];  dC[l][k][d] = sum_p W[l*Lc+k][p] dF[l][d][p] */
int lsv2o_decode_backward(int levels, int Lc, int D, int64_t HW, const float* wmap,
                          const float* codebooks, const float* dF, double* dW, double* dC)
{
    for (int l = 0; l < levels; ++l)
        for (int k = 0; k < Lc; ++k) {
            const float* cb = codebooks + ((size_t)l * Lc + k) * D;
#pragma omp parallel for schedule(static)
            for (int64_t p = 0; p < HW; ++p) {
                double s = 0.0;
                for (int d = 0; d < D; ++d) s += (double)cb[d] * (double)dF[((size_t)l * D + d) * HW + p];
                dW[((size_t)l * Lc + k) * HW + p] = s;
            }
#pragma omp parallel for schedule(static)
            for (int d = 0; d < D; ++d) {
                double s = 0.0;
                for (int64_t p = 0; p < HW; ++p)
                    s += (double)wmap[((size_t)l * Lc + k) * HW + p] * (double)dF[((size_t)l * D + d) * HW + p];
                dC[((size_t)l * Lc + k) * D + d] = s;
            }
        }
    return 0;
}

/* ------------------------------------------------------------------------- */
/* R4/R5 sparse coefficients: softmax -> top-k -> renormalise                 */
/*   utils/vq_utils.py:9-24 (dense soft code), :26-40 (weights + indices in    */
/*   ASCENDING CHANNEL order, indices returned as float).                      */
/* Ties in top-k: lower channel index wins.                                     */
/* ------------------------------------------------------------------------- */
int lsv2o_topk_soft_code(int64_t N, int C, int k, const float* logits,
                         float* dense /*[N][C] or NULL*/, float* weights /*[N][k] or NULL*/,
                         float* indices /*[N][k] or NULL*/)
{
#pragma omp parallel
    {
        float* y = (float*)malloc(sizeof(float) * (size_t)C);
        uint8_t* m = (uint8_t*)malloc((size_t)C);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < N; ++i) {
            const float* x = logits + (size_t)i * C;
            float mx = x[0];
            for (int c = 1; c < C; ++c) mx = fmaxf(mx, x[c]);
            float s = 0.f;
            for (int c = 0; c < C; ++c) { y[c] = expf(x[c] - mx); s += y[c]; }
            for (int c = 0; c < C; ++c) { y[c] = y[c] / s; m[c] = 0; }
            for (int j = 0; j < k; ++j) {
                int best = -1;
                for (int c = 0; c < C; ++c)
                    if (!m[c] && (best < 0 || y[c] > y[best])) best = c;
                m[best] = 1;
            }
            float ts = 0.f;
            for (int c = 0; c < C; ++c) if (m[c]) ts += y[c];
            const float den = ts + 1e-10f;
            int o = 0;
            for (int c = 0; c < C; ++c) {
                const float v = m[c] ? y[c] / den : 0.f;
                if (dense) dense[(size_t)i * C + c] = v;
                if (m[c] && o < k) {
                    if (weights) weights[(size_t)i * k + o] = v;
                    if (indices) indices[(size_t)i * k + o] = (float)c;
                    ++o;
                }
            }
        }
        free(y); free(m);
    }
    return 0;
}

int lsv2o_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
