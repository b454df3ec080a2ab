// geom_backward.cu -- the RGB stage's backward: gradients of the rendered image with respect to the Gaussians'
// geometry, opacity and colour.  Replaces renderCUDA (backward) + preprocessCUDA (backward) of the reference's
// rasterizer for `train.py` WITHOUT --include_feature (/root/reference/train.py:169-172,247-258: l1 + ssim on the RGB
// render, then optimizer / densification on means, scales, rotations, opacity, SH).
//
// Two kernels:
//   k_render_backward_rgb   one CTA per 16x16 tile; replays the tile's depth-sorted list FRONT to back with the
//                           forward's exact arithmetic (same alpha, same T, same contributor set), and reduces nine
//                           per-(pixel, instance) terms over the tile's 256 pixels into per-Gaussian accumulators;
//   k_preprocess_backward   one thread per Gaussian; pulls the accumulators back through conic -> cov2D -> (J, W) ->
//                           cov3D -> scale / rotation, the projection of the mean, and SH -> RGB.
//
// Front to back instead of the usual back to front: with C the pixel's final colour (saved by the forward) and
// P_i the colour accumulated before instance i,
//     dC/dalpha_i = T_i c_i - (C - P_i - c_i alpha_i T_i) / (1 - alpha_i)
// because every later term (and the background's T_final bg) carries the factor (1 - alpha_i).  T_i and P_i are
// what the forward computes anyway, so no division rebuilds T and the replay stops at n_contrib like the forward.
// The clamp alpha = min(0.99, o G) is treated as the identity in the derivative ([LINEAGE]: the 3DGS backward the
// fork chain shares does the same), skipped pairs (power > 0, alpha < 1/255) have no gradient.
//
// Reduction over pixels: a warp = 32 pixels sums its nine terms with a transposing butterfly (14 shuffles instead of
// 45) into a PRIVATE row of shared memory (one plain store per (warp, instance): no shared atomics); the CTA then adds
// the eight warps' rows in a fixed order and issues ONE global atomic per (instance, term, tile) -- "warp-aggregated
// rather than per-thread global atomics" (BASELINE.json north_star (e)).
#include "common.cuh"

namespace lsv2 {

constexpr int GB_THREADS = TILE_PIX;          // thread = pixel
constexpr int GB_WARPS = GB_THREADS / 32;
constexpr int GB_BATCH = 128;                 // instances staged per round
constexpr int GB_TERMS = 9;                   // rgb(3) | q | q dx | q dy | q dx^2 | q dx dy | q dy^2
constexpr int GB_ACC = LSV2_GEOM_ACC;         // floats per Gaussian in the accumulator array (padded to 12)

struct RgbBackArgs {
    int W, H, gx;
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;            // blend record, 3 x float4 per Gaussian (preprocess.cu)
    const uint32_t* n_contrib;
    const float* color;           // the forward's output [3][H][W] (background included)
    const float* dL_dcolor;       // [3][H][W]
    float* acc;                   // [P][GB_ACC], zero-initialised by the caller
};

// Sums a[0..7] over the warp; lane l ends up with the sum of a[idx(l)], idx = 4*bit4(l) + 2*bit3(l) + bit2(l).
__device__ __forceinline__ float butterfly8(const float (&a)[8], int lane)
{
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
    float b[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float keep = h16 ? a[k + 4] : a[k], send = h16 ? a[k] : a[k + 4];
        b[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float c[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float keep = h8 ? b[k + 2] : b[k], send = h8 ? b[k] : b[k + 2];
        c[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    float v = (h4 ? c[1] : c[0]) + __shfl_xor_sync(0xffffffffu, h4 ? c[0] : c[1], 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

__global__ void __launch_bounds__(GB_THREADS) k_render_backward_rgb(RgbBackArgs a)
{
    __shared__ float4 s_rec[GB_BATCH][3];
    __shared__ uint32_t s_gid[GB_BATCH];
    __shared__ float s_part[GB_WARPS][GB_BATCH][GB_TERMS];      // private per warp: plain stores
    __shared__ uint32_t s_mask[GB_WARPS][GB_BATCH / 32];        // which rows of s_part a warp wrote this round
    __shared__ uint32_t s_max;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tx = tile % a.gx, ty = tile / a.gx;
    const int px = tx * TILE + (tid & (TILE - 1));
    const int py = ty * TILE + (tid >> 4);
    const bool inside = px < a.W && py < a.H;
    const float pxf = (float)px, pyf = (float)py;
    const float wy0 = (float)(ty * TILE + 2 * warp), wy1 = wy0 + 1.0f;      // the warp's two pixel rows
    const uint2 range = a.ranges[tile];
    const size_t HW = (size_t)a.H * a.W;
    const size_t pix = inside ? (size_t)py * a.W + px : 0;

    if (tid == 0) s_max = 0;
    __syncthreads();
    const uint32_t last = inside ? a.n_contrib[pix] : 0u;
    atomicMax(&s_max, last);
    float dC[3] = {0.f, 0.f, 0.f}, Cf[3] = {0.f, 0.f, 0.f}, pre[3] = {0.f, 0.f, 0.f};
    if (inside) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { dC[c] = a.dL_dcolor[c * HW + pix]; Cf[c] = a.color[c * HW + pix]; }
    }
    __syncthreads();
    const uint32_t n_replay = min(s_max, range.y - range.x);

    float T = 1.0f;
    for (uint32_t off = 0; off < n_replay; off += GB_BATCH) {
        const int batch_n = (int)min((uint32_t)GB_BATCH, n_replay - off);
        __syncthreads();                                        // previous round's flush has read s_part / s_mask
        if (tid < batch_n) {
            const uint32_t g = a.point_list[range.x + off + tid];
            s_gid[tid] = g;
            s_rec[tid][0] = a.rec[3 * (size_t)g + 0];
            s_rec[tid][1] = a.rec[3 * (size_t)g + 1];
            s_rec[tid][2] = a.rec[3 * (size_t)g + 2];
        }
        __syncthreads();

        for (int jb = 0; jb < batch_n; jb += 32) {
        uint32_t wrote = 0u;
        const int j_end = min(jb + 32, batch_n);
#pragma unroll 1
        for (int j = jb; j < j_end; ++j) {
            const float4 r2 = s_rec[j][2];                      // (g, b, ext_y, ext_x)
            const float4 r0 = s_rec[j][0];                      // (x, y, ax, b')
            // rows this warp owns against the splat's alpha >= 1/255 extent (rounded outwards in the forward)
            if (r2.z < 0.0f || r0.y + r2.z < wy0 || r0.y - r2.z > wy1) continue;
            const float4 r1 = s_rec[j][1];                      // (az, power cut-off, opacity, r)
            float v[8];
            float v8 = 0.0f;
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = 0.0f;
            bool hit = false;
            if (off + j < last) {
                const float dx = r0.x - pxf, dy = r0.y - pyf;
                const float power = fmaf(r0.z * dx, dx, fmaf(r1.x * dy, dy, (r0.w * dx) * dy));
                if (!(power > 0.0f || power < r1.y)) {
                    const float G = repro_expf(power);
                    const float alpha = fminf(ALPHA_MAX, r1.z * G);
                    if (!(alpha < ALPHA_MIN)) {
                        hit = true;
                        const float w = alpha * T;
                        const float inv1m = 1.0f / (1.0f - alpha);
                        const float col[3] = {r1.w, r2.x, r2.y};
                        float dLda = 0.0f;
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float cw = col[c] * w;
                            const float suffix = (Cf[c] - pre[c]) - cw;
                            dLda = fmaf(dC[c], col[c] * T - suffix * inv1m, dLda);
                            pre[c] += cw;
                            v[c] = w * dC[c];
                        }
                        const float q = G * dLda;               // d/d opacity; times opacity = d/d power
                        v[3] = q;
                        v[4] = q * dx;
                        v[5] = q * dy;
                        v[6] = v[4] * dx;
                        v[7] = v[4] * dy;
                        v8 = v[5] * dy;
                        T = T * (1.0f - alpha);
                    }
                }
            }
            if (__ballot_sync(0xffffffffu, hit) == 0u) continue;
            const float s = butterfly8(v, lane);
            float s8 = v8;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s8 += __shfl_xor_sync(0xffffffffu, s8, o);
            if ((lane & 3) == 0) s_part[warp][j][lane >> 2] = s;
            if (lane == 0) s_part[warp][j][8] = s8;
            wrote |= 1u << (j & 31);
        }
        if (lane == 0) s_mask[warp][jb >> 5] = wrote;                   // warp-uniform value
        }
        __syncthreads();
        // fixed-order sum of the warps' rows, one global atomic per (instance, term)
        for (int e = tid; e < batch_n * GB_TERMS; e += GB_THREADS) {
            const int j = e / GB_TERMS, k = e - j * GB_TERMS;
            float sum = 0.0f;
            bool any = false;
#pragma unroll
            for (int w = 0; w < GB_WARPS; ++w)
                if ((s_mask[w][j >> 5] >> (j & 31)) & 1u) { sum += s_part[w][j][k]; any = true; }
            if (any && sum != 0.0f) atomicAdd(a.acc + (size_t)s_gid[j] * GB_ACC + k, sum);
        }
    }
}

int launch_render_backward_rgb(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_image& img,
                               const float* dL_dcolor, float* acc, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec || !b.ranges || !b.point_list || !img.color || !img.n_contrib || !dL_dcolor || !acc)
        return LSV2_EINVAL;
    RgbBackArgs a;
    a.W = v.width; a.H = v.height; a.gx = gx;
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.rec = reinterpret_cast<const float4*>(g.blend_rec);
    a.n_contrib = img.n_contrib;
    a.color = img.color;
    a.dL_dcolor = dL_dcolor;
    a.acc = acc;
    k_render_backward_rgb<<<gx * gy, GB_THREADS, 0, st>>>(a);
    LSV2_LAUNCH_CHECK("k_render_backward_rgb");
    return LSV2_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// per-Gaussian chain: accumulators -> (mean2D, conic, opacity, rgb) -> inputs of the op
// ---------------------------------------------------------------------------------------------------------------
__constant__ float c_GB_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__constant__ float c_GB_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

struct PreBackArgs {
    int P, W, H, deg, M;
    float tanfovx, tanfovy, fx, fy, mod;
    const float *vm, *pm, *campos;
    const float *means, *scales, *rots, *cov3d_pre, *shs;
    const int32_t* radii;
    const float4* conic_opacity;
    const uint8_t* clamped;
    const float* acc;
    float *d_means, *d_means2d, *d_opac, *d_scales, *d_rots, *d_cov3d, *d_shs, *d_colors;
};

constexpr int PB_THREADS = 128;
constexpr int PB_SH_FLOATS = 48;                 // degree-3 row: 16 coefficients x RGB
constexpr int PB_SH_STRIDE = PB_SH_FLOATS + 1;    // odd stride: conflict-free per-thread rows

__global__ void __launch_bounds__(PB_THREADS) k_preprocess_backward(PreBackArgs a)
{
    __shared__ float s_m[32];
    __shared__ float s_cam[3];
    const int tid = threadIdx.x;
    if (tid < 16) s_m[tid] = a.vm[tid];
    else if (tid < 32) s_m[tid] = a.pm[tid - 16];
    else if (tid < 35) s_cam[tid - 32] = a.campos[tid - 32];
    __syncthreads();
    const int i = blockIdx.x * PB_THREADS + tid;
    const bool live = i < a.P;                              // no early exit: the SH rows go through block-wide staging
    const float* vm = s_m;
    const float* pm = s_m + 16;
    const size_t si = (size_t)(live ? i : 0);

    float dmean[3] = {0.f, 0.f, 0.f}, dm2[2] = {0.f, 0.f}, dop = 0.f;
    float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float drgb[3] = {0.f, 0.f, 0.f};
    const bool vis = live && a.radii[i] > 0;
    const int ncoef = (a.deg + 1) * (a.deg + 1);

    if (vis) {
        const float* acc = a.acc + si * GB_ACC;
        const float4 co = a.conic_opacity[i];
        const float op = co.w;
        // power = ax dx^2 + az dy^2 + b' dx dy with ax = -a/2, az = -c/2, b' = -b;  dL/dpower summed = op * acc[3..]
        const float ax = -0.5f * co.x, bp = -co.y, az = -0.5f * co.z;
        const float S1x = op * acc[4], S1y = op * acc[5];
        const float dA = -0.5f * (op * acc[6]);                 // dL/d conic a
        const float dB = -(op * acc[7]);                        // dL/d conic b
        const float dCc = -0.5f * (op * acc[8]);                // dL/d conic c
        const float dpx = (2.0f * ax) * S1x + bp * S1y;         // dL/d pixel-space mean
        const float dpy = (2.0f * az) * S1y + bp * S1x;
        dop = acc[3];
        drgb[0] = acc[0]; drgb[1] = acc[1]; drgb[2] = acc[2];
        // means2D gradient in the reference's convention: with respect to the NDC-scaled point
        // ([LINEAGE] ddelx_dx = W/2; consumed as a densification statistic, scene/gaussian_model.py:506-508)
        dm2[0] = dpx * (0.5f * (float)a.W);
        dm2[1] = dpy * (0.5f * (float)a.H);

        const float p0 = a.means[3 * si + 0], p1 = a.means[3 * si + 1], p2 = a.means[3 * si + 2];
        const float vx = ((vm[0] * p0 + vm[4] * p1) + vm[8] * p2) + vm[12];
        const float vy = ((vm[1] * p0 + vm[5] * p1) + vm[9] * p2) + vm[13];
        const float vz = ((vm[2] * p0 + vm[6] * p1) + vm[10] * p2) + vm[14];

        // ---- projection of the mean: px = ((hx pw + 1) W - 1) / 2, pw = 1 / (hw + 1e-7)
        {
            const float hx = ((pm[0] * p0 + pm[4] * p1) + pm[8] * p2) + pm[12];
            const float hy = ((pm[1] * p0 + pm[5] * p1) + pm[9] * p2) + pm[13];
            const float hw = ((pm[3] * p0 + pm[7] * p1) + pm[11] * p2) + pm[15];
            const float pw = 1.0f / (hw + W_GUARD);
            const float dhx = dm2[0] * pw, dhy = dm2[1] * pw;
            const float dhw = -((dm2[0] * hx + dm2[1] * hy) * pw) * pw;
#pragma unroll
            for (int j = 0; j < 3; ++j) dmean[j] += (dhx * pm[4 * j + 0] + dhy * pm[4 * j + 1]) + dhw * pm[4 * j + 3];
        }

        // ---- cov3D (recomputed as the forward does)
        float c6[6];
        float R[3][3], sc[3] = {0.f, 0.f, 0.f};
        float qr = 0.f, qx = 0.f, qy = 0.f, qz = 0.f;
        if (a.cov3d_pre) {
#pragma unroll
            for (int k = 0; k < 6; ++k) c6[k] = a.cov3d_pre[6 * si + k];
        } else {
            const float4 q = reinterpret_cast<const float4*>(a.rots)[i];
            qr = q.x; qx = q.y; qy = q.z; qz = q.w;
            R[0][0] = 1.f - 2.f * (qy * qy + qz * qz);
            R[0][1] = 2.f * (qx * qy - qr * qz);
            R[0][2] = 2.f * (qx * qz + qr * qy);
            R[1][0] = 2.f * (qx * qy + qr * qz);
            R[1][1] = 1.f - 2.f * (qx * qx + qz * qz);
            R[1][2] = 2.f * (qy * qz - qr * qx);
            R[2][0] = 2.f * (qx * qz - qr * qy);
            R[2][1] = 2.f * (qy * qz + qr * qx);
            R[2][2] = 1.f - 2.f * (qx * qx + qy * qy);
#pragma unroll
            for (int k = 0; k < 3; ++k) sc[k] = a.mod * a.scales[3 * si + k];
            int o = 0;
#pragma unroll
            for (int r_ = 0; r_ < 3; ++r_)
#pragma unroll
                for (int c_ = r_; c_ < 3; ++c_)
                    c6[o++] = ((R[r_][0] * sc[0]) * (R[c_][0] * sc[0]) + (R[r_][1] * sc[1]) * (R[c_][1] * sc[1])) +
                              (R[r_][2] * sc[2]) * (R[c_][2] * sc[2]);
        }

        // ---- EWA projection, forward values
        const float limx = FRUSTUM_CLAMP * a.tanfovx, limy = FRUSTUM_CLAMP * a.tanfovy;
        const float txtz = vx / vz, tytz = vy / vz;
        const bool clx = txtz < -limx || txtz > limx, cly = tytz < -limy || tytz > limy;
        const float tx = fminf(limx, fmaxf(-limx, txtz)) * vz;
        const float ty = fminf(limy, fmaxf(-limy, tytz)) * vz;
        const float iz = 1.0f / vz, iz2 = iz * iz, iz3 = iz2 * iz;
        const float J00 = a.fx * iz, J02 = -(a.fx * tx) * iz2, J11 = a.fy * iz, J12 = -(a.fy * ty) * iz2;
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
        const float i2 = 1.0f / (det * det);
        // conic = (cv2, -cv1, cv0) / det
        const float dcv0 = i2 * ((-(cv2 * cv2) * dA + (cv1 * cv2) * dB) - (cv1 * cv1) * dCc);
        const float dcv1 = i2 * (((2.0f * cv1 * cv2) * dA - (cv0 * cv2 + cv1 * cv1) * dB) + (2.0f * cv0 * cv1) * dCc);
        const float dcv2 = i2 * ((-(cv1 * cv1) * dA + (cv0 * cv1) * dB) - (cv0 * cv0) * dCc);

        // dL/dSigma as a full symmetric matrix Gs (entries treated as independent), dL/dM rows
        float Gs[3][3];
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int k = 0; k < 3; ++k)
                Gs[j][k] = (dcv0 * (M0[j] * M0[k]) + (0.5f * dcv1) * (M0[j] * M1[k] + M0[k] * M1[j])) + dcv2 * (M1[j] * M1[k]);
        float dM0[3], dM1[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            dM0[j] = (2.0f * dcv0) * A0[j] + dcv1 * A1[j];
            dM1[j] = dcv1 * A0[j] + (2.0f * dcv2) * A1[j];
        }
        float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            dJ00 += dM0[j] * vm[4 * j + 0];
            dJ02 += dM0[j] * vm[4 * j + 2];
            dJ11 += dM1[j] * vm[4 * j + 1];
            dJ12 += dM1[j] * vm[4 * j + 2];
        }
        // J00 = fx/vz, J02 = -fx tx/vz^2 (tx = vx inside the frustum clamp, = +-limx vz outside), same for y
        float dvx = 0.f, dvy = 0.f, dvz = 0.f;
        dvz += -(a.fx * iz2) * dJ00 - (a.fy * iz2) * dJ11;
        dvz += (2.0f * (a.fx * tx) * iz3) * dJ02 + (2.0f * (a.fy * ty) * iz3) * dJ12;
        if (clx) dvz += -(a.fx * iz2) * (tx * iz) * dJ02; else dvx += -(a.fx * iz2) * dJ02;
        if (cly) dvz += -(a.fy * iz2) * (ty * iz) * dJ12; else dvy += -(a.fy * iz2) * dJ12;
#pragma unroll
        for (int j = 0; j < 3; ++j) dmean[j] += (dvx * vm[4 * j + 0] + dvy * vm[4 * j + 1]) + dvz * vm[4 * j + 2];

        if (a.cov3d_pre) {
            // the six stored entries: off-diagonal ones stand for two matrix entries each
            dcov[0] = Gs[0][0]; dcov[1] = 2.0f * Gs[0][1]; dcov[2] = 2.0f * Gs[0][2];
            dcov[3] = Gs[1][1]; dcov[4] = 2.0f * Gs[1][2]; dcov[5] = Gs[2][2];
        } else {
            // Sigma = Lm Lm^T, Lm = R diag(s):  dL/dLm = 2 Gs Lm
            float dR[3][3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float ds = 0.f;
#pragma unroll
                for (int r_ = 0; r_ < 3; ++r_) {
                    const float dLm = 2.0f * ((Gs[r_][0] * (R[0][k] * sc[k]) + Gs[r_][1] * (R[1][k] * sc[k])) +
                                              Gs[r_][2] * (R[2][k] * sc[k]));
                    ds += dLm * R[r_][k];
                    dR[r_][k] = dLm * sc[k];
                }
                dscale[k] = a.mod * ds;
            }
            drot[0] = 2.0f * ((((-qz * dR[0][1] + qy * dR[0][2]) + qz * dR[1][0]) - qx * dR[1][2]) +
                              (-qy * dR[2][0] + qx * dR[2][1]));
            drot[1] = 2.0f * ((((qy * dR[0][1] + qz * dR[0][2]) + qy * dR[1][0]) - 2.0f * qx * dR[1][1]) +
                              (((-qr * dR[1][2] + qz * dR[2][0]) + qr * dR[2][1]) - 2.0f * qx * dR[2][2]));
            drot[2] = 2.0f * ((((-2.0f * qy * dR[0][0] + qx * dR[0][1]) + qr * dR[0][2]) + qx * dR[1][0]) +
                              (((qz * dR[1][2] - qr * dR[2][0]) + qz * dR[2][1]) - 2.0f * qy * dR[2][2]));
            drot[3] = 2.0f * ((((-2.0f * qz * dR[0][0] - qr * dR[0][1]) + qx * dR[0][2]) + qr * dR[1][0]) +
                              (((-2.0f * qz * dR[1][1] + qy * dR[1][2]) + qx * dR[2][0]) + qy * dR[2][1]));
        }
    }

    // ---- colour: SH coefficients and the view direction's dependence on the mean
    // The SH block is the bulk of this kernel's bytes (192 B read + 192 B written per Gaussian).  A CTA's rows are one
    // contiguous block: they are staged through shared memory with 16-byte coalesced accesses (per-thread row accesses
    // touch 32 different lines per instruction), only rows of visible Gaussians are read, and the gradient row
    // overwrites the staged coefficients in place (element (k, c) is read before it is written).
    __shared__ float s_sh[PB_THREADS * PB_SH_STRIDE];
    __shared__ int s_visf[PB_THREADS];
    const int row_f = a.M * 3;
    const bool staged = a.shs && a.d_shs && row_f == PB_SH_FLOATS &&
                        ((reinterpret_cast<uintptr_t>(a.shs) | reinterpret_cast<uintptr_t>(a.d_shs)) & 15) == 0;
    const size_t blk_base = (size_t)blockIdx.x * PB_THREADS;
    const int blk_rows = min(PB_THREADS, a.P - (int)blk_base);
    if (staged) {
        s_visf[tid] = vis ? 1 : 0;
        __syncthreads();
        const float4* src4 = reinterpret_cast<const float4*>(a.shs + blk_base * PB_SH_FLOATS);
        for (int e = tid; e < blk_rows * (PB_SH_FLOATS / 4); e += PB_THREADS) {
            const int r = e / (PB_SH_FLOATS / 4), q = e - r * (PB_SH_FLOATS / 4);
            float* d = s_sh + r * PB_SH_STRIDE + 4 * q;
            if (s_visf[r]) {
                const float4 v = __ldg(src4 + e);
                d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
            } else {
                d[0] = 0.f; d[1] = 0.f; d[2] = 0.f; d[3] = 0.f;         // culled: the gradient row is zero
            }
        }
        __syncthreads();
    }
    if (a.shs) {
        float* dsh = staged ? s_sh + tid * PB_SH_STRIDE : (a.d_shs && live ? a.d_shs + si * a.M * 3 : nullptr);
        if (vis) {
            const float* sh = staged ? s_sh + tid * PB_SH_STRIDE : a.shs + si * a.M * 3;
            const float p0 = a.means[3 * si + 0], p1 = a.means[3 * si + 1], p2 = a.means[3 * si + 2];
            const float d0 = p0 - s_cam[0], d1 = p1 - s_cam[1], d2 = p2 - s_cam[2];
            const float len = sqrtf((d0 * d0 + d1 * d1) + d2 * d2);
            const float x = d0 / len, y = d1 / len, z = d2 / len;
            const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
            float gx = 0.f, gy = 0.f, gz = 0.f;                 // dL/d(unit direction)
            float basis[16];
            basis[0] = C0;
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            basis[1] = -C1 * y; basis[2] = C1 * z; basis[3] = -C1 * x;
            basis[4] = c_GB_C2[0] * xy; basis[5] = c_GB_C2[1] * yz; basis[6] = c_GB_C2[2] * ((2.0f * zz - xx) - yy);
            basis[7] = c_GB_C2[3] * xz; basis[8] = c_GB_C2[4] * (xx - yy);
            basis[9] = (c_GB_C3[0] * y) * (3.0f * xx - yy);
            basis[10] = (c_GB_C3[1] * xy) * z;
            basis[11] = (c_GB_C3[2] * y) * ((4.0f * zz - xx) - yy);
            basis[12] = (c_GB_C3[3] * z) * ((2.0f * zz - 3.0f * xx) - 3.0f * yy);
            basis[13] = (c_GB_C3[4] * x) * ((4.0f * zz - xx) - yy);
            basis[14] = (c_GB_C3[5] * z) * (xx - yy);
            basis[15] = (c_GB_C3[6] * x) * (xx - 3.0f * yy);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float dr = (a.clamped && a.clamped[3 * si + c]) ? 0.0f : drgb[c];
#define SH(k) sh[(k) * 3 + c]
                if (a.deg > 0) {
                    float ex = -C1 * SH(3), ey = -C1 * SH(1), ez = C1 * SH(2);
                    if (a.deg > 1) {
                        ex += (c_GB_C2[0] * y) * SH(4) - (2.0f * c_GB_C2[2] * x) * SH(6) + (c_GB_C2[3] * z) * SH(7) +
                              (2.0f * c_GB_C2[4] * x) * SH(8);
                        ey += (c_GB_C2[0] * x) * SH(4) + (c_GB_C2[1] * z) * SH(5) - (2.0f * c_GB_C2[2] * y) * SH(6) -
                              (2.0f * c_GB_C2[4] * y) * SH(8);
                        ez += (c_GB_C2[1] * y) * SH(5) + (4.0f * c_GB_C2[2] * z) * SH(6) + (c_GB_C2[3] * x) * SH(7);
                        if (a.deg > 2) {
                            ex += (6.0f * c_GB_C3[0] * xy) * SH(9) + (c_GB_C3[1] * yz) * SH(10) -
                                  (2.0f * c_GB_C3[2] * xy) * SH(11) - (6.0f * c_GB_C3[3] * xz) * SH(12) +
                                  (c_GB_C3[4] * ((4.0f * zz - 3.0f * xx) - yy)) * SH(13) +
                                  (2.0f * c_GB_C3[5] * xz) * SH(14) + (3.0f * c_GB_C3[6] * (xx - yy)) * SH(15);
                            ey += (3.0f * c_GB_C3[0] * (xx - yy)) * SH(9) + (c_GB_C3[1] * xz) * SH(10) +
                                  (c_GB_C3[2] * ((4.0f * zz - xx) - 3.0f * yy)) * SH(11) -
                                  (6.0f * c_GB_C3[3] * yz) * SH(12) - (2.0f * c_GB_C3[4] * xy) * SH(13) -
                                  (2.0f * c_GB_C3[5] * yz) * SH(14) - (6.0f * c_GB_C3[6] * xy) * SH(15);
                            ez += (c_GB_C3[1] * xy) * SH(10) + (8.0f * c_GB_C3[2] * yz) * SH(11) +
                                  (c_GB_C3[3] * ((6.0f * zz - 3.0f * xx) - 3.0f * yy)) * SH(12) +
                                  (8.0f * c_GB_C3[4] * xz) * SH(13) + (c_GB_C3[5] * (xx - yy)) * SH(14);
                        }
                    }
                    gx += dr * ex; gy += dr * ey; gz += dr * ez;
                }
#undef SH
                if (dsh) {
                    for (int k = 0; k < a.M; ++k) dsh[k * 3 + c] = k < ncoef ? basis[k] * dr : 0.0f;
                }
            }
            // direction = d / |d|:  dL/dd = (g - dir (dir . g)) / |d|
            const float dot = (x * gx + y * gy) + z * gz;
            dmean[0] += (gx - x * dot) / len;
            dmean[1] += (gy - y * dot) / len;
            dmean[2] += (gz - z * dot) / len;
        } else if (dsh && !staged) {
            for (int k = 0; k < a.M * 3; ++k) dsh[k] = 0.0f;
        }
    } else if (a.d_colors && live) {
#pragma unroll
        for (int c = 0; c < 3; ++c) a.d_colors[3 * si + c] = drgb[c];
    }
    if (staged) {
        __syncthreads();
        float4* dst4 = reinterpret_cast<float4*>(a.d_shs + blk_base * PB_SH_FLOATS);
        for (int e = tid; e < blk_rows * (PB_SH_FLOATS / 4); e += PB_THREADS) {
            const int r = e / (PB_SH_FLOATS / 4), q = e - r * (PB_SH_FLOATS / 4);
            const float* d = s_sh + r * PB_SH_STRIDE + 4 * q;
            dst4[e] = make_float4(d[0], d[1], d[2], d[3]);
        }
    }
    if (!live) return;

    if (a.d_means) { a.d_means[3 * si + 0] = dmean[0]; a.d_means[3 * si + 1] = dmean[1]; a.d_means[3 * si + 2] = dmean[2]; }
    if (a.d_means2d) { a.d_means2d[3 * si + 0] = dm2[0]; a.d_means2d[3 * si + 1] = dm2[1]; a.d_means2d[3 * si + 2] = 0.0f; }
    if (a.d_opac) a.d_opac[si] = dop;
    if (a.d_scales) { a.d_scales[3 * si + 0] = dscale[0]; a.d_scales[3 * si + 1] = dscale[1]; a.d_scales[3 * si + 2] = dscale[2]; }
    if (a.d_rots) {
        reinterpret_cast<float4*>(a.d_rots)[i] = make_float4(drot[0], drot[1], drot[2], drot[3]);
    }
    if (a.d_cov3d) {
#pragma unroll
        for (int k = 0; k < 6; ++k) a.d_cov3d[6 * si + k] = dcov[k];
    }
}

int launch_preprocess_backward(const lsv2_view& v, const lsv2_scene& s, const lsv2_geom& g, const float* acc,
                               const lsv2_geom_grads& out, cudaStream_t st)
{
    if (s.P == 0) return LSV2_OK;
    if (!acc || !g.radii || !g.conic_opacity || !s.means3d) return LSV2_EINVAL;
    if (!s.cov3d_precomp && (!s.scales || !s.rotations)) return LSV2_EINVAL;
    if (s.shs && s.sh_coeffs > 16) return LSV2_EUNSUPPORTED;
    if (out.rotations && (reinterpret_cast<uintptr_t>(out.rotations) & 15)) return LSV2_EINVAL;
    PreBackArgs a;
    a.P = s.P; a.W = v.width; a.H = v.height; a.deg = v.sh_degree; a.M = s.sh_coeffs;
    a.tanfovx = v.tanfovx; a.tanfovy = v.tanfovy;
    a.fx = (float)v.width / (2.0f * v.tanfovx);
    a.fy = (float)v.height / (2.0f * v.tanfovy);
    a.mod = v.scale_modifier;
    a.vm = v.viewmatrix; a.pm = v.projmatrix; a.campos = v.campos;
    a.means = s.means3d; a.scales = s.scales; a.rots = s.rotations; a.cov3d_pre = s.cov3d_precomp; a.shs = s.shs;
    a.radii = g.radii;
    a.conic_opacity = reinterpret_cast<const float4*>(g.conic_opacity);
    a.clamped = g.clamped;
    a.acc = acc;
    a.d_means = out.means3d; a.d_means2d = out.means2d; a.d_opac = out.opacities; a.d_scales = out.scales;
    a.d_rots = out.rotations; a.d_cov3d = out.cov3d; a.d_shs = out.shs; a.d_colors = out.colors;
    k_preprocess_backward<<<(s.P + PB_THREADS - 1) / PB_THREADS, PB_THREADS, 0, st>>>(a);
    LSV2_LAUNCH_CHECK("k_preprocess_backward");
    return LSV2_OK;
}

}  // namespace lsv2
