// abi.cu -- extern "C" entry points of liblsv2.so (see include/lsv2.h for the contract and
// the reference interfaces each one replaces).
#include <stdio.h>
#include <string.h>

#include "common.cuh"

namespace lsv2 {

static thread_local char g_cuda_error[512] = "";

void set_cuda_error(cudaError_t e, const char* what)
{
    snprintf(g_cuda_error, sizeof g_cuda_error, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}

int sort_num_passes(int nbits);
int launch_segment_sum(int Lc, int64_t HW, int nseg, const float* x, const int32_t* seg, float* out, cudaStream_t st);
int launch_decode_fp32(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                       int normalize, float* out, cudaStream_t st);
int launch_decode_tc(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                     int normalize, int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st);
size_t decode_tc_scratch_bytes(int levels, int Lc, int D);
int launch_decode_backward(int levels, int Lc, int D, int64_t HW, const float* wmap, const float* codebooks,
                           const float* dF, float* dW, float* dC, cudaStream_t st);
int launch_cos_loss_forward(int D, int64_t HW, const float* F, const float* gt, const float* mask, int normalize,
                            float* alpha_beta, float* partial, cudaStream_t st);
int launch_cos_loss_backward(int D, int64_t HW, const float* F, const float* gt, const float* alpha_beta,
                             const float* grad_out, float* dF, cudaStream_t st);
int launch_gram_loss_forward(int Lc, int64_t HW, const float* wmap, const float* h, const float* g2, const float* mask,
                             const float* G, int normalize, float* u, float* alpha_beta, float* partial,
                             const int32_t* seg, int nseg, cudaStream_t st);
int launch_gram_loss_backward(int Lc, int64_t HW, const float* wmap, const float* h, const float* u,
                              const float* alpha_beta, const float* grad_out, float* dw, float* aw, float* bw,
                              const int32_t* seg, int nseg, cudaStream_t st);
int launch_colsumsq(int D, int64_t HW, const float* x, float* out, cudaStream_t st);
int launch_topk_soft_code_bwd(int64_t N, int C, const float* logits, const float* dense, const float* d_dense,
                              float* d_logits, cudaStream_t st);
int launch_topk_soft_code(int64_t N, int C, int k, const float* logits, float* dense, float* weights,
                          float* indices, cudaStream_t st);

int launch_adam(float* p, const float* g, float* m, float* v, int64_t n, double lr, double b1, double b2, double eps,
                int64_t step, int zero_grad, const float* dev_scalars, cudaStream_t st);
size_t render_decode_scratch_bytes(int levels);
int launch_pack_coeff(int P, const lsv2_geom& g, const lsv2_features& f, uint32_t* status, cudaStream_t st);
int launch_render_wmap_tc(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                          const lsv2_image& img, void* scratch, size_t scratch_bytes, cudaStream_t st);
int launch_render_decode(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                         const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                         int out_dtype, int out_hwc, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st);
size_t render_query_scratch_bytes(int levels);
int launch_render_query(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                        const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, const float* embeds,
                        int n_pos, int n_neg, float* out, void* scratch, size_t scratch_bytes, cudaStream_t st);

static int tile_bits_for(const lsv2_view& v)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    const int64_t tiles = (int64_t)gx * gy;
    int bits = 1;
    while ((1ll << bits) < tiles) ++bits;
    return bits;
}

// depth keys are positive floats (z > NEAR_Z), so bit 31 of the low word is always clear
constexpr int DEPTH_KEY_BITS = 31;

static bool view_ok(const lsv2_view* v)
{
    return v && v->width > 0 && v->height > 0 && v->viewmatrix && v->projmatrix && v->campos && v->bg &&
           v->tanfovx > 0.f && v->tanfovy > 0.f;
}

}  // namespace lsv2

using namespace lsv2;

extern "C" {

int lsv2_abi_version(void) { return LSV2_ABI_VERSION; }

const char* lsv2_strerror(int code)
{
    switch (code) {
        case LSV2_OK: return "ok";
        case LSV2_EINVAL: return "invalid argument combination";
        case LSV2_ECUDA: return "CUDA error";
        case LSV2_ENOSPC: return "scratch buffer too small (or, for lsv2_decode_static, more levels than its built-in scratch holds)";
        case LSV2_EUNSUPPORTED: return "shape not supported by the sm_100a kernels";
        default: return "unknown error";
    }
}

const char* lsv2_last_cuda_error(void) { return g_cuda_error; }

size_t lsv2_sort_scratch_bytes(uint64_t capacity, int32_t P)
{
    const size_t a = sort_scratch_bytes(capacity);
    const size_t b = scan_scratch_bytes(P);
    const size_t c = depth_order_scratch_bytes(P);
    return (a > b ? (a > c ? a : c) : (b > c ? b : c)) + 256;
}

int lsv2_preprocess(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom,
                    const lsv2_binning* binning, void* stream)
{
    if (!view_ok(view) || !scene || !geom || !binning || !binning->status || scene->P < 0) return LSV2_EINVAL;
    if (scene->P > 0 && (!scene->means3d || !scene->opacities || !geom->radii || !geom->xy || !geom->depth ||
                         !geom->conic_opacity || !geom->rgb || !geom->tiles_touched || !geom->offsets))
        return LSV2_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    LSV2_CUDA_TRY(cudaMemsetAsync(binning->status, 0, LSV2_STATUS_WORDS * sizeof(uint32_t), st));
    int rc = launch_preprocess(*view, *scene, *geom, st);
    if (rc) return rc;
    return launch_scan(scene->P, geom->tiles_touched, geom->offsets, binning->status + LSV2_STATUS_NUM_RENDERED,
                       binning->scratch, binning->scratch_bytes, st);
}

int lsv2_bin(const lsv2_view* view, int32_t P, const lsv2_geom* geom, const lsv2_binning* binning, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !binning->status || !binning->ranges) return LSV2_EINVAL;
    if (binning->capacity > 0 && (!binning->keys || !binning->point_list || !binning->keys_alt))
        return LSV2_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const int tb = tile_bits_for(*view);
    // Gaussians in depth order first (P pairs), then the instances on the tile word alone (R pairs): the same
    // list as one stable sort on (tile | depth), see radix_sort.cu
    const uint64_t* order = nullptr;
    const uint32_t* offsets = geom->offsets;
    int rc = launch_depth_order(P, *geom, DEPTH_KEY_BITS, binning->scratch, binning->scratch_bytes, &order, &offsets,
                                st);
    if (rc) return rc;
    const int npass = sort_num_passes(tb);
    // emit the unsorted pairs into the buffer from which an npass ping-pong ends in (keys, point_list)
    lsv2_binning emit = *binning;
    if (!(npass & 1)) emit.keys_alt = binning->keys;
    rc = launch_duplicate(*view, P, *geom, emit, order, offsets, st);
    if (rc) return rc;
    const int n_tiles = ((view->width + TILE - 1) / TILE) * ((view->height + TILE - 1) / TILE);
    bool ranges_done = false;
    rc = launch_sort(*binning, tb, n_tiles, &ranges_done, binning->scratch, binning->scratch_bytes, st);
    if (rc) return rc;
    return ranges_done ? LSV2_OK : launch_ranges(*view, *binning, st);
}

int lsv2_compact_coeff(int32_t P, const lsv2_geom* geom, const lsv2_features* feat, void* stream)
{
    if (!geom || !feat) return LSV2_EINVAL;
    return launch_compact_coeff(P, *geom, *feat, (cudaStream_t)stream);
}

int lsv2_render_forward(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                        const lsv2_features* feat, const lsv2_image* image, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image) return LSV2_EINVAL;
    if (!image->color || !image->final_T || !image->n_contrib) return LSV2_EINVAL;
    return launch_render_forward(*view, *geom, *binning, *feat, *image, (cudaStream_t)stream);
}

int lsv2_pack_coeff(int32_t P, const lsv2_geom* geom, const lsv2_features* feat, uint32_t* status, void* stream)
{
    if (P < 0 || !geom || !feat) return LSV2_EINVAL;
    return launch_pack_coeff(P, *geom, *feat, status, (cudaStream_t)stream);
}

int lsv2_render_forward_tc(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                           const lsv2_features* feat, const lsv2_image* image, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image) return LSV2_EINVAL;
    if (!image->color || !image->final_T || !image->n_contrib || !image->wmap) return LSV2_EINVAL;
    if (feat->mode != LSV2_MODE_QUICK && feat->mode != LSV2_MODE_DENSE) return LSV2_EUNSUPPORTED;
    if (!feat->packed) return LSV2_EUNSUPPORTED;
    return launch_render_wmap_tc(*view, *geom, *binning, *feat, *image, binning->scratch, binning->scratch_bytes,
                                 (cudaStream_t)stream);
}

int lsv2_forward(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom,
                 const lsv2_binning* binning, const lsv2_features* feat, const lsv2_image* image, void* stream)
{
    int rc = lsv2_preprocess(view, scene, geom, binning, stream);
    if (rc) return rc;
    rc = lsv2_bin(view, scene->P, geom, binning, stream);
    if (rc) return rc;
    if (feat && feat->mode == LSV2_MODE_DENSE && feat->sparse_count) {
        rc = lsv2_compact_coeff(scene->P, geom, feat, stream);
        if (rc) return rc;
    }
    if (feat && image && image->wmap && feat->packed) {
        rc = lsv2_pack_coeff(scene->P, geom, feat, binning->status, stream);
        if (rc) return rc;
        rc = lsv2_render_forward_tc(view, geom, binning, feat, image, stream);
        if (rc != LSV2_EUNSUPPORTED) return rc;
    }
    return lsv2_render_forward(view, geom, binning, feat, image, stream);
}

int lsv2_mark_visible(int32_t P, const float* means3d, const float* viewmatrix, uint8_t* present, void* stream)
{
    if (P < 0 || (P > 0 && (!means3d || !viewmatrix || !present))) return LSV2_EINVAL;
    return launch_mark_visible(P, means3d, viewmatrix, present, (cudaStream_t)stream);
}

int lsv2_render_backward(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                         const lsv2_features* feat, const lsv2_image* image, const float* dL_dwmap,
                         const float* dL_dcolor, int32_t dense_grad, float* dL_dcoeff, float* dL_dcolors,
                         void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image) return LSV2_EINVAL;
    return launch_render_backward(*view, *geom, *binning, *feat, *image, dL_dwmap, dL_dcolor, dense_grad,
                                  dL_dcoeff, dL_dcolors, (cudaStream_t)stream);
}

int lsv2_render_backward_rows(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                              const lsv2_features* feat, const lsv2_image* image, const float* dL_dwmap,
                              const float* dL_dcolor, int32_t row_stride, float* dL_drows, float* dL_dcolors,
                              void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image || row_stride <= 0) return LSV2_EINVAL;
    if (!feat->sparse_count || !feat->sparse_index) return LSV2_EINVAL;
    return launch_render_backward(*view, *geom, *binning, *feat, *image, dL_dwmap, dL_dcolor, 0, dL_drows,
                                  dL_dcolors, (cudaStream_t)stream, row_stride);
}

int lsv2_topk_soft_code_backward_rows(int64_t N, const float* logits, const float* dense,
                                      const uint32_t* sparse_count, const uint16_t* sparse_index,
                                      const float* d_rows, int32_t row_stride, float* d_logits, void* stream)
{
    if (N < 0) return LSV2_EINVAL;
    return launch_topk_soft_code_bwd_rows(N, logits, dense, sparse_count, sparse_index, d_rows, row_stride, d_logits,
                                          (cudaStream_t)stream);
}

int lsv2_render_backward_rgb(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                             const lsv2_image* image, const float* dL_dcolor, float* acc, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !image) return LSV2_EINVAL;
    return launch_render_backward_rgb(*view, *geom, *binning, *image, dL_dcolor, acc, (cudaStream_t)stream);
}

int lsv2_preprocess_backward(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom, const float* acc,
                             const lsv2_geom_grads* grads, void* stream)
{
    if (!view_ok(view) || !scene || !geom || !grads || scene->P < 0) return LSV2_EINVAL;
    return launch_preprocess_backward(*view, *scene, *geom, acc, *grads, (cudaStream_t)stream);
}

size_t lsv2_decode_scratch_bytes(int32_t levels, int32_t Lc, int32_t D)
{
    return decode_tc_scratch_bytes(levels, Lc, D);
}

int lsv2_decode(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap, const float* codebooks,
                int32_t normalize, int32_t precision, int32_t out_dtype, void* out, void* scratch,
                size_t scratch_bytes, void* stream)
{
    if (levels < 0 || Lc <= 0 || D <= 0 || HW < 0 || !wmap || !codebooks || !out) return LSV2_EINVAL;
    if (precision == LSV2_DECODE_FP32) {
        if (out_dtype != 0) return LSV2_EUNSUPPORTED;
        return launch_decode_fp32(levels, Lc, D, HW, wmap, codebooks, normalize, (float*)out, (cudaStream_t)stream);
    }
    if (precision == LSV2_DECODE_BF16)
        return launch_decode_tc(levels, Lc, D, HW, wmap, codebooks, normalize, out_dtype, out, scratch,
                                scratch_bytes, (cudaStream_t)stream);
    return LSV2_EINVAL;
}

int lsv2_decode_static(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap, const float* codebooks,
                       int32_t normalize, int32_t out_dtype, void* out, void* stream)
{
    if (levels < 0 || Lc <= 0 || D <= 0 || HW < 0 || !wmap || !codebooks || !out) return LSV2_EINVAL;
    if (out_dtype != 0 && out_dtype != 1) return LSV2_EINVAL;
    return launch_decode_tc_static(levels, Lc, D, HW, wmap, codebooks, normalize, out_dtype, out, (cudaStream_t)stream);
}

size_t lsv2_render_decode_scratch_bytes(int32_t levels) { return render_decode_scratch_bytes(levels); }

int lsv2_render_decode(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                       const lsv2_features* feat, const lsv2_image* image, const float* codebooks, int32_t levels,
                       int32_t Lc, int32_t D, int32_t normalize, int32_t out_dtype, void* out, void* scratch,
                       size_t scratch_bytes, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image || !codebooks || !out) return LSV2_EINVAL;
    if (!image->color || !image->final_T || !image->n_contrib || !binning->ranges || !binning->status) return LSV2_EINVAL;
    if ((out_dtype & ~LSV2_OUT_HWC) != 0 && (out_dtype & ~LSV2_OUT_HWC) != 1) return LSV2_EINVAL;
    return launch_render_decode(*view, *geom, *binning, *feat, *image, codebooks, levels, Lc, D, normalize,
                                out_dtype & 1, (out_dtype & LSV2_OUT_HWC) ? 1 : 0, out, scratch, scratch_bytes,
                                (cudaStream_t)stream);
}

size_t lsv2_render_query_scratch_bytes(int32_t levels) { return render_query_scratch_bytes(levels); }

int lsv2_render_query(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                      const lsv2_features* feat, const lsv2_image* image, const float* codebooks, int32_t levels,
                      int32_t Lc, int32_t D, const float* phrase_embeds, int32_t n_pos, int32_t n_neg, float* out,
                      void* scratch, size_t scratch_bytes, void* stream)
{
    if (!view_ok(view) || !geom || !binning || !feat || !image || !codebooks || !phrase_embeds || !out) return LSV2_EINVAL;
    if (!image->color || !image->final_T || !image->n_contrib || !binning->ranges || !binning->status) return LSV2_EINVAL;
    return launch_render_query(*view, *geom, *binning, *feat, *image, codebooks, levels, Lc, D, phrase_embeds, n_pos,
                               n_neg, out, scratch, scratch_bytes, (cudaStream_t)stream);
}

int lsv2_decode_backward(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap,
                         const float* codebooks, const float* dF, float* dW, float* dC, void* stream)
{
    if (levels < 0 || Lc <= 0 || D <= 0 || HW < 0 || !dF) return LSV2_EINVAL;
    if ((dW && !codebooks) || (dC && !wmap)) return LSV2_EINVAL;
    return launch_decode_backward(levels, Lc, D, HW, wmap, codebooks, dF, dW, dC, (cudaStream_t)stream);
}

int lsv2_topk_soft_code(int64_t N, int32_t C, int32_t k, const float* logits, float* dense, float* weights,
                        float* indices, void* stream)
{
    if (N < 0 || !logits) return LSV2_EINVAL;
    return launch_topk_soft_code(N, C, k, logits, dense, weights, indices, (cudaStream_t)stream);
}

size_t lsv2_knn_scratch_bytes(int32_t P) { return knn_scratch_bytes(P); }

int lsv2_dist2_knn3(int32_t P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                    void* stream)
{
    if (P < 0) return LSV2_EINVAL;
    return launch_dist2_knn3(P, points, mean_dist2, scratch, scratch_bytes, (cudaStream_t)stream);
}

int lsv2_topk_soft_code_sparse(int64_t N, int32_t k, const float* logits, float* dense, uint32_t* sparse_count,
                               uint16_t* sparse_index, float* sparse_weight, uint32_t* packed, void* stream)
{
    if (N < 0) return LSV2_EINVAL;
    return launch_topk_soft_code_sparse(N, k, logits, dense, sparse_count, sparse_index, sparse_weight, packed,
                                        (cudaStream_t)stream);
}

int lsv2_topk_soft_code_backward(int64_t N, int32_t C, const float* logits, const float* dense,
                                 const float* d_dense, float* d_logits, void* stream)
{
    if (N < 0 || (N > 0 && (!logits || !dense || !d_dense || !d_logits))) return LSV2_EINVAL;
    return launch_topk_soft_code_bwd(N, C, logits, dense, d_dense, d_logits, (cudaStream_t)stream);
}

int32_t lsv2_cos_loss_blocks(int64_t HW) { return (int32_t)((HW + 255) / 256); }

int lsv2_cos_loss_forward(int32_t D, int64_t HW, const float* F, const float* gt, const float* mask,
                          int32_t normalize, float* alpha_beta, float* partial, void* stream)
{
    if (D <= 0 || HW < 0 || (HW > 0 && (!F || !gt || !alpha_beta || !partial))) return LSV2_EINVAL;
    return launch_cos_loss_forward(D, HW, F, gt, mask, normalize, alpha_beta, partial, (cudaStream_t)stream);
}

int lsv2_cos_loss_backward(int32_t D, int64_t HW, const float* F, const float* gt, const float* alpha_beta,
                           const float* grad_out, float* dF, void* stream)
{
    if (D <= 0 || HW < 0 || (HW > 0 && (!F || !gt || !alpha_beta || !dF))) return LSV2_EINVAL;
    return launch_cos_loss_backward(D, HW, F, gt, alpha_beta, grad_out, dF, (cudaStream_t)stream);
}

int lsv2_gram_loss_forward(int32_t Lc, int64_t HW, const float* wmap, const float* h, const float* g2,
                           const float* mask, const float* G, int32_t normalize, float* u, float* alpha_beta,
                           float* partial, void* stream)
{
    if (HW < 0 || (HW > 0 && (!wmap || !h || !g2 || !G || !u || !alpha_beta || !partial))) return LSV2_EINVAL;
    return launch_gram_loss_forward(Lc, HW, wmap, h, g2, mask, G, normalize, u, alpha_beta, partial, nullptr, 0,
                                    (cudaStream_t)stream);
}

int lsv2_gram_loss_forward_seg(int32_t Lc, int64_t HW, int32_t nseg, const float* wmap, const float* CE, const float* e2,
                               const int32_t* seg, const float* G, int32_t normalize, float* u, float* alpha_beta,
                               float* partial, void* stream)
{
    if (HW < 0 || nseg <= 0 || (HW > 0 && (!wmap || !CE || !e2 || !seg || !G || !u || !alpha_beta || !partial)))
        return LSV2_EINVAL;
    return launch_gram_loss_forward(Lc, HW, wmap, CE, e2, nullptr, G, normalize, u, alpha_beta, partial, seg, nseg,
                                    (cudaStream_t)stream);
}

int lsv2_gram_loss_backward(int32_t Lc, int64_t HW, const float* wmap, const float* h, const float* u,
                            const float* alpha_beta, const float* grad_out, float* dw, float* aw, float* bw, void* stream)
{
    if (HW < 0 || (HW > 0 && (!wmap || !h || !u || !alpha_beta || !dw || !aw || !bw))) return LSV2_EINVAL;
    return launch_gram_loss_backward(Lc, HW, wmap, h, u, alpha_beta, grad_out, dw, aw, bw, nullptr, 0,
                                     (cudaStream_t)stream);
}

int lsv2_gram_loss_backward_seg(int32_t Lc, int64_t HW, int32_t nseg, const float* wmap, const float* CE,
                                const int32_t* seg, const float* u, const float* alpha_beta, const float* grad_out,
                                float* dw, float* aw, float* bw, void* stream)
{
    if (HW < 0 || nseg <= 0 || (HW > 0 && (!wmap || !CE || !seg || !u || !alpha_beta || !dw || !aw || !bw)))
        return LSV2_EINVAL;
    return launch_gram_loss_backward(Lc, HW, wmap, CE, u, alpha_beta, grad_out, dw, aw, bw, seg, nseg,
                                     (cudaStream_t)stream);
}

int lsv2_segment_sum(int32_t Lc, int64_t HW, int32_t nseg, const float* x, const int32_t* seg, float* out, void* stream)
{
    if (Lc < 0 || HW < 0 || nseg < 0 || (HW > 0 && Lc > 0 && nseg > 0 && (!x || !seg || !out))) return LSV2_EINVAL;
    return launch_segment_sum(Lc, HW, nseg, x, seg, out, (cudaStream_t)stream);
}

int lsv2_colsumsq(int32_t D, int64_t HW, const float* x, float* out, void* stream)
{
    if (D <= 0 || HW < 0 || (HW > 0 && (!x || !out))) return LSV2_EINVAL;
    return launch_colsumsq(D, HW, x, out, (cudaStream_t)stream);
}

int lsv2_adam_step(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2, double eps,
                   int64_t step, int32_t zero_grad, const float* step_scalars, void* stream)
{
    if (n < 0 || (step < 1 && !step_scalars) || (n > 0 && (!p || !g || !m || !v))) return LSV2_EINVAL;
    return launch_adam(p, g, m, v, n, lr, beta1, beta2, eps, step < 1 ? 1 : step, zero_grad, step_scalars,
                       (cudaStream_t)stream);
}

}  // extern "C"
