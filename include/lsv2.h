/*
 * lsv2.h -- C ABI of liblsv2.so: the B200 (sm_100a) implementation of the LangSplatV2
 * rasterization hot path (preprocess -> binning -> blend with sparse-coefficient
 * accumulation -> global codebook decode, and the coefficient/codebook backward).
 *
 * This is the drop-in boundary.  It replaces what the reference binds through
 *   `from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`
 *   (/root/reference/gaussian_renderer/__init__.py:15, settings :37-52, call :108-119),
 * i.e. the `_C.rasterize_gaussians` / `_C.rasterize_gaussians_backward` / `_C.mark_visible`
 * pybind entry points of the (un-vendored) submodule
 *   submodules/efficient-langsplat-rasterization @ 1305b7c23ff64092cd6c4b0e4996d6557e56c76a,
 * plus the Python decode helpers that every caller runs right after the op
 *   (eval_lerf.py:210-220, scene/gaussian_model.py:520-549, train.py:38-58).
 *
 * Conventions
 *   - plain C, no torch / pybind types; every pointer is a DEVICE pointer unless the name
 *     ends in `_host`; all tensors are dense, row-major, fp32 unless stated.
 *   - every call is asynchronous on `stream` (pass torch.cuda.current_stream().cuda_stream),
 *     allocates nothing and never synchronises; scratch comes from the caller.
 *   - return value: 0 on success, otherwise an LSV2_E* code (lsv2_strerror()).
 *   - instance buffers are capacity-bounded: `capacity` instances are addressable; the
 *     true count is written to status[LSV2_STATUS_NUM_RENDERED] on the device and
 *     status[LSV2_STATUS_OVERFLOW] is set when it exceeds `capacity` (the frame is then
 *     rendered from the first `capacity` emitted instances and must be re-run).  The
 *     reference instead copies the count to the host between two kernels (one blocking
 *     sync per frame); lsv2_preprocess + a host read + lsv2_bin reproduces that.
 */
#ifndef LSV2_H
#define LSV2_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSV2_ABI_VERSION 2
#define LSV2_TILE 16

/* error codes */
enum {
    LSV2_OK = 0,
    LSV2_EINVAL = 1,      /* bad argument combination (reference: Python Exception)          */
    LSV2_ECUDA = 2,       /* a CUDA launch / API call failed (lsv2_last_cuda_error())         */
    LSV2_ENOSPC = 3,      /* scratch too small                                                */
    LSV2_EUNSUPPORTED = 4 /* shape outside what the kernels are built for                     */
};

/* feature modes (reference: settings.include_feature / settings.quick_render,
 * gaussian_renderer/__init__.py:50-51,87-103) */
enum {
    LSV2_MODE_RGB = 0,     /* RGB only; weight map not produced                               */
    LSV2_MODE_DENSE = 1,   /* include_feature: language_feature_precomp f32[P][L]             */
    LSV2_MODE_QUICK = 2    /* quick_render: weights f32[P][K] + float-typed indices f32[P][K] */
};

/* status words (uint32 each) */
enum {
    LSV2_STATUS_NUM_RENDERED = 0,
    LSV2_STATUS_OVERFLOW = 1,
    LSV2_STATUS_BAD_INDEX = 2,   /* quick mode: an index outside [0,L) was seen and skipped */
    LSV2_STATUS_WORDS = 8
};

/* Per-view settings == GaussianRasterizationSettings (gaussian_renderer/__init__.py:37-52).
 * The four matrices/vectors stay on the device exactly as the reference passes them
 * (transposed / row-vector 4x4s, scene/cameras.py:55-58). */
typedef struct lsv2_view {
    int32_t width, height;
    float tanfovx, tanfovy;
    float scale_modifier;
    int32_t sh_degree;        /* active degree 0..3 */
    const float* viewmatrix;  /* f32[16] world_view_transform                 */
    const float* projmatrix;  /* f32[16] full_proj_transform                  */
    const float* campos;      /* f32[3]  camera_center                        */
    const float* bg;          /* f32[3]                                       */
} lsv2_view;

/* Borrowed Gaussian parameters == the tensor kwargs of GaussianRasterizer.forward
 * (gaussian_renderer/__init__.py:108-119).  Exactly one of shs / colors_precomp and exactly
 * one of (scales, rotations) / cov3d_precomp must be non-NULL. */
typedef struct lsv2_scene {
    int32_t P;
    int32_t sh_coeffs;           /* M in shs f32[P][M][3]; 0 when colors_precomp is used */
    const float* means3d;        /* f32[P][3] */
    const float* opacities;      /* f32[P]    */
    const float* scales;         /* f32[P][3] */
    const float* rotations;      /* f32[P][4] (w,x,y,z), used as given */
    const float* cov3d_precomp;  /* f32[P][6] */
    const float* shs;            /* f32[P][M][3] */
    const float* colors_precomp; /* f32[P][3] */
} lsv2_scene;

/* Per-Gaussian state written by lsv2_preprocess (the reference's opaque geomBuffer). */
typedef struct lsv2_geom {
    int32_t* radii;          /* i32[P]   (also the op's 3rd return value)         */
    float* xy;               /* f32[P][2] pixel-space mean                         */
    float* depth;            /* f32[P]                                             */
    float* conic_opacity;    /* f32[P][4] (conic a,b,c, opacity)                   */
    float* rgb;              /* f32[P][3]                                          */
    uint8_t* clamped;        /* u8[P][3]  SH colour clamped at 0 (may be NULL)     */
    uint32_t* tiles_touched; /* u32[P]                                             */
    uint32_t* offsets;       /* u32[P]    inclusive scan of tiles_touched          */
    float* cov3d;            /* f32[P][6] or NULL                                  */
    int32_t* rect;           /* i32[P][4] (minx,miny,maxx,maxy) or NULL            */
    float* blend_rec;        /* f32[P][12] packed per-Gaussian record the blend reads:
                                (x, y, -a/2, -b), (-c/2, power cut-off, opacity, r), (g, b, dy, dx) with (dy, dx) the
                                half-extents of the alpha >= 1/255 ellipse's bounding box (dy < 0: never contributes) */
} lsv2_geom;

/* Per-instance state (the reference's binningBuffer) + per-tile ranges. */
typedef struct lsv2_binning {
    uint64_t capacity;       /* instances addressable in each array below          */
    uint64_t* keys;          /* u64[capacity]  sorted (tile<<32 | Gaussian id), in the order of the reference's
                                (tile<<32 | depth bits) key == (keys[i] & ~0xffffffff) | bits(depth[id])       */
    uint32_t* point_list;    /* u32[capacity]  sorted Gaussian ids                 */
    uint64_t* keys_alt;      /* u64[capacity]  ping-pong buffer of the sort        */
    uint32_t* vals_alt;      /* u32[capacity]  unused since the packed sort; may be NULL */
    uint32_t* ranges;        /* u32[tiles][2]  (start,end) per 16x16 tile          */
    void* scratch;           /* lsv2_sort_scratch_bytes(capacity, P)               */
    size_t scratch_bytes;
    uint32_t* status;        /* u32[LSV2_STATUS_WORDS]                             */
} lsv2_binning;

/* Sparse-coefficient inputs of the blend. */
typedef struct lsv2_features {
    int32_t mode;                /* LSV2_MODE_*                                            */
    int32_t L;                   /* weight-map channels (64 train, 192 quick)              */
    int32_t K;                   /* quick: entries per Gaussian                            */
    const float* coeff_dense;    /* DENSE: f32[P][L]                                       */
    const float* weights_quick;  /* QUICK: f32[P][K]                                       */
    const float* indices_quick;  /* QUICK: f32[P][K] float-typed integers in [0,L)         */
    /* DENSE only: compacted rows written by lsv2_compact_coeff (scratch, may be NULL to
     * let lsv2_render_forward read the dense rows directly) */
    uint32_t* sparse_count;      /* u32[P]  non-zeros per row, 0xFFFFFFFF = row too dense  */
    uint16_t* sparse_index;      /* u16[P][LSV2_SPARSE_CAP]                                */
    float* sparse_weight;        /* f32[P][LSV2_SPARSE_CAP]                                */
    /* Tensor-core blend only (lsv2_render_forward_tc, lsv2_render_decode): the coefficients as the kernel's B operand
     * wants them, written by lsv2_pack_coeff -- per Gaussian LSV2_PACKED_WORDS u32: 16 x (fp16 head | fp16 remainder
     * << 16), 16 channel bytes, the entry count (0xFFFFFFFF: read the dense row).  Duplicate channels are merged and
     * out-of-range ones dropped (and flagged) at pack time. */
    uint32_t* packed;            /* u32[P][LSV2_PACKED_WORDS]                              */
} lsv2_features;
#define LSV2_SPARSE_CAP 16
#define LSV2_PACKED_WORDS 24

/* Per-pixel outputs (rendered image, weight map, and the reference's imgBuffer). */
typedef struct lsv2_image {
    float* color;         /* f32[3][H][W]  = sum c*alpha*T + T*bg               */
    float* wmap;          /* f32[L][H][W]  no background term; NULL in RGB mode  */
    float* final_T;       /* f32[H][W]                                           */
    uint32_t* n_contrib;  /* u32[H][W]     index (1-based) of the last contributor in the tile list */
} lsv2_image;

/* ---- library ------------------------------------------------------------------- */
int lsv2_abi_version(void);
const char* lsv2_strerror(int code);
const char* lsv2_last_cuda_error(void);   /* text of the last CUDA failure on this thread */

/* ---- sizes --------------------------------------------------------------------- */
size_t lsv2_sort_scratch_bytes(uint64_t capacity, int32_t P);

/* ---- forward stages (each replaces a stage of _C.rasterize_gaussians) ------------ */

/* preprocessCUDA + InclusiveSum: fills geom, status[NUM_RENDERED].  `scratch` is
 * binning->scratch (only scratch/scratch_bytes/status of `binning` are used). */
int lsv2_preprocess(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom,
                    const lsv2_binning* binning, void* stream);

/* duplicateWithKeys + stable radix sort on the (tile|depth) key + identifyTileRanges. */
int lsv2_bin(const lsv2_view* view, int32_t P, const lsv2_geom* geom,
             const lsv2_binning* binning, void* stream);

/* DENSE mode helper: compact rows of coeff_dense for Gaussians with radii > 0. */
int lsv2_compact_coeff(int32_t P, const lsv2_geom* geom, const lsv2_features* feat, void* stream);

/* renderCUDA forward. */
int lsv2_render_forward(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                        const lsv2_features* feat, const lsv2_image* image, void* stream);

/* renderCUDA forward with the sparse-coefficient accumulation on the tensor cores (the blend half of
 * lsv2_render_decode, render_decode.cu): per batch of 16 depth-sorted instances the per-pixel blending weights
 * alpha*T and the instances' coefficients are split into an fp16 head and an fp16 remainder and
 * W += A_hi.S_hi + A_lo.S_hi + A_hi.S_lo accumulates in TMEM (fp32); the fp32 [L][H][W] map is written once per
 * tile.  color / final_T / n_contrib are bit-identical to lsv2_render_forward (same per-pixel arithmetic); the
 * weight map agrees with it to <= 1e-4 of its per-pixel norm (stated tolerance of the fp32 blend).  QUICK or DENSE
 * with the compacted rows of lsv2_compact_coeff; L a multiple of 64, <= 192, K <= LSV2_SPARSE_CAP; otherwise
 * LSV2_EUNSUPPORTED (use lsv2_render_forward).  Uses the first 256 bytes of binning->scratch. */
int lsv2_render_forward_tc(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                           const lsv2_features* feat, const lsv2_image* image, void* stream);

/* Fills feat->packed from the QUICK inputs (weights_quick / indices_quick) or, in DENSE mode, from the compacted rows
 * of lsv2_compact_coeff (call it first; geom->radii selects the rows that exist).  An index outside [0, L) sets
 * status[LSV2_STATUS_BAD_INDEX] (status may be NULL).  The table depends only on the coefficients, not on the view:
 * a resident scene packs once. */
int lsv2_pack_coeff(int32_t P, const lsv2_geom* geom, const lsv2_features* feat, uint32_t* status, void* stream);

/* preprocess + bin + compact + blend back to back, no host involvement (CUDA-graph capturable).  The blend is
 * lsv2_render_forward_tc where it applies, lsv2_render_forward otherwise. */
int lsv2_forward(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom,
                 const lsv2_binning* binning, const lsv2_features* feat, const lsv2_image* image,
                 void* stream);

/* _C.mark_visible: present[i] = (view-space z > 0.2). */
int lsv2_mark_visible(int32_t P, const float* means3d, const float* viewmatrix, uint8_t* present,
                      void* stream);

/* ---- backward (replaces _C.rasterize_gaussians_backward for the frozen-geometry
 *      feature stage: scene/gaussian_model.py:238-243, train.py:173) ------------------
 * dL_dcoeff f32[P][L] must be zero-initialised by the caller; it receives
 *   sum_pix alpha*T * dL_dwmap[c][pix]
 * on every channel (dense_grad != 0) or only on each row's non-zero support (== what
 * survives autograd through torch.where in utils/vq_utils.py:20).  dL_dcolors f32[P][3]
 * (zero-initialised) receives the colour gradient when dL_dcolor is non-NULL. */
int lsv2_render_backward(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                         const lsv2_features* feat, const lsv2_image* image,
                         const float* dL_dwmap, const float* dL_dcolor, int32_t dense_grad,
                         float* dL_dcoeff, float* dL_dcolors, void* stream);

/* The same with the gradient in the order of the COMPACTED rows (support-only by construction): dL_drows
 * f32[P][row_stride], zero-initialised by the caller, entry e of row g receives the sum for channel
 * sparse_index[g][e], e < sparse_count[g] (entries e >= row_stride are dropped; rows with more than 16 non-zeros have
 * no row form and get nothing -- lsv2_topk_soft_code_sparse never emits one).  For top-4 that is 16 B per Gaussian
 * instead of 256 B: less to clear, less for lsv2_topk_soft_code_backward_rows to read, and all a data-parallel
 * trainer has to exchange (the top-k backward is linear in this gradient, so ranks sum the rows and run it once). */
int lsv2_render_backward_rows(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                              const lsv2_features* feat, const lsv2_image* image, const float* dL_dwmap,
                              const float* dL_dcolor, int32_t row_stride, float* dL_drows, float* dL_dcolors,
                              void* stream);

/* ---- RGB-stage backward (replaces _C.rasterize_gaussians_backward for train.py WITHOUT --include_feature:
 *      train.py:169-172 l1 + ssim on the RGB render, :247-258 optimizer / densification;
 *      gradients reach means3D, means2D, shs / colors_precomp, opacities, scales, rotations / cov3D_precomp
 *      exactly as gaussian_renderer/__init__.py:108-119 passes them) ----------------------------------------
 * Step 1, lsv2_render_backward_rgb: per-Gaussian sums over all pixels the Gaussian contributed to,
 *   acc f32[P][LSV2_GEOM_ACC], ZERO-initialised by the caller:
 *   [0..2] sum w dL/dC_rgb            (w = alpha T)            -> dL/d rgb
 *   [3]    sum q                      (q = G dL/dalpha)        -> dL/d opacity
 *   [4..5] sum q dx, sum q dy         (d = mean2D - pixel)
 *   [6..8] sum q dx^2, q dx dy, q dy^2
 *   image->color must be the forward's output (background included), image->n_contrib its contributor counts.
 * Step 2, lsv2_preprocess_backward: pulls acc through conic -> cov2D -> cov3D -> scale / rotation, the projection
 *   and SH -> RGB.  Every non-NULL array of lsv2_geom_grads is fully written (zeros for culled Gaussians).
 *   means2d receives (dL/dpx * W/2, dL/dpy * H/2, 0): the gradient with respect to the NDC-scaled screen point,
 *   which is what the callers read as a densification statistic (scene/gaussian_model.py:506-508). */
#define LSV2_GEOM_ACC 12
typedef struct lsv2_geom_grads {
    float* means3d;     /* f32[P][3]     */
    float* means2d;     /* f32[P][3]     */
    float* opacities;   /* f32[P]        */
    float* scales;      /* f32[P][3]     (scales / rotations path)  */
    float* rotations;   /* f32[P][4]     16-byte aligned            */
    float* cov3d;       /* f32[P][6]     (cov3d_precomp path)       */
    float* shs;         /* f32[P][M][3]  (shs path)                 */
    float* colors;      /* f32[P][3]     (colors_precomp path)      */
} lsv2_geom_grads;
int lsv2_render_backward_rgb(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                             const lsv2_image* image, const float* dL_dcolor, float* acc, void* stream);
int lsv2_preprocess_backward(const lsv2_view* view, const lsv2_scene* scene, const lsv2_geom* geom,
                             const float* acc, const lsv2_geom_grads* grads, void* stream);

/* ---- global codebook decode (eval_lerf.py:214-218; scene/gaussian_model.py:533-549) ---
 * wmap f32[levels*Lc][HW], codebooks f32[levels][Lc][D] -> out [levels][D][HW]
 * out = C_l^T W_l, optionally divided by (||.||_2 over D + 1e-10).
 * precision: LSV2_DECODE_FP32 (CUDA cores, fp32 FMA) or LSV2_DECODE_BF16 (tcgen05 tensor
 * cores, bf16 operands, fp32 accumulate in TMEM).  out_dtype: 0 = f32, 1 = bf16. */
enum { LSV2_DECODE_FP32 = 0, LSV2_DECODE_BF16 = 1 };
size_t lsv2_decode_scratch_bytes(int32_t levels, int32_t Lc, int32_t D);
int lsv2_decode(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap,
                const float* codebooks, int32_t normalize, int32_t precision, int32_t out_dtype,
                void* out, void* scratch, size_t scratch_bytes, void* stream);

/* ---- lsv2_decode without caller scratch and without any allocation -------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same computation as lsv2_decode(..., precision = LSV2_DECODE_BF16, ...): tensor-core decode, Lc == 64, D == 512,
 * out_dtype 0 = f32 / 1 = bf16; the caller still passes wmap, codebooks and out.  The codebook image the kernel
 * needs (64 KB per level) lives in a static device array inside the library, LSV2_DECODE_STATIC_MAX_LEVELS levels
 * = 192 KB per GPU, held from library load to process exit.  levels > LSV2_DECODE_STATIC_MAX_LEVELS returns
 * LSV2_ENOSPC; other shapes LSV2_EUNSUPPORTED.  Calls no cudaMalloc / cudaFree / allocating library routine.
 * The array is shared by every call on a device and keeps the last call's contents (each call rewrites all it
 * reads, nothing needs clearing): calls that overlap on the GPU -- different streams with no ordering between
 * them -- race on it; serialise them (one stream, or events). */
#define LSV2_DECODE_STATIC_MAX_LEVELS 3
int lsv2_decode_static(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap,
                       const float* codebooks, int32_t normalize, int32_t out_dtype, void* out, void* stream);

/* ---- fused blend + decode (the serving path: eval_lerf.py:321-352 renders, then decodes) ----
 * renderCUDA forward and the codebook decode in ONE kernel: the blend's sparse-coefficient
 * accumulation runs on the tensor cores (alpha*T and coefficients as fp16 head + fp16 remainder,
 * three MMAs per batch, fp32 accumulate in TMEM: fp32-blend accuracy) and the L-channel weight map
 * goes straight from TMEM into the decode contraction (fp16 operands, fp32 accumulate; the L2 norm
 * comes from the Gram matrix C C^T before the decode), so it never exists in HBM.  The decode of
 * tile i runs under the blend of tile i+1; features leave through TMA tensor stores when the row
 * pitch W*sizeof(out) is a multiple of 16 bytes (per-thread stores otherwise).  Same inputs as lsv2_render_forward (QUICK, or DENSE with the
 * compacted rows of lsv2_compact_coeff) + codebooks f32[levels][Lc][D]; feat->L must equal
 * levels*Lc.  image->color / final_T / n_contrib are written as by lsv2_render_forward
 * (bit-identical); image->wmap may be NULL (if not, it receives the TMEM weight map, fp32).
 * out: [levels][D][H][W], out_dtype 0 = f32, 1 = bf16; OR-ing LSV2_OUT_HWC gives the consumers' channels-last
 * layout [levels][H][W][D] (eval/openclip_encoder.py:116 takes the map that way).  Supported: Lc = 64, D = 512,
 * levels <= 3, K <= LSV2_SPARSE_CAP, feat->packed set; otherwise LSV2_EUNSUPPORTED (use the two-call path). */
#define LSV2_OUT_HWC 2
size_t lsv2_render_decode_scratch_bytes(int32_t levels);
int lsv2_render_decode(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                       const lsv2_features* feat, const lsv2_image* image, const float* codebooks,
                       int32_t levels, int32_t Lc, int32_t D, int32_t normalize, int32_t out_dtype,
                       void* out, void* scratch, size_t scratch_bytes, void* stream);

/* ---- fused blend + relevancy query (eval_lerf.py:351-352 / backend_renderer.py:204-210 call
 *      OpenCLIPNetwork.get_max_across_quick, eval/openclip_encoder.py:114-138, on the decoded map) -------------
 * Same blend; instead of the 512-d map the kernel produces, per level and positive phrase,
 *   min over negatives of softmax(10 * [sim_pos, sim_neg])[0],  sim = <F / (|F| + 1e-10), e>,
 * from sim = w^T (C_l E^T) / |C_l^T w|: the feature map is never formed or written (levels*n_pos*H*W*4 bytes leave the
 * chip instead of levels*512*H*W*4).  phrase_embeds f32[n_pos + n_neg][D], positives first, n_pos + n_neg <= 64;
 * out f32[levels][n_pos][H][W].  image->color / final_T / n_contrib as lsv2_render_forward. */
size_t lsv2_render_query_scratch_bytes(int32_t levels);
int lsv2_render_query(const lsv2_view* view, const lsv2_geom* geom, const lsv2_binning* binning,
                      const lsv2_features* feat, const lsv2_image* image, const float* codebooks, int32_t levels,
                      int32_t Lc, int32_t D, const float* phrase_embeds, int32_t n_pos, int32_t n_neg, float* out,
                      void* scratch, size_t scratch_bytes, void* stream);

/* autograd of the un-normalised decode (scene/gaussian_model.py:538):
 *   dW[l*Lc+k][p] = sum_d C[l][k][d] dF[l][d][p]   (f32[levels*Lc][HW])
 *   dC[l][k][d]  += sum_p W[l*Lc+k][p] dF[l][d][p] (f32[levels][Lc][D], zero-initialised) */
int lsv2_decode_backward(int32_t levels, int32_t Lc, int32_t D, int64_t HW, const float* wmap,
                         const float* codebooks, const float* dF, float* dW, float* dC,
                         void* stream);

/* ---- sparse-coefficient producer (utils/vq_utils.py:9-40) ---------------------------
 * logits f32[N][C] -> softmax, keep top-k, renormalise by (sum + 1e-10).
 * dense f32[N][C] (may be NULL); weights f32[N][k] / indices f32[N][k] in ascending channel
 * order (may be NULL). */
int lsv2_topk_soft_code(int64_t N, int32_t C, int32_t k, const float* logits, float* dense,
                        float* weights, float* indices, void* stream);

/* simple_knn._C.distCUDA2 (submodules/simple-knn/spatial.cu:15-25 -> simple_knn.cu:148-220; the one call site is
 * scene/gaussian_model.py:194): mean_dist2[i] = mean of the squared distances from points[i] (f32[P][3]) to its three
 * nearest other points (fewer than three: FLT_MAX stays in the sum, as there).  Same values as the reference bit for bit (the search
 * only prunes).  No host synchronisation; scratch >= lsv2_knn_scratch_bytes(P). */
size_t lsv2_knn_scratch_bytes(int32_t P);
int lsv2_dist2_knn3(int32_t P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes,
                    void* stream);

/* The same producer for 64 codes per row, emitting in the same pass what the blend reads (replaces
 * utils/vq_utils.py:9-40 + the compaction / packing of a training iteration): the dense row (may be NULL), the
 * compacted row of lsv2_compact_coeff (sparse_count u32[N], sparse_index u16[N][16], sparse_weight f32[N][16]) and the
 * packed row of lsv2_pack_coeff (packed u32[N][LSV2_PACKED_WORDS]).  k <= 16. */
int lsv2_topk_soft_code_sparse(int64_t N, int32_t k, const float* logits, float* dense, uint32_t* sparse_count,
                               uint16_t* sparse_index, float* sparse_weight, uint32_t* packed, void* stream);

/* Backward of the dense form (autograd of utils/vq_utils.py:9-24, reached from
 * scene/gaussian_model.py:510-518 during train.py:173): `dense` is the forward's output
 * (its non-zeros are the top-k mask), d_dense = dL/d dense, d_logits = dL/d logits, all f32[N][C]. */
int lsv2_topk_soft_code_backward(int64_t N, int32_t C, const float* logits, const float* dense,
                                 const float* d_dense, float* d_logits, void* stream);

/* ... with the upstream gradient in row form (lsv2_render_backward_rows): 64 codes per row; logits / dense /
 * d_logits f32[N][64], sparse_count u32[N], sparse_index u16[N][16], d_rows f32[N][row_stride]. */
int lsv2_topk_soft_code_backward_rows(int64_t N, const float* logits, const float* dense,
                                      const uint32_t* sparse_count, const uint16_t* sparse_index,
                                      const float* d_rows, int32_t row_stride, float* d_logits, void* stream);

/* ---- feature-stage loss (train.py:157-164, utils/loss_utils.py:24-25) -----------------
 * loss = 1 - mean_p cos(a_p, b_p),  a = mask * F / (|F| + 1e-10) (normalize != 0) or mask * F,  b = mask * gt,
 * cos as torch.nn.functional.cosine_similarity (each norm clamped at 1e-8).  F, gt: f32[D][HW]; mask f32[HW]
 * or NULL (all ones).  forward: partial f32[lsv2_cos_loss_blocks(HW)] receives per-block sums of cos_p
 * (loss = 1 - sum(partial) / HW) and alpha_beta f32[2][HW] the per-pixel scalars of the gradient.
 * backward: dF = grad_out * dL/dF (grad_out: device scalar or NULL = 1). */
int32_t lsv2_cos_loss_blocks(int64_t HW);
int lsv2_cos_loss_forward(int32_t D, int64_t HW, const float* F, const float* gt, const float* mask,
                          int32_t normalize, float* alpha_beta, float* partial, void* stream);
int lsv2_cos_loss_backward(int32_t D, int64_t HW, const float* F, const float* gt, const float* alpha_beta,
                           const float* grad_out, float* dF, void* stream);

/* The same loss straight from the weight map, the feature map never materialised (decode + normalise + cos loss of
 * train.py:156-164 in one step; scene/gaussian_model.py:533-543 for one layer).  With F = C^T w:
 *   h = C gt  f32[Lc][HW]  (lsv2_decode_backward(..., dF = gt, dW = h)),  G = C C^T f32[Lc][Lc],  g2 = |gt_p|^2 f32[HW]
 *   (lsv2_colsumsq).  forward: u = G w f32[Lc][HW], alpha_beta f32[2][HW], partial as lsv2_cos_loss_forward.
 *   backward: dw = dL/d wmap, aw = (go alpha) w, bw = (go beta) w, all f32[Lc][HW];
 *   dL/dC = aw gt^T + (bw w^T) C  (two calls of lsv2_decode_backward's dC form).  Lc must be 64. */
int lsv2_gram_loss_forward(int32_t Lc, int64_t HW, const float* wmap, const float* h, const float* g2,
                           const float* mask, const float* G, int32_t normalize, float* u, float* alpha_beta,
                           float* partial, void* stream);
int lsv2_gram_loss_backward(int32_t Lc, int64_t HW, const float* wmap, const float* h, const float* u,
                            const float* alpha_beta, const float* grad_out, float* dw, float* aw, float* bw, void* stream);
int lsv2_colsumsq(int32_t D, int64_t HW, const float* x, float* out, void* stream);

/* The same two calls for a ground truth given as gt = table[seg] (/root/reference/scene/cameras.py:59-96: the
 * loader's form before it expands it; seg < 0 = unlabelled = masked out): CE = C table^T f32[Lc][nseg] stands in for
 * h (h_p = CE[:, seg_p]), e2 f32[nseg] = |table_s|^2 for g2, seg i32[HW] for the mask. */
int lsv2_gram_loss_forward_seg(int32_t Lc, int64_t HW, int32_t nseg, const float* wmap, const float* CE, const float* e2,
                               const int32_t* seg, const float* G, int32_t normalize, float* u, float* alpha_beta,
                               float* partial, void* stream);
int lsv2_gram_loss_backward_seg(int32_t Lc, int64_t HW, int32_t nseg, const float* wmap, const float* CE,
                                const int32_t* seg, const float* u, const float* alpha_beta, const float* grad_out,
                                float* dw, float* aw, float* bw, void* stream);

/* out[k][s] += sum_{p : seg[p] == s} x[k][p]   (x f32[Lc][HW], seg i32[HW] (< 0: skipped), out f32[Lc][nseg], caller
 * zeroes out).  The feature stage's ground truth is table[seg] (/root/reference/scene/cameras.py:59-96), so the
 * codebook gradient (alpha w) gt^T of lsv2_gram_loss_backward equals (segment sums of alpha w) . table. */
int lsv2_segment_sum(int32_t Lc, int64_t HW, int32_t nseg, const float* x, const int32_t* seg, float* out, void* stream);

/* ---- optimizer step (scene/gaussian_model.py:255 torch.optim.Adam(l, lr=0.0, eps=1e-15); train.py:261-263) ------
 * One Adam step on the flat range p[0, n) with gradient g and moments m, v (all f32, 16-byte aligned), in torch's
 * single-tensor operation order, no weight decay / amsgrad; `step` is the 1-based step count of this range.
 * zero_grad != 0 clears g in the same pass.  A group of 4 elements whose g, m and v are all zero is skipped, which
 * is exactly what the dense update does to it.  step_scalars (device f32[2], may be NULL): when given, the kernel
 * reads {lr / (1 - beta1^step), sqrt(1 - beta2^step)} from it instead of deriving them from lr / step on the host, so
 * a captured CUDA graph can be replayed with the next step's values. */
int lsv2_adam_step(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2, double eps,
                   int64_t step, int32_t zero_grad, const float* step_scalars, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LSV2_H */
