// render_decode.cu -- K6+K7 fused: alpha blending whose sparse-coefficient accumulation runs on the 5th-gen tensor
// cores, overlapped in the same CTA with the global codebook decode + L2 normalisation of the previous work unit,
// whose 512-d features go from TMEM through registers straight to HBM as full 128-byte lines.
// The L-channel weight map never exists in HBM (nor in shared memory as read-modify-write accumulators).
// Replaces renderCUDA (forward) + the einsum/norm/divide every caller runs after it
// (/root/reference/gaussian_renderer/__init__.py:108-119 ; eval_lerf.py:214-218 ; SURVEY.md rows R8+R9).
// Three paths share the kernel (template PATH): weight map only (the 3-tuple op), decode, relevancy query.
//
// Persistent CTAs (one per SM); work unit = a 32 x 8 pixel band (the 8-row halves of two horizontally adjacent 16x16
// tiles, each half with its own tile's depth-sorted list), handed out by an atomic counter.  Warp roles (fd_common.cuh):
//   warps 0-7   blend: one thread per pixel; warp = one 8 x 4 block of a half band; thread = A-tile row = TMEM lane of
//               the blend MMA.  At the end of a unit the same threads move the weight tile W (TMEM, fp32) into the
//               decode's fp16 operand (pass A) and compute the per-pixel 1/norm from the Gram product (pass B).
//   warps 12-13 producers, one per tile list: gather the packed blend records and packed coefficient rows of 16
//               depth-sorted instances per batch (ids 8 batches ahead, L2 prefetch 6 ahead, registers 2 ahead), write
//               the header + per-block cut-off masks, scatter each instance's (channel, fp16 head | fp16 remainder)
//               pairs into a zeroed B tile  S[L channels][hi 16 | lo 16]  (K-major, 64-byte swizzle).
//   warp 14     MB: issues the blend MMAs (converged warp, elect.sync); owns the TMEM allocation.
//   warp 15     MD: issues the Gram and decode MMAs and the bulk copies of the codebook chunks.
//   warps 8-11  epilogue: F slots (TMEM) -> registers -> x 1/norm -> st.global.cs, lane = x: every store instruction
//               writes one full 128-byte line of one feature plane (or 128 contiguous bytes of a pixel, channels-last).
//
// Blend, per batch:  the pixel threads evaluate alpha and T exactly as k_render_forward does (same fp32 operations,
// same thresholds => identical n_contrib / final_T / RGB), walking only the instances whose alpha >= 1/255 box reaches
// their warp's block, and store the blending weight alpha*T of each hit as fp16 head + fp16 remainder into the A tile
// [128 px][hi 16 | lo 16] of their half;
//   W[128 x L] += A_hi.S_hi + A_lo.S_hi + A_hi.S_lo        (M = 128, N = L, three K = 16 MMAs, fp32 in TMEM)
// which carries ~22 significant bits per factor: the weight map meets the fp32-blend tolerance (1e-4 of its norm).
// End of unit:  W -> fp16 rows of Wdec[level][256 px][64 codes];  U_l = Wdec_l.(G_hi + G_lo) with G_l = s C_l C_l^T
// gives |F_l|^2 = w^T G w per pixel; the blend warps move on to the next unit while
// Decode:  F[128 px x 64 features] = Wdec_l[128 px x 64] . C_l[64 x 64 features]  runs chunk by chunk into two TMEM
// slots (the 128 columns the two W halves leave free), codebook chunks streamed from an L2-resident fp16 image.
// The history of this kernel (what was measured and dropped: TMA tensor stores, transposed decode, unrolled role
// loops, divergent MMA issue, look-ahead across units) is in DESIGN.md 5.1 / 5.2 and profiles/r02*.
#include <string.h>

#include "fd_common.cuh"

// -DFD_WATCHDOG (debug library liblsv2_wd.so only, `make wd`): every mbarrier wait that spins for more than ~0.2 s
// writes (line, block, warp, parity) once into a host-mapped buffer (lsv2_debug_set_watchdog) and keeps waiting, so a
// deadlocked launch can be read out from the host while it hangs: scripts/fused_watchdog.py.
#ifdef FD_WATCHDOG
namespace lsv2 {
__device__ unsigned int* g_fd_wd = nullptr;          // [0] = records written, then 4 words per record
__device__ __noinline__ void fd_wd_report(int line, uint32_t parity)
{
    unsigned int* w = g_fd_wd;
    if (!w) return;
    const unsigned int k = atomicAdd(w, 1u);
    if (k < 1000u) {
        w[4 + 4 * k + 0] = (unsigned int)line;
        w[4 + 4 * k + 1] = blockIdx.x;
        w[4 + 4 * k + 2] = threadIdx.x >> 5;
        w[4 + 4 * k + 3] = parity;
        __threadfence_system();
    }
}
__device__ __forceinline__ void fd_wait_wd(uint64_t* bar, uint32_t parity, int line)
{
    long long t0 = clock64();
    bool told = false;
    while (!mbar_try_wait(bar, parity)) {
        if (!told && clock64() - t0 > 400000000LL) { told = true; if ((threadIdx.x & 31) == 0) fd_wd_report(line, parity); }
        __nanosleep(64);
    }
}
}  // namespace lsv2
extern "C" int lsv2_debug_set_watchdog(void* mapped_words)
{
    return cudaMemcpyToSymbol(lsv2::g_fd_wd, &mapped_words, sizeof(void*)) == cudaSuccess ? 0 : 2;
}
#define FD_WAIT(bar, par) fd_wait_wd(bar, par, __LINE__)
#define FD_WAIT_SLEEP(bar, par) fd_wait_wd(bar, par, __LINE__)
#else
#define FD_WAIT(bar, par) mbar_wait(bar, par)
#define FD_WAIT_SLEEP(bar, par) mbar_wait_sleep(bar, par)
#endif

// -DFD_PROF: per-CTA cycle counters of where each warp role waits (debug library liblsv2_prof.so only; the
// counters land behind the Gram images in the scratch buffer, see scripts/fused_prof.py)
#ifdef FD_PROF
#define FD_PROF_DECL long long prof_t0 = 0; long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define FD_TIC() prof_t0 = clock64()
#define FD_TOC(i) do { const long long _t = clock64(); prof_acc[i] += _t - prof_t0; prof_t0 = _t; } while (0)
#define FD_PROF_FLUSH(role) do { if (a.prof) for (int _i = 0; _i < 8; ++_i) a.prof[((size_t)blockIdx.x * 6 + (role)) * 8 + _i] = prof_acc[_i]; } while (0)
#define FD_PROFB_DECL long long profb_t0 = 0; long long profb_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define FD_TICB() profb_t0 = clock64()
#define FD_TOCB(i) do { const long long _t = clock64(); profb_acc[i] += _t - profb_t0; profb_t0 = _t; } while (0)
#define FD_PROFB_FLUSH(role) do { if (a.prof) for (int _i = 0; _i < 8; ++_i) a.prof[((size_t)blockIdx.x * 6 + (role)) * 8 + _i] = profb_acc[_i]; } while (0)
#elif defined(FD_WATCHDOG)
// heartbeat: the last phase marker every warp passed, [CTA][warp], readable from the host while the kernel hangs
#define FD_WD_MARK(code) do { unsigned int* _w = g_fd_wd; if (_w && (threadIdx.x & 31) == 0) \
        *(volatile unsigned int*)(_w + 4100 + blockIdx.x * 16 + (threadIdx.x >> 5)) = (unsigned int)(code); } while (0)
#define FD_PROFB_DECL
#define FD_TICB() FD_WD_MARK(100)
#define FD_TOCB(i) FD_WD_MARK(101 + (i))
#define FD_PROFB_FLUSH(role)
#define FD_PROF_DECL
#define FD_TIC() FD_WD_MARK(1)
#define FD_TOC(i) FD_WD_MARK(2 + (i))
#define FD_PROF_FLUSH(role) FD_WD_MARK(99)
#else
#define FD_PROFB_DECL
#define FD_TICB()
#define FD_TOCB(i)
#define FD_PROFB_FLUSH(role)
#define FD_PROF_DECL
#define FD_TIC()
#define FD_TOC(i)
#define FD_PROF_FLUSH(role)
#endif

namespace lsv2 {

// codebooks f32 [levels][64][512] -> fp16 image [levels][512 rows n][64 k], 128B-swizzled (A operand of the decode)
__global__ void k_fd_prep_b(int levels, const float* __restrict__ codebooks, uint8_t* __restrict__ bimg)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= levels * FD_D * 8) return;
    const int chunk = t & 7, n = (t >> 3) % FD_D, l = t / (FD_D * 8);
    const float* src = codebooks + ((size_t)l * FD_LC + chunk * 8) * FD_D + n;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = pack_f16x2(src[(size_t)(2 * i) * FD_D], src[(size_t)(2 * i + 1) * FD_D]);
    *reinterpret_cast<uint4*>(bimg + (size_t)l * (FD_D * 128) + sw128_offset(n, chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
}

// Gram images.  block (n, l), thread k:  G[n][k] = s * <C16_n, C16_k> with C16 the fp16-rounded codebook (what the
// decode MMA multiplies by), s = 1024 / max_n |C16_n|^2; stored as fp16 head (rows 0-63) + remainder (rows 64-127).
__global__ void __launch_bounds__(64) k_fd_prep_g(const float* __restrict__ codebooks, uint8_t* __restrict__ gimg,
                                                 float* __restrict__ gscale_inv)
{
    __shared__ float s_row[FD_D];
    __shared__ float s_v[FD_LC];
    __shared__ float s_red[2];
    const int n = blockIdx.x, l = blockIdx.y, k = threadIdx.x;
    const float* cb = codebooks + (size_t)l * FD_LC * FD_D;
    for (int d = k; d < FD_D; d += 64) s_row[d] = __half2float(__float2half_rn(cb[(size_t)n * FD_D + d]));
    __syncthreads();
    float dot = 0.0f, nrm = 0.0f;
    const float4* rk = reinterpret_cast<const float4*>(cb + (size_t)k * FD_D);
    for (int d4 = 0; d4 < FD_D / 4; ++d4) {
        const float4 v = __ldg(rk + d4);
        const float c0 = __half2float(__float2half_rn(v.x)), c1 = __half2float(__float2half_rn(v.y));
        const float c2 = __half2float(__float2half_rn(v.z)), c3 = __half2float(__float2half_rn(v.w));
        dot = fmaf(c0, s_row[4 * d4], dot); dot = fmaf(c1, s_row[4 * d4 + 1], dot);
        dot = fmaf(c2, s_row[4 * d4 + 2], dot); dot = fmaf(c3, s_row[4 * d4 + 3], dot);
        nrm = fmaf(c0, c0, nrm); nrm = fmaf(c1, c1, nrm); nrm = fmaf(c2, c2, nrm); nrm = fmaf(c3, c3, nrm);
    }
    float mx = nrm;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((k & 31) == 0) s_red[k >> 5] = mx;
    __syncthreads();
    mx = fmaxf(s_red[0], s_red[1]);
    const float s = mx > 0.0f ? 1024.0f / mx : 1.0f;
    if (n == 0 && k == 0) gscale_inv[l] = 1.0f / s;
    s_v[k] = dot * s;
    __syncthreads();
    if (k < 8) {                                  // thread k packs 16-byte chunk k of row n (head and remainder)
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_f16x2(s_v[k * 8 + 2 * i], s_v[k * 8 + 2 * i + 1], hi[i], lo[i]);
        uint8_t* base = gimg + (size_t)l * FD_CB_BYTES;
        *reinterpret_cast<uint4*>(base + sw128_offset(n, k)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(base + sw128_offset(64 + n, k)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

// Query images.  block (k, l), thread q:  M_l[k][q] = <C_l[k], E[q]> in fp32, stored as the B operand of
// sim = Wdec_l . M_l  (row n = phrase q, column k = code) as fp16 head (rows 0-63) + remainder (rows 64-127).
__global__ void __launch_bounds__(64) k_fd_prep_q(const float* __restrict__ codebooks, const float* __restrict__ embeds, int Q,
                                                 uint8_t* __restrict__ qimg)
{
    __shared__ float s_row[FD_D];
    const int k = blockIdx.x, l = blockIdx.y, q = threadIdx.x;
    const float* cb = codebooks + ((size_t)l * FD_LC + k) * FD_D;
    for (int d = q; d < FD_D; d += 64) s_row[d] = cb[d];
    __syncthreads();
    float dot = 0.0f;
    if (q < Q) {
        const float4* e = reinterpret_cast<const float4*>(embeds + (size_t)q * FD_D);
        for (int d4 = 0; d4 < FD_D / 4; ++d4) {
            const float4 v = __ldg(e + d4);
            dot = fmaf(v.x, s_row[4 * d4], dot); dot = fmaf(v.y, s_row[4 * d4 + 1], dot);
            dot = fmaf(v.z, s_row[4 * d4 + 2], dot); dot = fmaf(v.w, s_row[4 * d4 + 3], dot);
        }
    }
    const __half h = __float2half_rn(dot);
    const __half lo = __float2half_rn(__fsub_rn(dot, __half2float(h)));
    uint8_t* base = qimg + (size_t)l * FD_CB_BYTES;
    *reinterpret_cast<__half*>(base + sw128_offset((uint32_t)q, (uint32_t)(k >> 3)) + (k & 7) * 2) = h;
    *reinterpret_cast<__half*>(base + sw128_offset((uint32_t)(64 + q), (uint32_t)(k >> 3)) + (k & 7) * 2) = lo;
}

// ---- packed coefficients (lsv2_pack_coeff) -------------------------------------------------------------------------
// One thread per Gaussian: entries -> (fp16 head | fp16 remainder << 16) words + channel bytes + count.  Duplicate
// channels are merged (their weights add up, as the blend would add them), channels outside [0, L) are dropped and
// flagged.  MODE_DENSE reads the compacted rows of lsv2_compact_coeff; a row too dense for them keeps count =
// FD_ROW_DENSE and the blend reads it from coeff_dense.
template <int MODE>
__global__ void __launch_bounds__(256) k_fd_pack(int P, int L, int K, const float* __restrict__ wq, const float* __restrict__ iq,
                                                 const int32_t* __restrict__ radii, const uint32_t* __restrict__ sp_count,
                                                 const uint16_t* __restrict__ sp_index, const float* __restrict__ sp_weight,
                                                 uint32_t* __restrict__ crec, uint32_t* __restrict__ status)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    float w[FD_KMAX];
    int ch[FD_KMAX];
    uint32_t n = 0;
    bool bad = false;
    if (MODE == LSV2_MODE_QUICK) {
        for (int k = 0; k < K; ++k) {
            const float wk = wq[(size_t)g * K + k];
            const int c = (int)iq[(size_t)g * K + k];
            if (c < 0 || c >= L) { bad = true; continue; }
            bool merged = false;
#pragma unroll
            for (int e = 0; e < FD_KMAX; ++e)
                if ((uint32_t)e < n && ch[e] == c) { w[e] += wk; merged = true; }
            if (!merged) {
#pragma unroll
                for (int e = 0; e < FD_KMAX; ++e) if ((uint32_t)e == n) { w[e] = wk; ch[e] = c; }
                ++n;
            }
        }
    } else {
        if (radii && radii[g] <= 0) return;                 // the compacted row of an invisible Gaussian does not exist
        n = sp_count[g];
        if (n != FD_ROW_DENSE) {
            n = min(n, (uint32_t)FD_KMAX);
#pragma unroll
            for (int e = 0; e < FD_KMAX; ++e)
                if ((uint32_t)e < n) { w[e] = sp_weight[(size_t)g * FD_KMAX + e]; ch[e] = (int)sp_index[(size_t)g * FD_KMAX + e]; }
        }
    }
    uint32_t out[FD_CREC_WORDS];
#pragma unroll
    for (int i = 0; i < FD_CREC_WORDS; ++i) out[i] = 0u;
    if (n != FD_ROW_DENSE) {
#pragma unroll
        for (int e = 0; e < FD_KMAX; ++e)
            if ((uint32_t)e < n) {
                const __half h = __float2half_rn(w[e]);
                const __half lo = __float2half_rn(__fsub_rn(w[e], __half2float(h)));
                out[e] = (uint32_t)__half_as_ushort(h) | ((uint32_t)__half_as_ushort(lo) << 16);
                out[16 + (e >> 2)] |= (uint32_t)(ch[e] & 0xFF) << (8 * (e & 3));
            }
    }
    out[20] = n;
    uint4* dst = reinterpret_cast<uint4*>(crec + (size_t)g * FD_CREC_WORDS);
#pragma unroll
    for (int i = 0; i < FD_CREC_WORDS / 4; ++i) dst[i] = make_uint4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
    if (bad && status) status[LSV2_STATUS_BAD_INDEX] = 1u;
}

int launch_pack_coeff(int P, const lsv2_geom& g, const lsv2_features& f, uint32_t* status, cudaStream_t st)
{
    if (P <= 0) return LSV2_OK;
    if (!f.packed || f.L <= 0 || f.L > 256) return LSV2_EINVAL;
    if (f.mode == LSV2_MODE_QUICK) {
        if (f.K <= 0 || !f.weights_quick || !f.indices_quick) return LSV2_EINVAL;
        if (f.K > FD_KMAX) return LSV2_EUNSUPPORTED;
        k_fd_pack<LSV2_MODE_QUICK><<<ceil_div(P, 256), 256, 0, st>>>(P, f.L, f.K, f.weights_quick, f.indices_quick, nullptr,
                                                                     nullptr, nullptr, nullptr, f.packed, status);
    } else if (f.mode == LSV2_MODE_DENSE) {
        if (!f.sparse_count || !f.sparse_index || !f.sparse_weight) return LSV2_EINVAL;
        k_fd_pack<LSV2_MODE_DENSE><<<ceil_div(P, 256), 256, 0, st>>>(P, f.L, 0, nullptr, nullptr, g.radii, f.sparse_count,
                                                                     f.sparse_index, f.sparse_weight, f.packed, status);
    } else {
        return LSV2_EINVAL;
    }
    LSV2_LAUNCH_CHECK("k_fd_pack");
    return LSV2_OK;
}

// What one producer lane gathers (kept in registers one batch ahead).  Two lanes share an instance: lane part p
// (= lane >> 4) takes the packed entry pairs p, p + 2, p + 4, p + 6 (entries 2(p + 2i), 2(p + 2i) + 1), part 0 also
// the position / shape / colour words of the blend record, part 1 its cut-off / opacity word.
struct FdGather {
    uint32_t g, cnt;
    float4 ra, rb;                       // part 0: rec[0], rec[2] ; part 1: rec[1], -
    uint2 hl[4];
    uint4 ch;
};

__device__ __forceinline__ void fd_gather(const FdArgs& a, uint32_t g, int part, FdGather& o)
{
    o.g = g;
    o.cnt = 0;
    if (g == 0xFFFFFFFFu) return;
    if (part == 0) {
        o.ra = __ldg(a.rec + 3 * (size_t)g);
        o.rb = __ldg(a.rec + 3 * (size_t)g + 2);
    } else {
        o.ra = __ldg(a.rec + 3 * (size_t)g + 1);
    }
    const uint32_t* row = a.crec + (size_t)g * FD_CREC_WORDS;
#pragma unroll
    for (int i = 0; i < 4; ++i) o.hl[i] = __ldg(reinterpret_cast<const uint2*>(row) + part + 2 * i);
    o.ch = __ldg(reinterpret_cast<const uint4*>(row) + 4);
    o.cnt = __ldg(row + 20);
}

// pull what fd_gather will read for Gaussian g towards L2 (no registers held)
__device__ __forceinline__ void fd_prefetch(const FdArgs& a, uint32_t g, int part)
{
    if (g == 0xFFFFFFFFu) return;
#ifdef FD_NO_PREFETCH
    return;
#endif
    if (part == 0) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rec + 3 * (size_t)g));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rec + 3 * (size_t)g + 2));
    } else {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.crec + (size_t)g * FD_CREC_WORDS));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.crec + (size_t)g * FD_CREC_WORDS + 20));
    }
}

// one coefficient of a dense row into an S tile: head at k-slice 0, remainder at k-slice 1 of row `c` (the
// remainder's 16-byte chunk index is the head's with bit 1 flipped, i.e. byte offset ^ 32)
__device__ __forceinline__ void fd_put(unsigned char* S, uint32_t c, uint32_t inst, float w)
{
    const uint32_t off = sw64_offset(c, inst >> 3) + (inst & 7u) * 2u;
    const __half h = __float2half_rn(w);
    *reinterpret_cast<__half*>(S + off) = h;
    *reinterpret_cast<__half*>(S + (off ^ 32u)) = __float2half_rn(__fsub_rn(w, __half2float(h)));
}

// HWC: channels-last output (compile-time, so that neither variant carries the other's epilogue in its hot loop).
template <bool OUT_BF16, int PATH, bool HWC = false>
__global__ void __launch_bounds__(PATH != FD_PATH_WMAP ? FD_THREADS : FD_THREADS_NODECODE, 1) k_render_decode(FdArgs a)
{
    constexpr bool DECODE = PATH == FD_PATH_DECODE;     // codebook decode + epilogue warps
    constexpr bool QUERY = PATH == FD_PATH_QUERY;       // relevancy query, finished by the blend warps themselves
    constexpr bool WROWS = PATH != FD_PATH_WMAP;        // the blend's W goes into the fp16 operand rows
    extern __shared__ unsigned char smem_raw[];
    // offset arithmetic on the __shared__ symbol (not a uintptr_t round trip) keeps the shared address space
    // visible to the compiler: LDS/STS instead of generic LD/ST in the blend loop
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* ring = smem + FD_SMEM_RING;
    FdHeader* hdrs = reinterpret_cast<FdHeader*>(smem + FD_SMEM_HDR);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FD_SMEM_BAR);
    uint64_t* full = bars;                        // [2]  producer -> blend, mma          (32 arrivals)
    uint64_t* a_full = bars + 2;                  // [2]  blend -> mma                    (256 arrivals)
    uint64_t* stage_free = bars + 4;              // [2]  mma -> producer                 (commit / arrive)
    uint64_t* w_full = bars + 6;                  //      mma -> blend: W complete        (commit / arrive)
    uint64_t* wraw_full = bars + 7;               // [3]  blend -> MD: fp16 rows of level l written     (256)
    uint64_t* u_full = bars + 10;                 // [3]  MD -> blend: U_l = Wdec_l.G_l ready           (commit)
    uint64_t* inv_full = bars + 13;               //      blend -> epilogue: 1/norm of the unit written (256)
    uint64_t* dec_done = bars + 14;               //      MD -> blend: decode MMAs of the unit done     (commit)
    uint64_t* cb_full = bars + 15;                // [2]  bulk copy landed                 (tx)
    uint64_t* cb_free = bars + 17;                // [2]  MMAs done with the chunk         (commit)
    uint64_t* f_full = bars + 19;                 // [4]  F slot ready                     (commit)
    uint64_t* f_free = bars + 23;                 // [4]  F slot drained                   (4 arrivals)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 27);
    uint32_t* s_ndone = s_tmem + 1;               // [2]  saturated blend warps, by unit parity
    FdTileInfo* tinfo = reinterpret_cast<FdTileInfo*>(s_tmem + 4);     // [2] by unit parity
    uint32_t* s_misc = s_tmem + 8;                // [3]  producers: unit id, batches of list 0 / 1
    float* s_inv = reinterpret_cast<float*>(smem + FD_SMEM_INV);       // [2][3][256] by unit parity
    unsigned char* wdec = smem + FD_SMEM_WDEC;
    unsigned char* s_cb = smem + FD_SMEM_CB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = a.L, levels = a.levels;
    const size_t HW = (size_t)a.H * a.W;
    const bool normalize = QUERY || (DECODE && a.normalize != 0);

    if (tid == 0) {
        for (int s = 0; s < FD_STAGES; ++s) {
            mbar_init(full + s, 2);
            mbar_init(a_full + s, FD_BLEND_WARPS);
            mbar_init(stage_free + s, 1);
            mbar_init(cb_full + s, 1);
            mbar_init(cb_free + s, 1);
        }
        mbar_init(w_full, 1);
        for (int l = 0; l < 3; ++l) { mbar_init(wraw_full + l, FD_BLEND_WARPS); mbar_init(u_full + l, 1); }
        mbar_init(inv_full, FD_BLEND_WARPS);
        mbar_init(dec_done, 1);
        for (int q = 0; q < FD_FSLOTS; ++q) { mbar_init(f_full + q, 1); mbar_init(f_free + q, QUERY ? FD_BLEND_WARPS : FD_EPI_WARPS); }
        s_ndone[0] = 0;
        s_ndone[1] = 0;
        mbar_fence_init();
    }
    if (warp == FD_WARP_MB) tmem_alloc(s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp == FD_WARP_PROD || warp == FD_WARP_PROD + 1) {
        // ======================================= producers =======================================
        // Warp t serves tile list t (the left / right tile of the band): lanes i and i + 16 share instance slot i.
        // Depth of the software pipeline, per list: instance ids 8 batches ahead, an L2 prefetch of their records and
        // coefficient rows 6 batches ahead, the register gather 2 batches ahead, so the memory latency stays off the
        // per-batch path: wait stage -> zero the S tile -> scatter -> arrive.
        const int t = warp - FD_WARP_PROD;
        const int inst = lane & 15, part = lane >> 4;
        const int ptid = tid - 32 * FD_WARP_PROD;
        uint32_t use_g = 0;                             // stage uses so far (global)
        FD_PROF_DECL
        FD_TIC();
        // Units are claimed from a global counter.  (Dealing most of them round-robin without the atomic, and walking the
        // next unit's set-up chain -- range, ids, records -- during the current unit, with registers or with cp.async
        // into shared memory, were both measured on one box against this form: within +-1 %, profiles/r02m.)
        for (uint32_t u_iter = 0;; ++u_iter) {
            if (ptid == 0) s_misc[0] = atomicAdd(a.tile_counter, 1u);
            named_bar_sync(1, FD_PROD_THREADS);
            const int unit = (int)s_misc[0];
            const bool exit_now = unit >= a.n_units;
            const int band = unit / a.ux, pair = unit % a.ux;
            const int tx = 2 * pair + t, ty = band >> 1;
            uint2 range = make_uint2(0, 0);
            if (!exit_now && tx < a.gx) range = a.ranges[ty * a.gx + tx];
            const uint32_t n_inst = range.y - range.x;
            if (lane == 0) s_misc[1 + t] = (n_inst + FD_BATCH - 1) / FD_BATCH;
            named_bar_sync(1, FD_PROD_THREADS);
            const uint32_t n_batches = max(s_misc[1], s_misc[2]);
            const float y0 = (float)(band * FD_UNIT_H), x0 = (float)(tx * TILE);
            auto load_id = [&](uint32_t b) -> uint32_t {
                const uint32_t i = b * FD_BATCH + (uint32_t)inst;
                return i < n_inst ? __ldg(a.point_list + range.x + i) : 0xFFFFFFFFu;
            };
            FdGather cur, nxt, nx2;                     // batches b, b + 1 (gathers in flight / landed), b + 2 (id only)
            uint32_t idq[5];                            // ids of batches b+3 .. b+7
            {
                const uint32_t i0 = load_id(0), i1 = load_id(1), i2 = load_id(2);
#pragma unroll
                for (int k = 0; k < 5; ++k) idq[k] = load_id(3 + k);
                fd_gather(a, i0, part, cur);
                fd_gather(a, i1, part, nxt);
                nx2.g = i2;
#pragma unroll
                for (int k = 0; k < 3; ++k) fd_prefetch(a, idq[k], part);
            }
            unsigned char* const S_base = ring + 2 * FD_AHALF_BYTES + (size_t)t * FD_S_BYTES;
            // byte offset of this lane's instance column inside an S row, without the row / chunk-swizzle part
            const uint32_t col_off = (uint32_t)(inst & 7) * 2u;
            const uint32_t chunk0 = (uint32_t)inst >> 3;
            // One batch per iteration (single copy of the code: the roles' loops must share the instruction caches).
            // `cur` holds batch b; the register gather of batch b + 1 is issued at the top and lands while this batch
            // is assembled (its lines were pulled into L2 four batches earlier).
#pragma unroll 1
            for (uint32_t b = 0;; ++b) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                FD_TOC(2);
                if (use > 0) FD_WAIT_SLEEP(stage_free + s, (use - 1) & 1);
                FD_TOC(0);
                FdHeader& Hd = hdrs[s];
                // s_ndone[parity] only ever grows: every blend warp adds exactly one per unit (when its block saturates,
                // or at the end of the unit), so "all 8 blocks of this unit are done" is count - 8 * (u_iter / 2) >= 8
                const bool stop_l = exit_now || b >= n_batches ||
                                    (int32_t)(*(volatile uint32_t*)(s_ndone + (u_iter & 1)) - FD_BLEND_WARPS * (u_iter >> 1)) >= FD_BLEND_WARPS;
                // s_ndone may change between the 64 reads: OR-reduce so the decision is uniform over both warps
                if (named_bar_or(2, FD_PROD_THREADS, stop_l)) {
                    if (ptid == 0) { Hd.count = exit_now ? -2 : -1; Hd.unit = unit; Hd.batches = (int32_t)b; }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(full + s);
                    ++use_g;
                    break;
                }
                FD_TOC(3);
                unsigned char* S = S_base + (size_t)s * FD_STAGE_BYTES;
                // only the L channel rows the blend MMA reads (N = L): 4 KB at L = 64 (training), 12 KB at L = 192
                for (int q = lane; q < L * 4; q += 32) reinterpret_cast<uint4*>(S)[q] = make_uint4(0, 0, 0, 0);
                fd_gather(a, nx2.g, part, nx2);              // batch b + 2: under the write stream an L2 hit takes about one
                                                             // batch time, so the gather gets two batches to land
                fd_prefetch(a, idq[3], part);                // batch b + 6
                const uint32_t id_new = load_id(b + 8);
                __syncwarp();
                FD_TOC(4);
                uint32_t m4 = 0;
                if (cur.g != 0xFFFFFFFFu) {
                    if (part == 0) {
                        Hd.g0[t * FD_BATCH + inst] = cur.ra;
                        Hd.g2[t * FD_BATCH + inst] = make_float2(cur.rb.x, cur.rb.y);
                        m4 = fd_block_mask_ext(cur.ra.y - y0, cur.ra.x - x0, cur.rb.z, cur.rb.w);
                    } else {
                        Hd.g1[t * FD_BATCH + inst] = cur.ra;
                    }
                    if (cur.cnt != FD_ROW_DENSE) {
                        const uint32_t chw[4] = {cur.ch.x, cur.ch.y, cur.ch.z, cur.ch.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
#pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                const uint32_t e = 2u * (uint32_t)(part + 2 * i) + (uint32_t)j;
                                if (e < cur.cnt) {
                                    const uint32_t c = (chw[i] >> (8u * (2u * (uint32_t)part + (uint32_t)j))) & 0xFFu;
                                    const uint32_t v = j == 0 ? cur.hl[i].x : cur.hl[i].y;
                                    const uint32_t off = c * 64u + (((chunk0 ^ (c >> 1)) & 3u) << 4) + col_off;
                                    *reinterpret_cast<unsigned short*>(S + off) = (unsigned short)(v & 0xFFFFu);
                                    *reinterpret_cast<unsigned short*>(S + (off ^ 32u)) = (unsigned short)(v >> 16);
                                }
                            }
                        }
                    } else {
                        const float* row = a.coeff_dense + (size_t)cur.g * L;
                        for (int c = part; c < L; c += 2) {
                            const float v = __ldg(row + c);
                            if (v != 0.0f) fd_put(S, (uint32_t)c, (uint32_t)inst, v);
                        }
                    }
                }
                if (b == 0) FD_TOC(5); else FD_TOC(1);
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, (m4 >> w) & 1u);     // part-0 lanes: bit i = instance i
                    if (lane == 0) Hd.mask[4 * t + w] = bal & 0xFFFFu;
                }
                if (ptid == 0) { Hd.count = 1; Hd.unit = unit; }
                fence_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(full + s);
                // batch b + 1 becomes current; queue the id of b + 3
                cur = nxt;
                nxt = nx2;
                nx2.g = idq[0];
#pragma unroll
                for (int k = 0; k < 4; ++k) idq[k] = idq[k + 1];
                idq[4] = id_new;
                ++use_g;
                FD_TOC(6);
            }
            if (exit_now) break;
        }
        if (ptid == 0) FD_PROF_FLUSH(1);
    } else if (warp == FD_WARP_MB) {
        // ======================================= blend MMA issue =======================================
        // The whole warp runs the loop converged; one elected lane issues (see elect_one, tc_common.cuh).
        const uint32_t idesc_blend = umma_idesc_f16_f32(128, L);
        const uint32_t ring_addr = bcast0(smem_u32(ring));
        const uint32_t tmem_u = bcast0(tmem);
        uint32_t batch_in_unit = 0;
#pragma unroll 1
        for (uint32_t use_g = 0;; ++use_g) {
            const int s = use_g & 1;
            const uint32_t par = (use_g >> 1) & 1;
            FD_WAIT_SLEEP(full + s, par);
            const int count = (int)bcast0((uint32_t)hdrs[s].count);
            FD_WAIT(a_full + s, par);                  // batches: A tiles written; terminators: header read by all
            if (count == -2) break;
            if (count >= 0) {
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t sa = ring_addr + (uint32_t)s * FD_STAGE_BYTES;
                    const uint64_t dA = umma_desc_k_sw64(sa), dS = umma_desc_k_sw64(sa + 2 * FD_AHALF_BYTES);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint64_t a0 = dA + (uint64_t)(h * (FD_AHALF_BYTES >> 4)), s0 = dS + (uint64_t)(h * (FD_S_BYTES >> 4));
                        const uint32_t d = tmem_u + (uint32_t)(h * FD_MAX_L);
                        umma_bf16(d, a0, s0, idesc_blend, batch_in_unit > 0 ? 1u : 0u);
                        umma_bf16(d, a0 + 2, s0, idesc_blend, 1u);
                        umma_bf16(d, a0, s0 + 2, idesc_blend, 1u);
                    }
                    umma_commit(stage_free + s);
                }
                __syncwarp();
                ++batch_in_unit;
                continue;
            }
            // ---- end of unit ----
            if (elect_one()) {
                mbar_arrive(stage_free + s);
                if (batch_in_unit > 0) umma_commit(w_full); else mbar_arrive(w_full);
            }
            __syncwarp();
            batch_in_unit = 0;
        }
    } else if (WROWS && warp == FD_WARP_MD) {
        // ================================ Gram + decode MMA issue, codebook fetch ================================
        // Converged warp, one elected lane issues.
        constexpr uint32_t idesc_u = umma_idesc_f16_f32(128, FD_LC);
        constexpr uint32_t idesc_dec = umma_idesc_f16_f32(128, FD_NFEAT);
        static_assert(FD_NFEAT == 64, "the channels-last decode swaps the operands of an M = 128, N = 64 MMA");
        const uint32_t tmem_u = bcast0(tmem);
        // operand descriptors are built once; a k-step (32 bytes) or a row block only adds to the address field
        const uint64_t desc_w0 = umma_desc_k_sw128(bcast0(smem_u32(wdec))), desc_cb0 = umma_desc_k_sw128(bcast0(smem_u32(s_cb)));
        const int n_g = normalize ? levels : 0;
        const int jobs_per_unit = QUERY ? 2 * levels : n_g + levels * FD_CHUNKS;
        uint32_t cb_issued = 0, cb_used = 0, f_n = 0;
        FD_PROF_DECL
        FD_TIC();
        auto fetch_next = [&]() {                        // codebook-ring job `cb_issued` -> slot cb_issued & 1
            const int slot = (int)(cb_issued & 1u);
            if (cb_issued >= FD_CB_SLOTS) FD_WAIT(cb_free + slot, ((cb_issued >> 1) - 1) & 1);
            const int j = (int)(cb_issued % (uint32_t)jobs_per_unit);
            const uint8_t* src = QUERY ? ((j & 1) ? a.qimg : a.gimg) + (size_t)(j >> 1) * FD_CB_BYTES
                                       : (j < n_g ? a.gimg + (size_t)j * FD_CB_BYTES : a.bimg + (size_t)(j - n_g) * FD_CB_BYTES);
            if (elect_one()) {
                mbar_arrive_expect_tx(cb_full + slot, FD_CB_BYTES);
                bulk_g2s(s_cb + (size_t)slot * FD_CB_BYTES, src, FD_CB_BYTES, cb_full + slot);
            }
            __syncwarp();
            ++cb_issued;
        };
        fetch_next();
        fetch_next();
#pragma unroll 1
        for (uint32_t it = 0;; ++it) {
            bool stop = false;
#pragma unroll 1
            for (int l = 0; l < levels; ++l) {
                // rows of level l are in shared memory (all levels are needed before the decode starts anyway;
                // the per-level barrier lets the Gram products start while later levels are still converted)
                FD_TOC(3);
                FD_WAIT_SLEEP(wraw_full + l, it & 1);
                FD_TOC(0);
                if (l == 0 && (int)bcast0((uint32_t)tinfo[it & 1].unit) < 0) { stop = true; break; }
                if (!normalize) continue;
                tc_fence_after();
                const uint32_t n = cb_used;
                const int slot = (int)(n & 1u);
                FD_WAIT(cb_full + slot, (n >> 1) & 1);
                if (elect_one()) {
                    const uint64_t g_hi = desc_cb0 + (uint64_t)(slot * (FD_CB_BYTES >> 4)), g_lo = g_hi + (uint64_t)((64 * 128) >> 4);
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        const uint32_t d = tmem_u + (uint32_t)(m * FD_MAX_L + l * FD_LC);
                        const uint64_t da = desc_w0 + (uint64_t)(l * (FD_WDEC_LEVEL >> 4) + m * ((128 * 128) >> 4));
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_bf16(d, da + (uint64_t)(2 * ks), g_hi + (uint64_t)(2 * ks), idesc_u, ks > 0 ? 1u : 0u);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_bf16(d, da + (uint64_t)(2 * ks), g_lo + (uint64_t)(2 * ks), idesc_u, 1u);
                    }
                    umma_commit(u_full + l);
                    umma_commit(cb_free + slot);
                }
                __syncwarp();
                ++cb_used;
                if (n >= 1) fetch_next();
                if (QUERY) {
                    // sim_l = Wdec_l . (M_hi + M_lo) into the spare TMEM columns (one level at a time: the blend warps
                    // drain it before the next level's product may land)
                    const uint32_t nq = it * (uint32_t)levels + (uint32_t)l;
                    const uint32_t n2 = cb_used;
                    const int slot2 = (int)(n2 & 1u);
                    FD_WAIT(cb_full + slot2, (n2 >> 1) & 1);
                    if (nq >= 1) FD_WAIT(f_free, (nq - 1) & 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t q_hi = desc_cb0 + (uint64_t)(slot2 * (FD_CB_BYTES >> 4)), q_lo = q_hi + (uint64_t)((64 * 128) >> 4);
#pragma unroll
                        for (int m = 0; m < 2; ++m) {
                            const uint32_t d = tmem_u + (uint32_t)(FD_TMEM_F + m * 64);
                            const uint64_t da = desc_w0 + (uint64_t)(l * (FD_WDEC_LEVEL >> 4) + m * ((128 * 128) >> 4));
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(d, da + (uint64_t)(2 * ks), q_hi + (uint64_t)(2 * ks), idesc_u, ks > 0 ? 1u : 0u);
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(d, da + (uint64_t)(2 * ks), q_lo + (uint64_t)(2 * ks), idesc_u, 1u);
                        }
                        umma_commit(f_full);
                        umma_commit(cb_free + slot2);
                    }
                    __syncwarp();
                    ++cb_used;
                    fetch_next();
                }
            }
            if (stop) break;
            if (QUERY) continue;
            tc_fence_after();
#pragma unroll 1
            for (int l = 0; l < levels; ++l) {
#pragma unroll 1
                for (int c = 0; c < FD_CHUNKS; ++c) {
                    const uint32_t n = cb_used;
                    const int slot = (int)(n & 1u);
                    FD_TOC(3);
                    FD_WAIT(cb_full + slot, (n >> 1) & 1);
                    FD_TOC(1);
#pragma unroll 1
                    for (int gm = 0; gm < 2 * FD_FGROUPS; ++gm) {
                        const int g = gm >> 1, m = gm & 1;
                        const int fs = (int)(f_n % FD_FSLOTS);
                        FD_TOC(3);
                        if (f_n >= FD_FSLOTS) FD_WAIT(f_free + fs, ((f_n / FD_FSLOTS) - 1) & 1);
                        FD_TOC(2);
                        tc_fence_after();
                        if (elect_one()) {
                            // CHW: D[128 px (half m)][64 features (group g)] -- lane = pixel, so a store instruction
                            // writes 32 pixels of one feature plane.  Channels-last: the operands swap roles,
                            // D[128 features (the chunk)][64 px (decode rows 64 gm ..)] -- lane = feature, so a store
                            // instruction writes 32 consecutive features of one pixel.  Same bytes in shared memory, same
                            // descriptor layout (both K-major, 128-byte swizzle), same instruction shape M = 128, N = 64.
                            const uint64_t db = HWC ? desc_w0 + (uint64_t)(l * (FD_WDEC_LEVEL >> 4) + gm * ((64 * 128) >> 4))
                                                      : desc_cb0 + (uint64_t)(slot * (FD_CB_BYTES >> 4) + g * ((FD_NFEAT * 128) >> 4));
                            const uint64_t da = HWC ? desc_cb0 + (uint64_t)(slot * (FD_CB_BYTES >> 4))
                                                      : desc_w0 + (uint64_t)(l * (FD_WDEC_LEVEL >> 4) + m * ((128 * 128) >> 4));
                            const uint32_t d = tmem_u + (uint32_t)(FD_TMEM_F + fs * FD_NFEAT);
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(d, da + (uint64_t)(2 * ks), db + (uint64_t)(2 * ks), idesc_dec, ks > 0 ? 1u : 0u);
                            umma_commit(f_full + fs);
                        }
                        __syncwarp();
                        ++f_n;
                        if (gm == 1 && n >= 1) fetch_next();
                    }
                    if (elect_one()) umma_commit(cb_free + slot);
                    __syncwarp();
                    ++cb_used;
                }
            }
            if (elect_one()) umma_commit(dec_done);
            __syncwarp();
        }
        if (lane == 0) FD_PROF_FLUSH(2);
        // drain the codebook prefetch that no unit consumed
        while (cb_used < cb_issued) { FD_WAIT(cb_full + (cb_used & 1u), (cb_used >> 1) & 1); ++cb_used; }
    } else if (DECODE && warp >= FD_BLEND_WARPS && warp < FD_BLEND_WARPS + FD_EPI_WARPS) {
        // ==================================== epilogue: F slots -> global ====================================
        // Decode MMA block m covers band rows 4m .. 4m+3; TMEM lane quadrant q = this warp = band row 4m + q,
        // lane = x: one store instruction writes 32 consecutive pixels of one feature plane (a 128-byte line).
        const int q = warp & 3;
        const uint32_t lane_addr = tmem + ((uint32_t)(q * 32) << 16) + FD_TMEM_F;
        uint32_t f_n = 0;
        FD_PROF_DECL
        FD_TIC();
        for (uint32_t it = 0;; ++it) {
            FD_TOC(2);
            FD_WAIT_SLEEP(inv_full, it & 1);
            FD_TOC(0);
            const int unit = tinfo[it & 1].unit;
            if (unit < 0) break;
            const int band = unit / a.ux, pair = unit % a.ux;
            const int px = pair * FD_UNIT_W + lane;
            const float* inv_u = s_inv + (size_t)(it & 1) * 3 * 256;
            const size_t esz = OUT_BF16 ? 2 : 4;
            for (int l = 0; l < levels; ++l) {
                float inv[2];
                inv[0] = normalize ? inv_u[l * 256 + q * 32 + lane] : 1.0f;
                inv[1] = normalize ? inv_u[l * 256 + 128 + q * 32 + lane] : 1.0f;
                for (int cg = 0; cg < FD_CHUNKS * FD_FGROUPS; ++cg) {
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        const int fs = (int)(f_n % FD_FSLOTS);
                        FD_TOC(2);
                        FD_WAIT(f_full + fs, (f_n / FD_FSLOTS) & 1);
                        FD_TOC(1);
                        ++f_n;
                        tc_fence_after();
                        const int py = band * FD_UNIT_H + 4 * m + q;
                        const bool in = px < a.W && py < a.H;
                        const uint64_t stride = (uint64_t)HW * esz;
                        const float sc = inv[m];
#pragma unroll
                        for (int hh = 0; hh < FD_NFEAT / 32; ++hh) {
                            uint32_t r[32];
                            tmem_ld32(lane_addr + (uint32_t)(fs * FD_NFEAT + hh * 32), r);
                            tmem_wait_ld();
                            if (hh == FD_NFEAT / 32 - 1) {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(f_free + fs);     // slot drained: the next MMA may overwrite it
                            }
                            // channels-last (transposed decode): this lane = feature (cg >> 1) * 128 + 32 q + lane,
                            // r[j] = decode row 64 pg + 32 hh + j = the pixel (x = j, band row 2 pg + hh); one store
                            // instruction = 32 consecutive features of one pixel = one 128-byte line
                            if (HWC) {
                                const int pg = (cg & 1) * 2 + m;
                                const int yy = band * FD_UNIT_H + 2 * pg + hh;
                                const int rr0 = 64 * pg + 32 * hh;
                                const size_t feat = (size_t)(cg >> 1) * 128 + (size_t)(q * 32 + lane);
                                if (yy < a.H) {
                                    uint64_t pp = (uint64_t)(uintptr_t)a.out +
                                                  ((((size_t)l * a.H + yy) * a.W + (size_t)pair * FD_UNIT_W) * FD_D + feat) * esz;
                                    const int nx = min(32, a.W - pair * FD_UNIT_W);
                                    // 1/norm of pixel j: lane j fetches it once and the warp passes it round -- a
                                    // shared-memory load between the `asm volatile` stores would be ordered behind the
                                    // previous store by its "memory" clobber (32 serialised load latencies per chunk:
                                    // the first version of this path ran at 7.5 ms per frame)
                                    const float inv_l = normalize ? inv_u[l * 256 + rr0 + lane] : 1.0f;
                                    float invr[32];
#pragma unroll
                                    for (int j = 0; j < 32; ++j) invr[j] = __shfl_sync(0xffffffffu, inv_l, j);
#pragma unroll
                                    for (int j = 0; j < 32; ++j) {
                                        if (j < nx) {
                                            const float v = __uint_as_float(r[j]) * invr[j];
                                            if (OUT_BF16) {
                                                const unsigned short hv = __bfloat16_as_ushort(__float2bfloat16_rn(v));
                                                asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(pp), "h"(hv) : "memory");
                                            } else {
                                                asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(pp), "f"(v) : "memory");
                                            }
                                        }
                                        pp += (uint64_t)FD_D * esz;
                                    }
                                }
                            } else if (in) {
                                const size_t d0 = (size_t)l * FD_D + (size_t)cg * FD_NFEAT + (size_t)hh * 32;
                                uint64_t p[4];
                                p[0] = (uint64_t)(uintptr_t)a.out + (d0 * HW + (size_t)py * a.W + px) * esz;
                                p[1] = p[0] + stride; p[2] = p[1] + stride; p[3] = p[2] + stride;
                                const uint64_t s4 = 4 * stride;
#pragma unroll
                                for (int i = 0; i < 32; i += 4) {
#pragma unroll
                                    for (int u = 0; u < 4; ++u) {
                                        const float v = __uint_as_float(r[i + u]) * sc;
                                        if (OUT_BF16) {
                                            const unsigned short hv = __bfloat16_as_ushort(__float2bfloat16_rn(v));
                                            asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p[u]), "h"(hv) : "memory");
                                        } else {
                                            asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p[u]), "f"(v) : "memory");
                                        }
                                        p[u] += s4;
                                    }
                                }
                            }
                        }
                    }
                }
            }
        }
        if (warp == FD_BLEND_WARPS && lane == 0) FD_PROF_FLUSH(3);
    } else if (warp < FD_BLEND_WARPS) {
        // ==================================== blend ====================================
        FD_PROF_DECL
        FD_PROFB_DECL
        FD_TIC();
        const int half = warp >> 2;                                  // 0: left tile's list, 1: right tile's list
        const int strip = warp & 3;                                  // 8 x 4 pixel block of the half band: 2 * by + bx
        const int row = strip * 32 + lane;                           // A-tile row == TMEM lane of the blend MMA
        const uint32_t lane_base = tmem + ((uint32_t)(strip * 32) << 16);
        const int lx = (strip & 1) * 8 + (lane & 7), ly = (strip >> 1) * 4 + (lane >> 3);
        // the decode sees the band as 8 rows of 32 pixels: row of the 256-row decode operand this pixel owns
        const uint32_t prow = (uint32_t)(ly * 32 + half * 16 + lx);
        // ... and the decode row whose norm this thread computes (TMEM lane quadrant = strip, MMA block = half)
        const uint32_t nrow = (uint32_t)(half * 128 + strip * 32 + lane);
        uint32_t units_done = 0;
        uint32_t use_g = 0;
        for (;;) {
            // ---------------- blend phase ----------------
            float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
            uint32_t last = 0;
            bool done = false, warp_done = false, first = true;
            int unit = 0, px = 0, py = 0;
            float pxf = 0.f, pyf = 0.f;
            bool inside = false;
            int count;
            int nb = 0;
            auto locate = [&]() {
                const int band = unit / a.ux, pair = unit % a.ux;
                px = pair * FD_UNIT_W + half * 16 + lx;
                py = band * FD_UNIT_H + ly;
                pxf = (float)px; pyf = (float)py;
                inside = px < a.W && py < a.H;
            };
            for (;; ++use_g) {
                const int s = use_g & 1;
                FD_TOC(1);
                FD_WAIT(full + s, (use_g >> 1) & 1);
                FD_TOC(0);
                const FdHeader& Hd = hdrs[s];
                count = Hd.count;
                if (count < 0) {
                    // acknowledge the terminator: its header may be overwritten only after every thread has read it
                    unit = Hd.unit; nb = Hd.batches;
                    __syncwarp();
                    if (lane == 0) {
                        if (count == -1 && !warp_done) atomicAdd(s_ndone + (units_done & 1), 1u);   // exactly one per warp per unit
                        mbar_arrive(a_full + s);
                    }
                    ++use_g;
                    break;
                }
                FD_TICB();
                if (first) {
                    first = false;
                    unit = Hd.unit;
                    locate();
                    done = !inside;
                    warp_done = __all_sync(0xffffffffu, done);
                    if (warp_done && lane == 0) atomicAdd(s_ndone + (units_done & 1), 1u);
                }
                unsigned char* A = ring + (size_t)s * FD_STAGE_BYTES + (size_t)half * FD_AHALF_BYTES;
                const uint32_t m = warp_done ? 0u : Hd.mask[warp];
                FD_TOCB(0);
                // The hot loop is NOT unrolled: it stays a few hundred instructions, so the five warp roles' loops share
                // the instruction caches (L0 6 KB, L1.5 32 KB) instead of thrashing them -- with the 16 instances
                // unrolled (15 KB) every role ran ~5x slower.
                // The mask is warp-uniform: the warp walks only the instances whose alpha >= 1/255 box reaches its 8 x 4
                // block, four set bits per iteration (the four quadratic forms / exps interleave; a short tail repeats
                // its last instance with the weight forced to 0).  Skipped instances would have added w = 0 exactly, so
                // colour, T, n_contrib and the weight tile are bit-identical to walking all sixteen.  The A row is
                // zeroed first; each (pixel, instance) hit then stores its fp16 head / remainder at the instance's column.
                {
                    const uint32_t off0 = sw64_offset((uint32_t)row, 0u), off1 = sw64_offset((uint32_t)row, 1u);
                    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
                    *reinterpret_cast<uint4*>(A + off0) = z4;
                    *reinterpret_cast<uint4*>(A + (off0 ^ 32u)) = z4;
                    *reinterpret_cast<uint4*>(A + off1) = z4;
                    *reinterpret_cast<uint4*>(A + (off1 ^ 32u)) = z4;
                }
                uint32_t mm = __shfl_sync(0xffffffffu, m, 0);     // uniform by construction; make it so for the compiler
#pragma unroll 1
                while (mm != 0u) {
                    int jj[4];
                    bool vk[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        vk[u] = mm != 0u;
                        jj[u] = vk[u] ? (__ffs((int)mm) - 1) : (u > 0 ? jj[u - 1] : 0);
                        mm &= mm - 1u;
                    }
                    float pw_[4], op_[4], cr_[4];
                    bool hit[4];
                    bool any = false;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int j = half * FD_BATCH + jj[u];
                        const float4 g0 = Hd.g0[j];
                        const float4 g1 = Hd.g1[j];
                        const float dx = g0.x - pxf, dy = g0.y - pyf;
                        pw_[u] = fmaf(g0.z * dx, dx, fmaf(g1.x * dy, dy, (g0.w * dx) * dy));
                        hit[u] = vk[u] && !(pw_[u] > 0.0f || pw_[u] < g1.y);
                        op_[u] = g1.z;
                        cr_[u] = g1.w;
                        any |= hit[u];
                    }
                    FD_TOCB(1);
                    if (any && !done) {
                        // a hit has cut-off <= power <= 0 (cut-off > -6), so the unchecked exp is exact there;
                        // evaluated for all four so the chains interleave, discarded where !hit
                        float al[4], xe[4], e4[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) xe[u] = fmaxf(fminf(pw_[u], 0.0f), -80.0f);
                        repro_expf_inrange4(xe, e4);
#pragma unroll
                        for (int u = 0; u < 4; ++u) al[u] = hit[u] ? fminf(ALPHA_MAX, op_[u] * e4[u]) : 0.0f;
                        FD_TOCB(2);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            // select-only form of: skip if alpha < 1/255 or done; stop if T' < 1e-4; else blend.
                            // A skipped instance adds w = 0 (exact), so the sums match the branching form bit for bit.
                            const bool valid = !(al[u] < ALPHA_MIN) && !done;
                            const float test_T = T * (1.0f - al[u]);
                            const bool stop = valid && (test_T < T_MIN);
                            const bool use = valid && !stop;
                            done = done || stop;
                            const float w = use ? al[u] * T : 0.0f;
                            const float2 g2 = Hd.g2[half * FD_BATCH + jj[u]];
                            C0 = fmaf(cr_[u], w, C0);
                            C1 = fmaf(g2.x, w, C1);
                            C2 = fmaf(g2.y, w, C2);
                            T = use ? test_T : T;
                            last = use ? (uint32_t)nb * FD_BATCH + (uint32_t)jj[u] + 1u : last;
                            if (use) {
                                // fp16 head + fp16 remainder at column jj[u] of this pixel's row (k-slice 0 | k-slice 1)
                                const __half hh = __float2half_rn(w);
                                const __half ll = __float2half_rn(__fsub_rn(w, __half2float(hh)));
                                const uint32_t o = sw64_offset((uint32_t)row, (uint32_t)(jj[u] >> 3)) + (uint32_t)(jj[u] & 7) * 2u;
                                *reinterpret_cast<__half*>(A + o) = hh;
                                *reinterpret_cast<__half*>(A + (o ^ 32u)) = ll;
                            }
                        }
                        FD_TOCB(3);
                    }
                    FD_TOCB(5);
                }
                ++nb;
                fence_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full + s);     // one arrival per warp
                if (!warp_done && __all_sync(0xffffffffu, done)) {
                    warp_done = true;
                    if (lane == 0) atomicAdd(s_ndone + (units_done & 1), 1u);
                }
                FD_TOCB(6);
            }
            if (count == -2) {
                if (WROWS) {
                    // The terminating arrivals open a NEW phase of inv_full / wraw_full[0]: a warp that is through with the
                    // last unit must not arrive while slower warps still owe their arrival of that unit's phase (8
                    // arrivals complete a phase no matter who made them -- the epilogue would be released before the
                    // slow warps' 1/norm values exist).
                    // ... and it must not arrive before the epilogue has CONSUMED the last unit's inv_full phase: a waiter
                    // that falls two phases behind sees the parity it waits for as "not yet completed" and waits
                    // forever.  Every other phase is protected by the data flow (a unit's pass A waits for the previous
                    // unit's dec_done, which implies the epilogue passed that unit's inv_full); the terminator is not --
                    // it follows the last pass B within a microsecond.  dec_done of the last unit (decode MMAs complete
                    // => all but the last F slots drained => the epilogue is inside that unit) closes it.  Found with the
                    // watchdog build (scripts/fused_watchdog.py) when a slower epilogue variant lost the race every time.
                    if (units_done > 0) {
                        FD_WAIT(wraw_full, (units_done - 1) & 1);
                        if (DECODE) FD_WAIT(inv_full, (units_done - 1) & 1);
                        if (DECODE) FD_WAIT(dec_done, (units_done - 1) & 1);
                    }
                    if (tid == 0) { tinfo[units_done & 1].unit = -1; tinfo[units_done & 1].nb = 0; }
                    __syncwarp();
                    if (lane == 0) { mbar_arrive(wraw_full); if (DECODE) mbar_arrive(inv_full); }
                }
                break;
            }
            if (first) locate();                           // a unit with two empty lists
            const size_t pix = (size_t)py * a.W + px;

            // ---------------- W: TMEM -> fp16 rows of the decode operand ----------------
            FD_TOC(6);
            FD_WAIT(w_full, units_done & 1);
            FD_TOC(2);
            tc_fence_after();
            if (WROWS) {
                // the previous unit's MMAs are done with Wdec (query: its last product was waited for by every thread)
                if (DECODE && units_done > 0) FD_WAIT(dec_done, (units_done - 1) & 1);
                FD_TOC(3);
                if (tid == 0) { tinfo[units_done & 1].unit = unit; tinfo[units_done & 1].nb = nb; }
            }
            for (int l = 0; l < levels; ++l) {
                unsigned char* Wl = wdec + (size_t)l * FD_WDEC_LEVEL;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t r[32];
                    if (nb > 0) {
                        tmem_ld32(lane_base + (uint32_t)(half * FD_MAX_L + l * FD_LC + hh * 32), r);
                        tmem_wait_ld();
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = 0u;
                    }
                    if (a.wmap != nullptr && inside) {
                        float* o = a.wmap + (size_t)(l * FD_LC + hh * 32) * HW + pix;
#pragma unroll
                        for (int i = 0; i < 32; ++i) st_stream_f32(o + (size_t)i * HW, __uint_as_float(r[i]));
                    }
                    if (WROWS) {
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) {
                            uint32_t p[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                p[i] = pack_f16x2(__uint_as_float(r[ch * 8 + 2 * i]), __uint_as_float(r[ch * 8 + 2 * i + 1]));
                            *reinterpret_cast<uint4*>(Wl + sw128_offset(prow, (uint32_t)(hh * 4 + ch))) =
                                make_uint4(p[0], p[1], p[2], p[3]);
                        }
                    }
                }
                if (WROWS) {
                    fence_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(wraw_full + l);
                }
            }
            if (!WROWS) tc_fence_before();
            FD_TOC(4);

            if (inside) {
                a.final_T[pix] = T;
                a.n_contrib[pix] = last;
                a.color[pix] = fmaf(T, a.bg[0], C0);
                a.color[HW + pix] = fmaf(T, a.bg[1], C1);
                a.color[2 * HW + pix] = fmaf(T, a.bg[2], C2);
            }

            // ---------------- 1 / (|F_l| + 1e-10) per pixel from the Gram products ----------------
            // |F|^2 = w^T (C C^T) w with w the fp16 row the decode multiplies: U = Wdec.G comes back in TMEM block
            // `half`, lane quadrant `strip`, i.e. for decode row nrow (not this thread's own pixel -- any thread may
            // serve any row, the rows are in shared memory).  The scale itself is applied in fp32 by the epilogue.
            if (WROWS) {
                if (normalize) {
                    float* inv_u = s_inv + (size_t)(units_done & 1) * 3 * 256;
                    // the decode row nrow is band row (nrow >> 5), column (nrow & 31)
                    const int qx = (unit % a.ux) * FD_UNIT_W + (int)(nrow & 31u), qy = (unit / a.ux) * FD_UNIT_H + (int)(nrow >> 5);
                    for (int l = 0; l < levels; ++l) {
                        const unsigned char* Wl = wdec + (size_t)l * FD_WDEC_LEVEL;
                        FD_WAIT(u_full + l, units_done & 1);
                        tc_fence_after();
                        float ss = 0.0f;
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            uint32_t r[32];
                            tmem_ld32(lane_base + (uint32_t)(half * FD_MAX_L + l * FD_LC + hh * 32), r);
                            tmem_wait_ld();
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch) {
                                const uint4 w8 = *reinterpret_cast<const uint4*>(Wl + sw128_offset(nrow, (uint32_t)(hh * 4 + ch)));
                                const uint32_t ww[4] = {w8.x, w8.y, w8.z, w8.w};
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const float2 wf = __half22float2(*reinterpret_cast<const __half2*>(&ww[i]));
                                    ss = fmaf(__uint_as_float(r[ch * 8 + 2 * i]), wf.x, ss);
                                    ss = fmaf(__uint_as_float(r[ch * 8 + 2 * i + 1]), wf.y, ss);
                                }
                            }
                        }
                        const float inv = 1.0f / (sqrtf(fmaxf(ss, 0.0f) * a.gscale_inv[l]) + 1e-10f);
                        if (DECODE) inv_u[l * 256 + nrow] = inv;
                        if (QUERY) {
                            // similarities of the normalised feature with every phrase: sim = (w . M) / |F|; then
                            // min over the negatives of softmax(10 [pos, neg])[0] = 1 / (1 + exp(10 (max neg - pos)))
                            // (eval/openclip_encoder.py:126-137: the pairwise softmax is monotone in the negative)
                            const uint32_t nq = units_done * (uint32_t)levels + (uint32_t)l;
                            FD_WAIT(f_full, nq & 1);
                            tc_fence_after();
                            uint32_t s0[32], s1[32];
                            tmem_ld32(lane_base + (uint32_t)(FD_TMEM_F + half * 64), s0);
                            tmem_ld32(lane_base + (uint32_t)(FD_TMEM_F + half * 64 + 32), s1);
                            tmem_wait_ld();
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(f_free);             // the next level's product may land
                            float maxneg = -1e30f;
#pragma unroll
                            for (int i = 0; i < 64; ++i) {
                                const float v = __uint_as_float(i < 32 ? s0[i] : s1[i - 32]) * inv;
                                if (i >= a.n_pos && i < a.n_pos + a.n_neg) maxneg = fmaxf(maxneg, v);
                            }
                            if (qx < a.W && qy < a.H) {
                                float* o = a.qout + ((size_t)l * a.n_pos * a.H + qy) * a.W + qx;
#pragma unroll
                                for (int i = 0; i < 64; ++i) {
                                    if (i < a.n_pos) {
                                        const float v = __uint_as_float(i < 32 ? s0[i] : s1[i - 32]) * inv;
                                        st_stream_f32(o + (size_t)i * HW, 1.0f / (1.0f + expf(10.0f * (maxneg - v))));
                                    }
                                }
                            }
                        }
                    }
                    tc_fence_before();
                }
                if (DECODE) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(inv_full);
                }
            }
            FD_TOC(5);
            ++units_done;
        }
        if (tid == 0) { FD_PROF_FLUSH(0); FD_PROFB_FLUSH(4); }
    }
#ifdef FD_PROF
    // (role 0 = blend warp 0 and warp 5: two samples of the pixel warps)
#endif
    tc_fence_before();
    __syncthreads();
    if (warp == FD_WARP_MB) tmem_dealloc(tmem, 512);
}

// ---- host side ------------------------------------------------------------------------------------------------
size_t render_decode_scratch_bytes(int levels)
{
    const size_t lv = (size_t)(levels > 0 ? levels : 1);
    return 512 + lv * FD_D * 128 + lv * FD_CB_BYTES + 512 + (size_t)NUM_SMS * 6 * 8 * sizeof(long long);
}

static int fd_check_features(const lsv2_features& f)
{
    if (f.L <= 0 || f.L > FD_MAX_L || (f.L % FD_LC) != 0) return LSV2_EUNSUPPORTED;
    if (f.mode == LSV2_MODE_QUICK) {
        if (f.K <= 0) return LSV2_EINVAL;
        if (f.K > FD_KMAX) return LSV2_EUNSUPPORTED;
    } else if (f.mode == LSV2_MODE_DENSE) {
        if (!f.coeff_dense) return LSV2_EINVAL;
    } else {
        return LSV2_EINVAL;
    }
    if (!f.packed) return LSV2_EINVAL;              // lsv2_pack_coeff first
    return LSV2_OK;
}

static void fd_fill_args(FdArgs& a, const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                         const lsv2_image& img, int gx, int gy)
{
    a.W = v.width; a.H = v.height; a.gx = gx; a.n_tiles = gx * gy;
    a.ux = (v.width + FD_UNIT_W - 1) / FD_UNIT_W;
    a.n_units = a.ux * ((v.height + FD_UNIT_H - 1) / FD_UNIT_H);
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.rec = reinterpret_cast<const float4*>(g.blend_rec);
    a.bg = v.bg;
    a.mode = f.mode; a.L = f.L; a.K = f.K; a.levels = f.L / FD_LC;
    a.coeff_dense = f.coeff_dense; a.crec = f.packed;
    a.color = img.color; a.wmap = img.wmap; a.final_T = img.final_T; a.n_contrib = img.n_contrib;
    a.status = b.status;
    a.bimg = nullptr; a.gimg = nullptr; a.gscale_inv = nullptr; a.out = nullptr; a.normalize = 0;
    a.hwc = 0; a.qimg = nullptr; a.qout = nullptr; a.n_pos = 0; a.n_neg = 0;
    a.prof = nullptr;
    a.tile_counter = nullptr;
}

int launch_render_decode(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                         const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                         int out_dtype, int out_hwc, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec || !codebooks || !out) return LSV2_EINVAL;
    if (Lc != FD_LC || D != FD_D || levels < 1 || levels * Lc > FD_MAX_L) return LSV2_EUNSUPPORTED;
    if (f.L != levels * Lc) return LSV2_EINVAL;
    { const int rc = fd_check_features(f); if (rc) return rc; }
    if (!scratch || scratch_bytes < render_decode_scratch_bytes(levels)) return LSV2_ENOSPC;
    uint8_t* base = reinterpret_cast<uint8_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    uint32_t* counter = reinterpret_cast<uint32_t*>(base);
    float* gscale_inv = reinterpret_cast<float*>(base + 128);
    uint8_t* bimg = base + 256;
    uint8_t* gimg = bimg + (size_t)levels * FD_D * 128;
    LSV2_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
    k_fd_prep_b<<<ceil_div(levels * FD_D * 8, 256), 256, 0, st>>>(levels, codebooks, bimg);
    LSV2_LAUNCH_CHECK("k_fd_prep_b");
    if (normalize) {
        k_fd_prep_g<<<dim3(FD_LC, levels), 64, 0, st>>>(codebooks, gimg, gscale_inv);
        LSV2_LAUNCH_CHECK("k_fd_prep_g");
    }

    FdArgs a;
    fd_fill_args(a, v, g, b, f, img, gx, gy);
    a.bimg = bimg; a.gimg = gimg; a.gscale_inv = gscale_inv; a.out = out; a.normalize = normalize; a.hwc = out_hwc;
    a.tile_counter = counter;
#ifdef FD_PROF
    a.prof = reinterpret_cast<long long*>(gimg + (size_t)levels * FD_CB_BYTES);
#endif
    const int grid = min(NUM_SMS, a.n_units);
#define FD_LAUNCH(O, C)                                                                                         \
    do {                                                                                                        \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode<O, FD_PATH_DECODE, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           (int)FD_SMEM_TOTAL));                                                \
        k_render_decode<O, FD_PATH_DECODE, C><<<grid, FD_THREADS, FD_SMEM_TOTAL, st>>>(a);                      \
    } while (0)
    if (out_hwc) { if (out_dtype == 0) FD_LAUNCH(false, true); else FD_LAUNCH(true, true); }
    else         { if (out_dtype == 0) FD_LAUNCH(false, false); else FD_LAUNCH(true, false); }
#undef FD_LAUNCH
    LSV2_LAUNCH_CHECK("k_render_decode");
    return LSV2_OK;
}

// Blend + relevancy query (eval/openclip_encoder.py:114-138 on top of eval_lerf.py:214-218): the 512-d map is never
// formed -- sim = w^T (C E^T) / |C^T w| per level and phrase, then the pairwise softmax / min over the negatives.
// scratch: counter | 1/s per level | Gram images | query images.
size_t render_query_scratch_bytes(int levels)
{
    const size_t lv = (size_t)(levels > 0 ? levels : 1);
    return 512 + 2 * lv * FD_CB_BYTES + 512;
}

int launch_render_query(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                        const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, const float* embeds,
                        int n_pos, int n_neg, float* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec || !codebooks || !embeds || !out) return LSV2_EINVAL;
    if (Lc != FD_LC || D != FD_D || levels < 1 || levels * Lc > FD_MAX_L) return LSV2_EUNSUPPORTED;
    if (n_pos < 1 || n_neg < 1 || n_pos + n_neg > 64) return LSV2_EUNSUPPORTED;
    if (f.L != levels * Lc) return LSV2_EINVAL;
    { const int rc = fd_check_features(f); if (rc) return rc; }
    if (!scratch || scratch_bytes < render_query_scratch_bytes(levels)) return LSV2_ENOSPC;
    uint8_t* base = reinterpret_cast<uint8_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    uint32_t* counter = reinterpret_cast<uint32_t*>(base);
    float* gscale_inv = reinterpret_cast<float*>(base + 128);
    uint8_t* gimg = base + 256;
    uint8_t* qimg = gimg + (size_t)levels * FD_CB_BYTES;
    LSV2_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
    k_fd_prep_g<<<dim3(FD_LC, levels), 64, 0, st>>>(codebooks, gimg, gscale_inv);
    LSV2_LAUNCH_CHECK("k_fd_prep_g");
    k_fd_prep_q<<<dim3(FD_LC, levels), 64, 0, st>>>(codebooks, embeds, n_pos + n_neg, qimg);
    LSV2_LAUNCH_CHECK("k_fd_prep_q");
    FdArgs a;
    fd_fill_args(a, v, g, b, f, img, gx, gy);
    a.gimg = gimg; a.gscale_inv = gscale_inv; a.qimg = qimg; a.qout = out; a.n_pos = n_pos; a.n_neg = n_neg;
    a.normalize = 1;
    a.tile_counter = counter;
    const int grid = min(NUM_SMS, a.n_units);
    LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode<false, FD_PATH_QUERY>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)FD_SMEM_TOTAL));
    k_render_decode<false, FD_PATH_QUERY><<<grid, FD_THREADS, FD_SMEM_TOTAL, st>>>(a);
    LSV2_LAUNCH_CHECK("k_render_decode<query>");
    return LSV2_OK;
}

// The same blend with the weight map as its product (the 3-tuple of the drop-in op): W leaves TMEM as fp32
// [L][H][W], nothing is decoded.  scratch: 128 bytes (the tile counter).
int launch_render_wmap_tc(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                          const lsv2_image& img, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec || !img.wmap) return LSV2_EINVAL;
    { const int rc = fd_check_features(f); if (rc) return rc; }
    if (!scratch || scratch_bytes < 256) return LSV2_ENOSPC;
    uint32_t* counter = reinterpret_cast<uint32_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    LSV2_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
    FdArgs a;
    fd_fill_args(a, v, g, b, f, img, gx, gy);
    a.tile_counter = counter;
    const int grid = min(NUM_SMS, a.n_units);
    LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode<false, FD_PATH_WMAP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)FD_SMEM_NODECODE_TOTAL));
    k_render_decode<false, FD_PATH_WMAP><<<grid, FD_THREADS_NODECODE, FD_SMEM_NODECODE_TOTAL, st>>>(a);
    LSV2_LAUNCH_CHECK("k_render_decode<wmap>");
    return LSV2_OK;
}

}  // namespace lsv2
