// render_decode.cu -- K6+K7 fused: per-tile alpha blending whose sparse-coefficient accumulation runs on the
// 5th-gen tensor cores, followed in the same CTA by the global codebook decode + L2 normalisation.
// The L-channel weight map never exists in HBM (nor in shared memory as read-modify-write accumulators).
// Replaces renderCUDA (forward) + the einsum/norm/divide every caller runs after it
// (/root/reference/gaussian_renderer/__init__.py:108-119 ; eval_lerf.py:214-218 ; SURVEY.md rows R8+R9).
//
// Persistent CTAs (one per SM), tiles handed out by an atomic counter.  Warp roles:
//   warps 0-7  blend + epilogue: one thread per pixel of the 16x16 tile (warp = 16x2 strip).  Pixels 0..127 are
//              rows of MMA half 0, pixels 128..255 of half 1; thread = TMEM lane.
//   warps 8-9  producer: gathers the packed blend records of 64 depth-sorted instances per batch, builds the
//              per-strip cut-off masks, and scatters each instance's (channel, weight) pairs into a zeroed
//              fp16 B tile  S[L channels][64 instances]  (K-major, 128B-swizzled).
//   warp 10    one thread issues every tcgen05.mma and the bulk copies of the codebook tiles.
//
// Blend phase, per batch:  the pixel threads evaluate alpha and T exactly as k_render_forward does (same fp32
// operations, same thresholds => identical n_contrib / final_T / RGB) and store the blending weight alpha*T of
// each (pixel, instance) as fp16 into the A tile [128 px][64 inst] of their half;  W[128 x L] += A . S  (M=128,
// N=L, K=64, fp32 accumulators in TMEM columns [h*L, h*L+L)).
// Decode phase, per tile:  W is read once from TMEM, rounded to fp16 into A operands [128 px][64 codes] per level
// (aliasing the blend ring, idle by then), and  F[128 x 128] = W_l . C_l[:, chunk]  runs chunk by chunk into two
// ping-pong TMEM slots per half; pass 0 sums squares (thread = pixel => thread-local L2 norm), pass 1 re-issues the
// MMAs, scales and streams the features to HBM.  One level's codebook (64 KB fp16) is resident in shared memory and
// re-fetched level by level from L2 with bulk async copies.
#include "fd_common.cuh"

namespace lsv2 {

// codebooks f32 [levels][64][512] -> fp16 B image [levels][512 rows n][64 k], 128B-swizzled
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

// everything the producer gathers for one instance (kept in registers one batch ahead)
struct FdGather {
    uint32_t g, cnt;
    float4 r0, r1, r2;
    float wv[FD_KMAX], iv[FD_KMAX];
};

template <int MODE>
__device__ __forceinline__ void fd_gather(const FdArgs& a, uint32_t g, FdGather& o)
{
    o.g = g;
    o.cnt = 0;
    if (g == 0xFFFFFFFFu) return;
    o.r0 = __ldg(a.rec + 3 * (size_t)g);
    o.r1 = __ldg(a.rec + 3 * (size_t)g + 1);
    o.r2 = __ldg(a.rec + 3 * (size_t)g + 2);
    if (MODE == LSV2_MODE_QUICK) {
        o.cnt = (uint32_t)a.K;
        const float* wsrc = a.wq + (size_t)g * a.K;
        const float* isrc = a.iq + (size_t)g * a.K;
        if ((a.K & 3) == 0) {
#pragma unroll
            for (int k4 = 0; k4 < FD_KMAX / 4; ++k4)
                if (k4 * 4 < a.K) {
                    const float4 w4 = __ldg(reinterpret_cast<const float4*>(wsrc) + k4);
                    const float4 i4 = __ldg(reinterpret_cast<const float4*>(isrc) + k4);
                    o.wv[4 * k4] = w4.x; o.wv[4 * k4 + 1] = w4.y; o.wv[4 * k4 + 2] = w4.z; o.wv[4 * k4 + 3] = w4.w;
                    o.iv[4 * k4] = i4.x; o.iv[4 * k4 + 1] = i4.y; o.iv[4 * k4 + 2] = i4.z; o.iv[4 * k4 + 3] = i4.w;
                }
        } else {
#pragma unroll
            for (int k = 0; k < FD_KMAX; ++k)
                if (k < a.K) { o.wv[k] = __ldg(wsrc + k); o.iv[k] = __ldg(isrc + k); }
        }
    } else {
        o.cnt = a.sp_count[g];
        if (o.cnt != FD_ROW_DENSE) {
#pragma unroll
            for (int k = 0; k < FD_KMAX; ++k)
                if ((uint32_t)k < o.cnt) {
                    o.wv[k] = a.sp_weight[(size_t)g * FD_KMAX + k];
                    o.iv[k] = (float)a.sp_index[(size_t)g * FD_KMAX + k];
                }
        }
    }
}

template <int MODE, bool OUT_BF16>
__global__ void __launch_bounds__(FD_THREADS, 1) k_render_decode(FdArgs a)
{
    extern __shared__ unsigned char smem_raw[];
    // offset arithmetic on the __shared__ symbol (not a uintptr_t round trip) keeps the shared address space
    // visible to the compiler: LDS/STS instead of generic LD/ST in the blend loop
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* ring = smem + FD_SMEM_RING;
    unsigned char* s_b = smem + FD_SMEM_B;
    FdHeader* hdrs = reinterpret_cast<FdHeader*>(smem + FD_SMEM_HDR);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FD_SMEM_BAR);
    uint64_t* full = bars;                        // [2]  producer -> blend, mma     (64 arrivals)
    uint64_t* a_full = bars + 2;                  // [2]  blend -> mma               (256 arrivals)
    uint64_t* stage_free = bars + 4;              // [2]  mma -> producer            (commit / arrive)
    uint64_t* w_full = bars + 6;                  //      mma -> blend: W complete   (commit / arrive)
    uint64_t* adec_full = bars + 7;               //      blend -> mma               (256 arrivals)
    uint64_t* ring_free = bars + 8;               //      mma -> producer            (commit / arrive)
    uint64_t* b_full = bars + 9;                  // [4]  bulk copy landed           (tx)
    uint64_t* b_free = bars + 13;                 // [4]  mma done with chunk        (commit)
    uint64_t* f_full = bars + 17;                 // [2]  F slot ready               (commit)
    uint64_t* f_free = bars + 19;                 // [2]  F slot drained             (256 arrivals)
    uint64_t* g_full = bars + 21;                 //      Gram images landed         (tx)
    uint64_t* u_full = bars + 22;                 //      U = W.G ready              (commit)
    uint64_t* adec2_full = bars + 23;             //      blend -> mma: W rescaled   (256 arrivals)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 24);
    uint32_t* s_ndone = s_tmem + 1;               // [2]  saturated strips, by tile parity
    int32_t* s_tile = reinterpret_cast<int32_t*>(s_tmem + 3);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = a.L, levels = a.levels;
    const size_t HW = (size_t)a.H * a.W;

    if (tid == 0) {
        for (int s = 0; s < FD_STAGES; ++s) {
            mbar_init(full + s, FD_PROD_THREADS);
            mbar_init(a_full + s, 32 * FD_BLEND_WARPS);
            mbar_init(stage_free + s, 1);
            mbar_init(f_full + s, 1);
            mbar_init(f_free + s, 32 * FD_BLEND_WARPS);
        }
        mbar_init(w_full, 1);
        mbar_init(adec_full, 32 * FD_BLEND_WARPS);
        mbar_init(ring_free, 1);
        mbar_init(g_full, 1);
        mbar_init(u_full, 1);
        mbar_init(adec2_full, 32 * FD_BLEND_WARPS);
        for (int c = 0; c < FD_CHUNKS; ++c) { mbar_init(b_full + c, 1); mbar_init(b_free + c, 1); }
        s_ndone[0] = 0;
        s_ndone[1] = 0;
        mbar_fence_init();
    }
    if (warp == 10) tmem_alloc(s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp >= FD_BLEND_WARPS && warp < FD_BLEND_WARPS + 2) {
        // ======================================= producer =======================================
        // The gathers of batch b+1 (ids two batches ahead) are issued before batch b is assembled, so the
        // global-memory latency is off the per-batch critical path: wait stage -> zero B tile -> scatter -> arrive.
        const int pt = tid - 32 * FD_BLEND_WARPS;       // 0..63 = instance slot of the batch
        const int pw = pt >> 5;
        uint32_t use_g = 0;                             // stage uses so far (global)
        for (uint32_t t_iter = 0;; ++t_iter) {
            if (pt == 0) *s_tile = (int32_t)atomicAdd(a.tile_counter, 1u);
            named_bar_sync(1, FD_PROD_THREADS);
            const int tile = *s_tile;
            named_bar_sync(1, FD_PROD_THREADS);
            const bool exit_now = tile >= a.n_tiles;
            uint2 range = make_uint2(0, 0);
            if (!exit_now) range = a.ranges[tile];
            const uint32_t n_inst = range.y - range.x;
            const uint32_t n_batches = (n_inst + FD_BATCH - 1) / FD_BATCH;
            const float y0 = (float)((tile / a.gx) * TILE), x0 = (float)((tile % a.gx) * TILE);
            auto load_id = [&](uint32_t b) -> uint32_t {
                const uint32_t i = b * FD_BATCH + pt;
                return i < n_inst ? __ldg(a.point_list + range.x + i) : 0xFFFFFFFFu;
            };
            FdGather cur, nxt;
            fd_gather<MODE>(a, load_id(0), cur);
            uint32_t gid_next = load_id(1);
            if (t_iter > 0 && !exit_now) mbar_wait_sleep(ring_free, (t_iter - 1) & 1);
            for (uint32_t b = 0;; ++b, ++use_g) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                if (use > 0) mbar_wait_sleep(stage_free + s, (use - 1) & 1);
                FdHeader& Hd = hdrs[s];
                const bool stop = exit_now || b >= n_batches ||
                                  *(volatile uint32_t*)(s_ndone + (t_iter & 1)) >= FD_BLEND_WARPS;
                // s_ndone may change between the 64 reads: OR-reduce so the decision is uniform
                if (named_bar_or(1, FD_PROD_THREADS, stop)) {
                    if (pt == 0) { Hd.count = exit_now ? -2 : -1; Hd.tile = tile; Hd.batches = (int32_t)b; }
                    mbar_arrive(full + s);
                    ++use_g;
                    break;
                }
                unsigned char* S = ring + (size_t)s * FD_STAGE_BYTES + 2 * FD_TILE16K;
                for (int q = pt; q < L * 8; q += FD_PROD_THREADS)
                    reinterpret_cast<uint4*>(S)[q] = make_uint4(0, 0, 0, 0);
                fd_gather<MODE>(a, gid_next, nxt);          // in flight while this batch is assembled
                gid_next = load_id(b + 2);
                named_bar_sync(1, FD_PROD_THREADS);
                uint32_t m8 = 0;
                if (cur.g != 0xFFFFFFFFu) {
                    Hd.g0[pt] = cur.r0;
                    Hd.g1[pt] = cur.r1;
                    Hd.g2[pt] = make_float2(cur.r2.x, cur.r2.y);
                    m8 = fd_strip_mask(cur.r0.y - y0, cur.r0.z, cur.r0.w, cur.r1.x, cur.r1.y, cur.r0.x - x0);
                    const uint32_t col = (uint32_t)(pt & 7) * 2u, chunk = (uint32_t)pt >> 3;
                    if (cur.cnt != FD_ROW_DENSE) {
                        bool ascending = true;              // distinct channels: plain stores, no read-modify-write
#pragma unroll
                        for (int k = 1; k < FD_KMAX; ++k)
                            if ((uint32_t)k < cur.cnt) ascending &= cur.iv[k] > cur.iv[k - 1];
#pragma unroll
                        for (int k = 0; k < FD_KMAX; ++k) {
                            if ((uint32_t)k < cur.cnt) {
                                const int c = (int)cur.iv[k];
                                if (c >= 0 && c < L) {
                                    __half* p = reinterpret_cast<__half*>(S + sw128_offset((uint32_t)c, chunk) + col);
                                    if (ascending) {
                                        *p = __float2half_rn(cur.wv[k]);
                                    } else {
                                        const float old = __half2float(*p);    // duplicates of a channel add up
                                        *p = __float2half_rn(old + cur.wv[k]);
                                    }
                                } else {
                                    a.status[LSV2_STATUS_BAD_INDEX] = 1u;
                                }
                            }
                        }
                    } else {
                        const float* row = a.coeff_dense + (size_t)cur.g * L;
                        for (int c = 0; c < L; ++c) {
                            const float v = __ldg(row + c);
                            if (v != 0.0f)
                                *reinterpret_cast<__half*>(S + sw128_offset((uint32_t)c, chunk) + col) = __float2half_rn(v);
                        }
                    }
                }
#pragma unroll
                for (int w = 0; w < FD_BLEND_WARPS; ++w) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, (m8 >> w) & 1u);
                    if (lane == 0) Hd.mask[w][pw] = bal;
                }
                if (pt == 0) {
                    Hd.count = (int32_t)min((uint32_t)FD_BATCH, n_inst - b * FD_BATCH);
                    Hd.tile = tile;
                }
                fence_async_smem();
                mbar_arrive(full + s);
                cur = nxt;
            }
            if (exit_now) break;
        }
    } else if (warp == 10) {
        // ================================ MMA issue + codebook fetch ================================
        if (lane == 0) {
            const uint32_t idesc_blend = umma_idesc_f16_f32(128, L);
            constexpr uint32_t idesc_dec = umma_idesc_f16_f32(128, FD_CHUNK_N);
            const uint32_t ring_addr = smem_u32(ring), b_addr = smem_u32(s_b);
            const int passes = a.normalize ? 2 : 1;
            uint32_t b_issued[FD_CHUNKS] = {0, 0, 0, 0};     // bulk copies issued per chunk slot
            uint32_t b_used[FD_CHUNKS] = {0, 0, 0, 0};       // b_full phases consumed per chunk slot
            uint32_t f_n = 0;                                // F slot fills so far (slots alternate)
            int next_level = 0;                              // level the next codebook fetch belongs to
            auto fetch = [&](int c) {                        // chunk c of level `lvl`
                mbar_arrive_expect_tx(b_full + c, FD_TILE16K);
                bulk_g2s(s_b + (size_t)c * FD_TILE16K, a.bimg + ((size_t)next_level * FD_CHUNKS + c) * FD_TILE16K,
                         FD_TILE16K, b_full + c);
                ++b_issued[c];
            };
            for (int c = 0; c < FD_CHUNKS; ++c) fetch(c);    // level 0, ready before the first tile needs it
            next_level = (next_level + 1) % levels;
            uint32_t tiles_done = 0;
            uint32_t batch_in_tile = 0;
            uint32_t a_uses[FD_STAGES] = {0, 0};             // a_full phases consumed (batches only)
            for (uint32_t use_g = 0;; ++use_g) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                mbar_wait_sleep(full + s, use & 1);
                const FdHeader& Hd = hdrs[s];
                const int count = Hd.count;
                if (count == -2) break;
                if (count >= 0) {
                    mbar_wait(a_full + s, a_uses[s] & 1);
                    ++a_uses[s];
                    tc_fence_after();
                    const uint32_t sa = ring_addr + (uint32_t)s * FD_STAGE_BYTES;
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_bf16(tmem + (uint32_t)(h * L), umma_desc_k_sw128(sa + h * FD_TILE16K + ks * 32),
                                      umma_desc_k_sw128(sa + 2 * FD_TILE16K + ks * 32), idesc_blend,
                                      (batch_in_tile > 0 || ks > 0) ? 1u : 0u);
                    umma_commit(stage_free + s);
                    ++batch_in_tile;
                    continue;
                }
                // ---- end of tile ----
                const uint32_t nb = batch_in_tile;
                batch_in_tile = 0;
                mbar_arrive(stage_free + s);
                if (nb > 0) umma_commit(w_full); else mbar_arrive(w_full);
                mbar_wait_sleep(adec_full, tiles_done & 1);
                tc_fence_after();
                if (nb == 0) {
                    mbar_arrive(ring_free);
                    ++tiles_done;
                    continue;
                }
                for (int l = 0; l < levels; ++l) {
                    for (int pass = 0; pass < passes; ++pass) {
                        for (int c = 0; c < FD_CHUNKS; ++c) {
                            const int q = (int)(f_n & 1u);
                            if (f_n >= 2) mbar_wait(f_free + q, ((f_n >> 1) - 1) & 1);
                            if (pass == 0) { mbar_wait(b_full + c, b_used[c] & 1); ++b_used[c]; }
                            tc_fence_after();
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const uint32_t aa = ring_addr + (uint32_t)(h * levels + l) * FD_TILE16K;
#pragma unroll
                                for (int ks = 0; ks < 4; ++ks)
                                    umma_bf16(tmem + (uint32_t)(h * 256 + q * FD_CHUNK_N), umma_desc_k_sw128(aa + ks * 32),
                                              umma_desc_k_sw128(b_addr + c * FD_TILE16K + ks * 32), idesc_dec,
                                              ks > 0 ? 1u : 0u);
                            }
                            umma_commit(f_full + q);
                            ++f_n;
                            if (pass == passes - 1) {
                                umma_commit(b_free + c);
                                // refill the chunk slot released one step ago with the next level's tile
                                const int pc = c - 1;
                                if (pc >= 0) {
                                    mbar_wait(b_free + pc, (b_issued[pc] - 1) & 1);
                                    fetch(pc);
                                }
                            }
                        }
                    }
                    // last chunk slot of this level
                    mbar_wait(b_free + (FD_CHUNKS - 1), (b_issued[FD_CHUNKS - 1] - 1) & 1);
                    fetch(FD_CHUNKS - 1);
                    next_level = (next_level + 1) % levels;
                }
                umma_commit(ring_free);
                ++tiles_done;
            }
            // drain the codebook prefetch that no tile consumed
            for (int c = 0; c < FD_CHUNKS; ++c)
                while (b_used[c] < b_issued[c]) { mbar_wait(b_full + c, b_used[c] & 1); ++b_used[c]; }
        }
    } else {
        // ==================================== blend + epilogue ====================================
        const int half = warp >> 2;                                  // MMA half this strip belongs to
        const int row = (warp & 3) * 32 + lane;                      // A-tile row == TMEM lane
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int lx = lane & 15, ly = warp * 2 + (lane >> 4);
        uint32_t f_n = 0;                                            // F slots consumed so far (slots alternate)
        uint32_t tiles_done = 0;
        uint32_t use_g = 0;
        for (;;) {
            // ---------------- blend phase ----------------
            float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
            uint32_t last = 0;
            bool done = false, warp_done = false, first = true;
            int tile = 0, px = 0, py = 0;
            float pxf = 0.f, pyf = 0.f;
            bool inside = false;
            int count;
            int nb = 0;
            for (;; ++use_g) {
                const int s = use_g & 1;
                mbar_wait(full + s, (use_g >> 1) & 1);
                const FdHeader& Hd = hdrs[s];
                count = Hd.count;
                if (count < 0) { tile = Hd.tile; nb = Hd.batches; ++use_g; break; }
                if (first) {
                    first = false;
                    tile = Hd.tile;
                    px = (tile % a.gx) * TILE + lx;
                    py = (tile / a.gx) * TILE + ly;
                    pxf = (float)px; pyf = (float)py;
                    inside = px < a.W && py < a.H;
                    done = !inside;
                    warp_done = __all_sync(0xffffffffu, done);
                    if (warp_done && lane == 0) atomicAdd(s_ndone + (tiles_done & 1), 1u);
                    if (warp == 0 && lane == 0) s_ndone[(tiles_done + 1) & 1] = 0;
                }
                unsigned char* A = ring + (size_t)s * FD_STAGE_BYTES + (size_t)half * FD_TILE16K;
                const uint64_t m = warp_done ? 0ull : (((uint64_t)Hd.mask[warp][1] << 32) | Hd.mask[warp][0]);
#pragma unroll 1
                for (int c = 0; c < 8; ++c) {
                    const uint32_t m8 = (uint32_t)(m >> (8 * c)) & 0xFFu;
                    uint32_t hw[4] = {0u, 0u, 0u, 0u};
                    if (m8 != 0u && !(a.dbg & 2)) {
#pragma unroll
                        for (int q4 = 0; q4 < 2; ++q4) {
                            const uint32_t m4 = (m8 >> (4 * q4)) & 0xFu;
                            if (m4 == 0u) continue;
                            // stage 1: the four quadratic forms and alphas are independent (ILP); alpha does not
                            // depend on T.  stage 2: the short sequential transmittance chain.
                            float pw_[4], op_[4], cr_[4];
                            bool hit[4];
                            bool any = false;
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const int j = 8 * c + 4 * q4 + u;
                                const float4 g0 = Hd.g0[j];
                                const float4 g1 = Hd.g1[j];
                                const float dx = g0.x - pxf, dy = g0.y - pyf;
                                pw_[u] = fmaf(g0.z * dx, dx, fmaf(g1.x * dy, dy, (g0.w * dx) * dy));
                                hit[u] = ((m4 >> u) & 1u) && !(pw_[u] > 0.0f || pw_[u] < g1.y);
                                op_[u] = g1.z;
                                cr_[u] = g1.w;
                                any |= hit[u];
                            }
                            float wgt[4] = {0.f, 0.f, 0.f, 0.f};
                            if (any && !done) {
                                float al[4];
#pragma unroll
                                for (int u = 0; u < 4; ++u)
                                {   // a hit has cut-off <= power <= 0 (cut-off > -6), so the unchecked exp is exact there;
                                    // evaluated for all four so the chains interleave, discarded where !hit
                                    const float e = repro_expf_inrange(fmaxf(fminf(pw_[u], 0.0f), -80.0f));
                                    al[u] = hit[u] ? fminf(ALPHA_MAX, op_[u] * e) : 0.0f;
                                }
#pragma unroll
                                for (int u = 0; u < 4; ++u) {
                                    // select-only form of: skip if alpha < 1/255 or done; stop if T' < 1e-4; else blend.
                                    // A skipped instance adds w = 0 (exact), so the sums match the branching form bit for bit.
                                    const int j = 8 * c + 4 * q4 + u;
                                    const bool valid = !(al[u] < ALPHA_MIN) && !done;
                                    const float test_T = T * (1.0f - al[u]);
                                    const bool stop = valid && (test_T < T_MIN);
                                    const bool use = valid && !stop;
                                    done = done || stop;
                                    const float w = use ? al[u] * T : 0.0f;
                                    const float2 g2 = Hd.g2[j];
                                    C0 = fmaf(cr_[u], w, C0);
                                    C1 = fmaf(g2.x, w, C1);
                                    C2 = fmaf(g2.y, w, C2);
                                    wgt[u] = w;
                                    T = use ? test_T : T;
                                    last = use ? (uint32_t)nb * FD_BATCH + (uint32_t)j + 1u : last;
                                }
                            }
                            hw[2 * q4] = pack_f16x2(wgt[0], wgt[1]);
                            hw[2 * q4 + 1] = pack_f16x2(wgt[2], wgt[3]);
                        }
                    }
                    *reinterpret_cast<uint4*>(A + sw128_offset((uint32_t)row, (uint32_t)c)) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
                }
                ++nb;
                fence_async_smem();
                tc_fence_before();
                mbar_arrive(a_full + s);
                if (!warp_done && __all_sync(0xffffffffu, done)) {
                    warp_done = true;
                    if (lane == 0) atomicAdd(s_ndone + (tiles_done & 1), 1u);
                }
            }
            if (count == -2) break;
            if (first) {                                   // a tile with an empty list
                px = (tile % a.gx) * TILE + lx;
                py = (tile / a.gx) * TILE + ly;
                inside = px < a.W && py < a.H;
                if (warp == 0 && lane == 0) s_ndone[(tiles_done + 1) & 1] = 0;
            }
            const size_t pix = (size_t)py * a.W + px;

            // ---------------- W: TMEM -> fp16 A operands of the decode ----------------
            mbar_wait(w_full, tiles_done & 1);
            tc_fence_after();
            if (nb > 0) {
                for (int l = 0; l < levels; ++l) {
                    unsigned char* Ad = ring + (size_t)(half * levels + l) * FD_TILE16K;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        uint32_t r[32];
                        tmem_ld32(lane_base + (uint32_t)(half * L + l * FD_LC + hh * 32), r);
                        tmem_wait_ld();
                        if (a.wmap != nullptr && inside) {
                            float* o = a.wmap + (size_t)(l * FD_LC + hh * 32) * HW + pix;
#pragma unroll
                            for (int i = 0; i < 32; ++i) st_stream_f32(o + (size_t)i * HW, __uint_as_float(r[i]));
                        }
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) {
                            uint32_t p[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                p[i] = pack_f16x2(__uint_as_float(r[ch * 8 + 2 * i]), __uint_as_float(r[ch * 8 + 2 * i + 1]));
                            *reinterpret_cast<uint4*>(Ad + sw128_offset((uint32_t)row, (uint32_t)(hh * 4 + ch))) =
                                make_uint4(p[0], p[1], p[2], p[3]);
                        }
                    }
                }
                fence_async_smem();
            } else if (a.wmap != nullptr && inside) {
                for (int c = 0; c < L; ++c) st_stream_f32(a.wmap + (size_t)c * HW + pix, 0.0f);
            }
            tc_fence_before();
            mbar_arrive(adec_full);

            if (inside) {
                a.final_T[pix] = T;
                a.n_contrib[pix] = last;
                a.color[pix] = fmaf(T, a.bg[0], C0);
                a.color[HW + pix] = fmaf(T, a.bg[1], C1);
                a.color[2 * HW + pix] = fmaf(T, a.bg[2], C2);
            }

            // ---------------- decode epilogue ----------------
            if (nb == 0) {
                if (inside) {
                    if (OUT_BF16) {
                        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(a.out) + pix;
                        for (int d = 0; d < levels * FD_D; ++d) o[(size_t)d * HW] = __float2bfloat16_rn(0.0f);
                    } else {
                        float* o = reinterpret_cast<float*>(a.out) + pix;
                        for (int d = 0; d < levels * FD_D; ++d) st_stream_f32(o + (size_t)d * HW, 0.0f);
                    }
                }
            } else {
                const int passes = a.normalize ? 2 : 1;
                for (int l = 0; l < levels; ++l) {
                    float ss = 0.0f, inv = 1.0f;
                    for (int pass = 0; pass < passes; ++pass) {
                        const bool store = pass == passes - 1;
                        if (store && a.normalize) inv = 1.0f / (sqrtf(ss) + 1e-10f);
                        for (int c = 0; c < FD_CHUNKS; ++c) {
                            const int q = (int)(f_n & 1u);
                            mbar_wait(f_full + q, (f_n >> 1) & 1);
                            ++f_n;
                            tc_fence_after();
                            const uint32_t col0 = (uint32_t)(half * 256 + q * FD_CHUNK_N);
                            auto consume = [&](uint32_t (&r)[32], int cc) {
                                if (!store) {
#pragma unroll
                                    for (int i = 0; i < 32; ++i) ss = fmaf(__uint_as_float(r[i]), __uint_as_float(r[i]), ss);
                                } else if (inside && !(a.dbg & 1)) {
                                    // scale in place, then 32 stores through four independent running addresses
#pragma unroll
                                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * inv);
                                    const uint64_t esz = OUT_BF16 ? 2u : 4u;
                                    const uint64_t stride = (uint64_t)HW * esz;
                                    const size_t ch0 = (size_t)l * FD_D + c * FD_CHUNK_N + cc * 32;
                                    uint64_t p[4];
                                    p[0] = (uint64_t)(uintptr_t)a.out + (ch0 * HW + pix) * esz;
                                    p[1] = p[0] + stride; p[2] = p[1] + stride; p[3] = p[2] + stride;
                                    const uint64_t s4 = 4 * stride;
#pragma unroll
                                    for (int i = 0; i < 32; i += 4) {
#pragma unroll
                                        for (int u = 0; u < 4; ++u) {
                                            if (OUT_BF16) {
                                                const unsigned short v =
                                                    __bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(r[i + u])));
                                                asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p[u]), "h"(v) : "memory");
                                            } else {
                                                asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p[u]), "r"(r[i + u]) : "memory");
                                            }
                                            p[u] += s4;
                                        }
                                    }
                                }
                            };
                            // two register buffers: the next TMEM read is in flight while 32 columns are consumed
                            uint32_t ra[32], rb[32];
                            const uint32_t t0 = lane_base + col0;
                            tmem_ld32(t0, ra);
                            tmem_wait_ld();
                            tmem_ld32(t0 + 32, rb);
                            consume(ra, 0);
                            tmem_wait_ld();
                            tmem_ld32(t0 + 64, ra);
                            consume(rb, 1);
                            tmem_wait_ld();
                            tmem_ld32(t0 + 96, rb);
                            consume(ra, 2);
                            tmem_wait_ld();
                            tc_fence_before();
                            mbar_arrive_relaxed(f_free + q);      // slot drained: the next MMA may overwrite it
                            consume(rb, 3);
                        }
                    }
                }
            }
            ++tiles_done;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 10) tmem_dealloc(tmem, 512);
}

size_t render_decode2_scratch_bytes(int levels);
int launch_render_decode2(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                          const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                          int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st);

size_t render_decode3_scratch_bytes(int levels);
int launch_render_decode3(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                          const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                          int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st);

static size_t render_decode1_scratch_bytes(int levels)
{
    return (size_t)(levels > 0 ? levels : 1) * FD_D * 128 + 512;
}

size_t render_decode_scratch_bytes(int levels)
{
    const size_t s1 = render_decode1_scratch_bytes(levels), s2 = render_decode2_scratch_bytes(levels);
    return s1 > s2 ? s1 : s2;
}

static int launch_render_decode1(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                                 const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                                 int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st);

// Default: the first-generation kernel (blend and decode phases alternate inside a CTA).  LSV2_FD_V2=1 selects
// render_decode2.cu (blend of tile i+1 overlapped with the decode of tile i; parity-green but, as measured in
// DESIGN.md 5.2, not yet faster: its decode side stalls on L2 latency while the HBM write stream is saturated).
int launch_render_decode(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                         const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                         int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    const char* e3 = getenv("LSV2_FD_V3");
    if (e3 && atoi(e3) != 0)
        return launch_render_decode3(v, g, b, f, img, codebooks, levels, Lc, D, normalize, out_dtype, out, scratch,
                                     scratch_bytes, st);
    const char* e = getenv("LSV2_FD_V2");
    if (e && atoi(e) != 0)
        return launch_render_decode2(v, g, b, f, img, codebooks, levels, Lc, D, normalize, out_dtype, out, scratch,
                                     scratch_bytes, st);
    return launch_render_decode1(v, g, b, f, img, codebooks, levels, Lc, D, normalize, out_dtype, out, scratch,
                                 scratch_bytes, st);
}

static int launch_render_decode1(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
                         const lsv2_image& img, const float* codebooks, int levels, int Lc, int D, int normalize,
                         int out_dtype, void* out, void* scratch, size_t scratch_bytes, cudaStream_t st)
{
    const int gx = (v.width + TILE - 1) / TILE, gy = (v.height + TILE - 1) / TILE;
    if (gx * gy == 0) return LSV2_OK;
    if (!g.blend_rec || !codebooks || !out) return LSV2_EINVAL;
    if (Lc != FD_LC || D != FD_D || levels < 1 || levels * Lc > FD_MAX_L) return LSV2_EUNSUPPORTED;
    if (f.L != levels * Lc) return LSV2_EINVAL;
    if (f.mode == LSV2_MODE_QUICK) {
        if (f.K <= 0 || !f.weights_quick || !f.indices_quick) return LSV2_EINVAL;
        if (f.K > FD_KMAX) return LSV2_EUNSUPPORTED;
    } else if (f.mode == LSV2_MODE_DENSE) {
        if (!f.coeff_dense || !f.sparse_count || !f.sparse_index || !f.sparse_weight) return LSV2_EINVAL;
    } else {
        return LSV2_EINVAL;
    }
    if (!scratch || scratch_bytes < render_decode1_scratch_bytes(levels)) return LSV2_ENOSPC;
    uint8_t* base = reinterpret_cast<uint8_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    uint32_t* counter = reinterpret_cast<uint32_t*>(base);
    uint8_t* bimg = base + 128;
    LSV2_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
    k_fd_prep_b<<<ceil_div(levels * FD_D * 8, 256), 256, 0, st>>>(levels, codebooks, bimg);
    LSV2_LAUNCH_CHECK("k_fd_prep_b");

    FdArgs a;
    a.W = v.width; a.H = v.height; a.gx = gx; a.n_tiles = gx * gy;
    a.ranges = reinterpret_cast<const uint2*>(b.ranges);
    a.point_list = b.point_list;
    a.rec = reinterpret_cast<const float4*>(g.blend_rec);
    a.bg = v.bg;
    a.mode = f.mode; a.L = f.L; a.K = f.K; a.levels = levels;
    a.coeff_dense = f.coeff_dense; a.wq = f.weights_quick; a.iq = f.indices_quick;
    a.sp_count = f.sparse_count; a.sp_index = f.sparse_index; a.sp_weight = f.sparse_weight;
    a.color = img.color; a.wmap = img.wmap; a.final_T = img.final_T; a.n_contrib = img.n_contrib;
    a.status = b.status;
    a.bimg = bimg; a.out = out; a.normalize = normalize; a.tile_counter = counter;
    { const char* e = getenv("LSV2_FD_DEBUG"); a.dbg = e ? atoi(e) : 0; }
    const int grid = min(NUM_SMS, gx * gy);
#define FD_LAUNCH(M, O)                                                                                          \
    do {                                                                                                         \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode<M, O>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                           (int)FD_SMEM_TOTAL));                                                 \
        k_render_decode<M, O><<<grid, FD_THREADS, FD_SMEM_TOTAL, st>>>(a);                                       \
    } while (0)
    if (f.mode == LSV2_MODE_QUICK) {
        if (out_dtype == 0) FD_LAUNCH(LSV2_MODE_QUICK, false); else FD_LAUNCH(LSV2_MODE_QUICK, true);
    } else {
        if (out_dtype == 0) FD_LAUNCH(LSV2_MODE_DENSE, false); else FD_LAUNCH(LSV2_MODE_DENSE, true);
    }
#undef FD_LAUNCH
    LSV2_LAUNCH_CHECK("k_render_decode");
    return LSV2_OK;
}

}  // namespace lsv2
