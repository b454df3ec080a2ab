// render_decode2.cu -- fused blend + codebook decode, second generation: the blend of tile i+1 overlaps the
// decode and the 3 MB of feature stores of tile i inside one persistent CTA.
// Replaces renderCUDA (forward) + the einsum / norm / divide every caller runs after it
// (/root/reference/gaussian_renderer/__init__.py:108-119 ; eval_lerf.py:214-218 ; SURVEY.md rows R8+R9).
//
// What changed against render_decode.cu (same arithmetic for alpha, T, n_contrib, RGB and the weight tile W):
//   * the pixel warps no longer run the decode epilogue.  They blend, and at the end of a tile round W
//     (fp32, TMEM) to fp16 A operands in a dedicated shared-memory buffer, then start the next tile at once;
//   * eight epilogue warps (one per half and TMEM lane quarter, thread = pixel) drain the decode
//     accumulators and stream the normalised features to HBM while the pixel warps blend;
//   * two MMA issuers: one for the blend accumulation, one for the decode (+ the codebook chunk ring), so neither
//     pipeline waits behind the other's barriers;
//   * the L2 norm comes from the Gram matrix:  |F|^2 = w^T (C C^T) w  =  sum_k U[k] w[k]  with  U = W . G  one tiny MMA per
//     level (G split hi + lo in fp16 so the product is exact to ~2^-21), which removes the second decode pass;
//   * codebooks stream through a 3-slot ring of 32-feature chunks (4 KB each) instead of a resident 64 KB level.
//
// Warps: 0-7 blend (thread = pixel, warp = 16x2 strip; pixels 0..127 = MMA half 0, 128..255 = half 1),
//        8-15 epilogue, 16-17 producer, 18 blend-MMA issuer, 19 decode-MMA issuer + chunk fetch.
// TMEM:  columns [0, 2L) W (half h at h*L);  [384, 512) decode: four 32-column F slots (half h, parity j at
//        384 + (2h+j)*32), aliased per level by U (half h at 384 + 64h) before that level's F chunks are issued.
#include "fd2_common.cuh"

namespace lsv2 {

constexpr int F2_THREADS = 640;
constexpr int F2_EPI_WARP0 = 8, F2_EPI_WARPS = 8, F2_PROD_WARP0 = 16, F2_BMMA_WARP = 18, F2_DMMA_WARP = 19;
constexpr uint32_t F2_TMEM_F = 384;

constexpr size_t F2_SMEM_RING = 0;
constexpr size_t F2_SMEM_ADEC = F2_SMEM_RING + (size_t)FD_STAGES * FD_STAGE_BYTES;                // 112 KB
constexpr size_t F2_SMEM_CB = F2_SMEM_ADEC + (size_t)2 * (FD_MAX_L / FD_LC) * FD_TILE16K;       // + 96 KB
constexpr size_t F2_SMEM_HDR = F2_SMEM_CB + (size_t)F2_CB_SLOTS * F2_CB_BYTES;                   // + 12 KB
constexpr size_t F2_SMEM_BAR = F2_SMEM_HDR + FD_STAGES * sizeof(FdHeader);
constexpr size_t F2_SMEM_TOTAL = F2_SMEM_BAR + 40 * sizeof(uint64_t) + 64 + 1024;                // + alignment slack
static_assert(F2_SMEM_TOTAL <= 227 * 1024, "render_decode2 shared memory budget");

// ---- codebook images -------------------------------------------------------------------------------------
// per level 640 rows of 64 fp16 (128 B, 128B-swizzled, 8-row pattern): rows [0,64) G_hi, [64,128) G_lo,
// [128,640) the codebook transposed (row n = feature, k = code).  G = s * C16 C16^T with C16 the fp16-rounded codebook
// (what the decode MMA multiplies by), s = 1024 / max_n |C16_n|^2.
__global__ void k_f2_prep_c(int levels, const float* __restrict__ codebooks, uint8_t* __restrict__ img)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= levels * FD_D * 8) return;
    const int chunk = t & 7, n = (t >> 3) % FD_D, l = t / (FD_D * 8);
    const float* src = codebooks + ((size_t)l * FD_LC + chunk * 8) * FD_D + n;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = pack_f16x2(src[(size_t)(2 * i) * FD_D], src[(size_t)(2 * i + 1) * FD_D]);
    *reinterpret_cast<uint4*>(img + (size_t)l * F2_LVL_BYTES + sw128_offset(128 + n, chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
}

// block (n, l), thread k:  G[n][k] = s * <C16_n, C16_k>;  every block recomputes s from the 64 squared norms
__global__ void __launch_bounds__(64) k_f2_prep_g(const float* __restrict__ codebooks, uint8_t* __restrict__ img,
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
    if (k < 8) {                                  // thread k packs 16-byte chunk k of row n (hi and lo images)
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float v0 = s_v[k * 8 + 2 * i], v1 = s_v[k * 8 + 2 * i + 1];
            const __half h0 = __float2half_rn(v0), h1 = __float2half_rn(v1);
            hi[i] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
            lo[i] = pack_f16x2(v0 - __half2float(h0), v1 - __half2float(h1));
        }
        uint8_t* base = img + (size_t)l * F2_LVL_BYTES;
        *reinterpret_cast<uint4*>(base + sw128_offset(n, k)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(base + sw128_offset(64 + n, k)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

template <int MODE, bool OUT_BF16>
__global__ void __launch_bounds__(F2_THREADS, 1) k_render_decode2(FdArgs a)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* ring = smem + F2_SMEM_RING;
    unsigned char* s_adec = smem + F2_SMEM_ADEC;
    unsigned char* s_cb = smem + F2_SMEM_CB;
    FdHeader* hdrs = reinterpret_cast<FdHeader*>(smem + F2_SMEM_HDR);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + F2_SMEM_BAR);
    uint64_t* full = bars;                        // [2]  producer -> blend, blend-mma   (64 arrivals)
    uint64_t* a_full = bars + 2;                  // [2]  blend -> blend-mma             (256 arrivals)
    uint64_t* stage_free = bars + 4;              // [2]  blend-mma -> producer          (commit / arrive)
    uint64_t* w_full = bars + 6;                  //      blend-mma -> blend: W complete (commit / arrive)
    uint64_t* adec_full = bars + 7;               //      blend -> decode-mma, epilogue  (256 arrivals)
    uint64_t* adec_free = bars + 8;               //      decode-mma -> blend            (commit)
    uint64_t* cb_full = bars + 9;                 // [3]  chunk landed                   (tx)
    uint64_t* cb_free = bars + 12;                // [3]  decode-mma done with the chunk (commit)
    uint64_t* f_full = bars + 15;                 // [4]  F slot ready                   (commit)
    uint64_t* f_free = bars + 19;                 // [4]  F slot drained                 (32 arrivals: one warp)
    uint64_t* u_full = bars + 23;                 //      U = W.G ready                  (commit)
    uint64_t* u_free = bars + 24;                 //      U read                         (128 arrivals)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 25);
    uint32_t* s_ndone = s_tmem + 1;               // [2]  saturated strips, by tile parity
    int32_t* s_tile = reinterpret_cast<int32_t*>(s_tmem + 3);
    F2TileInfo* tinfo = reinterpret_cast<F2TileInfo*>(bars + 28);   // [2]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = a.L, levels = a.levels;
    const size_t HW = (size_t)a.H * a.W;

    if (tid == 0) {
        for (int s = 0; s < FD_STAGES; ++s) {
            mbar_init(full + s, FD_PROD_THREADS);
            mbar_init(a_full + s, 32 * FD_BLEND_WARPS);
            mbar_init(stage_free + s, 1);
        }
        mbar_init(w_full, 1);
        mbar_init(adec_full, 32 * FD_BLEND_WARPS);
        mbar_init(adec_free, 1);
        for (int c = 0; c < F2_CB_SLOTS; ++c) { mbar_init(cb_full + c, 1); mbar_init(cb_free + c, 1); }
        for (int q = 0; q < 4; ++q) { mbar_init(f_full + q, 1); mbar_init(f_free + q, 128); }
        mbar_init(u_full, 1);
        mbar_init(u_free, 32 * F2_EPI_WARPS);
        s_ndone[0] = 0;
        s_ndone[1] = 0;
        mbar_fence_init();
    }
    if (warp == F2_BMMA_WARP) tmem_alloc(s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp >= F2_PROD_WARP0 && warp < F2_PROD_WARP0 + 2) {
        // ======================================= producer =======================================
        const int pt = tid - 32 * F2_PROD_WARP0;        // 0..63 = instance slot of the batch
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
            F2Packed cur;
            F2Raw nxt;
            f2_gather<MODE>(a, load_id(0), nxt);
            f2_pack(nxt, L, cur);
            uint32_t gid_next = load_id(1);
            for (uint32_t b = 0;; ++b, ++use_g) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                if (use > 0) mbar_wait_sleep(stage_free + s, (use - 1) & 1);
                FdHeader& Hd = hdrs[s];
                const bool stop = exit_now || b >= n_batches ||
                                  *(volatile uint32_t*)(s_ndone + (t_iter & 1)) >= FD_BLEND_WARPS;
                if (named_bar_or(1, FD_PROD_THREADS, stop)) {
                    if (pt == 0) { Hd.count = exit_now ? -2 : -1; Hd.tile = tile; Hd.batches = (int32_t)b; }
                    mbar_arrive(full + s);
                    ++use_g;
                    break;
                }
                unsigned char* S = ring + (size_t)s * FD_STAGE_BYTES + 2 * FD_TILE16K;
                for (int q = pt; q < L * 8; q += FD_PROD_THREADS)
                    reinterpret_cast<uint4*>(S)[q] = make_uint4(0, 0, 0, 0);
                f2_gather<MODE>(a, gid_next, nxt);          // in flight while this batch is assembled
                gid_next = load_id(b + 2);
                named_bar_sync(1, FD_PROD_THREADS);
                uint32_t m8 = 0;
                if (cur.g != 0xFFFFFFFFu) {
                    Hd.g0[pt] = cur.r0;
                    Hd.g1[pt] = cur.r1;
                    Hd.g2[pt] = make_float2(cur.r2.x, cur.r2.y);
                    m8 = fd_strip_mask(cur.r0.y - y0, cur.r0.z, cur.r0.w, cur.r1.x, cur.r1.y, cur.r0.x - x0);
                    const uint32_t col = (uint32_t)(pt & 7) * 2u, chunk = (uint32_t)pt >> 3;
                    if (cur.bad) a.status[LSV2_STATUS_BAD_INDEX] = 1u;
                    if (cur.cnt != FD_ROW_DENSE) {
#pragma unroll
                        for (int k = 0; k < FD_KMAX; ++k) {
                            if ((uint32_t)k < cur.cnt) {
                                const uint32_t c = cur.wc[k] >> 16;
                                if (c != 0xFFFFu) {
                                    unsigned short* p = reinterpret_cast<unsigned short*>(S + sw128_offset(c, chunk) + col);
                                    if (cur.ascending) {
                                        *p = (unsigned short)(cur.wc[k] & 0xFFFFu);
                                    } else {                 // duplicates of a channel add up
                                        const float old = __half2float(__ushort_as_half(*p));
                                        const float add = __half2float(__ushort_as_half((unsigned short)(cur.wc[k] & 0xFFFFu)));
                                        *p = __half_as_ushort(__float2half_rn(old + add));
                                    }
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
                f2_pack(nxt, L, cur);
            }
            if (exit_now) break;
        }
    } else if (warp == F2_BMMA_WARP) {
        // ================================ blend MMA issue ================================
        if (lane == 0) {
            const uint32_t idesc_blend = umma_idesc_f16_f32(128, L);
            const uint32_t ring_addr = smem_u32(ring);
            uint32_t batch_in_tile = 0;
            uint32_t a_uses[FD_STAGES] = {0, 0};             // a_full phases consumed (batches only)
            for (uint32_t use_g = 0;; ++use_g) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                mbar_wait_sleep(full + s, use & 1);
                const int count = hdrs[s].count;
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
                // ---- end of tile: W is final once every MMA issued so far has completed ----
                if (batch_in_tile > 0) umma_commit(w_full); else mbar_arrive(w_full);
                batch_in_tile = 0;
                mbar_arrive(stage_free + s);
            }
        }
    } else if (warp == F2_DMMA_WARP) {
        // ========================= decode MMA issue + codebook chunk ring =========================
        if (lane == 0) {
            constexpr uint32_t idesc_dec = umma_idesc_f16_f32(128, F2_NCH);
            const uint32_t adec_addr = smem_u32(s_adec), cb_addr = smem_u32(s_cb);
            const int first_chunk = a.normalize ? 0 : F2_G_CHUNKS;
            uint32_t ld_n = 0, use_n = 0;                    // chunk loads issued / chunks consumed
            int ld_level = 0, ld_chunk = first_chunk;
            auto issue_load = [&]() {                        // the chunk sequence is the same for every tile
                const uint32_t slot = ld_n % F2_CB_SLOTS;
                if (ld_n >= F2_CB_SLOTS) mbar_wait(cb_free + slot, ((ld_n / F2_CB_SLOTS) - 1) & 1);
                mbar_arrive_expect_tx(cb_full + slot, F2_CB_BYTES);
                bulk_g2s(s_cb + (size_t)slot * F2_CB_BYTES,
                         a.bimg + (size_t)ld_level * F2_LVL_BYTES + (size_t)ld_chunk * F2_CB_BYTES, F2_CB_BYTES,
                         cb_full + slot);
                ++ld_n;
                if (++ld_chunk == F2_LVL_CHUNKS) { ld_chunk = first_chunk; ld_level = (ld_level + 1) % levels; }
            };
            auto take_chunk = [&]() -> uint32_t {            // -> shared address of the next chunk, landed
                const uint32_t slot = use_n % F2_CB_SLOTS;
                mbar_wait(cb_full + slot, (use_n / F2_CB_SLOTS) & 1);
                tc_fence_after();
                return cb_addr + slot * F2_CB_BYTES;
            };
            auto release_chunk = [&]() {
                umma_commit(cb_free + (use_n % F2_CB_SLOTS));
                ++use_n;
                issue_load();                                // refills the slot released one chunk ago
            };
            issue_load();
            issue_load();
            uint32_t R = 0;                                  // fills of every F slot so far
            uint32_t u_n = 0;
            for (uint32_t t = 0;; ++t) {
                mbar_wait_sleep(adec_full, t & 1);
                if (tinfo[t & 1].tile < 0) break;
                tc_fence_after();
                for (int l = 0; l < levels; ++l) {
                    if (a.normalize) {
                        if (R > 0)
                            for (int q = 0; q < 4; ++q) mbar_wait(f_free + q, (R - 1) & 1);
                        tc_fence_after();
                        for (int gc = 0; gc < F2_G_CHUNKS; ++gc) {
                            const uint32_t cb = take_chunk();
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const uint32_t aa = adec_addr + (uint32_t)(h * levels + l) * FD_TILE16K;
#pragma unroll
                                for (int ks = 0; ks < 4; ++ks)
                                    umma_bf16(tmem + F2_TMEM_F + (uint32_t)(h * 64 + (gc & 1) * 32),
                                              umma_desc_k_sw128(aa + ks * 32), umma_desc_k_sw128(cb + ks * 32), idesc_dec,
                                              (gc >= 2 || ks > 0) ? 1u : 0u);
                            }
                            release_chunk();
                        }
                        umma_commit(u_full);
                        mbar_wait(u_free, u_n & 1);
                        ++u_n;
                        tc_fence_after();
                    }
                    for (int c = 0; c < F2_C_CHUNKS; ++c) {
                        const uint32_t cb = take_chunk();
                        const uint32_t fills = R + (uint32_t)(c >> 1);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const int q = h * 2 + (c & 1);
                            if (fills > 0) mbar_wait(f_free + q, (fills - 1) & 1);
                            tc_fence_after();
                            const uint32_t aa = adec_addr + (uint32_t)(h * levels + l) * FD_TILE16K;
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(tmem + F2_TMEM_F + (uint32_t)(q * F2_NCH), umma_desc_k_sw128(aa + ks * 32),
                                          umma_desc_k_sw128(cb + ks * 32), idesc_dec, ks > 0 ? 1u : 0u);
                            umma_commit(f_full + q);
                        }
                        release_chunk();
                    }
                    R += F2_C_CHUNKS / 2;
                }
                umma_commit(adec_free);
            }
            while (use_n < ld_n) {                           // drain the prefetch no tile consumed
                mbar_wait(cb_full + (use_n % F2_CB_SLOTS), (use_n / F2_CB_SLOTS) & 1);
                ++use_n;
            }
        }
    } else if (warp >= F2_EPI_WARP0 && warp < F2_EPI_WARP0 + F2_EPI_WARPS) {
        // ==================================== decode epilogue ====================================
        // eight warps: warp -> (half h, TMEM lane quarter q4); thread = one pixel = one TMEM lane of half h
        const int q4 = warp & 3, h = (warp - F2_EPI_WARP0) >> 2;
        const int row = q4 * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(q4 * 32) << 16) + F2_TMEM_F;
        const int lx = lane & 15;
        uint32_t R = 0, u_n = 0;
        for (uint32_t t = 0;; ++t) {
            mbar_wait(adec_full, t & 1);
            const int tile = tinfo[t & 1].tile;
            if (tile < 0) break;
            const int px = (tile % a.gx) * TILE + lx;
            const int py = (tile / a.gx) * TILE + (h * 4 + q4) * 2 + (lane >> 4);
            const bool inside = px < a.W && py < a.H;
            const size_t pix = (size_t)py * a.W + px;
            for (int l = 0; l < levels; ++l) {
                float inv = 1.0f;
                if (a.normalize) {
                    mbar_wait(u_full, u_n & 1);
                    ++u_n;
                    tc_fence_after();
                    const unsigned char* Ad = s_adec + (size_t)(h * levels + l) * FD_TILE16K;
                    float ss = 0.0f;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        uint32_t r[32];
                        tmem_ld32(lane_base + (uint32_t)(h * 64 + hh * 32), r);
                        tmem_wait_ld();
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) {
                            const uint4 w8 = *reinterpret_cast<const uint4*>(Ad + sw128_offset((uint32_t)row, (uint32_t)(hh * 4 + ch)));
                            const uint32_t ww[4] = {w8.x, w8.y, w8.z, w8.w};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float2 wf = __half22float2(*reinterpret_cast<const __half2*>(&ww[i]));
                                ss = fmaf(__uint_as_float(r[ch * 8 + 2 * i]), wf.x, ss);
                                ss = fmaf(__uint_as_float(r[ch * 8 + 2 * i + 1]), wf.y, ss);
                            }
                        }
                    }
                    inv = 1.0f / (sqrtf(fmaxf(ss, 0.0f) * a.gscale_inv[l]) + 1e-10f);
                    tc_fence_before();
                    mbar_arrive_relaxed(u_free);
                }
                // F chunks: even chunks through ra, odd through rb; the next TMEM read is in flight during the stores
                uint32_t ra[32], rb[32];
                const size_t ch0 = (size_t)l * FD_D;
                auto drain_wait = [&](int c) {
                    mbar_wait(f_full + (h * 2 + (c & 1)), (R + (uint32_t)(c >> 1)) & 1);
                    tc_fence_after();
                };
                auto release = [&](int c) {
                    tmem_wait_ld();
                    tc_fence_before();
                    mbar_arrive_relaxed(f_free + (h * 2 + (c & 1)));      // relaxed: must not wait for the feature stores
                };
                // scale in place, then 32 stores through four independent running addresses
                auto store = [&](uint32_t (&r)[32], int c) {
                    if (!inside || (a.dbg & 1)) return;
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * inv);
                    const uint64_t esz = OUT_BF16 ? 2u : 4u;
                    // dbg & 4 (timing experiment): every channel lands on plane 0 -> same instructions, L2-resident footprint
                    const uint64_t stride = (a.dbg & 4) ? 0 : (uint64_t)HW * esz;
                    uint64_t p[4];
                    p[0] = (uint64_t)(uintptr_t)a.out + (((a.dbg & 4) ? 0 : (ch0 + (size_t)c * F2_NCH) * HW) +
                                                            ((a.dbg & 8) ? (pix & 4095) : pix)) * esz;   // dbg & 8: same pages, 25 MB footprint
                    p[1] = p[0] + stride; p[2] = p[1] + stride; p[3] = p[2] + stride;
                    const uint64_t s4 = 4 * stride;
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (OUT_BF16) {
                                const unsigned short v = __bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(r[i + u])));
                                asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p[u]), "h"(v) : "memory");
                            } else {
                                asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p[u]), "r"(r[i + u]) : "memory");
                            }
                            p[u] += s4;
                        }
                    }
                };
                const uint32_t slot0 = lane_base + (uint32_t)(h * 2 * F2_NCH);
                drain_wait(0);
                tmem_ld32(slot0, ra);
#pragma unroll 1
                for (int c = 0; c < F2_C_CHUNKS; c += 2) {
                    release(c);                                      // ra holds chunk c
                    drain_wait(c + 1);
                    tmem_ld32(slot0 + F2_NCH, rb);
                    store(ra, c);
                    release(c + 1);                                  // rb holds chunk c + 1
                    if (c + 2 < F2_C_CHUNKS) {
                        drain_wait(c + 2);
                        tmem_ld32(slot0, ra);
                    }
                    store(rb, c + 1);
                }
                R += F2_C_CHUNKS / 2;
            }
        }
    } else {
        // ========================================= blend =========================================
        const int half = warp >> 2;                                  // MMA half this strip belongs to
        const int row = (warp & 3) * 32 + lane;                      // A-tile row == TMEM lane
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int lx = lane & 15, ly = warp * 2 + (lane >> 4);
        uint32_t tiles_done = 0;
        uint32_t use_g = 0;
        for (;;) {
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
                                for (int u = 0; u < 4; ++u) {
                                    const float e = repro_expf_inrange(fmaxf(fminf(pw_[u], 0.0f), -80.0f));
                                    al[u] = hit[u] ? fminf(ALPHA_MAX, op_[u] * e) : 0.0f;
                                }
#pragma unroll
                                for (int u = 0; u < 4; ++u) {
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
            if (count == -2) {                             // no more tiles: tell the decode side and leave
                // never run two adec_full phases ahead of the decode side (a parity wait cannot tell them apart)
                if (tiles_done > 0) mbar_wait(adec_free, (tiles_done - 1) & 1);
                if (tid == 0) tinfo[tiles_done & 1].tile = -1;
                mbar_arrive(adec_full);
                break;
            }
            if (first) {                                   // a tile with an empty list
                px = (tile % a.gx) * TILE + lx;
                py = (tile / a.gx) * TILE + ly;
                inside = px < a.W && py < a.H;
                if (warp == 0 && lane == 0) s_ndone[(tiles_done + 1) & 1] = 0;
            }
            const size_t pix = (size_t)py * a.W + px;

            // ---------------- W: TMEM -> fp16 A operands of the decode ----------------
            mbar_wait(w_full, tiles_done & 1);
            if (tiles_done > 0) mbar_wait(adec_free, (tiles_done - 1) & 1);      // previous tile's decode MMAs done
            tc_fence_after();
            for (int l = 0; l < levels; ++l) {
                unsigned char* Ad = s_adec + (size_t)(half * levels + l) * FD_TILE16K;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t r[32];
                    if (nb > 0) {
                        tmem_ld32(lane_base + (uint32_t)(half * L + l * FD_LC + hh * 32), r);
                        tmem_wait_ld();
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = 0u;
                    }
                    if (a.wmap != nullptr && inside) {
                        float* o = a.wmap + (size_t)(l * FD_LC + hh * 32) * HW + pix;
#pragma unroll
                        for (int i = 0; i < 32; ++i, o += HW) st_stream_f32(o, __uint_as_float(r[i]));
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
            if (tid == 0) { tinfo[tiles_done & 1].tile = tile; tinfo[tiles_done & 1].nb = nb; }
            fence_async_smem();
            tc_fence_before();
            mbar_arrive(adec_full);

            if (inside) {
                a.final_T[pix] = T;
                a.n_contrib[pix] = last;
                a.color[pix] = fmaf(T, a.bg[0], C0);
                a.color[HW + pix] = fmaf(T, a.bg[1], C1);
                a.color[2 * HW + pix] = fmaf(T, a.bg[2], C2);
            }
            ++tiles_done;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == F2_BMMA_WARP) tmem_dealloc(tmem, 512);
}

int f2_prepare_images(int levels, const float* codebooks, int normalize, uint8_t* cimg, float* gscale_inv,
                      cudaStream_t st)
{
    k_f2_prep_c<<<ceil_div(levels * FD_D * 8, 256), 256, 0, st>>>(levels, codebooks, cimg);
    LSV2_LAUNCH_CHECK("k_f2_prep_c");
    if (normalize) {
        k_f2_prep_g<<<dim3(FD_LC, levels), 64, 0, st>>>(codebooks, cimg, gscale_inv);
        LSV2_LAUNCH_CHECK("k_f2_prep_g");
    }
    return LSV2_OK;
}

size_t render_decode2_scratch_bytes(int levels) { return f2_scratch_bytes(levels); }

int launch_render_decode2(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
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
    if (!scratch || scratch_bytes < render_decode2_scratch_bytes(levels)) return LSV2_ENOSPC;
    uint8_t* base = reinterpret_cast<uint8_t*>(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    uint32_t* counter = reinterpret_cast<uint32_t*>(base);
    float* gscale_inv = reinterpret_cast<float*>(base + 128);
    uint8_t* cimg = base + 256;
    LSV2_CUDA_TRY(cudaMemsetAsync(counter, 0, sizeof(uint32_t), st));
    { const int rc = f2_prepare_images(levels, codebooks, normalize, cimg, gscale_inv, st); if (rc) return rc; }

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
    a.bimg = cimg; a.gimg = nullptr; a.gscale_inv = gscale_inv;
    a.out = out; a.normalize = normalize; a.tile_counter = counter;
    { const char* e = getenv("LSV2_FD_DEBUG"); a.dbg = e ? atoi(e) : 0; }
    int grid = min(NUM_SMS, gx * gy);
    { const char* e = getenv("LSV2_FD_GRID"); if (e && atoi(e) > 0) grid = min(grid, atoi(e)); }   // debugging aid
#define F2_LAUNCH(M, O)                                                                                          \
    do {                                                                                                         \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode2<M, O>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                           (int)F2_SMEM_TOTAL));                                                 \
        k_render_decode2<M, O><<<grid, F2_THREADS, F2_SMEM_TOTAL, st>>>(a);                                      \
    } while (0)
    if (f.mode == LSV2_MODE_QUICK) {
        if (out_dtype == 0) F2_LAUNCH(LSV2_MODE_QUICK, false); else F2_LAUNCH(LSV2_MODE_QUICK, true);
    } else {
        if (out_dtype == 0) F2_LAUNCH(LSV2_MODE_DENSE, false); else F2_LAUNCH(LSV2_MODE_DENSE, true);
    }
#undef F2_LAUNCH
    LSV2_LAUNCH_CHECK("k_render_decode2");
    return LSV2_OK;
}

}  // namespace lsv2
