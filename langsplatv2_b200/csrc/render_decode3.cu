// render_decode3.cu -- fused blend + codebook decode, third generation: HALF a tile (16 x 8 pixels) per CTA,
// TWO CTAs per SM.  Each CTA alternates blend and decode phases like render_decode.cu, but needs only 256 TMEM
// columns and ~99 KB of shared memory, so the hardware co-schedules a blending CTA with a storing CTA on every SM:
// the latency-bound blend of one fills the issue slots the other's HBM-bound feature stores leave idle, with no
// cross-role hand-off inside a CTA to stall on.  Decode side as render_decode2.cu (32-feature codebook chunks
// through a 3-slot ring, L2 norm from the Gram matrix).
// Replaces renderCUDA (forward) + the einsum / norm / divide every caller runs after it
// (/root/reference/gaussian_renderer/__init__.py:108-119 ; eval_lerf.py:214-218 ; SURVEY.md rows R8+R9).
//
// Warps: 0-3 blend + epilogue (thread = pixel = TMEM lane; warp = 16x2 strip of the half tile),
//        4-5 producer, 6 MMA issuer (blend and decode MMAs, codebook chunk ring).
// TMEM (256 columns): [0, L) W ; [192, 256) two 32-column F slots, aliased per level by U = W.G (64 columns).
// Shared memory: two stages of { A [128 px][64 inst] fp16 | S [L][64 inst] fp16 } = 80 KB, reused between the
// phases as the decode's A operands [levels][128 px][64] fp16; 12 KB chunk ring; headers; barriers.
#include "fd2_common.cuh"

namespace lsv2 {

constexpr int F3_BLEND_WARPS = 4;
constexpr int F3_THREADS = 32 * F3_BLEND_WARPS + FD_PROD_THREADS + 32;      // 224
constexpr int F3_PROD_WARP0 = 4, F3_MMA_WARP = 6;
constexpr int F3_STAGE_BYTES = FD_TILE16K + FD_MAX_L * 128;                 // A | S = 40 KB
constexpr uint32_t F3_TMEM_COLS = 256;
constexpr uint32_t F3_TMEM_F = 192;

constexpr size_t F3_SMEM_RING = 0;
constexpr size_t F3_SMEM_CB = F3_SMEM_RING + (size_t)FD_STAGES * F3_STAGE_BYTES;                 // 80 KB
constexpr size_t F3_SMEM_HDR = F3_SMEM_CB + (size_t)F2_CB_SLOTS * F2_CB_BYTES;                   // + 12 KB
constexpr size_t F3_SMEM_BAR = F3_SMEM_HDR + FD_STAGES * sizeof(FdHeader);
constexpr size_t F3_SMEM_TOTAL = F3_SMEM_BAR + 32 * sizeof(uint64_t) + 64 + 1024;                // + alignment slack
static_assert(2 * (F3_SMEM_TOTAL + 1024) <= 227 * 1024, "two CTAs of render_decode3 must fit one SM");
static_assert((FD_MAX_L / FD_LC) * FD_TILE16K <= FD_STAGES * F3_STAGE_BYTES, "decode operands alias the blend ring");

template <int MODE, bool OUT_BF16>
__global__ void __launch_bounds__(F3_THREADS, 2) k_render_decode3(FdArgs a)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* ring = smem + F3_SMEM_RING;
    unsigned char* s_cb = smem + F3_SMEM_CB;
    FdHeader* hdrs = reinterpret_cast<FdHeader*>(smem + F3_SMEM_HDR);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + F3_SMEM_BAR);
    uint64_t* full = bars;                        // [2]  producer -> blend, mma       (64 arrivals)
    uint64_t* a_full = bars + 2;                  // [2]  blend -> mma                 (128 arrivals)
    uint64_t* stage_free = bars + 4;              // [2]  mma -> producer              (commit / arrive)
    uint64_t* w_full = bars + 6;                  //      mma -> blend: W complete     (commit / arrive)
    uint64_t* adec_full = bars + 7;               //      blend -> mma                 (128 arrivals)
    uint64_t* ring_free = bars + 8;               //      mma -> producer              (commit)
    uint64_t* cb_full = bars + 9;                 // [3]  chunk landed                 (tx)
    uint64_t* cb_free = bars + 12;                // [3]  mma done with the chunk      (commit)
    uint64_t* f_full = bars + 15;                 // [2]  F slot ready                 (commit)
    uint64_t* f_free = bars + 17;                 // [2]  F slot drained               (128 arrivals)
    uint64_t* u_full = bars + 19;                 //      U = W.G ready                (commit)
    uint64_t* u_free = bars + 20;                 //      U read                       (128 arrivals)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 21);
    uint32_t* s_ndone = s_tmem + 1;               // [2]  saturated strips, by unit parity
    int32_t* s_unit = reinterpret_cast<int32_t*>(s_tmem + 3);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int L = a.L, levels = a.levels;
    const size_t HW = (size_t)a.H * a.W;
    const int n_units = 2 * a.n_tiles;            // unit u: tile u >> 1, rows 8 * (u & 1) .. + 7 of it

    if (tid == 0) {
        for (int s = 0; s < FD_STAGES; ++s) {
            mbar_init(full + s, FD_PROD_THREADS);
            mbar_init(a_full + s, 32 * F3_BLEND_WARPS);
            mbar_init(stage_free + s, 1);
            mbar_init(f_full + s, 1);
            mbar_init(f_free + s, 32 * F3_BLEND_WARPS);
        }
        mbar_init(w_full, 1);
        mbar_init(adec_full, 32 * F3_BLEND_WARPS);
        mbar_init(ring_free, 1);
        for (int c = 0; c < F2_CB_SLOTS; ++c) { mbar_init(cb_full + c, 1); mbar_init(cb_free + c, 1); }
        mbar_init(u_full, 1);
        mbar_init(u_free, 32 * F3_BLEND_WARPS);
        s_ndone[0] = 0;
        s_ndone[1] = 0;
        mbar_fence_init();
    }
    if (warp == F3_MMA_WARP) tmem_alloc(s_tmem, F3_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if (warp >= F3_PROD_WARP0 && warp < F3_PROD_WARP0 + 2) {
        // ======================================= producer =======================================
        const int pt = tid - 32 * F3_PROD_WARP0;        // 0..63 = instance slot of the batch
        const int pw = pt >> 5;
        uint32_t use_g = 0;                             // stage uses so far (global)
        for (uint32_t t_iter = 0;; ++t_iter) {
            if (pt == 0) *s_unit = (int32_t)atomicAdd(a.tile_counter, 1u);
            named_bar_sync(1, FD_PROD_THREADS);
            const int unit = *s_unit;
            named_bar_sync(1, FD_PROD_THREADS);
            const bool exit_now = unit >= n_units;
            const int tile = unit >> 1, hh = unit & 1;
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
            // the previous unit's decode reads its A operands from the ring: wait until those MMAs are done
            if (t_iter > 0 && !exit_now) mbar_wait_sleep(ring_free, (t_iter - 1) & 1);
            for (uint32_t b = 0;; ++b, ++use_g) {
                const int s = use_g & 1;
                const uint32_t use = use_g >> 1;
                if (use > 0) mbar_wait_sleep(stage_free + s, (use - 1) & 1);
                FdHeader& Hd = hdrs[s];
                const bool stop = exit_now || b >= n_batches ||
                                  *(volatile uint32_t*)(s_ndone + (t_iter & 1)) >= F3_BLEND_WARPS;
                if (named_bar_or(1, FD_PROD_THREADS, stop)) {
                    if (pt == 0) { Hd.count = exit_now ? -2 : -1; Hd.tile = unit; Hd.batches = (int32_t)b; }
                    mbar_arrive(full + s);
                    ++use_g;
                    break;
                }
                unsigned char* S = ring + (size_t)s * F3_STAGE_BYTES + FD_TILE16K;
                for (int q = pt; q < L * 8; q += FD_PROD_THREADS)
                    reinterpret_cast<uint4*>(S)[q] = make_uint4(0, 0, 0, 0);
                f2_gather<MODE>(a, gid_next, nxt);          // in flight while this batch is assembled
                gid_next = load_id(b + 2);
                named_bar_sync(1, FD_PROD_THREADS);
                uint32_t m4 = 0;
                if (cur.g != 0xFFFFFFFFu) {
                    Hd.g0[pt] = cur.r0;
                    Hd.g1[pt] = cur.r1;
                    Hd.g2[pt] = make_float2(cur.r2.x, cur.r2.y);
                    m4 = (fd_strip_mask(cur.r0.y - y0, cur.r0.z, cur.r0.w, cur.r1.x, cur.r1.y, cur.r0.x - x0) >> (4 * hh)) & 0xFu;
                    const uint32_t col = (uint32_t)(pt & 7) * 2u, chunk = (uint32_t)pt >> 3;
                    if (cur.bad) a.status[LSV2_STATUS_BAD_INDEX] = 1u;
                    if (m4 != 0u) {                          // an instance no strip of this half can see stays a zero column
                        if (cur.cnt != FD_ROW_DENSE) {
#pragma unroll
                            for (int k = 0; k < FD_KMAX; ++k) {
                                if ((uint32_t)k < cur.cnt) {
                                    const uint32_t c = cur.wc[k] >> 16;
                                    if (c != 0xFFFFu) {
                                        unsigned short* p = reinterpret_cast<unsigned short*>(S + sw128_offset(c, chunk) + col);
                                        if (cur.ascending) {
                                            *p = (unsigned short)(cur.wc[k] & 0xFFFFu);
                                        } else {             // duplicates of a channel add up
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
                }
#pragma unroll
                for (int w = 0; w < F3_BLEND_WARPS; ++w) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, (m4 >> w) & 1u);
                    if (lane == 0) Hd.mask[w][pw] = bal;
                }
                if (pt == 0) {
                    Hd.count = (int32_t)min((uint32_t)FD_BATCH, n_inst - b * FD_BATCH);
                    Hd.tile = unit;
                }
                fence_async_smem();
                mbar_arrive(full + s);
                f2_pack(nxt, L, cur);
            }
            if (exit_now) break;
        }
    } else if (warp == F3_MMA_WARP) {
        // ===================== MMA issue (blend, then decode) + codebook chunk ring =====================
        if (lane == 0) {
            const uint32_t idesc_blend = umma_idesc_f16_f32(128, L);
            constexpr uint32_t idesc_dec = umma_idesc_f16_f32(128, F2_NCH);
            const uint32_t ring_addr = smem_u32(ring), cb_addr = smem_u32(s_cb);
            const int first_chunk = a.normalize ? 0 : F2_G_CHUNKS;
            uint32_t ld_n = 0, use_n = 0;                    // chunk loads issued / chunks consumed
            int ld_level = 0, ld_chunk = first_chunk;
            auto issue_load = [&]() {                        // the chunk sequence is the same for every unit
                const uint32_t slot = ld_n % F2_CB_SLOTS;
                if (ld_n >= F2_CB_SLOTS) mbar_wait(cb_free + slot, ((ld_n / F2_CB_SLOTS) - 1) & 1);
                mbar_arrive_expect_tx(cb_full + slot, F2_CB_BYTES);
                bulk_g2s(s_cb + (size_t)slot * F2_CB_BYTES,
                         a.bimg + (size_t)ld_level * F2_LVL_BYTES + (size_t)ld_chunk * F2_CB_BYTES, F2_CB_BYTES,
                         cb_full + slot);
                ++ld_n;
                if (++ld_chunk == F2_LVL_CHUNKS) { ld_chunk = first_chunk; ld_level = (ld_level + 1) % levels; }
            };
            auto take_chunk = [&]() -> uint32_t {
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
            uint32_t R = 0, u_n = 0, units_done = 0, batch_in_unit = 0;
            uint32_t a_uses[FD_STAGES] = {0, 0};
            for (uint32_t use_g = 0;; ++use_g) {
                const int s = use_g & 1;
                mbar_wait_sleep(full + s, (use_g >> 1) & 1);
                const int count = hdrs[s].count;
                if (count == -2) break;
                if (count >= 0) {
                    mbar_wait(a_full + s, a_uses[s] & 1);
                    ++a_uses[s];
                    tc_fence_after();
                    const uint32_t sa = ring_addr + (uint32_t)s * F3_STAGE_BYTES;
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_bf16(tmem, umma_desc_k_sw128(sa + ks * 32), umma_desc_k_sw128(sa + FD_TILE16K + ks * 32),
                                  idesc_blend, (batch_in_unit > 0 || ks > 0) ? 1u : 0u);
                    umma_commit(stage_free + s);
                    ++batch_in_unit;
                    continue;
                }
                // ---- end of unit: W is final once every MMA issued so far has completed ----
                if (batch_in_unit > 0) umma_commit(w_full); else mbar_arrive(w_full);
                batch_in_unit = 0;
                mbar_arrive(stage_free + s);
                mbar_wait_sleep(adec_full, units_done & 1);
                tc_fence_after();
                for (int l = 0; l < levels; ++l) {
                    const uint32_t aa = ring_addr + (uint32_t)l * FD_TILE16K;
                    if (a.normalize) {
                        if (R > 0) { mbar_wait(f_free + 0, (R - 1) & 1); mbar_wait(f_free + 1, (R - 1) & 1); }
                        tc_fence_after();
                        for (int gc = 0; gc < F2_G_CHUNKS; ++gc) {
                            const uint32_t cb = take_chunk();
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(tmem + F3_TMEM_F + (uint32_t)((gc & 1) * 32), umma_desc_k_sw128(aa + ks * 32),
                                          umma_desc_k_sw128(cb + ks * 32), idesc_dec, (gc >= 2 || ks > 0) ? 1u : 0u);
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
                        const int q = c & 1;
                        if (fills > 0) mbar_wait(f_free + q, (fills - 1) & 1);
                        tc_fence_after();
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_bf16(tmem + F3_TMEM_F + (uint32_t)(q * F2_NCH), umma_desc_k_sw128(aa + ks * 32),
                                      umma_desc_k_sw128(cb + ks * 32), idesc_dec, ks > 0 ? 1u : 0u);
                        umma_commit(f_full + q);
                        release_chunk();
                    }
                    R += F2_C_CHUNKS / 2;
                }
                umma_commit(ring_free);
                ++units_done;
            }
            while (use_n < ld_n) {                           // drain the prefetch no unit consumed
                mbar_wait(cb_full + (use_n % F2_CB_SLOTS), (use_n / F2_CB_SLOTS) & 1);
                ++use_n;
            }
        }
    } else {
        // ==================================== blend + epilogue ====================================
        const int row = warp * 32 + lane;                            // A-tile row == TMEM lane
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        const int lx = lane & 15;
        uint32_t units_done = 0, use_g = 0, R = 0, u_n = 0;
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
                const int tile = unit >> 1;
                px = (tile % a.gx) * TILE + lx;
                py = (tile / a.gx) * TILE + (unit & 1) * 8 + warp * 2 + (lane >> 4);
                pxf = (float)px; pyf = (float)py;
                inside = px < a.W && py < a.H;
            };
            for (;; ++use_g) {
                const int s = use_g & 1;
                mbar_wait(full + s, (use_g >> 1) & 1);
                const FdHeader& Hd = hdrs[s];
                count = Hd.count;
                if (count < 0) { unit = Hd.tile; nb = Hd.batches; ++use_g; break; }
                if (first) {
                    first = false;
                    unit = Hd.tile;
                    locate();
                    done = !inside;
                    warp_done = __all_sync(0xffffffffu, done);
                    if (warp_done && lane == 0) atomicAdd(s_ndone + (units_done & 1), 1u);
                    if (warp == 0 && lane == 0) s_ndone[(units_done + 1) & 1] = 0;
                }
                unsigned char* A = ring + (size_t)s * F3_STAGE_BYTES;
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
                    if (lane == 0) atomicAdd(s_ndone + (units_done & 1), 1u);
                }
            }
            if (count == -2) break;
            if (first) {                                   // a unit with an empty list
                locate();
                if (warp == 0 && lane == 0) s_ndone[(units_done + 1) & 1] = 0;
            }
            const size_t pix = (size_t)py * a.W + px;

            // ---------------- W: TMEM -> fp16 A operands of the decode (over the idle blend ring) ----------------
            mbar_wait(w_full, units_done & 1);
            tc_fence_after();
            for (int l = 0; l < levels; ++l) {
                unsigned char* Ad = ring + (size_t)l * FD_TILE16K;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t r[32];
                    if (nb > 0) {
                        tmem_ld32(lane_base + (uint32_t)(l * FD_LC + hh * 32), r);
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

            // ---------------- decode epilogue ----------------
            for (int l = 0; l < levels; ++l) {
                float inv = 1.0f;
                if (a.normalize) {
                    mbar_wait(u_full, u_n & 1);
                    ++u_n;
                    tc_fence_after();
                    const unsigned char* Ad = ring + (size_t)l * FD_TILE16K;
                    float ss = 0.0f;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        uint32_t r[32];
                        tmem_ld32(lane_base + F3_TMEM_F + (uint32_t)(hh * 32), r);
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
                uint32_t ra[32], rb[32];
                const size_t ch0 = (size_t)l * FD_D;
                auto drain_wait = [&](int c) {
                    mbar_wait(f_full + (c & 1), (R + (uint32_t)(c >> 1)) & 1);
                    tc_fence_after();
                };
                auto release = [&](int c) {
                    tmem_wait_ld();
                    tc_fence_before();
                    mbar_arrive_relaxed(f_free + (c & 1));
                };
                auto store = [&](uint32_t (&r)[32], int c) {
                    if (!inside || (a.dbg & 1)) return;
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * inv);
                    const uint64_t esz = OUT_BF16 ? 2u : 4u;
                    const uint64_t stride = (uint64_t)HW * esz;
                    uint64_t p[4];
                    p[0] = (uint64_t)(uintptr_t)a.out + ((ch0 + (size_t)c * F2_NCH) * HW + pix) * esz;
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
                const uint32_t slot0 = lane_base + F3_TMEM_F;
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
            ++units_done;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == F3_MMA_WARP) tmem_dealloc(tmem, F3_TMEM_COLS);
}

size_t render_decode3_scratch_bytes(int levels) { return f2_scratch_bytes(levels); }

int launch_render_decode3(const lsv2_view& v, const lsv2_geom& g, const lsv2_binning& b, const lsv2_features& f,
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
    if (!scratch || scratch_bytes < render_decode3_scratch_bytes(levels)) return LSV2_ENOSPC;
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
    int grid = min(2 * NUM_SMS, 2 * gx * gy);
    { const char* e = getenv("LSV2_FD_GRID"); if (e && atoi(e) > 0) grid = min(grid, atoi(e)); }   // debugging aid
#define F3_LAUNCH(M, O)                                                                                          \
    do {                                                                                                         \
        LSV2_CUDA_TRY(cudaFuncSetAttribute(k_render_decode3<M, O>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                           (int)F3_SMEM_TOTAL));                                                 \
        k_render_decode3<M, O><<<grid, F3_THREADS, F3_SMEM_TOTAL, st>>>(a);                                      \
    } while (0)
    if (f.mode == LSV2_MODE_QUICK) {
        if (out_dtype == 0) F3_LAUNCH(LSV2_MODE_QUICK, false); else F3_LAUNCH(LSV2_MODE_QUICK, true);
    } else {
        if (out_dtype == 0) F3_LAUNCH(LSV2_MODE_DENSE, false); else F3_LAUNCH(LSV2_MODE_DENSE, true);
    }
#undef F3_LAUNCH
    LSV2_LAUNCH_CHECK("k_render_decode3");
    return LSV2_OK;
}

}  // namespace lsv2
