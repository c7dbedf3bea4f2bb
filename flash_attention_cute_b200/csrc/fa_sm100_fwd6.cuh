// FlashAttention-2 forward for sm_100a, generation 6: 128-key S blocks, P handed over in parts.
//
// Replaces the reference's flash_attention_v2 (csrc/flash_attention_template.cuh:138-564),
// Mask (csrc/mask.cuh:30-88) and launch_flash_attention (template.cuh:567-583).  Same math
// (fp32 S and O accumulators, exp2-domain online softmax, P rounded to T before P·V, row sums
// from unrounded P, bottom-right causal alignment, kv-head = q-head / G).
//
// Generation 5 (64-key half blocks, two S buffers per M-tile) keeps the softmax warps busy but
// pays for it in shared-memory bandwidth: Q·Kᵀ with N = 64 re-reads the Q operand twice as often
// and the kernel ends up bound by the 128 B/clk of shared memory (~160 KB per half step).  This
// generation goes back to one 128-column S buffer per M-tile (full-rate N = 128 MMAs) and
// shortens the serial chain  S(j) -> softmax -> P(j) -> P·V(j) -> S(j+1)  instead:
//   * P(j) is published in kParts equal parts; P·V of a part is issued while the softmax warps
//     still exponentiate the next one, so only the last part and S(j+1) follow the softmax;
//   * one MMA issuer thread per M-tile, blocking (hardware-suspended) mbarrier waits — a polling
//     issuer steals issue slots from the softmax warps of its SM sub-partition (measured);
//   * the row max is reduced in the shadow of the first 32 exp2 (speculative stale-max path);
//   * the rare lazy rescale of O_t happens in the softmax warps themselves (O_t is quiescent
//     whenever S_t(j) is complete and no part of P_t(j) has been published).
//
//   grid = min(#SMs, #work tiles) persistent CTAs of 512 threads; work tile = 2 M-tiles of 128
//   query rows sharing one K/V stream; LPT + L2-section order, dynamic scheduler.
//     warps 0-3   softmax warpgroup, M-tile 0 (one thread per query row, no shuffles)
//     warps 4-7   softmax warpgroup, M-tile 1
//     warps 8-11  epilogue warpgroup: O_t / l -> T -> global
//     warps 12,13 MMA issuers (one elected thread each), one per M-tile
//     warp  14    TMA producer (one elected thread)
//     warp  15    spare
//   TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [256+D,256+2D);
//     P_t (16-bit, two per column) aliases the first 64 columns of S_t.
#pragma once
#include "fa_sm100_fwd4.cuh"

namespace fa100 {
namespace v6 {

using v4::Tile;
using v4::decode_tile;
using v4::TileConsumer;
using v4::Tracer;
using v4::kSchedDepth;
using v4::kPanelBytes;
using v4::kBlockN;
using v4::kRescaleThreshold;

struct Cfg {
    static constexpr int kSoftmaxWarps = 8;
    static constexpr int kEpiWarp0 = 8;
    static constexpr int kMmaWarp = 12;  // 12: M-tile 0, 13: M-tile 1
    static constexpr int kLoadWarp = 14;
    static constexpr int kNumWarps = 16;
    static constexpr int kNumThreads = kNumWarps * 32;
    static constexpr int kRegsLaunch = 128;  // 64K / 512
    static constexpr int kRegsSoftmax = 192;
    static constexpr int kRegsEpi = 64;
    static constexpr int kRegsOther = 64;
    static_assert(kSoftmaxWarps * kRegsSoftmax + 4 * kRegsEpi + 4 * kRegsOther <= kNumWarps * kRegsLaunch,
                  "register budget");
};

template <int kD, int kStages>
struct Smem {
    static constexpr int kPanels = kD / 64;
    static constexpr int kTileBytes = kPanels * kPanelBytes;  // a Q M-tile or one K / V block
    static constexpr int kQ0 = 0;
    static constexpr int kQ1 = kTileBytes;
    static constexpr int kKV = 2 * kTileBytes;
    static constexpr int kRowSum = kKV + kStages * kTileBytes;  // float [2][128] final row sums
    static constexpr int kTileRing = kRowSum + 2 * 128 * 4;      // int [kSchedDepth]
    static constexpr int kBars = kTileRing + 16;
    // barrier slots (8 B each)
    static constexpr int kBarQFull = 0;                  // [2]     TMA -> MMA
    static constexpr int kBarQEmpty = 2;                 // [2]     MMA commit -> TMA
    static constexpr int kBarKVFull = 4;                 // [kStages]
    static constexpr int kBarKVEmpty = 4 + kStages;      // [kStages] one arrival per MMA issuer
    static constexpr int kBarSFull = 4 + 2 * kStages;    // [2]     MMA commit -> softmax
    static constexpr int kBarPFull = 6 + 2 * kStages;    // [2][4]  softmax (4 warps) -> MMA (tile, part)
    static constexpr int kBarRowSum = 14 + 2 * kStages;  // [2]     softmax (4 warps) -> epilogue
    static constexpr int kBarOFull = 16 + 2 * kStages;   // [2]     MMA commit -> epilogue
    static constexpr int kBarOEmpty = 18 + 2 * kStages;  // [2]     epilogue (4 warps) -> MMA
    static constexpr int kBarTileFull = 20 + 2 * kStages;   // [kSchedDepth] producer -> everyone
    static constexpr int kBarTileEmpty = 24 + 2 * kStages;  // [kSchedDepth] consumer warps -> producer
    static constexpr int kNumBars = 28 + 2 * kStages;
    static constexpr int kTmemPtr = kBars + kNumBars * 8;
    static constexpr int kTotal = kTmemPtr + 16;
    static constexpr int kDynamicBytes = kTotal;
};

// kEmu: of every 8 element pairs, how many take the FMA-pipe exp2 (0 = all MUFU).
// kParts: equal parts in which P is handed to the MMA issuer (1, 2 or 4).
template <int kD, bool kBf16, int kStages, int kEmu, int kParts, bool kTrace>
__global__ void __launch_bounds__(Cfg::kNumThreads, 1)
fa_fwd_sm100_v6(const __grid_constant__ FwdParams params) {
    using L = Smem<kD, kStages>;
    using C = Cfg;
    constexpr int kEpiWarp0 = C::kEpiWarp0, kMmaWarp = C::kMmaWarp, kLoadWarp = C::kLoadWarp;
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kTileBytes = L::kTileBytes;
    static_assert(kParts == 1 || kParts == 2 || kParts == 4, "kParts");
    static_assert(kEmu >= 0 && kEmu <= 8, "kEmu");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    uint8_t* smem_gen = smem_raw;
    if (smem_base & 1023u) {  // 128B-swizzled TMA/UMMA tiles need a 1024 B aligned base
        if (threadIdx.x == 0) atomicCAS(params.status, 0u, 0x80000063u);
        return;
    }

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    unsigned int* status = params.status;
    auto bar = [&](int slot) -> uint32_t { return smem_base + L::kBars + slot * 8; };

    // ---- one-time setup ----
    if (warp == kLoadWarp && lane == 0) {
        prefetch_tensormap(&params.tm_q);
        prefetch_tensormap(&params.tm_k);
        prefetch_tensormap(&params.tm_v);
        for (int t = 0; t < 2; ++t) {
            mbar_init(bar(L::kBarQFull + t), 1);
            mbar_init(bar(L::kBarQEmpty + t), 1);
            mbar_init(bar(L::kBarSFull + t), 1);
            for (int q = 0; q < 4; ++q) mbar_init(bar(L::kBarPFull + 4 * t + q), 4);
            mbar_init(bar(L::kBarRowSum + t), 4);
            mbar_init(bar(L::kBarOFull + t), 1);
            mbar_init(bar(L::kBarOEmpty + t), 4);
        }
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(L::kBarKVFull + s), 1);
            mbar_init(bar(L::kBarKVEmpty + s), 2);  // one arrival per MMA issuer
        }
        for (int s = 0; s < kSchedDepth; ++s) {
            mbar_init(bar(L::kBarTileFull + s), 1);
            mbar_init(bar(L::kBarTileEmpty + s), C::kSoftmaxWarps + 4 + 2);  // softmax + epilogue + 2 MMA
        }
        fence_mbar_init();
    }
    if (warp == kMmaWarp) {
        tmem_alloc(smem_base + L::kTmemPtr, 512);
        tmem_relinquish_alloc_permit();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + L::kTmemPtr);
    const int P = params.rows_per_head;
    const bool packed = params.pack_g > 1;
    Tracer<kTrace> tr;

    if (warp >= kMmaWarp) {
        reg_dealloc<C::kRegsOther>();
        // elect_one() rather than lane == 0: ptxas then knows the role runs on one thread with uniform
        // operands and emits back-to-back UTCHMMA / UTMALDG without a per-instruction elect loop.
        const bool elected = elect_one();
        if (warp == kLoadWarp && elected) {
            // =========================== TMA producer ===========================
            tr.init(params, 4, true);
            int slot = 0;
            uint32_t phase = 0;
            uint32_t qe_phase[2] = {0, 0};
            Tile w;
            int next_idx = blockIdx.x;  // first tile: no atomic
            for (int round = 0;; ++round) {
                // publish this round's tile index (or the end marker) to the other warps
                const int idx = next_idx < params.num_tiles ? next_idx : -1;
                const int ts = round % kSchedDepth;
                mbar_wait(bar(L::kBarTileEmpty + ts), ((round / kSchedDepth) & 1) ^ 1u, status, 21);
                asm volatile("st.shared.s32 [%0], %1;" ::"r"(smem_base + L::kTileRing + ts * 4), "r"(idx) : "memory");
                mbar_arrive(bar(L::kBarTileFull + ts));
                if (idx < 0) break;
                // prefetch the next index; the atomic's latency hides behind this tile's TMA issue
                next_idx = int(gridDim.x) + atomicAdd(params.sched, 1);
                decode_tile(params, idx, w);
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (w.nblk[t] > 0) {
                        mbar_wait(bar(L::kBarQEmpty + t), qe_phase[t] ^ 1u, status, 10);
                        qe_phase[t] ^= 1u;
                        const uint32_t qbar = bar(L::kBarQFull + t);
                        mbar_arrive_expect_tx(qbar, params.q_tile_bytes);
                        for (int pn = 0; pn < kPanels; ++pn)
                            tma_load_5d(smem_base + (t ? L::kQ1 : L::kQ0) + pn * kPanelBytes,
                                        &params.tm_q, qbar, pn * 64, w.pos0[t], w.g_coord, w.kv_head, w.b);
                    }
                }
                tr.ev(1);
                for (int it = 0; it < 2 * w.nmax; ++it) {  // K0 V0 K1 V1 ...
                    const int j = it >> 1;
                    const CUtensorMap* tm = (it & 1) ? &params.tm_v : &params.tm_k;
                    mbar_wait(bar(L::kBarKVEmpty + slot), phase ^ 1u, status, 1);
                    const uint32_t fb = bar(L::kBarKVFull + slot);
                    mbar_arrive_expect_tx(fb, kTileBytes);
                    for (int pn = 0; pn < kPanels; ++pn)
                        tma_load_4d(smem_base + L::kKV + slot * kTileBytes + pn * kPanelBytes, tm, fb,
                                    pn * 64, j * kBlockN, w.kv_head, w.b);
                    if (++slot == kStages) { slot = 0; phase ^= 1u; }
                }
            }
        } else if ((warp == kMmaWarp || warp == kMmaWarp + 1) && elected) {
            // =========================== MMA issuers (one per M-tile) ===========================
            const int t = warp - kMmaWarp;
            tr.init(params, 3, t == 0);
            constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 128, false, false);
            constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
            constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO, version, SW128
            constexpr uint32_t kLboK = (16u >> 4) << 16;                          // Q, K: K-major panels
            constexpr uint32_t kLboV = (uint32_t(kPanelBytes) >> 4) << 16;        // V: MN-major panels
            auto desc = [](uint32_t lo) -> uint64_t {
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(kDescHi));
                return d;
            };
            const uint32_t q_lo = ((smem_base + (t ? L::kQ1 : L::kQ0)) >> 4) | kLboK;
            const uint32_t kv_lo = (smem_base + L::kKV) >> 4;  // + slot * (kTileBytes >> 4)
            const uint32_t tmem_s = tmem_base + t * 128;
            const uint32_t tmem_o = tmem_base + 256 + t * kD;

            auto issue_s = [&](int kslot_) {
                const uint32_t k_lo = (kv_lo + kslot_ * (kTileBytes >> 4)) | kLboK;
#pragma unroll
                for (int ks = 0; ks < kD / 16; ++ks) {
                    const uint32_t off = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
                    umma_ss(tmem_s, desc(q_lo + off), desc(k_lo + off), idesc_qk, ks > 0);
                }
            };
            // k-steps [ks0, ks1) of O_t += P_t · V_j (16 keys = 8 packed P columns each)
            auto issue_pv = [&](int vslot_, bool acc, int ks0, int ks1) {
                const uint32_t v_lo = (kv_lo + vslot_ * (kTileBytes >> 4)) | kLboV;
#pragma unroll
                for (int ks = ks0; ks < ks1; ++ks)
                    umma_ts(tmem_o, tmem_s + ks * 8, desc(v_lo + ks * (2048 >> 4)), idesc_pv,
                            (acc || ks > 0) ? 1u : 0u);
            };

            int slot = 0;           // ring position of the next K/V tile of this issuer
            uint32_t phase = 0;
            auto advance = [&]() { if (++slot == kStages) { slot = 0; phase ^= 1u; } };
            uint32_t q_phase = 0, oe_phase = 0;
            uint32_t blk_base = 0;  // KV blocks of earlier work tiles, mod 2 (P barrier parity)
            Tile w;
            TileConsumer<L> sched(smem_base, status);
            while (sched.next(params, w, true)) {
                if (w.nmax == 0) continue;
                const int n = w.nblk[t];
                if (n > 0) {
                    // prologue: S(0) = Q · K_0ᵀ
                    mbar_wait(bar(L::kBarQFull + t), q_phase, status, 3);
                    q_phase ^= 1u;
                    mbar_wait(bar(L::kBarKVFull + slot), phase, status, 2);
                    issue_s(slot);
                    tc_commit(bar(L::kBarSFull + t));
                    tc_commit(bar(L::kBarKVEmpty + slot));
                    if (n == 1) tc_commit(bar(L::kBarQEmpty + t));
                    advance();
                    tr.ev(2);
                    for (int j = 0; j < n; ++j) {
                        const int vslot = slot; const uint32_t vphase = phase; advance();
                        const uint32_t pp = (blk_base + j) & 1u;
#pragma unroll
                        for (int q = 0; q < kParts; ++q) {
                            mbar_wait(bar(L::kBarPFull + 4 * t + q), pp, status, 5);
                            if (q == 0) {
                                if (j == 0) {  // O_t of the previous work tile must have been drained
                                    mbar_wait(bar(L::kBarOEmpty + t), oe_phase ^ 1u, status, 11);
                                    oe_phase ^= 1u;
                                }
                                mbar_wait(bar(L::kBarKVFull + vslot), vphase, status, 4);
                                tr.ev(3);
                            }
                            issue_pv(vslot, j > 0, q * (8 / kParts), (q + 1) * (8 / kParts));
                        }
                        tc_commit(bar(L::kBarKVEmpty + vslot));
                        if (j + 1 < n) {
                            mbar_wait(bar(L::kBarKVFull + slot), phase, status, 6);
                            issue_s(slot);
                            tc_commit(bar(L::kBarSFull + t));
                            tc_commit(bar(L::kBarKVEmpty + slot));
                            if (j + 2 == n) tc_commit(bar(L::kBarQEmpty + t));
                            advance();
                        } else {
                            tc_commit(bar(L::kBarOFull + t));
                        }
                        tr.ev(5);
                    }
                }
                // K/V tiles only the other (longer) M-tile uses: stay in lockstep with the ring
                for (int it = (n > 0 ? 2 * n : 0); it < 2 * w.nmax; ++it) {
                    mbar_wait(bar(L::kBarKVFull + slot), phase, status, 16);
                    mbar_arrive(bar(L::kBarKVEmpty + slot));
                    advance();
                }
                blk_base += static_cast<uint32_t>(n);
            }
        }
    } else if (warp >= kEpiWarp0) {
        // =========================== epilogue warpgroup ===========================
        reg_dealloc<C::kRegsEpi>();
        tr.init(params, 2, warp == kEpiWarp0 && lane == 0);
        const int r = threadIdx.x & 127;  // row inside an M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t rs_phase[2] = {0, 0}, of_phase[2] = {0, 0};
        Tile w;
        TileConsumer<L> sched(smem_base, status);
        while (sched.next(params, w, lane == 0)) {
            // O / l -> T -> global (row-contiguous 16 B stores)
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                if (w.pos0[t] >= params.seqlen_q) continue;
                const int n = w.nblk[t];
                const int g = r / P;
                const int pos = w.pos0[t] + (r - g * P);
                const bool row_valid = (g < params.pack_g) && (pos < params.seqlen_q);
                const int head = packed ? w.hu * params.group + g : w.hu;
                uint8_t* orow = reinterpret_cast<uint8_t*>(params.o) +
                                2 * (static_cast<int64_t>(w.b) * params.o_stride_b +
                                     static_cast<int64_t>(head) * params.o_stride_h +
                                     static_cast<int64_t>(pos) * params.o_stride_s);
                float inv = 0.f;
                if (n > 0) {
                    mbar_wait(bar(L::kBarRowSum + t), rs_phase[t], status, 14);
                    rs_phase[t] ^= 1u;
                    const float l = ld_shared_f32(smem_base + L::kRowSum + (t * 128 + r) * 4);
                    inv = (l > 0.f) ? __frcp_rn(l) : 0.f;
                    mbar_wait(bar(L::kBarOFull + t), of_phase[t], status, 8);
                    of_phase[t] ^= 1u;
                    tc_fence_after_sync();
                    tr.ev(8);
                }
                const uint32_t tmem_o = tmem_base + lane_base + 256 + t * kD;
#pragma unroll
                for (int c = 0; c < kD / 32; ++c) {
                    uint32_t o[32];
                    if (n > 0) {
                        tmem_ld32(tmem_o + c * 32, o);
                        tmem_wait_ld();
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = 0u;
                    }
                    if (c == kD / 32 - 1 && n > 0) {  // O_t fully read: the next work tile may overwrite it
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar(L::kBarOEmpty + t));
                    }
                    if (row_valid) {
#pragma unroll
                        for (int q4 = 0; q4 < 4; ++q4) {
                            const int col = c * 32 + q4 * 8;
                            if (col < params.headdim) {
                                uint4 v;
                                v.x = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 0]) * inv,
                                                   __uint_as_float(o[q4 * 8 + 1]) * inv);
                                v.y = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 2]) * inv,
                                                   __uint_as_float(o[q4 * 8 + 3]) * inv);
                                v.z = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 4]) * inv,
                                                   __uint_as_float(o[q4 * 8 + 5]) * inv);
                                v.w = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 6]) * inv,
                                                   __uint_as_float(o[q4 * 8 + 7]) * inv);
                                *reinterpret_cast<uint4*>(orow + col * 2) = v;
                            }
                        }
                    }
                }
                tr.ev(9);
            }
        }
    } else {
        // =========================== softmax warpgroups ===========================
        reg_alloc<C::kRegsSoftmax>();
        const int t = warp >> 2;          // M-tile handled by this warpgroup
        tr.init(params, t, (warp & 3) == 0 && lane == 0);
        const int r = threadIdx.x & 127;  // query row inside the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t tmem_s = tmem_base + lane_base + t * 128;
        const uint32_t tmem_o = tmem_base + lane_base + 256 + t * kD;
        const float scale_log2 = params.scale_log2;
        const int causal_off = params.seqlen_k - params.seqlen_q;
        const uint32_t rowsum_addr = smem_base + L::kRowSum + (t * 128 + r) * 4;
        const uint64_t sc2 = pack_f32x2(scale_log2, scale_log2);
        uint32_t sf_phase = 0;
        Tile w;
        TileConsumer<L> sched(smem_base, status);
        while (sched.next(params, w, lane == 0)) {
            const int n = w.nblk[t];
            if (n == 0) continue;
            const int my_pos0 = w.pos0[t];
            const int g = r / P;
            const int pos = my_pos0 + (r - g * P);
            int row_limit = params.seqlen_k - 1;  // last visible column
            if (params.causal) row_limit = min(row_limit, pos + causal_off);

            float m = -INFINITY;  // running row max (raw score units), possibly stale (lazy rescale)
            float l = 0.f;        // running row sum of exp2, relative to m
            float msc = 0.f;      // m * scale_log2 (0 while m == -inf)
            uint64_t nm2 = pack_f32x2(0.f, 0.f);  // (-msc, -msc)

            for (int j = 0; j < n; ++j) {
                mbar_wait(bar(L::kBarSFull + t), sf_phase, status, 7);
                sf_phase ^= 1u;
                tc_fence_after_sync();
                tr.ev(1);

                uint32_t s[128];
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    tmem_ld32(tmem_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
                tmem_wait_ld();
                tr.ev(2);

                // exp2 of 32 keys -> 16 packed columns of P, stored straight back over S
                auto exp_chunk = [&](const uint32_t* sv, int c, uint64_t& sa, uint64_t& sb) {
                    uint32_t pk[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const uint64_t x2 = fma_f32x2(
                            pack_f32x2(__uint_as_float(sv[2 * q]), __uint_as_float(sv[2 * q + 1])), sc2, nm2);
                        float p0, p1;
                        if (((q * kEmu) & 7) < kEmu) {
                            ex2_emu_x2(x2, p0, p1);
                        } else {
                            float x0, x1;
                            unpack_f32x2(x2, x0, x1);
                            p0 = ex2_approx(x0);
                            p1 = ex2_approx(x1);
                        }
                        if (q & 1) sb = add_f32x2(sb, pack_f32x2(p0, p1));
                        else sa = add_f32x2(sa, pack_f32x2(p0, p1));
                        pk[q] = pack2<kBf16>(p0, p1);
                    }
                    tmem_st16(tmem_s + c * 16, pk);
                };
                auto row_max = [&]() -> float {
                    float mx0 = max3(__uint_as_float(s[0]), __uint_as_float(s[1]), __uint_as_float(s[2]));
                    float mx1 = max3(__uint_as_float(s[3]), __uint_as_float(s[4]), __uint_as_float(s[5]));
                    float mx2 = max3(__uint_as_float(s[6]), __uint_as_float(s[7]), __uint_as_float(s[8]));
                    float mx3 = max3(__uint_as_float(s[9]), __uint_as_float(s[10]), __uint_as_float(s[11]));
#pragma unroll
                    for (int i = 12; i < 124; i += 8) {
                        mx0 = max3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
                        mx1 = max3(mx1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
                        mx2 = max3(mx2, __uint_as_float(s[i + 4]), __uint_as_float(s[i + 5]));
                        mx3 = max3(mx3, __uint_as_float(s[i + 6]), __uint_as_float(s[i + 7]));
                    }
                    mx0 = max3(mx0, __uint_as_float(s[124]), __uint_as_float(s[125]));
                    mx1 = max3(mx1, __uint_as_float(s[126]), __uint_as_float(s[127]));
                    return fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
                };
                auto publish = [&](int part) {
                    tmem_wait_st();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(L::kBarPFull + 4 * t + part));
                };

                // element mask only on blocks that cross the diagonal or the end of K
                const int col0 = j * kBlockN;
                const bool need_mask = (col0 + kBlockN > params.seqlen_k) ||
                                       (params.causal && (col0 + kBlockN - 1 > my_pos0 + causal_off));
                uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);

                // Fast path (every block but the first and the masked ones): the first 32 keys are
                // exponentiated against the stale running max while the row max of the whole block
                // is reduced in the shadow of those MUFU ops; the result stands unless some row of
                // this warp grew past the lazy-rescale threshold (then the block is redone below —
                // the scores are still in registers, P chunk 0 is simply overwritten; nothing has
                // been published yet).
                bool fast = false;
                if (j > 0 && !need_mask) {
                    uint64_t sa = pack_f32x2(0.f, 0.f), sb = pack_f32x2(0.f, 0.f);
                    exp_chunk(&s[0], 0, sa, sb);
                    const float mx = row_max();
                    fast = !__any_sync(0xffffffffu, (mx - m) * scale_log2 > kRescaleThreshold);
                    if (fast) { la = sa; lb = sb; }
                }
                if (!fast) {
                    if (need_mask) {
                        const int rel = row_limit - col0;
#pragma unroll
                        for (int i = 0; i < 128; ++i)
                            if (i > rel) s[i] = 0xff800000u;  // -inf
                    }
                    const float m_new = fmaxf(m, row_max());
                    // lazy rescale: keep the stale max unless this row grew by more than 2^8
                    float alpha = 1.f;
                    if ((m_new - m) * scale_log2 > kRescaleThreshold) {  // false for NaN (-inf - -inf)
                        alpha = ex2_approx((m - m_new) * scale_log2);
                        m = m_new;
                        msc = m * scale_log2;
                        nm2 = pack_f32x2(-msc, -msc);
                        l *= alpha;
                    }
                    if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
                        // O_t is quiescent here: S_t(j) complete implies P·V_t(j-1) complete, and
                        // P·V_t(j) cannot start before this warp publishes part 0 of P_t(j).
                        tr.ev(7);
#pragma unroll 1
                        for (int c = 0; c < kD / 8; ++c) {  // small chunks: the 128 scores stay in registers
                            uint32_t o[8];
                            tmem_ld8(tmem_o + c * 8, o);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                            tmem_st8(tmem_o + c * 8, o);
                        }
                        tmem_wait_st();
                    }
                    exp_chunk(&s[0], 0, la, lb);
                }
                tr.ev(3);
                if constexpr (kParts == 4) publish(0);
                exp_chunk(&s[32], 1, la, lb);
                if constexpr (kParts == 4) publish(1);
                if constexpr (kParts == 2) publish(0);
                exp_chunk(&s[64], 2, la, lb);
                if constexpr (kParts == 4) publish(2);
                exp_chunk(&s[96], 3, la, lb);
                publish(kParts - 1);
                tr.ev(4);
                float l0, l1;
                unpack_f32x2(add_f32x2(la, lb), l0, l1);
                l += l0 + l1;
            }
            st_shared_f32(rowsum_addr, l);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(L::kBarRowSum + t));
        }
    }

    // ---- teardown ----
    tc_fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
    if (threadIdx.x == 0) {
        // the last CTA out re-arms the scheduler for the next launch that uses this slot
        __threadfence();
        if (atomicAdd(params.sched + 1, 1) == int(gridDim.x) - 1) {
            params.sched[0] = 0;
            params.sched[1] = 0;
            __threadfence();
        }
    }
}

}  // namespace v6
}  // namespace fa100
