// FlashAttention-2 forward for sm_100a (B200): persistent, warp-specialised tcgen05 / TMEM / TMA
// kernel with a half-block software pipeline.
//
// Replaces the reference's flash_attention_v2 (csrc/flash_attention_template.cuh:138-564),
// Mask (csrc/mask.cuh:30-88) and launch_flash_attention (template.cuh:567-583).  Same math
// (fp32 S and O accumulators, exp2-domain online softmax, P rounded to T before P·V, row sums
// from unrounded P, bottom-right causal alignment, kv-head = q-head / G).
//
//   TMEM: tile t, buffer b: S/P at columns [128 t + 64 b, +64); O_t at [256 + D t, +D).
//   Every 128-key KV block is visited as two 64-key half blocks h = 0, 1, 2, ... that alternate
//   between the two S buffers of their M-tile, so that S(h+1) is already waiting when softmax(h)
//   ends and S(h+2) is issued right behind P·V(h): the softmax warpgroups run back to back and
//   the tensor pipe always has the other buffer's or the other tile's MMAs queued.
//
//   grid = min(#SMs, #work tiles) persistent CTAs of 512 threads; work tile = 2 M-tiles of 128
//   query rows sharing one K/V stream; LPT + L2-section order, dynamic scheduler
//   (fa_sm100_sched.cuh).  An M-tile's 128 rows are (head-in-group g, position p) pairs:
//   row r -> g = r / P, p = r % P with P = rows_per_head (unpacked: P = 128, one q head;
//   GQA-packed: P = 128 / G and the Q TMA box is (64, P, G) of the 5-D tensor map).
//     warps 0-3   softmax warpgroup, M-tile 0 (one thread per query row, no shuffles); also
//                 performs the (rare) lazy rescale of O_t in TMEM
//     warps 4-7   softmax warpgroup, M-tile 1
//     warps 8-11  epilogue warpgroup: O_t / l -> T -> global; zeroes the V rows that belong to
//                 the next sequence of a packed variable-length batch
//     warps 12,13 MMA issuers (one elected thread each), one per M-tile: S_t = Q_t·Khᵀ (SS),
//                 O_t += P_t·Vh (TS).  A K/V slot is released when both have committed on it.
//     warp  14    TMA producer of the K/V ring + tile scheduler (one elected thread)
//     warp  15    Q loader (one elected thread)
//
// Every mbarrier wait is bounded (fa_sm100_ptx.cuh): when one gives up, every role leaves its
// loop without touching global memory again and the CTA tears down normally.
#pragma once
#include "fa_sm100_sched.cuh"

namespace fa100 {
namespace prefill {

struct Cfg {
    static constexpr int kSoftmaxWarps = 8;
    static constexpr int kEpiWarp0 = 8;
    static constexpr int kMmaWarp = 12;   // 12: M-tile 0, 13: M-tile 1
    static constexpr int kLoadWarp = 14;
    static constexpr int kQWarp = 15;
    static constexpr int kNumWarps = 16;
    static constexpr int kNumThreads = kNumWarps * 32;
    static constexpr int kRegsLaunch = 128;  // 64K / 512
    static constexpr int kRegsSoftmax = 168;
    static constexpr int kRegsEpi = 112;
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
    static constexpr int kHalfBytes = kTileBytes / 2;           // 64 keys of K or V
    static constexpr int kHalfPanel = kPanelBytes / 2;          // 64 rows x 64 16-bit elements
    static constexpr int kStageBytes = 2 * kHalfBytes;          // ring stage = two half tiles
    static constexpr int kRowSum = kKV + kStages * kStageBytes;  // float [2][128] final row sums
    static constexpr int kTileRec = 64;                           // decoded work tile: 16 ints
    static constexpr int kTileRing = kRowSum + 2 * 128 * 4;      // [kSchedDepth] records
    static constexpr int kBars = kTileRing + kSchedDepth * kTileRec;
    // barrier slots (8 B each)
    static constexpr int kBarQFull = 0;                  // [2]     TMA -> MMA
    static constexpr int kBarQEmpty = 2;                 // [2]     MMA commit -> TMA
    static constexpr int kBarKVFull = 4;                 // [kStages]
    static constexpr int kBarKVEmpty = 4 + kStages;      // [kStages]
    static constexpr int kBarSFull = 4 + 2 * kStages;    // [2][2]  MMA commit -> softmax   (tile, buffer)
    static constexpr int kBarPFull = 8 + 2 * kStages;    // [2][2]  softmax (4 warps) -> MMA (tile, buffer)
    static constexpr int kBarPVDone = 12 + 2 * kStages;  // [2]     MMA commit after the last two P·V of a tile (rescale guard)
    static constexpr int kBarRowSum = 14 + 2 * kStages;  // [2]     softmax (4 warps) -> epilogue
    static constexpr int kBarOFull = 16 + 2 * kStages;   // [2]     MMA commit -> epilogue
    static constexpr int kBarOEmpty = 18 + 2 * kStages;  // [2]     epilogue (4 warps) -> MMA, softmax
    static constexpr int kBarTileFull = 20 + 2 * kStages;   // [kSchedDepth] producer -> everyone
    static constexpr int kBarTileEmpty = 24 + 2 * kStages;  // [kSchedDepth] consumer warps -> producer
    static constexpr int kBarVFix = 28 + 2 * kStages;       // [1]  epilogue (4 warps) -> MMA, producer: V tail rows zeroed
    static constexpr int kBarVFixReq = 29 + 2 * kStages;    // [1]  producer -> epilogue: the tail stage's TMA is issued
    static constexpr int kNumBars = 30 + 2 * kStages;
    static constexpr int kTmemPtr = kBars + kNumBars * 8;
    static constexpr int kAbort = kTmemPtr + 8;  // int: the launch found the device status word set
    static constexpr int kTotal = kTmemPtr + 16;
    // No slack: the kernel has no static shared memory, so the dynamic window starts 1024 B aligned
    // (checked at run time: a misaligned base aborts with status code 0x80000063).
    static constexpr int kDynamicBytes = kTotal;
};

// Work-tile hand-off.  The producer thread decodes each tile index once (a chain of integer
// divisions: some 3000 cycles for a lone thread on a busy scheduler, measured) and publishes the
// decoded record through the shared-memory ring, one tile ahead of the consumers; they only copy
// it out.  mblock < 0 ends the CTA.
template <typename L>
__device__ __forceinline__ void tile_store(uint32_t smem_base, int slot, const Tile& t) {
    const uint32_t a = smem_base + L::kTileRing + slot * L::kTileRec;
    asm volatile("st.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(t.mblock), "r"(t.hu), "r"(t.b), "r"(t.kv_head) : "memory");
    asm volatile("st.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a + 16), "r"(t.g_coord), "r"(t.pos0[0]), "r"(t.pos0[1]), "r"(t.nblk[0]) : "memory");
    asm volatile("st.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a + 32), "r"(t.nblk[1]), "r"(t.nmax), "r"(t.seqlen_q), "r"(t.seqlen_k) : "memory");
    asm volatile("st.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a + 48), "r"(t.q_row0), "r"(t.k_row0), "r"(t.flags), "r"(0) : "memory");
}
template <typename L>
struct TileConsumer {
    uint32_t smem_base;
    int round = 0;
    const WaitCtx& ctx;
    __device__ __forceinline__ TileConsumer(uint32_t base, const WaitCtx& c) : smem_base(base), ctx(c) {}
    // All lanes of a consumer warp call this together (or the single active thread of a role).
    // Returns 1 = tile, 0 = end of work, -1 = the launch was aborted.
    __device__ __forceinline__ int next(Tile& t, bool elected) {
        const int slot = round % kSchedDepth;
        const uint32_t parity = (round / kSchedDepth) & 1;
        if (!mbar_wait(smem_base + L::kBars + (L::kBarTileFull + slot) * 8, parity, ctx, 20)) return -1;
        const uint32_t a = smem_base + L::kTileRing + slot * L::kTileRec;
        int pad;
        asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(t.mblock), "=r"(t.hu), "=r"(t.b), "=r"(t.kv_head) : "r"(a) : "memory");
        asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(t.g_coord), "=r"(t.pos0[0]), "=r"(t.pos0[1]), "=r"(t.nblk[0]) : "r"(a + 16) : "memory");
        asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(t.nblk[1]), "=r"(t.nmax), "=r"(t.seqlen_q), "=r"(t.seqlen_k) : "r"(a + 32) : "memory");
        asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(t.q_row0), "=r"(t.k_row0), "=r"(t.flags), "=r"(pad) : "r"(a + 48) : "memory");
        __syncwarp(__activemask());
        if (elected) mbar_arrive(smem_base + L::kBars + (L::kBarTileEmpty + slot) * 8);
        ++round;
        return t.mblock >= 0 ? 1 : 0;
    }
};

// kEmu: of every 8 element pairs, how many take the FMA-pipe exp2 (0 = all MUFU).
// kSpec: speculative stale-max fast path in the softmax (row max reduced under the first exp2 chunk).
// kFence: tcgen05.fence::after_thread_sync between the P-full wait and the P·V MMA it orders.
template <int kD, bool kBf16, int kStages, int kEmu, bool kSpec, bool kFence, bool kTrace>
__global__ void __launch_bounds__(Cfg::kNumThreads, 1)
fa_fwd_sm100(const __grid_constant__ FwdParams params) {
    using L = Smem<kD, kStages>;
    using C = Cfg;
    constexpr int kEpiWarp0 = C::kEpiWarp0, kMmaWarp = C::kMmaWarp, kLoadWarp = C::kLoadWarp, kQWarp = C::kQWarp;
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kHalfBytes = L::kHalfBytes;
    constexpr uint32_t kHalfPanel = L::kHalfPanel;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    uint8_t* smem_gen = smem_raw;
    const WaitCtx& ctx = params.wait;
    if (smem_base & 1023u) {  // 128B-swizzled TMA/UMMA tiles need a 1024 B aligned base
        if (threadIdx.x == 0 && atomicCAS(ctx.status, 0u, 0x80000063u) == 0u && ctx.status_host) {
            *reinterpret_cast<volatile unsigned int*>(ctx.status_host) = 0x80000063u;
            __threadfence_system();
        }
        return;
    }

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    auto bar = [&](int slot) -> uint32_t { return smem_base + L::kBars + slot * 8; };
#define FA_WAIT(slot_, parity_, code_) \
    do { if (!mbar_wait(bar(slot_), (parity_), ctx, (code_))) return; } while (0)

    // ---- one-time setup ----
    if (warp == kLoadWarp && lane == 0) {
        // A launch that finds the abort flag of an earlier one still set does nothing at all.
        *reinterpret_cast<volatile int*>(smem_gen + L::kAbort) =
            (*reinterpret_cast<volatile unsigned int*>(ctx.status) != 0u) ? 1 : 0;
        prefetch_tensormap(&params.tm_q);
        prefetch_tensormap(&params.tm_k);
        prefetch_tensormap(&params.tm_v);
        for (int t = 0; t < 2; ++t) {
            mbar_init(bar(L::kBarQFull + t), 1);
            mbar_init(bar(L::kBarQEmpty + t), 1);
            for (int b = 0; b < 2; ++b) {
                mbar_init(bar(L::kBarSFull + 2 * t + b), 1);
                mbar_init(bar(L::kBarPFull + 2 * t + b), 4);
            }
            mbar_init(bar(L::kBarPVDone + t), 1);
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
            mbar_init(bar(L::kBarTileEmpty + s), C::kSoftmaxWarps + 4 + 3);  // softmax + epilogue + 2 MMA + Q loader
        }
        mbar_init(bar(L::kBarVFix), 4);
        mbar_init(bar(L::kBarVFixReq), 1);
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
    const bool launch_aborted = *reinterpret_cast<volatile int*>(smem_gen + L::kAbort) != 0;
    const int P = params.rows_per_head;
    const bool packed = params.pack_g > 1;

    if (launch_aborted) {
        // fall through to the teardown
    } else if (warp >= kMmaWarp) {
        reg_dealloc<C::kRegsOther>();
        // elect_one() rather than lane == 0: ptxas then knows the role runs on one thread with uniform
        // operands and emits back-to-back UTCHMMA / UTMALDG without a per-instruction elect loop.
        const bool elected = elect_one();
        if (warp == kLoadWarp && elected) {
            // =========================== TMA producer + tile scheduler ===========================
            [&]() {
                Tracer<kTrace> tr;
                tr.init(params, 4, true);
                int slot = 0;
                uint32_t phase = 0;
                uint32_t vfix_count = 0;  // packed batches: work tiles so far whose V tail needs zeroing
                Tile w, wn;
                const bool varlen = params.cu_q != nullptr;
                const int num_tiles = varlen ? params.plan[0] : params.num_tiles;
                VarlenCursor cur;
                // decode tile index idx (or the end marker) and hand the record to the other warps
                auto publish = [&](int round, int idx, Tile& t) -> bool {
                    const int ts = round % kSchedDepth;
                    if (idx < num_tiles) {
                        if (varlen) decode_tile_varlen(params, idx, cur, t);
                        else decode_tile(params, idx, t);
                    } else {
                        t.mblock = -1;
                        t.nmax = 0;
                    }
                    if (!mbar_wait(bar(L::kBarTileEmpty + ts), ((round / kSchedDepth) & 1) ^ 1u, ctx, 21)) return false;
                    tile_store<L>(smem_base, ts, t);
                    mbar_arrive(bar(L::kBarTileFull + ts));
                    return true;
                };
                if (!publish(0, blockIdx.x, w)) return;  // first tile: no atomic
                // The index after that is always in flight: the atomic's round trip (about 1000 cycles
                // with every CTA on one counter) must not sit between two work tiles.
                int next_idx = int(gridDim.x) + atomicAdd(params.sched, 1);
                for (int round = 0; w.mblock >= 0; ++round) {
                    bool next_published = false;
                    // The next tile is decoded and published while this tile's stages wait for ring
                    // slots: the Q loader needs it before this tile ends, and a consumer that finishes
                    // early finds it ready.
                    auto publish_next = [&]() -> bool {
                        if (!publish(round + 1, next_idx, wn)) return false;
                        next_published = true;
                        if (wn.mblock >= 0) next_idx = int(gridDim.x) + atomicAdd(params.sched, 1);
                        return true;
                    };
                    tr.ev(1);
                    // Ring stages of two 64-key half tiles, in the order the MMA issuers consume them
                    // (one full/empty barrier pair per stage, so an issuer does one wait and one
                    // commit per half-block):
                    //   {K0,K1} {V0,K2} {V1,K3} ... {V(nh-3),K(nh-1)} {V(nh-2),V(nh-1)}
                    const int nh = 2 * w.nmax;  // half-blocks of the longer M-tile
                    for (int i = 0; i < nh; ++i) {
                        const CUtensorMap* tm0 = (i == 0) ? &params.tm_k : &params.tm_v;
                        const CUtensorMap* tm1 = (i == nh - 1) ? &params.tm_v : &params.tm_k;
                        const int h0 = (i == 0) ? 0 : i - 1;
                        const int h1 = (i == nh - 1) ? nh - 1 : i + 1;
                        FA_WAIT(L::kBarKVEmpty + slot, phase ^ 1u, 1);
                        const uint32_t fb = bar(L::kBarKVFull + slot);
                        mbar_arrive_expect_tx(fb, 2 * kHalfBytes);
                        const uint32_t dst = smem_base + L::kKV + slot * (2 * kHalfBytes);
                        for (int pn = 0; pn < kPanels; ++pn)
                            tma_load_4d(dst + pn * kHalfPanel, tm0, fb, pn * 64, w.k_row0 + h0 * 64, w.kv_head, w.b);
                        for (int pn = 0; pn < kPanels; ++pn)
                            tma_load_4d(dst + kHalfBytes + pn * kHalfPanel, tm1, fb, pn * 64, w.k_row0 + h1 * 64, w.kv_head, w.b);
                        if (++slot == kStages) { slot = 0; phase ^= 1u; }
                        tr.ev(6);
                        if (i == 1 && !publish_next()) return;
                    }
                    if (!next_published && !publish_next()) return;
                    if (w.flags & kTileVFix) {
                        // Tell the epilogue that the tail stage's loads are on their way: only now is its
                        // wait on that stage's full barrier unambiguous (the ring is kStages deep and
                        // the barriers carry one parity bit; a waiter that runs a whole tile ahead of the
                        // ring would otherwise mistake an older fill of the slot for this one).  One
                        // request at a time: the previous one must have been answered.
                        if (vfix_count > 0) FA_WAIT(L::kBarVFix, (vfix_count - 1) & 1u, 22);
                        mbar_arrive(bar(L::kBarVFixReq));
                        ++vfix_count;
                    }
                    w = wn;
                }
            }();
        } else if (warp == kQWarp && elected) {
            // =========================== Q loader ===========================
            // Q_t of a work tile goes out as soon as M-tile t of the previous one has issued its last
            // S (QEmpty) — independently for the two M-tiles and of the K/V ring, so the tile that
            // finishes first restarts without waiting for the other.
            [&]() {
                uint32_t qe_phase0 = 0, qe_phase1 = 0;
                Tile w;
                TileConsumer<L> sched(smem_base, ctx);
                while (sched.next(w, true) > 0) {
                    bool pend0 = w.nblk[0] > 0, pend1 = w.nblk[1] > 0;
                    auto load_q = [&](int t) {
                        const uint32_t qbar = bar(L::kBarQFull + t);
                        mbar_arrive_expect_tx(qbar, params.q_tile_bytes);
                        for (int pn = 0; pn < kPanels; ++pn)
                            tma_load_5d(smem_base + (t ? L::kQ1 : L::kQ0) + pn * kPanelBytes, &params.tm_q, qbar,
                                        pn * 64, w.q_row0 + (t ? w.pos0[1] : w.pos0[0]), w.g_coord, w.kv_head, w.b);
                    };
                    for (uint32_t spins = 0; pend0 || pend1;) {
                        if (pend0 && mbar_try_wait(bar(L::kBarQEmpty + 0), qe_phase0 ^ 1u)) {
                            qe_phase0 ^= 1u; pend0 = false; load_q(0);
                        }
                        if (pend1 && mbar_try_wait(bar(L::kBarQEmpty + 1), qe_phase1 ^ 1u)) {
                            qe_phase1 ^= 1u; pend1 = false; load_q(1);
                        }
                        if ((++spins & 0xffu) == 0 && *reinterpret_cast<volatile unsigned int*>(ctx.status) != 0u) return;
                    }
                }
            }();
        } else if ((warp == kMmaWarp || warp == kMmaWarp + 1) && elected) {
            // =========================== MMA issuer ===========================
            [&]() {
                Tracer<kTrace> tr;
                tr.init(params, 3, warp == kMmaWarp);
                constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 64, false, false);
                constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
                // UMMA smem descriptors are rebuilt per instruction from a 32-bit base (start address
                // field) plus a compile-time offset.
                constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO, version, SW128
                constexpr uint32_t kLboK = (16u >> 4) << 16;                          // Q, K: K-major panels
                constexpr uint32_t kLboV = (kHalfPanel >> 4) << 16;                   // V: MN-major panels
                auto desc = [](uint32_t lo) -> uint64_t {
                    uint64_t d;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(kDescHi));
                    return d;
                };
                const uint32_t q_lo0 = ((smem_base + L::kQ0) >> 4) | kLboK, q_lo1 = ((smem_base + L::kQ1) >> 4) | kLboK;
                const uint32_t kv_lo = (smem_base + L::kKV) >> 4;  // + slot * (kHalfBytes >> 4)

                // S(t, buffer b) = Q_t · Khᵀ, Kh = the 64-key half tile in ring slot kslot_
                auto issue_s = [&](int t, int b, int kslot_) {
                    const uint32_t k_lo = (kv_lo + kslot_ * (kHalfBytes >> 4)) | kLboK;
                    const uint32_t d = tmem_base + t * 128 + b * 64;
#pragma unroll
                    for (int ks = 0; ks < kD / 16; ++ks) {
                        const uint32_t qoff = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
                        const uint32_t koff = ((ks >> 2) * kHalfPanel + (ks & 3) * 32) >> 4;
                        umma_ss(d, desc((t ? q_lo1 : q_lo0) + qoff), desc(k_lo + koff), idesc_qk, ks > 0);
                    }
                };
                // O_t += P(t, buffer b) · Vh, Vh = the 64-key half tile in ring slot vslot_
                auto issue_pv = [&](int t, int b, int vslot_, bool acc) {
                    const uint32_t v_lo = (kv_lo + vslot_ * (kHalfBytes >> 4)) | kLboV;
                    const uint32_t a = tmem_base + t * 128 + b * 64;
                    const uint32_t d = tmem_base + 256 + t * kD;
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_ts(d, a + ks * 8, desc(v_lo + ks * (2048 >> 4)), idesc_pv, (acc || ks > 0) ? 1u : 0u);
                };

                // Ring position of this work tile's first half tile (the producer's order is
                // K0 K1 | V0 K2 | V1 K3 | ... | V(nh-2) | V(nh-1), nh = half-blocks of the longer M-tile).
                const int t = warp - kMmaWarp;
                int slot0 = 0;
                uint32_t phase0 = 0;
                uint32_t q_phase = 0, oe_phase = 0, vfix_phase = 0;
                uint32_t blk_base = 0;  // KV blocks of earlier work tiles, mod 2 (P barrier parity)
                Tile w;
                Tile wn;
                TileConsumer<L> sched(smem_base, ctx);
                // The next tile's record is published one work tile ahead; it is read in front of the
                // wait for the last P of this tile (idle time) so that nothing but the Q/K waits sits
                // between the last P·V of one tile and the first S of the next.
                int have = sched.next(w, true);
                while (have > 0) {
                    bool fetched = false;
                    int have_next = 0;
                    if (w.nmax > 0) {
                        auto ring = [&](int it, int& s_, uint32_t& ph_) {
                            const int x = slot0 + it;
                            const int q = x / kStages;
                            s_ = x - q * kStages;
                            ph_ = phase0 ^ static_cast<uint32_t>(q & 1);
                        };
                        const int nh = 2 * (t ? w.nblk[1] : w.nblk[0]);
                        const int nhmax = 2 * w.nmax;
                        const bool vfix = (w.flags & kTileVFix) != 0;
                        // wait for ring stage i of this work tile; returns its slot, -1 = aborted
                        auto acquire = [&](int i, uint32_t code) -> int {
                            int sl; uint32_t ph;
                            ring(i, sl, ph);
                            if (!mbar_wait(bar(L::kBarKVFull + sl), ph, ctx, code)) return -1;
                            return sl;
                        };
                        if (nh > 0) {
                            // prologue: S(0), S(1) from stage 0 = {K0, K1}
                            FA_WAIT(L::kBarQFull + t, q_phase, 3);
                            q_phase ^= 1u;
                            const int s0 = acquire(0, 2);
                            if (s0 < 0) return;
                            issue_s(t, 0, 2 * s0);
                            tc_commit(bar(L::kBarSFull + 2 * t + 0));
                            issue_s(t, 1, 2 * s0 + 1);
                            tc_commit(bar(L::kBarSFull + 2 * t + 1));
                            tc_commit(bar(L::kBarKVEmpty + s0));
                            if (nh == 2) tc_commit(bar(L::kBarQEmpty + t));
                            tr.ev(2);
                            int st = 0;  // stage holding V(h): stage h+1 {V(h), K(h+2)}, or the last one {V(nh-2), V(nh-1)}
                            for (int h = 0; h < nh; ++h) {
                                const int b = h & 1;
                                if (h == nh - 1) {
                                    have_next = sched.next(wn, true);
                                    if (have_next < 0) return;
                                    fetched = true;
                                }
                                FA_WAIT(L::kBarPFull + 2 * t + b, (blk_base + (h >> 1)) & 1u, 5);
                                if (h == 0) {  // O_t of the previous work tile must have been drained
                                    FA_WAIT(L::kBarOEmpty + t, oe_phase ^ 1u, 11);
                                    oe_phase ^= 1u;
                                }
                                const bool tail = (h == nhmax - 1);  // V(nhmax-1) sits in the second half of the last stage
                                if (!tail) {
                                    st = acquire(h + 1, 4);
                                    if (st < 0) return;
                                    // the last stage of a packed sequence: its foreign V rows must be zero first
                                    if (vfix && h + 1 == nhmax - 1) FA_WAIT(L::kBarVFix, vfix_phase, 18);
                                }
                                if constexpr (kFence) tc_fence_after_sync();
                                tr.ev(3);
                                issue_pv(t, b, 2 * st + (tail ? 1 : 0), h > 0);
                                if (h + 2 < nh) {
                                    issue_s(t, b, 2 * st + 1);
                                    tc_commit(bar(L::kBarSFull + 2 * t + b));
                                    if (h + 3 == nh) tc_commit(bar(L::kBarQEmpty + t));
                                } else {
                                    tc_commit(bar(L::kBarPVDone + t));  // no S follows: the rescale guard of the last steps
                                }
                                if (h + 1 < nhmax - 1 || tail) tc_commit(bar(L::kBarKVEmpty + st));
                                if (h + 1 == nh) tc_commit(bar(L::kBarOFull + t));
                                tr.ev(5);
                            }
                        }
                        // stages only the other (longer) M-tile uses: stay in lockstep with the ring
                        tr.ev(23);
                        for (int i = (nh > 0 ? nh + 1 : 0); i < nhmax; ++i) {
                            const int sl = acquire(i, 16);
                            if (sl < 0) return;
                            mbar_arrive(bar(L::kBarKVEmpty + sl));
                        }
                        if (vfix) vfix_phase ^= 1u;
                        blk_base += static_cast<uint32_t>((t ? w.nblk[1] : w.nblk[0]));
                        {
                            const int x = slot0 + 2 * w.nmax;  // nhmax stages
                            const int q = x / kStages;
                            slot0 = x - q * kStages;
                            phase0 ^= static_cast<uint32_t>(q & 1);
                        }
                    }  // w.nmax > 0
                    if (!fetched) have_next = sched.next(wn, true);
                    w = wn;
                    have = have_next;
                }
            }();
        }
    } else if (warp >= kEpiWarp0) {
        // =========================== epilogue warpgroup ===========================
        reg_dealloc<C::kRegsEpi>();
        [&]() {
            Tracer<kTrace> tr;
            tr.init(params, 2, warp == kEpiWarp0 && lane == 0);
            const int r = threadIdx.x & 127;  // row inside an M-tile == TMEM lane
            const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
            uint32_t rs_phase[2] = {0, 0}, of_phase[2] = {0, 0};
            int slot0 = 0;        // ring position of the work tile's first stage (as the issuers track it)
            uint32_t phase0 = 0;
            uint32_t vfix_phase = 0;
            Tile w;
            TileConsumer<L> sched(smem_base, ctx);
            while (sched.next(w, lane == 0) > 0) {
                if (w.nmax > 0) {
                    const int x = slot0 + 2 * w.nmax;
                    if (w.flags & kTileVFix) {
                        // Packed batch: the last stage {V(nh-2), V(nh-1)} reaches past the sequence's last
                        // key.  Zero those V rows (whole 128 B lines of every panel; the swizzle only
                        // permutes 16 B chunks inside a line) before the issuers feed them to P·V.
                        const int xl = x - 1;
                        const int ql = xl / kStages;
                        const int sl = xl - ql * kStages;
                        FA_WAIT(L::kBarVFixReq, vfix_phase, 24);  // the producer has issued this very fill
                        vfix_phase ^= 1u;
                        FA_WAIT(L::kBarKVFull + sl, phase0 ^ static_cast<uint32_t>(ql & 1), 19);
                        const uint32_t stage = smem_base + L::kKV + sl * (2 * kHalfBytes);
#pragma unroll
                        for (int half = 0; half < 2; ++half) {
                            int valid = w.seqlen_k - (2 * w.nmax - 2 + half) * 64;
                            valid = valid < 0 ? 0 : (valid > 64 ? 64 : valid);
                            const int chunks = (64 - valid) * 8;  // 16 B chunks per panel
                            for (int pn = 0; pn < kPanels; ++pn) {
                                const uint32_t base = stage + half * kHalfBytes + pn * kHalfPanel + valid * 128;
                                for (int c = r; c < chunks; c += 128)
                                    asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + c * 16), "r"(0) : "memory");
                            }
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar(L::kBarVFix));
                    }
                    const int q = x / kStages;
                    slot0 = x - q * kStages;
                    phase0 ^= static_cast<uint32_t>(q & 1);
                }
                // O / l -> T -> global (row-contiguous 16 B stores)
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if ((t ? w.pos0[1] : w.pos0[0]) >= w.seqlen_q) continue;
                    const int n = (t ? w.nblk[1] : w.nblk[0]);
                    const int g = r / P;
                    const int pos = (t ? w.pos0[1] : w.pos0[0]) + (r - g * P);
                    const bool row_valid = (g < params.pack_g) && (pos < w.seqlen_q);
                    const int head = packed ? w.hu * params.group + g : w.hu;
                    uint8_t* orow = reinterpret_cast<uint8_t*>(params.o) +
                                    2 * (static_cast<int64_t>(w.b) * params.o_stride_b +
                                         static_cast<int64_t>(head) * params.o_stride_h +
                                         static_cast<int64_t>(w.q_row0 + pos) * params.o_stride_s);
                    float inv = 0.f;
                    if (n > 0) {
                        FA_WAIT(L::kBarRowSum + t, rs_phase[t], 14);
                        rs_phase[t] ^= 1u;
                        const float l = ld_shared_f32(smem_base + L::kRowSum + (t * 128 + r) * 4);
                        inv = (l > 0.f) ? __frcp_rn(l) : 0.f;
                        FA_WAIT(L::kBarOFull + t, of_phase[t], 8);
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
                        if (c == kD / 32 - 1 && n > 0) {  // O_t and l fully read: the next work tile may overwrite them
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
        }();
    } else {
        // =========================== softmax warpgroups ===========================
        reg_alloc<C::kRegsSoftmax>();
        [&]() {
            Tracer<kTrace> tr;
            const int t = warp >> 2;          // M-tile handled by this warpgroup
            tr.init(params, t, (warp & 3) == 0 && lane == 0);
            const int r = threadIdx.x & 127;  // query row inside the M-tile == TMEM lane
            const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
            const uint32_t tmem_s = tmem_base + lane_base + t * 128;
            const uint32_t tmem_o = tmem_base + lane_base + 256 + t * kD;
            const float scale_log2 = params.scale_log2;
            const uint32_t rowsum_addr = smem_base + L::kRowSum + (t * 128 + r) * 4;
            const uint64_t sc2 = pack_f32x2(scale_log2, scale_log2);
            uint32_t blk_base = 0;  // KV blocks of earlier work tiles, mod 2 (S barrier parity)
            uint32_t oe_phase = 0;  // work tiles with n > 0 so far, mod 2 (row-sum slot hand-back)
            Tile w;
            TileConsumer<L> sched(smem_base, ctx);
            while (sched.next(w, lane == 0) > 0) {
                const int n = (t ? w.nblk[1] : w.nblk[0]);
                if (n == 0) continue;
                const int my_pos0 = (t ? w.pos0[1] : w.pos0[0]);
                const int g = r / P;
                const int pos = my_pos0 + (r - g * P);
                const int causal_off = w.seqlen_k - w.seqlen_q;
                int row_limit = w.seqlen_k - 1;  // last visible column
                if (params.causal) row_limit = min(row_limit, pos + causal_off);

                float m = -INFINITY;  // running row max (raw score units), possibly stale (lazy rescale)
                float l = 0.f;        // running row sum of exp2, relative to m
                float msc = 0.f;      // m * scale_log2 (0 while m == -inf)
                uint64_t nm2 = pack_f32x2(0.f, 0.f);  // (-msc, -msc)

                for (int h = 0; h < 2 * n; ++h) {
                    const int b = h & 1;
                    FA_WAIT(L::kBarSFull + 2 * t + b, (blk_base + (h >> 1)) & 1u, 7);
                    tc_fence_after_sync();
                    tr.ev(1);

                    uint32_t s[64];
                    tmem_ld32(tmem_s + b * 64, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
                    tmem_ld32(tmem_s + b * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
                    tmem_wait_ld();
                    tr.ev(2);

                    // exp2 of 32 keys -> 16 packed columns of P, stored straight back over S
                    auto exp_chunk = [&](int c, uint64_t& sa, uint64_t& sb) {
                        uint32_t pk[16];
#pragma unroll
                        for (int q = 0; q < 16; ++q) {
                            const uint64_t x2 = fma_f32x2(
                                pack_f32x2(__uint_as_float(s[c * 32 + 2 * q]), __uint_as_float(s[c * 32 + 2 * q + 1])),
                                sc2, nm2);
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
                        tmem_st16(tmem_s + b * 64 + c * 16, pk);
                    };
                    auto row_max = [&]() -> float {
                        float mx0 = max3(__uint_as_float(s[0]), __uint_as_float(s[1]), __uint_as_float(s[2]));
                        float mx1 = max3(__uint_as_float(s[3]), __uint_as_float(s[4]), __uint_as_float(s[5]));
                        float mx2 = max3(__uint_as_float(s[6]), __uint_as_float(s[7]), __uint_as_float(s[8]));
                        float mx3 = max3(__uint_as_float(s[9]), __uint_as_float(s[10]), __uint_as_float(s[11]));
#pragma unroll
                        for (int i = 12; i < 60; i += 8) {
                            mx0 = max3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
                            mx1 = max3(mx1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
                            mx2 = max3(mx2, __uint_as_float(s[i + 4]), __uint_as_float(s[i + 5]));
                            mx3 = max3(mx3, __uint_as_float(s[i + 6]), __uint_as_float(s[i + 7]));
                        }
                        mx0 = max3(mx0, __uint_as_float(s[60]), __uint_as_float(s[61]));
                        mx1 = max3(mx1, __uint_as_float(s[62]), __uint_as_float(s[63]));
                        return fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
                    };

                    // element mask only on half-blocks that cross the diagonal or the end of K
                    const int col0 = h * 64;
                    const bool need_mask = (col0 + 64 > w.seqlen_k) ||
                                           (params.causal && (col0 + 63 > my_pos0 + causal_off));
                    uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);

                    // Fast path (all but the first and the masked half-blocks): the first 32 keys are
                    // exponentiated against the stale running max while the row max of the half-block
                    // is reduced in the shadow of those MUFU ops; the result stands unless some row of
                    // this warp grew past the lazy-rescale threshold (then the half-block is redone —
                    // the scores are still in registers, P chunk 0 is overwritten, nothing is published).
                    bool fast = false;
                    if (kSpec && h > 0 && !need_mask) {
                        uint64_t sa = pack_f32x2(0.f, 0.f), sb = pack_f32x2(0.f, 0.f);
                        exp_chunk(0, sa, sb);
                        const float mx = row_max();
                        fast = !__any_sync(0xffffffffu, (mx - m) * scale_log2 > kRescaleThreshold);
                        if (fast) { la = sa; lb = sb; }
                    }
                    if (!fast) {
                        if (need_mask) {
                            const int rel = row_limit - col0;
#pragma unroll
                            for (int i = 0; i < 64; ++i)
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
                        if (h > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
                            // O_t must be quiescent: P·V(h-1) complete (P·V(h) cannot have been issued,
                            // P(h) is not published yet).
                            tr.ev(7);
                            if (h + 1 < 2 * n)  // S(h+1) was committed behind P·V(h-1): wait without consuming
                                FA_WAIT(L::kBarSFull + 2 * t + (b ^ 1), (blk_base + ((h + 1) >> 1)) & 1u, 15);
                            else                // last half-block: P·V(2n-2) is the first of this tile's two PVDone commits
                                FA_WAIT(L::kBarPVDone + t, 0u, 15);
                            tc_fence_after_sync();
#pragma unroll 1
                            for (int c = 0; c < kD / 8; ++c) {
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
                        exp_chunk(0, la, lb);
                    }
                    tr.ev(3);
                    exp_chunk(1, la, lb);
                    tmem_wait_st();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(L::kBarPFull + 2 * t + b));
                    tr.ev(4);
                    float l0, l1;
                    unpack_f32x2(add_f32x2(la, lb), l0, l1);
                    l += l0 + l1;
                }
                blk_base += static_cast<uint32_t>(n);
                // The row-sum slot (and its barrier) is handed back by the epilogue together with O_t:
                // a one-block tile (S(0), S(1) need no O_t) can finish its softmax before the epilogue
                // has read the previous tile's sums — without this wait the sums are overwritten and
                // the RowSum barrier runs two phases ahead of its waiter (deadlock).
                FA_WAIT(L::kBarOEmpty + t, oe_phase ^ 1u, 17);
                oe_phase ^= 1u;
                st_shared_f32(rowsum_addr, l);
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(L::kBarRowSum + t));
            }
        }();
    }
#undef FA_WAIT

    // ---- teardown ----
    tc_fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        // An aborted launch may still have TMA / MMA work in flight towards this CTA's memories.
        if (*reinterpret_cast<volatile unsigned int*>(ctx.status) != 0u) __nanosleep(200000);
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

}  // namespace prefill
}  // namespace fa100
