#pragma once
#include "fa_sm100_fwd.cuh"

namespace fa100 {
namespace v4 {

constexpr int kBlockM = 128;
constexpr int kBlockN = 128;
constexpr int kPanelBytes = 128 * 128;  // 128 rows x 64 16-bit elements
// Role layout (512 threads).  setmaxnreg budget: the CTA is launched with 64K / 512 = 128 registers
// per thread; the sum over warps of the per-role counts must not exceed that pool or
// setmaxnreg.inc blocks forever.
struct Cfg {
    static constexpr int kSoftmaxWarps = 8;
    static constexpr int kCorrWarp0 = 8;
    static constexpr int kMmaWarp = 12;
    static constexpr int kLoadWarp = 13;
    static constexpr int kNumWarps = 16;
    static constexpr int kNumThreads = kNumWarps * 32;
    static constexpr int kRegsLaunch = 128;
    static constexpr int kRegsSoftmax = 184;
    static constexpr int kRegsCorr = 80;
    static constexpr int kRegsOther = 64;
    static_assert(kSoftmaxWarps * kRegsSoftmax + 4 * kRegsCorr + 4 * kRegsOther <= kNumWarps * kRegsLaunch,
                  "register budget");
};
constexpr float kRescaleThreshold = 8.0f;  // lazy rescale: only when the max grows by > 2^8

template <int kD, int kStages>
struct Smem {
    static constexpr int kPanels = kD / 64;
    static constexpr int kTileBytes = kPanels * kPanelBytes;  // a Q M-tile or one K / V block
    static constexpr int kQ0 = 0;
    static constexpr int kQ1 = kTileBytes;
    static constexpr int kKV = 2 * kTileBytes;
    static constexpr int kScale = kKV + kStages * kTileBytes;  // float [2][128] rescale factors
    static constexpr int kRowSum = kScale + 2 * 128 * 4;       // float [2][128] final row sums
    static constexpr int kTileRing = kRowSum + 2 * 128 * 4;    // int [kSchedDepth]
    static constexpr int kBars = kTileRing + 16;
    // barrier slots (8 B each)
    static constexpr int kBarQFull = 0;                  // [2]  TMA -> MMA
    static constexpr int kBarQEmpty = 2;                 // [2]  MMA commit -> TMA
    static constexpr int kBarKVFull = 4;                 // [kStages]
    static constexpr int kBarKVEmpty = 4 + kStages;      // [kStages]
    static constexpr int kBarSFull = 4 + 2 * kStages;    // [2]  MMA commit -> softmax
    static constexpr int kBarP1Full = 6 + 2 * kStages;   // [2]  softmax (4 warps) + correction (4 warps) -> MMA
    static constexpr int kBarP2Full = 8 + 2 * kStages;   // [2]  softmax (4 warps) -> MMA
    static constexpr int kBarScale = 10 + 2 * kStages;   // [2]  softmax (4 warps) -> correction
    static constexpr int kBarRowSum = 12 + 2 * kStages;  // [2]  softmax (4 warps) -> correction
    static constexpr int kBarOFull = 14 + 2 * kStages;   // [2]  MMA commit -> correction
    static constexpr int kBarOEmpty = 16 + 2 * kStages;  // [2]  correction (4 warps) -> MMA
    static constexpr int kBarTileFull = 18 + 2 * kStages;   // [kSchedDepth] producer -> everyone
    static constexpr int kBarTileEmpty = 22 + 2 * kStages;  // [kSchedDepth] consumer warps -> producer
    static constexpr int kNumBars = 26 + 2 * kStages;
    static constexpr int kTmemPtr = kBars + kNumBars * 8;
    static constexpr int kTotal = kTmemPtr + 16;
    // No slack: the kernel has no static shared memory, so the dynamic window starts 1024 B aligned
    // (checked at run time: a misaligned base aborts with status code 0x80000063).
    static constexpr int kDynamicBytes = kTotal;
};

// Number of 128-wide KV blocks an M-tile starting at query position `tile_pos0` has to visit.
__device__ __forceinline__ int tile_num_blocks(const FwdParams& p, int tile_pos0) {
    if (tile_pos0 >= p.seqlen_q) return 0;
    int kv_end = p.seqlen_k;
    if (p.causal) {
        int maxpos = min(tile_pos0 + p.rows_per_head, p.seqlen_q) - 1;
        kv_end = min(kv_end, maxpos + (p.seqlen_k - p.seqlen_q) + 1);
    }
    return kv_end <= 0 ? 0 : (kv_end + kBlockN - 1) / kBlockN;
}

// One CTA's view of a work tile.
struct Tile {
    int mblock, hu, b;   // m-block, head unit (q head or kv head when packed), batch
    int kv_head, g_coord;
    int pos0[2];         // first query position of each M-tile
    int nblk[2];         // KV blocks each M-tile visits
    int nmax;
};

// Work-tile order.  Linear index L -> (section, m-rank, head): sections of `section_heads`
// consecutive (batch, head) units whose K/V fit in L2 together; inside a section the m-blocks go
// heaviest (last) first and the heads vary fastest, so CTAs running at the same time read the
// same few K/V heads.
__device__ __forceinline__ void decode_tile(const FwdParams& p, int L, Tile& t) {
    const int nm = p.num_m_blocks;
    const int sec_tiles = p.section_heads * nm;
    const int sec = L / sec_tiles;
    const int rem = L - sec * sec_tiles;
    const int h0 = sec * p.section_heads;
    const int nh = min(p.section_heads, p.total_heads - h0);
    const int mrank = rem / nh;
    const int hflat = h0 + (rem - mrank * nh);
    t.mblock = nm - 1 - mrank;
    t.hu = hflat % p.head_units;
    t.b = hflat / p.head_units;
    const bool packed = p.pack_g > 1;
    t.kv_head = packed ? t.hu : t.hu / p.group;
    t.g_coord = packed ? 0 : t.hu % p.group;
    const int P = p.rows_per_head;
    t.pos0[0] = t.mblock * 2 * P;
    t.pos0[1] = t.pos0[0] + P;
    t.nblk[0] = tile_num_blocks(p, t.pos0[0]);
    t.nblk[1] = tile_num_blocks(p, t.pos0[1]);
    t.nmax = max(t.nblk[0], t.nblk[1]);
}

// Dynamic tile scheduler.  The persistent CTAs pull tiles in the order above (longest processing
// time first) from one global counter: CTA c starts with tile c, every further tile is
// gridDim.x + atomicAdd(counter, 1).  The TMA producer thread is the only one that touches the
// counter (it prefetches the next index while it issues the current tile's loads); it hands each
// index to the other warps through a kSchedDepth-deep shared-memory ring guarded by mbarriers
// (TileFull: 1 arrival; TileEmpty: one arrival per consumer warp).  -1 ends the CTA.
constexpr int kSchedDepth = 4;
template <typename L>
struct TileConsumer {
    uint32_t smem_base;
    int round = 0;
    unsigned int* status;
    __device__ __forceinline__ TileConsumer(uint32_t base, unsigned int* st) : smem_base(base), status(st) {}
    // all lanes of a consumer warp call this together (or the single active thread of a role)
    __device__ __forceinline__ bool next(const FwdParams& p, Tile& t, bool elected) {
        const int slot = round % kSchedDepth;
        const uint32_t parity = (round / kSchedDepth) & 1;
        mbar_wait(smem_base + L::kBars + (L::kBarTileFull + slot) * 8, parity, status, 20);
        int idx;
        asm volatile("ld.shared.s32 %0, [%1];" : "=r"(idx) : "r"(smem_base + L::kTileRing + slot * 4) : "memory");
        __syncwarp(__activemask());
        if (elected) mbar_arrive(smem_base + L::kBars + (L::kBarTileEmpty + slot) * 8);
        ++round;
        if (idx < 0) return false;
        decode_tile(p, idx, t);
        return true;
    }
};

// Debug timeline: CTA 0 records (event code << 56 | clock64) per role; see tools/trace_view.py.
constexpr int kTraceRoles = 5;
constexpr int kTraceMax = 4096;
template <bool kTrace>
struct Tracer {
    long long* buf;
    int n;
    __device__ __forceinline__ void init(const FwdParams& p, int role, bool on) {
        if constexpr (kTrace) {
            buf = (on && p.trace && blockIdx.x == 0) ? p.trace + role * kTraceMax : nullptr;
            n = 0;
        }
    }
    __device__ __forceinline__ void ev(int code) {
        if constexpr (kTrace) {
            if (buf && n < kTraceMax) buf[n++] = (static_cast<long long>(code) << 56) | (clock64_now() & 0xffffffffffffffll);
        }
    }
};

// kEmu: of every 8 element pairs, how many take the FMA-pipe exp2 (0 = all MUFU).
// kSplit: keys whose P is handed to the MMA warp first (64, 96 or 128 = no split).
template <int kD, bool kBf16, int kStages, int kEmu, int kSplit, bool kTrace>
__global__ void __launch_bounds__(Cfg::kNumThreads, 1)
fa_fwd_sm100_v4(const __grid_constant__ FwdParams params) {
    using L = Smem<kD, kStages>;
    using C = Cfg;
    constexpr int kCorrWarp0 = C::kCorrWarp0, kMmaWarp = C::kMmaWarp, kLoadWarp = C::kLoadWarp;
    constexpr int kRegsSoftmax = C::kRegsSoftmax, kRegsCorr = C::kRegsCorr, kRegsOther = C::kRegsOther;
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kTileBytes = L::kTileBytes;
    static_assert(kSplit == 64 || kSplit == 96 || kSplit == 128, "kSplit");
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
            mbar_init(bar(L::kBarP1Full + t), 4 + 4);
            mbar_init(bar(L::kBarP2Full + t), 4);
            mbar_init(bar(L::kBarScale + t), 4);
            mbar_init(bar(L::kBarRowSum + t), 4);
            mbar_init(bar(L::kBarOFull + t), 1);
            mbar_init(bar(L::kBarOEmpty + t), 4);
        }
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(L::kBarKVFull + s), 1);
            mbar_init(bar(L::kBarKVEmpty + s), 1);
        }
        for (int s = 0; s < kSchedDepth; ++s) {
            mbar_init(bar(L::kBarTileFull + s), 1);
            mbar_init(bar(L::kBarTileEmpty + s), C::kSoftmaxWarps + 4 + 1);  // softmax + correction + MMA
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
        reg_dealloc<kRegsOther>();
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
        } else if (warp == kMmaWarp && elected) {
            // =========================== MMA issuer ===========================
            tr.init(params, 3, true);
            constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 128, false, false);
            constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
            const uint32_t tmem_s[2] = {tmem_base, tmem_base + 128};
            const uint32_t tmem_o[2] = {tmem_base + 256, tmem_base + 256 + kD};
            // UMMA smem descriptors are rebuilt per instruction from a 32-bit base (start address
            // field) plus a compile-time offset: keeps the issuing thread inside its 64 registers.
            constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO, version, SW128
            constexpr uint32_t kLboK = (16u >> 4) << 16;           // Q, K: K-major panels
            constexpr uint32_t kLboV = (uint32_t(kPanelBytes) >> 4) << 16;  // V: MN-major panels
            auto desc = [](uint32_t lo) -> uint64_t {
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(kDescHi));
                return d;
            };
            const uint32_t q_lo[2] = {((smem_base + L::kQ0) >> 4) | kLboK, ((smem_base + L::kQ1) >> 4) | kLboK};
            const uint32_t kv_lo = (smem_base + L::kKV) >> 4;  // + slot * (kTileBytes >> 4)

            auto issue_s = [&](int t, int kslot_) {
                const uint32_t k_lo = (kv_lo + kslot_ * (kTileBytes >> 4)) | kLboK;
#pragma unroll
                for (int ks = 0; ks < kD / 16; ++ks) {
                    const uint32_t off = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
                    umma_ss(tmem_s[t], desc(q_lo[t] + off), desc(k_lo + off), idesc_qk, ks > 0);
                }
            };
            // P column of k-step ks (16 keys = 8 packed columns)
            auto p_col = [](int ks) -> uint32_t { return uint32_t(ks * 8); };
            auto issue_pv = [&](int t, int vslot_, bool acc, int ks0, int ks1) {
                const uint32_t v_lo = (kv_lo + vslot_ * (kTileBytes >> 4)) | kLboV;
#pragma unroll
                for (int ks = ks0; ks < ks1; ++ks) {
                    umma_ts(tmem_o[t], tmem_s[t] + p_col(ks), desc(v_lo + ks * (2048 >> 4)), idesc_pv,
                            (acc || ks > 0) ? 1u : 0u);
                }
            };

            // Ring position of the next K/V load (the producer issues K0 V0 K1 V1 ... per work tile).
            int slot0 = 0;
            uint32_t phase0 = 0;
            uint32_t q_phase[2] = {0, 0}, p_phase[2] = {0, 0}, oe_phase[2] = {0, 0};
            Tile w;
            TileConsumer<L> sched(smem_base, status);
            while (sched.next(params, w, true)) {
                if (w.nmax == 0) continue;
                auto ring = [&](int it, int& s_, uint32_t& ph_) {
                    const int x = slot0 + it;
                    const int q = x / kStages;
                    s_ = x - q * kStages;
                    ph_ = phase0 ^ static_cast<uint32_t>(q & 1);
                };
                // Blocks whose S (sdone) / P·V (pdone) MMAs have been issued, per M-tile.  A K/V slot is
                // released by whichever tile uses it last (tcgen05.commit covers both tiles' MMAs).
                int sdone[2] = {0, 0}, pdone[2] = {0, 0};
                bool part_a[2] = {false, false};
                // prologue: S_t(0) = Q_t · K_0ᵀ
                {
                    int ks; uint32_t kph;
                    ring(0, ks, kph);
                    mbar_wait(bar(L::kBarKVFull + ks), kph, status, 2);
                    tc_fence_after_sync();
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        if (w.nblk[t] > 0) {
                            mbar_wait(bar(L::kBarQFull + t), q_phase[t], status, 3);
                            q_phase[t] ^= 1u;
                            tc_fence_after_sync();
                            issue_s(t, ks);
                            tc_commit(bar(L::kBarSFull + t));
                            if (w.nblk[t] == 1) tc_commit(bar(L::kBarQEmpty + t));
                            sdone[t] = 1;
                        }
                    }
                    tc_commit(bar(L::kBarKVEmpty + ks));
                    tr.ev(2);
                }
                // Main loop: poll both M-tiles and issue whatever has its P ready, so that neither
                // tile's hand-off (in particular the second part of a split P) blocks the other.
                int remaining = w.nblk[0] + w.nblk[1];
                uint32_t idle = 0;
                while (remaining > 0) {
                    bool progressed = false;
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                    if (pdone[t] < w.nblk[t]) {
                        const int o = t ^ 1;
                        const int j = pdone[t];
                        int vs; uint32_t vph;
                        ring(2 * j + 1, vs, vph);
                        if (!part_a[t] && mbar_test_wait(bar(L::kBarP1Full + t), p_phase[t])) {
                            if (j == 0) {  // O_t of the previous work tile must have been drained
                                mbar_wait(bar(L::kBarOEmpty + t), oe_phase[t] ^ 1u, status, 11);
                                oe_phase[t] ^= 1u;
                            }
                            mbar_wait(bar(L::kBarKVFull + vs), vph, status, 4);
                            tc_fence_after_sync();
                            tr.ev(3 + t);
                            issue_pv(t, vs, j > 0, 0, kSplit / 16);
                            part_a[t] = true;
                            progressed = true;
                        }
                        if (part_a[t] && (kSplit == 128 || mbar_test_wait(bar(L::kBarP2Full + t), p_phase[t]))) {
                            if constexpr (kSplit < 128) {
                                tc_fence_after_sync();
                                issue_pv(t, vs, true, kSplit / 16, 8);
                            }
                            p_phase[t] ^= 1u;
                            part_a[t] = false;
                            progressed = true;
                            pdone[t] = j + 1;
                            --remaining;
                            if (j >= w.nblk[o] || pdone[o] > j) tc_commit(bar(L::kBarKVEmpty + vs));
                            if (j + 1 < w.nblk[t]) {
                                int ks; uint32_t kph;
                                ring(2 * j + 2, ks, kph);
                                mbar_wait(bar(L::kBarKVFull + ks), kph, status, 6);
                                tc_fence_after_sync();
                                issue_s(t, ks);
                                tc_commit(bar(L::kBarSFull + t));
                                if (j + 2 == w.nblk[t]) tc_commit(bar(L::kBarQEmpty + t));
                                sdone[t] = j + 2;
                                if (j + 1 >= w.nblk[o] || sdone[o] > j + 1) tc_commit(bar(L::kBarKVEmpty + ks));
                            } else {
                                tc_commit(bar(L::kBarOFull + t));
                            }
                            tr.ev(5 + t);
                        }
                    }
                    }
                    if (progressed) {
                        idle = 0;
                    } else if ((++idle & 0xfffu) == 0 &&
                               *reinterpret_cast<volatile unsigned int*>(status) != 0u) {
                        remaining = 0;  // another role's watchdog fired: fall out
                    }
                }
                {
                    const int x = slot0 + 2 * w.nmax;
                    const int q = x / kStages;
                    slot0 = x - q * kStages;
                    phase0 ^= static_cast<uint32_t>(q & 1);
                }
            }
        }
    } else if (warp >= kCorrWarp0) {
        // =========================== correction warpgroup ===========================
        reg_dealloc<kRegsCorr>();
        tr.init(params, 2, warp == kCorrWarp0 && lane == 0);
        const int r = threadIdx.x & 127;  // row inside an M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t sc_phase[2] = {0, 0}, rs_phase[2] = {0, 0}, of_phase[2] = {0, 0};
        Tile w;
        TileConsumer<L> sched(smem_base, status);
        while (sched.next(params, w, lane == 0)) {
            for (int j = 0; j < w.nmax; ++j) {
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (j < w.nblk[t]) {
                        mbar_wait(bar(L::kBarScale + t), sc_phase[t], status, 13);
                        sc_phase[t] ^= 1u;
                        if (j > 0) {
                            const float alpha = ld_shared_f32(smem_base + L::kScale + (t * 128 + r) * 4);
                            if (__any_sync(0xffffffffu, alpha != 1.f)) {
                                tr.ev(7);
                                tc_fence_after_sync();
                                const uint32_t tmem_o = tmem_base + lane_base + 256 + t * kD;
#pragma unroll
                                for (int c = 0; c < kD / 32; ++c) {
                                    uint32_t o[32];
                                    tmem_ld32(tmem_o + c * 32, o);
                                    tmem_wait_ld();
#pragma unroll
                                    for (int i = 0; i < 32; ++i)
                                        o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                                    tmem_st32(tmem_o + c * 32, o);
                                }
                                tmem_wait_st();
                                tc_fence_before_sync();
                            }
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar(L::kBarP1Full + t));
                    }
                }
            }
            // ---- epilogue: O / l -> T -> global (row-contiguous 16 B stores) ----
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
        reg_alloc<kRegsSoftmax>();
        const int t = warp >> 2;          // M-tile handled by this warpgroup
        tr.init(params, t, (warp & 3) == 0 && lane == 0);
        const int r = threadIdx.x & 127;  // query row inside the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t tmem_s = tmem_base + lane_base + t * 128;
        const float scale_log2 = params.scale_log2;
        const int causal_off = params.seqlen_k - params.seqlen_q;
        const uint32_t scale_addr = smem_base + L::kScale + (t * 128 + r) * 4;
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

                // element mask only on blocks that cross the diagonal or the end of K
                const int col0 = j * kBlockN;
                const bool need_mask = (col0 + kBlockN > params.seqlen_k) ||
                                       (params.causal && (col0 + kBlockN - 1 > my_pos0 + causal_off));
                uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);
                float alpha = 1.f;

                // Fast path (every block but the first and the masked ones): the first 32 keys are
                // exponentiated against the stale running max while the row max of the whole block
                // is reduced in the shadow of those MUFU ops; the result stands unless some row of
                // this warp grew past the lazy-rescale threshold (then the block is redone below —
                // the scores are still in registers, P chunk 0 is simply overwritten).
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
                    if ((m_new - m) * scale_log2 > kRescaleThreshold) {  // false for NaN (-inf - -inf)
                        alpha = ex2_approx((m - m_new) * scale_log2);
                        m = m_new;
                        msc = m * scale_log2;
                        nm2 = pack_f32x2(-msc, -msc);
                    }
                }
                st_shared_f32(scale_addr, alpha);
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(L::kBarScale + t));
                tr.ev(3);

                auto publish = [&](int slot) {
                    tmem_wait_st();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(slot + t));
                };
                if (!fast) exp_chunk(&s[0], 0, la, lb);
#pragma unroll
                for (int c = 1; c < kSplit / 32; ++c) exp_chunk(&s[c * 32], c, la, lb);
                publish(L::kBarP1Full);
                tr.ev(4);
                if constexpr (kSplit < 128) {
#pragma unroll
                    for (int c = kSplit / 32; c < 4; ++c) exp_chunk(&s[c * 32], c, la, lb);
                    publish(L::kBarP2Full);
                    tr.ev(5);
                }
                float l0, l1;
                unpack_f32x2(add_f32x2(la, lb), l0, l1);
                l = fmaf(l, alpha, l0 + l1);
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

}  // namespace v4
}  // namespace fa100
