// FlashAttention-2 forward for sm_100a, generation 2: persistent, warp-specialised
// tcgen05 / TMEM / TMA kernel with a correction warpgroup, split-P hand-off and FMA-pipe exp2.
//
// Replaces the reference's flash_attention_v2 (csrc/flash_attention_template.cuh:138-564),
// Mask (csrc/mask.cuh:30-88) and launch_flash_attention (template.cuh:567-583).  Same math
// (fp32 S and O accumulators, exp2-domain online softmax, P rounded to T before P·V, row sums
// from unrounded P, bottom-right causal alignment, kv-head = q-head / G); different machine:
//
//   grid = min(#SMs, #work tiles) persistent CTAs of 512 threads; a work tile = 2 M-tiles of 128
//   query rows sharing one K/V stream.  Work tiles are ordered for L2 (sections of heads whose
//   K/V fit in L2) and LPT (heaviest causal tiles first inside a section) and dealt to the CTAs
//   in snake order, so every CTA gets the same amount of work without any atomics.
//
//     warps 0-3   softmax warpgroup, M-tile 0 (one thread per query row, no shuffles)
//     warps 4-7   softmax warpgroup, M-tile 1
//     warps 8-11  correction warpgroup: lazy O rescale in TMEM, final 1/l, O -> global
//     warp  12    MMA issuer (one thread): S_t = Q_t·K_jᵀ (SS), O_t += P_t·V_j (TS)
//     warp  13    TMA producer: Q tiles per work tile, K_j / V_j through an mbarrier ring
//     warps 14-15 spare (their registers go to the softmax warpgroups via setmaxnreg)
//
//   TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [256+D,256+2D);
//     P_t (16-bit, two per column) aliases the first 64 columns of S_t.
//   SMEM: Q0, Q1, kStages x (K or V) tiles, all 128B-swizzled [rows][64] panels as written by
//     TMA; K-major UMMA descriptors for Q and K, MN-major for V (no transpose anywhere).
//
// Per KV block j and M-tile t the dependency chain is
//   SFull[t] -> softmax: ld S, row max, (scale -> correction), exp2, P part A -> P1Full[t]
//            -> P part B -> P2Full[t];  MMA: PV part A | PV part B | S_t(j+1) -> SFull[t]
// and the two M-tiles run it half a period apart so the tensor pipe always has the other tile's
// MMAs to do.  The exp2 of a fraction of the elements runs on the FMA pipe (ex2_emu_x2).
//
// An M-tile's 128 rows are (head-in-group g, position p) pairs: row r -> g = r / P, p = r % P,
// with P = rows_per_head.  Unpacked: P = 128, one q head.  GQA-packed: P = 128 / G and the G
// q heads of one kv head share the tile (the Q TMA box is (64, P, G) of a 5-D tensor map).
#pragma once
#include "fa_sm100_ptx.cuh"

namespace fa100 {

struct FwdParams {
    CUtensorMap tm_q;  // 5-D (d, s, g, hkv, b); box (64, rows_per_head, pack_g, 1, 1)
    CUtensorMap tm_k;  // 4-D (d, s, hkv, b);    box (64, 128, 1, 1)
    CUtensorMap tm_v;  // 4-D (d, s, hkv, b);    box (64, 128, 1, 1)
    void* o;
    int64_t o_stride_b, o_stride_h, o_stride_s;  // elements
    int32_t batch, heads_q, heads_kv, group;     // group = heads_q / heads_kv
    int32_t seqlen_q, seqlen_k, headdim;
    int32_t pack_g;         // q heads packed into one M tile (1 = unpacked, else == group)
    int32_t rows_per_head;  // P
    int32_t num_m_blocks;   // ceil(seqlen_q / (2 P))
    int32_t head_units;     // pack ? heads_kv : heads_q
    int32_t total_heads;    // head_units * batch
    int32_t section_heads;  // heads per L2 section (tile order)
    int32_t num_tiles;      // num_m_blocks * total_heads
    int32_t q_tile_bytes;   // bytes TMA delivers for one Q M-tile (all panels)
    int32_t causal;
    float scale_log2;       // softmax_scale * log2(e)
    unsigned int* status;   // [0] abort flag/code, [1..3] detail
    long long* trace;       // debug timeline (kTrace builds only), else nullptr
    int* sched;             // [0] next-tile counter, [1] finished-CTA counter (both zero between launches)
};

namespace v2 {

constexpr int kBlockM = 128;
constexpr int kBlockN = 128;
constexpr int kPanelBytes = 128 * 128;  // 128 rows x 64 16-bit elements
// Role layout.  kHalves = threads per query row in the softmax: 1 = one thread owns a whole
// 128-key row of S (2 softmax warpgroups, 512 threads); 2 = two threads own 64 keys each
// (4 softmax warpgroups, 768 threads) and exchange their row maxima through shared memory —
// half the serial softmax latency per KV block, which is what bounds the S -> P -> O chain.
//   setmaxnreg budget: the CTA is launched with 64K / kNumThreads registers per thread; the sum
//   over warps of the per-role counts must not exceed that pool or setmaxnreg.inc blocks forever.
template <int kHalves>
struct Cfg {
    static constexpr int kSoftmaxWarps = 8 * kHalves;
    static constexpr int kCorrWarp0 = kSoftmaxWarps;
    static constexpr int kMmaWarp = kCorrWarp0 + 4;
    static constexpr int kLoadWarp = kMmaWarp + 1;
    static constexpr int kNumWarps = kSoftmaxWarps + 8;
    static constexpr int kNumThreads = kNumWarps * 32;
    static constexpr int kRegsLaunch = kHalves == 1 ? 128 : 80;  // ptxas cap for kNumThreads
    static constexpr int kRegsSoftmax = kHalves == 1 ? 184 : 88;
    static constexpr int kRegsCorr = kHalves == 1 ? 80 : 64;
    static constexpr int kRegsOther = 64;
    static_assert(kSoftmaxWarps * kRegsSoftmax + 4 * kRegsCorr + 4 * kRegsOther <= kNumWarps * kRegsLaunch,
                  "register budget");
};
constexpr float kRescaleThreshold = 8.0f;  // lazy rescale: only when the max grows by > 2^8

template <int kD, int kStages, int kHalves>
struct Smem {
    static constexpr int kPanels = kD / 64;
    static constexpr int kTileBytes = kPanels * kPanelBytes;  // a Q M-tile or one K / V block
    static constexpr int kQ0 = 0;
    static constexpr int kQ1 = kTileBytes;
    static constexpr int kKV = 2 * kTileBytes;
    static constexpr int kScale = kKV + kStages * kTileBytes;  // float [2][128] rescale factors
    static constexpr int kRowSum = kScale + 2 * 128 * 4;       // float [2][kHalves][128] final row sums
    static constexpr int kXchg = kRowSum + 2 * kHalves * 128 * 4;  // float [2][2 parity][2][128] row-max exchange
    static constexpr int kTileRing = kXchg + (kHalves == 2 ? 2 * 2 * 2 * 128 * 4 : 0);  // int [kSchedDepth]
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

// kEmu: of every 4 element pairs, how many take the FMA-pipe exp2 (0 = all MUFU).
// kSplit: keys whose P is handed to the MMA warp first (64, 96 or 128 = no split).
// kHalves: threads per softmax row (see Cfg).  With kHalves == 2 the split is implicit (each half
// publishes its own 64 keys) and kSplit must be 128.
template <int kD, bool kBf16, int kStages, int kEmu, int kSplit, int kHalves, bool kTrace>
__global__ void __launch_bounds__(Cfg<kHalves>::kNumThreads, 1)
fa_fwd_sm100_v2(const __grid_constant__ FwdParams params) {
    using L = Smem<kD, kStages, kHalves>;
    using C = Cfg<kHalves>;
    constexpr int kCorrWarp0 = C::kCorrWarp0, kMmaWarp = C::kMmaWarp, kLoadWarp = C::kLoadWarp;
    constexpr int kRegsSoftmax = C::kRegsSoftmax, kRegsCorr = C::kRegsCorr, kRegsOther = C::kRegsOther;
    static_assert(kHalves == 1 || kSplit == 128, "kHalves == 2 implies no explicit P split");
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kTileBytes = L::kTileBytes;
    static_assert(kSplit == 64 || kSplit == 96 || kSplit == 128, "kSplit");

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
            mbar_init(bar(L::kBarP1Full + t), 4 * kHalves + 4);
            mbar_init(bar(L::kBarP2Full + t), 4);
            mbar_init(bar(L::kBarScale + t), 4);
            mbar_init(bar(L::kBarRowSum + t), 4 * kHalves);
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
        if (warp == kLoadWarp && lane == 0) {
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
        } else if (warp == kMmaWarp && lane == 0) {
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
            // P column of k-step ks (16 keys = 8 packed columns).  kHalves == 2: each half writes
            // its P over the first half of its own S columns, so keys 64.. start at column 64.
            auto p_col = [](int ks) -> uint32_t {
                return kHalves == 2 ? uint32_t((ks >> 2) * 64 + (ks & 3) * 8) : uint32_t(ks * 8);
            };
            auto issue_pv = [&](int t, int vslot_, bool acc, int ks0, int ks1) {
                const uint32_t v_lo = (kv_lo + vslot_ * (kTileBytes >> 4)) | kLboV;
#pragma unroll
                for (int ks = ks0; ks < ks1; ++ks) {
                    umma_ts(tmem_o[t], tmem_s[t] + p_col(ks), desc(v_lo + ks * (2048 >> 4)), idesc_pv,
                            (acc || ks > 0) ? 1u : 0u);
                }
            };

            int slot = 0;
            uint32_t phase = 0;
            auto advance = [&]() { if (++slot == kStages) { slot = 0; phase ^= 1u; } };
            uint32_t q_phase[2] = {0, 0}, p_phase[2] = {0, 0}, oe_phase[2] = {0, 0};
            Tile w;
            TileConsumer<L> sched(smem_base, status);
            while (sched.next(params, w, true)) {
                if (w.nmax == 0) continue;
                // prologue: S_t(0) = Q_t · K_0ᵀ
                int kslot = slot; uint32_t kphase = phase; advance();
                mbar_wait(bar(L::kBarKVFull + kslot), kphase, status, 2);
                tc_fence_after_sync();
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (w.nblk[t] > 0) {
                        mbar_wait(bar(L::kBarQFull + t), q_phase[t], status, 3);
                        q_phase[t] ^= 1u;
                        tc_fence_after_sync();
                        issue_s(t, kslot);
                        tc_commit(bar(L::kBarSFull + t));
                        if (w.nblk[t] == 1) tc_commit(bar(L::kBarQEmpty + t));
                    }
                }
                tc_commit(bar(L::kBarKVEmpty + kslot));
                tr.ev(2);

                for (int j = 0; j < w.nmax; ++j) {
                    const int vslot = slot; const uint32_t vphase = phase; advance();
                    const bool has_next = (j + 1 < w.nmax);
                    if (has_next) { kslot = slot; kphase = phase; advance(); }
                    mbar_wait(bar(L::kBarKVFull + vslot), vphase, status, 4);
                    bool k_ready = false;
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        if (j < w.nblk[t]) {
                            if (j == 0) {  // O_t of the previous work tile must have been drained
                                mbar_wait(bar(L::kBarOEmpty + t), oe_phase[t] ^ 1u, status, 11);
                                oe_phase[t] ^= 1u;
                            }
                            mbar_wait(bar(L::kBarP1Full + t), p_phase[t], status, 5);
                            tc_fence_after_sync();
                            tr.ev(3 + t);
                            issue_pv(t, vslot, j > 0, 0, kSplit / 16);
                            if constexpr (kSplit < 128) {
                                mbar_wait(bar(L::kBarP2Full + t), p_phase[t], status, 12);
                                tc_fence_after_sync();
                                issue_pv(t, vslot, true, kSplit / 16, 8);
                            }
                            p_phase[t] ^= 1u;
                            if (j + 1 < w.nblk[t]) {
                                if (!k_ready) {
                                    mbar_wait(bar(L::kBarKVFull + kslot), kphase, status, 6);
                                    tc_fence_after_sync();
                                    k_ready = true;
                                }
                                issue_s(t, kslot);
                                tc_commit(bar(L::kBarSFull + t));
                                if (j + 2 == w.nblk[t]) tc_commit(bar(L::kBarQEmpty + t));
                            } else {
                                tc_commit(bar(L::kBarOFull + t));
                            }
                            tr.ev(5 + t);
                        }
                    }
                    tc_commit(bar(L::kBarKVEmpty + vslot));
                    if (has_next) tc_commit(bar(L::kBarKVEmpty + kslot));
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
                    float l = ld_shared_f32(smem_base + L::kRowSum + ((t * kHalves) * 128 + r) * 4);
                    if constexpr (kHalves == 2)
                        l += ld_shared_f32(smem_base + L::kRowSum + ((t * kHalves + 1) * 128 + r) * 4);
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
    } else if (kHalves == 2) {
        // ============ softmax, two threads per query row (4 warpgroups) ============
        // warp = t*8 + h*4 + q: M-tile t, key half h (keys [64h, 64h+64) of every KV block), TMEM
        // lane quarter q.  The two warps of a (t, q) pair see the same 32 rows; they exchange their
        // partial row maxima through shared memory behind a 64-thread named barrier and keep
        // separate partial row sums (added by the correction warpgroup at the end).  S is read from
        // TMEM twice (max pass, exp pass) so a thread never holds more than 32 scores; each half
        // writes its P over the first 32 columns of its own 64 S columns.
        reg_alloc<kRegsSoftmax>();
        const int t = warp >> 3;
        const int h = (warp >> 2) & 1;
        const int q4 = warp & 3;
        tr.init(params, t, h == 0 && q4 == 0 && lane == 0);
        const int r = q4 * 32 + lane;     // query row inside the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>(q4 * 32) << 16;
        const uint32_t tmem_s = tmem_base + lane_base + t * 128 + h * 64;
        const float scale_log2 = params.scale_log2;
        const int causal_off = params.seqlen_k - params.seqlen_q;
        const uint32_t scale_addr = smem_base + L::kScale + (t * 128 + r) * 4;
        const uint32_t rowsum_addr = smem_base + L::kRowSum + ((t * 2 + h) * 128 + r) * 4;
        const uint32_t xchg_base = smem_base + L::kXchg + t * (2 * 2 * 128 * 4) + r * 4;
        const int pair_bar = 1 + t * 4 + q4;  // named barrier of this (tile, lane quarter) pair
        const uint64_t sc2 = pack_f32x2(scale_log2, scale_log2);
        uint32_t sf_phase = 0;
        uint32_t xpar = 0;
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
            float m = -INFINITY;
            float l = 0.f;
            for (int j = 0; j < n; ++j) {
                mbar_wait(bar(L::kBarSFull + t), sf_phase, status, 7);
                sf_phase ^= 1u;
                tc_fence_after_sync();
                tr.ev(1);
                const int col0 = j * kBlockN;
                const bool need_mask = (col0 + kBlockN > params.seqlen_k) ||
                                       (params.causal && (col0 + kBlockN - 1 > my_pos0 + causal_off));
                const int rel0 = row_limit - col0 - h * 64;  // last visible index within my 64 keys

                // ---- pass 1: partial row max over my 64 keys ----
                float pmax;
                {
                    uint32_t s[64];
                    tmem_ld32(tmem_s, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
                    tmem_ld32(tmem_s + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
                    tmem_wait_ld();
                    tr.ev(2);
                    if (need_mask) {
#pragma unroll
                        for (int i = 0; i < 64; ++i)
                            if (i > rel0) s[i] = 0xff800000u;
                    }
                    float mx0 = max3(__uint_as_float(s[0]), __uint_as_float(s[1]), __uint_as_float(s[2]));
                    float mx1 = max3(__uint_as_float(s[3]), __uint_as_float(s[4]), __uint_as_float(s[5]));
#pragma unroll
                    for (int i = 6; i < 62; i += 4) {
                        mx0 = max3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
                        mx1 = max3(mx1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
                    }
                    pmax = fmaxf(max3(mx0, __uint_as_float(s[62]), __uint_as_float(s[63])), mx1);
                }
                st_shared_f32(xchg_base + (xpar * 2 + h) * 512, pmax);
                asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                const float other = ld_shared_f32(xchg_base + (xpar * 2 + (h ^ 1)) * 512);
                xpar ^= 1u;
                const float m_new = max3(m, pmax, other);

                // lazy rescale (both halves take the same decision: same m, same m_new)
                float alpha = 1.f;
                if ((m_new - m) * scale_log2 > kRescaleThreshold) {
                    alpha = ex2_approx((m - m_new) * scale_log2);
                    m = m_new;
                }
                if (h == 0) {
                    st_shared_f32(scale_addr, alpha);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(L::kBarScale + t));
                }
                tr.ev(3);

                // ---- pass 2: exp2, 32 keys at a time, P written back over consumed S columns ----
                const float msc = (m == -INFINITY) ? 0.f : m * scale_log2;
                const uint64_t nm2 = pack_f32x2(-msc, -msc);
                uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t sv[32];
                    tmem_ld32(tmem_s + c * 32, sv);
                    tmem_wait_ld();
                    if (need_mask) {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i > rel0) sv[i] = 0xff800000u;
                    }
                    uint32_t pk[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const uint64_t x2 = fma_f32x2(
                            pack_f32x2(__uint_as_float(sv[2 * q]), __uint_as_float(sv[2 * q + 1])), sc2, nm2);
                        float x0, x1, p0, p1;
                        unpack_f32x2(x2, x0, x1);
                        if ((q & 3) < kEmu) {
                            p0 = ex2_emu(x0);
                            p1 = ex2_emu(x1);
                        } else {
                            p0 = ex2_approx(x0);
                            p1 = ex2_approx(x1);
                        }
                        if (q & 1) lb = add_f32x2(lb, pack_f32x2(p0, p1));
                        else la = add_f32x2(la, pack_f32x2(p0, p1));
                        pk[q] = pack2<kBf16>(p0, p1);
                    }
                    tmem_st16(tmem_s + c * 16, pk);
                }
                tmem_wait_st();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(L::kBarP1Full + t));
                tr.ev(4);
                float l0, l1;
                unpack_f32x2(add_f32x2(la, lb), l0, l1);
                l = fmaf(l, alpha, l0 + l1);
            }
            st_shared_f32(rowsum_addr, l);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(L::kBarRowSum + t));
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

                // element mask only on blocks that cross the diagonal or the end of K
                const int col0 = j * kBlockN;
                const bool need_mask = (col0 + kBlockN > params.seqlen_k) ||
                                       (params.causal && (col0 + kBlockN - 1 > my_pos0 + causal_off));
                if (need_mask) {
                    const int rel = row_limit - col0;
#pragma unroll
                    for (int i = 0; i < 128; ++i)
                        if (i > rel) s[i] = 0xff800000u;  // -inf
                }

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
                const float m_new = fmaxf(max3(m, mx0, mx1), fmaxf(mx2, mx3));

                // lazy rescale: keep the stale max unless this row grew by more than 2^8
                float alpha = 1.f;
                if ((m_new - m) * scale_log2 > kRescaleThreshold) {  // false for NaN (-inf - -inf)
                    alpha = ex2_approx((m - m_new) * scale_log2);
                    m = m_new;
                }
                st_shared_f32(scale_addr, alpha);
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(L::kBarScale + t));
                tr.ev(3);

                const float msc = (m == -INFINITY) ? 0.f : m * scale_log2;
                const uint64_t nm2 = pack_f32x2(-msc, -msc);
                uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);
                // exp2 of 32 keys -> 16 packed columns of P, stored straight back over S
                auto exp_chunk = [&](const uint32_t* sv, int c) {
                    uint32_t pk[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const uint64_t x2 = fma_f32x2(
                            pack_f32x2(__uint_as_float(sv[2 * q]), __uint_as_float(sv[2 * q + 1])), sc2, nm2);
                        float p0, p1;
                        if ((q & 3) < kEmu) {
                            ex2_emu_x2(x2, p0, p1);
                        } else {
                            float x0, x1;
                            unpack_f32x2(x2, x0, x1);
                            p0 = ex2_approx(x0);
                            p1 = ex2_approx(x1);
                        }
                        if (q & 1) lb = add_f32x2(lb, pack_f32x2(p0, p1));
                        else la = add_f32x2(la, pack_f32x2(p0, p1));
                        pk[q] = pack2<kBf16>(p0, p1);
                    }
                    tmem_st16(tmem_s + c * 16, pk);
                };
                auto publish = [&](int slot) {
                    tmem_wait_st();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(slot + t));
                };
                // part A: the first kSplit keys, handed to the MMA warp at once
#pragma unroll
                for (int c = 0; c < kSplit / 32; ++c) exp_chunk(&s[c * 32], c);
                publish(L::kBarP1Full);
                tr.ev(4);
                if constexpr (kSplit < 128) {
                    // part B: re-read the remaining S columns from TMEM (P only overwrote columns
                    // [0, kSplit/2)).  The reload is what orders this work after the hand-off of
                    // part A, and it frees the registers of the second half during part A.
                    constexpr int kRest = 128 - kSplit;
                    uint32_t s2[kRest];
#pragma unroll
                    for (int c = 0; c < kRest / 32; ++c)
                        tmem_ld32(tmem_s + kSplit + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s2[c * 32]));
                    tmem_wait_ld();
                    if (need_mask) {
                        const int rel = row_limit - col0 - kSplit;
#pragma unroll
                        for (int i = 0; i < kRest; ++i)
                            if (i > rel) s2[i] = 0xff800000u;
                    }
#pragma unroll
                    for (int c = 0; c < kRest / 32; ++c) exp_chunk(&s2[c * 32], kSplit / 32 + c);
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

}  // namespace v2
}  // namespace fa100
