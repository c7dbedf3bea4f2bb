// FlashAttention-2 forward for sm_100a, generation 3: one 128-row M-tile per persistent CTA,
// THREE S/P buffers in TMEM and two softmax warpgroups that take alternate KV blocks.
//
// Why: with two M-tiles per CTA (fa_sm100_fwd.cuh, and every 2-Q-tile design) TMEM is full
// (2 x 128 S columns + 2 x 128 O columns), P has to alias S, and each tile runs the serial chain
//   S_j ready -> softmax (~1750 cycles) -> P_j -> P·V_j, Q·K_{j+1} (1024) -> S_{j+1} ready
// so the tensor pipe is busy 2 x 1024 cycles out of a ~3300-cycle period (measured, profiles/r01).
// With ONE tile, TMEM holds O (128 columns) and three S/P buffers (3 x 128): Q·K_{j+3} is issued
// before P_j exists, three blocks are in flight, and the softmax of block j+1 (other warpgroup)
// overlaps that of block j.  The kernel becomes throughput-bound (MUFU / issue slots / tensor
// pipe) instead of latency-bound.
//
//   warps 0-3   softmax warpgroup A: KV blocks 0, 2, 4, ...   (thread r <-> query row r)
//   warps 4-7   softmax warpgroup B: KV blocks 1, 3, 5, ...
//   warps 8-11  correction warpgroup: lazy O rescale, final normalisation, O -> global
//   warp  12    MMA issuer;  warp 13  TMA producer + tile scheduler;  warps 14-15 spare
//
// The running row max is shared by the two softmax warpgroups through shared memory (the owner of
// block j reads the reference max left by the owner of block j-1, MaxBar orders the two); each
// warpgroup keeps its own partial row sum relative to the reference max it last saw, and the
// correction warpgroup merges the two at the end.  Everything else (TMA ring, UMMA descriptors,
// lazy rescale threshold, masks, GQA packing, dynamic LPT tile scheduler) is as in generation 2.
#pragma once
#include "fa_sm100_fwd.cuh"

namespace fa100 {
namespace v3 {

using v2::kBlockN;
using v2::kPanelBytes;
using v2::kRescaleThreshold;
using v2::kSchedDepth;
using v2::Tracer;

constexpr int kNumThreads = 512;
constexpr int kCorrWarp0 = 8, kMmaWarp = 12, kLoadWarp = 13;
constexpr int kRegsSoftmax = 184, kRegsCorr = 80, kRegsOther = 64;
static_assert(8 * kRegsSoftmax + 4 * kRegsCorr + 4 * kRegsOther <= 16 * 128, "register budget");
constexpr int kSBufs = 3;

template <int kD, int kStages>
struct Smem {
    static constexpr int kPanels = kD / 64;
    static constexpr int kTileBytes = kPanels * kPanelBytes;  // the Q tile or one K / V block
    static constexpr int kQ = 0;
    static constexpr int kKV = kTileBytes;
    static constexpr int kScale = kKV + kStages * kTileBytes;   // float [3][128] rescale factor of block j (slot j % 3)
    static constexpr int kMax = kScale + 3 * 128 * 4;           // float [128] reference row max
    static constexpr int kFinal = kMax + 128 * 4;               // float [2 tile parity][2 wg][2 {l, m}][128]
    static constexpr int kTileRing = kFinal + 2 * 2 * 2 * 128 * 4;  // int [kSchedDepth]
    static constexpr int kBars = kTileRing + 16;
    static constexpr int kBarQFull = 0;                     // TMA -> MMA
    static constexpr int kBarQEmpty = 1;                    // MMA commit -> TMA
    static constexpr int kBarKVFull = 2;                    // [kStages]
    static constexpr int kBarKVEmpty = 2 + kStages;         // [kStages]
    static constexpr int kBarSFull = 2 + 2 * kStages;       // [3 buffers][2 owner wg] MMA commit -> softmax owner
    static constexpr int kBarPFull = 8 + 2 * kStages;       // [3] softmax (4 warps) + correction (4 warps) -> MMA
    static constexpr int kBarScale = 11 + 2 * kStages;      // [3] softmax owner (4 warps) -> correction
    static constexpr int kBarMax = 14 + 2 * kStages;        // [2] owner of block j -> owner of block j+1
    static constexpr int kBarPVDone = 16 + 2 * kStages;     // MMA commit after every P·V -> correction
    static constexpr int kBarFinal = 17 + 2 * kStages;      // [2 tile parity][2 wg] softmax warpgroup (4 warps) -> correction
    static constexpr int kBarOEmpty = 21 + 2 * kStages;     // correction (4 warps) -> MMA
    static constexpr int kBarTileFull = 22 + 2 * kStages;   // [kSchedDepth]
    static constexpr int kBarTileEmpty = 26 + 2 * kStages;  // [kSchedDepth]
    static constexpr int kNumBars = 30 + 2 * kStages;
    static constexpr int kTmemPtr = kBars + kNumBars * 8;
    static constexpr int kTotal = kTmemPtr + 16;
    static constexpr int kDynamicBytes = kTotal;  // base is 1024 B aligned (checked in the kernel)
};

struct Tile {
    int mblock, hu, b, kv_head, g_coord;
    int pos0;  // first query position of the M-tile
    int nblk;  // KV blocks to visit
};

// Same order as v2::decode_tile (L2 sections, heaviest m-blocks first), one M-tile per work tile.
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
    t.pos0 = t.mblock * p.rows_per_head;
    t.nblk = v2::tile_num_blocks(p, t.pos0);
}

template <typename L>
struct TileConsumer {
    uint32_t smem_base;
    unsigned int* status;
    int round = 0;
    int last_idx = -2;
    __device__ __forceinline__ TileConsumer(uint32_t base, unsigned int* st) : smem_base(base), status(st) {}
    __device__ __forceinline__ bool next(const FwdParams& p, Tile& t, bool elected) {
        const int slot = round % kSchedDepth;
        const uint32_t parity = (round / kSchedDepth) & 1;
        mbar_wait(smem_base + L::kBars + (L::kBarTileFull + slot) * 8, parity, status, 20);
        int idx;
        asm volatile("ld.shared.s32 %0, [%1];" : "=r"(idx) : "r"(smem_base + L::kTileRing + slot * 4) : "memory");
        __syncwarp(__activemask());
        if (elected) mbar_arrive(smem_base + L::kBars + (L::kBarTileEmpty + slot) * 8);
        ++round;
        last_idx = idx;
        if (idx < 0) return false;
        decode_tile(p, idx, t);
        return true;
    }
};

// Debug (kTrace builds): every warp leaves (tile index, block j, step code) behind in global memory
// so that a hung launch can be inspected afterwards (tools/hang_dump.py).
template <bool kTrace>
__device__ __forceinline__ void progress(const FwdParams& p, int idx, int j, int code) {
    if constexpr (kTrace) {
        if (p.trace && (threadIdx.x & 31) == 0 && *reinterpret_cast<volatile unsigned int*>(p.status) == 0u) {
            long long* q = p.trace + 5 * 4096 + (blockIdx.x * 16 + (threadIdx.x >> 5)) * 4;
            if (code == 999) {
                q[2] = j;  // wait site about to block on
            } else {
                q[0] = idx;
                q[1] = (static_cast<long long>(code) << 32) | static_cast<unsigned int>(j);
                q[2] = 0;
            }
        }
    }
}

#define WAITN(B, PH, CODE)                                   \
    do {                                                     \
        progress<kTrace>(params, -1000 - (CODE), (CODE), 999); \
        mbar_wait((B), (PH), status, (CODE));                \
    } while (0)

// kEmu: of every 4 element pairs, how many take the FMA-pipe exp2 (0 = all MUFU).
template <int kD, bool kBf16, int kStages, int kEmu, bool kTrace>
__global__ void __launch_bounds__(kNumThreads, 1)
fa_fwd_sm100_v3(const __grid_constant__ FwdParams params) {
    using L = Smem<kD, kStages>;
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kTileBytes = L::kTileBytes;

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
        mbar_init(bar(L::kBarQFull), 1);
        mbar_init(bar(L::kBarQEmpty), 1);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(L::kBarKVFull + s), 1);
            mbar_init(bar(L::kBarKVEmpty + s), 1);
        }
        for (int s = 0; s < kSBufs; ++s) {
            // one SFull per (buffer, owning warpgroup): a barrier with two alternating consumers lets
            // one of them get two phases ahead at a tile boundary, where the parity test aliases
            mbar_init(bar(L::kBarSFull + 2 * s), 1);
            mbar_init(bar(L::kBarSFull + 2 * s + 1), 1);
            mbar_init(bar(L::kBarPFull + s), 8);
            mbar_init(bar(L::kBarScale + s), 4);
        }
        mbar_init(bar(L::kBarMax + 0), 4);
        mbar_init(bar(L::kBarMax + 1), 4);
        mbar_init(bar(L::kBarPVDone), 1);
        for (int i = 0; i < 4; ++i) mbar_init(bar(L::kBarFinal + i), 4);
        mbar_init(bar(L::kBarOEmpty), 4);
        for (int s = 0; s < kSchedDepth; ++s) {
            mbar_init(bar(L::kBarTileFull + s), 1);
            mbar_init(bar(L::kBarTileEmpty + s), 8 + 4 + 1);  // softmax + correction + MMA
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
            // =========================== TMA producer + tile scheduler ===========================
            tr.init(params, 4, true);
            int slot = 0;
            uint32_t phase = 0;
            uint32_t qe_phase = 0;
            Tile w;
            int next_idx = blockIdx.x;
            auto load_kv = [&](const CUtensorMap* tm, int j) {
                WAITN(bar(L::kBarKVEmpty + slot), phase ^ 1u, 1);
                const uint32_t fb = bar(L::kBarKVFull + slot);
                mbar_arrive_expect_tx(fb, kTileBytes);
                for (int pn = 0; pn < kPanels; ++pn)
                    tma_load_4d(smem_base + L::kKV + slot * kTileBytes + pn * kPanelBytes, tm, fb, pn * 64,
                                j * kBlockN, w.kv_head, w.b);
                if (++slot == kStages) { slot = 0; phase ^= 1u; }
            };
            for (int round = 0;; ++round) {
                const int idx = next_idx < params.num_tiles ? next_idx : -1;
                const int ts = round % kSchedDepth;
                WAITN(bar(L::kBarTileEmpty + ts), ((round / kSchedDepth) & 1) ^ 1u, 21);
                asm volatile("st.shared.s32 [%0], %1;" ::"r"(smem_base + L::kTileRing + ts * 4), "r"(idx) : "memory");
                mbar_arrive(bar(L::kBarTileFull + ts));
                if (idx < 0) break;
                next_idx = int(gridDim.x) + atomicAdd(params.sched, 1);
                decode_tile(params, idx, w);
                const int n = w.nblk;
                progress<kTrace>(params, idx, round, 100);
                if (n == 0) continue;
                WAITN(bar(L::kBarQEmpty), qe_phase ^ 1u, 10);
                qe_phase ^= 1u;
                mbar_arrive_expect_tx(bar(L::kBarQFull), params.q_tile_bytes);
                for (int pn = 0; pn < kPanels; ++pn)
                    tma_load_5d(smem_base + L::kQ + pn * kPanelBytes, &params.tm_q, bar(L::kBarQFull), pn * 64,
                                w.pos0, w.g_coord, w.kv_head, w.b);
                tr.ev(1);
                // consumption order of the MMA warp: K0 K1 K2 | V0 K3 | V1 K4 | ...
                for (int i = 0; i < min(kSBufs, n); ++i) load_kv(&params.tm_k, i);
                for (int j = 0; j < n; ++j) {
                    progress<kTrace>(params, idx, j, 101);
                    load_kv(&params.tm_v, j);
                    if (j + kSBufs < n) load_kv(&params.tm_k, j + kSBufs);
                }
            }
        } else if (warp == kMmaWarp && lane == 0) {
            // =========================== MMA issuer ===========================
            tr.init(params, 3, true);
            constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 128, false, false);
            constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
            const uint32_t tmem_o = tmem_base + 384;
            constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO, version, SW128
            constexpr uint32_t kLboK = (16u >> 4) << 16;
            constexpr uint32_t kLboV = (uint32_t(kPanelBytes) >> 4) << 16;
            auto desc = [](uint32_t lo) -> uint64_t {
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(kDescHi));
                return d;
            };
            const uint32_t q_lo = ((smem_base + L::kQ) >> 4) | kLboK;
            const uint32_t kv_lo = (smem_base + L::kKV) >> 4;
            int slot = 0;
            uint32_t phase = 0;
            // S(i) = Q · K_iᵀ into S buffer b, from the next ring slot
            auto issue_s = [&](int b, int owner) {
                WAITN(bar(L::kBarKVFull + slot), phase, 2);
                tc_fence_after_sync();
                const uint32_t k_lo = (kv_lo + slot * (kTileBytes >> 4)) | kLboK;
#pragma unroll
                for (int ks = 0; ks < kD / 16; ++ks) {
                    const uint32_t off = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
                    umma_ss(tmem_base + b * 128, desc(q_lo + off), desc(k_lo + off), idesc_qk, ks > 0);
                }
                tc_commit(bar(L::kBarSFull + 2 * b + owner));
                tc_commit(bar(L::kBarKVEmpty + slot));
                if (++slot == kStages) { slot = 0; phase ^= 1u; }
            };
            uint32_t q_phase = 0, oe_phase = 0, pf_phase[kSBufs] = {0, 0, 0};
            Tile w;
            TileConsumer<L> sched(smem_base, status);
            while (sched.next(params, w, true)) {
                const int n = w.nblk;
                if (n == 0) continue;
                WAITN(bar(L::kBarQFull), q_phase, 3);
                q_phase ^= 1u;
                tc_fence_after_sync();
                for (int i = 0; i < min(kSBufs, n); ++i) {
                    issue_s(i, i & 1);
                    if (i == n - 1) tc_commit(bar(L::kBarQEmpty));
                }
                tr.ev(2);
                int b = 0;
                for (int j = 0; j < n; ++j) {
                    progress<kTrace>(params, sched.last_idx, j, 200);
                    WAITN(bar(L::kBarKVFull + slot), phase, 4);
                    const int vslot = slot;
                    if (++slot == kStages) { slot = 0; phase ^= 1u; }
                    if (j == 0) {  // O of the previous work tile must have been drained
                        WAITN(bar(L::kBarOEmpty), oe_phase ^ 1u, 11);
                        oe_phase ^= 1u;
                    }
                    WAITN(bar(L::kBarPFull + b), pf_phase[b], 5);
                    pf_phase[b] ^= 1u;
                    tc_fence_after_sync();
                    tr.ev(3);
                    const uint32_t v_lo = (kv_lo + vslot * (kTileBytes >> 4)) | kLboV;
#pragma unroll
                    for (int ks = 0; ks < kBlockN / 16; ++ks)
                        umma_ts(tmem_o, tmem_base + b * 128 + ks * 8, desc(v_lo + ks * (2048 >> 4)), idesc_pv,
                                (j > 0 || ks > 0) ? 1u : 0u);
                    tc_commit(bar(L::kBarKVEmpty + vslot));
                    tc_commit(bar(L::kBarPVDone));
                    if (j + kSBufs < n) {
                        issue_s(b, (j + kSBufs) & 1);
                        if (j + kSBufs == n - 1) tc_commit(bar(L::kBarQEmpty));
                    }
                    tr.ev(5);
                    if (++b == kSBufs) b = 0;
                }
            }
        }
    } else if (warp >= kCorrWarp0) {
        // =========================== correction warpgroup ===========================
        reg_dealloc<kRegsCorr>();
        tr.init(params, 2, warp == kCorrWarp0 && lane == 0);
        const int r = threadIdx.x & 127;  // row of the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t tmem_o = tmem_base + lane_base + 384;
        const float scale_log2 = params.scale_log2;
        uint32_t sc_phase[kSBufs] = {0, 0, 0}, pv_phase = 0, fin_phase[4] = {0, 0, 0, 0};
        int tile_count = 0;
        Tile w;
        TileConsumer<L> sched(smem_base, status);
        while (sched.next(params, w, lane == 0)) {
            if (w.pos0 >= params.seqlen_q) continue;
            const int n = w.nblk;
            int b = 0;
            for (int j = 0; j < n; ++j) {
                progress<kTrace>(params, sched.last_idx, j, 300);
                WAITN(bar(L::kBarScale + b), sc_phase[b], 13);
                sc_phase[b] ^= 1u;
                if (j > 0) {
                    const float alpha = ld_shared_f32(smem_base + L::kScale + (b * 128 + r) * 4);
                    // P·V_{j-1} must have landed in O before O may be touched (and before P·V_j is released)
                    WAITN(bar(L::kBarPVDone), pv_phase, 15);
                    pv_phase ^= 1u;
                    if (__any_sync(0xffffffffu, alpha != 1.f)) {
                        tr.ev(7);
                        tc_fence_after_sync();
#pragma unroll
                        for (int c = 0; c < kD / 32; ++c) {
                            uint32_t o[32];
                            tmem_ld32(tmem_o + c * 32, o);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                            tmem_st32(tmem_o + c * 32, o);
                        }
                        tmem_wait_st();
                        tc_fence_before_sync();
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(L::kBarPFull + b));
                if (++b == kSBufs) b = 0;
            }
            // ---- epilogue: merge the two partial row sums, O / l -> T -> global ----
            const int g = r / P;
            const int pos = w.pos0 + (r - g * P);
            const bool row_valid = (g < params.pack_g) && (pos < params.seqlen_q);
            const int head = packed ? w.hu * params.group + g : w.hu;
            uint8_t* orow = reinterpret_cast<uint8_t*>(params.o) +
                            2 * (static_cast<int64_t>(w.b) * params.o_stride_b +
                                 static_cast<int64_t>(head) * params.o_stride_h +
                                 static_cast<int64_t>(pos) * params.o_stride_s);
            float inv = 0.f;
            progress<kTrace>(params, sched.last_idx, n, 301);
            if (n > 0) {
                const int tp = tile_count & 1;
                ++tile_count;
                WAITN(bar(L::kBarFinal + tp * 2), fin_phase[tp * 2], 14);
                fin_phase[tp * 2] ^= 1u;
                const uint32_t fa = smem_base + L::kFinal + tp * (2 * 2 * 128 * 4) + r * 4;
                const float la = ld_shared_f32(fa), ma = ld_shared_f32(fa + 512);
                float lb = 0.f, mb = -INFINITY;
                if (n > 1) {  // warpgroup B only reports when it owned a block
                    WAITN(bar(L::kBarFinal + tp * 2 + 1), fin_phase[tp * 2 + 1], 17);
                    fin_phase[tp * 2 + 1] ^= 1u;
                    lb = ld_shared_f32(fa + 1024);
                    mb = ld_shared_f32(fa + 1536);
                }
                const float mf = fmaxf(ma, mb);  // the reference max O is relative to
                float l = 0.f;
                if (ma != -INFINITY) l += la * ex2_approx((ma - mf) * scale_log2);
                if (mb != -INFINITY) l += lb * ex2_approx((mb - mf) * scale_log2);
                inv = (l > 0.f) ? __frcp_rn(l) : 0.f;
                WAITN(bar(L::kBarPVDone), pv_phase, 8);
                pv_phase ^= 1u;
                tc_fence_after_sync();
                tr.ev(8);
            }
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
                if (c == kD / 32 - 1 && n > 0) {  // O fully read: the next work tile may overwrite it
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(L::kBarOEmpty));
                }
                if (row_valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        const int col = c * 32 + q4 * 8;
                        if (col < params.headdim) {
                            uint4 v;
                            v.x = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 0]) * inv, __uint_as_float(o[q4 * 8 + 1]) * inv);
                            v.y = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 2]) * inv, __uint_as_float(o[q4 * 8 + 3]) * inv);
                            v.z = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 4]) * inv, __uint_as_float(o[q4 * 8 + 5]) * inv);
                            v.w = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 6]) * inv, __uint_as_float(o[q4 * 8 + 7]) * inv);
                            *reinterpret_cast<uint4*>(orow + col * 2) = v;
                        }
                    }
                }
            }
            tr.ev(9);
        }
    } else {
        // =========================== softmax warpgroups A / B ===========================
        reg_alloc<kRegsSoftmax>();
        const int wg = warp >> 2;  // 0: even KV blocks, 1: odd KV blocks
        tr.init(params, wg, (warp & 3) == 0 && lane == 0);
        const int r = threadIdx.x & 127;  // query row of the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float scale_log2 = params.scale_log2;
        const int causal_off = params.seqlen_k - params.seqlen_q;
        const uint32_t max_addr = smem_base + L::kMax + r * 4;
        const uint64_t sc2 = pack_f32x2(scale_log2, scale_log2);
        uint32_t sf_phase[kSBufs] = {0, 0, 0};  // phases of my own SFull[b][wg] barriers
        uint32_t mx_phase = 0;  // phase of MaxBar[1 - wg], the one this warpgroup waits on
        int tile_count = 0;
        Tile w;
        TileConsumer<L> sched(smem_base, status);
        while (sched.next(params, w, lane == 0)) {
            const int n = w.nblk;
            if (n == 0) continue;
            const int tp = tile_count & 1;
            ++tile_count;
            if (n <= wg) {
                // No block of this tile is mine (n == 1, warpgroup B).  Touch no barrier: a warpgroup
                // that only arrived could run several tiles ahead and complete a later phase early.
                continue;
            }
            const int g = r / P;
            const int pos = w.pos0 + (r - g * P);
            int row_limit = params.seqlen_k - 1;  // last visible column
            if (params.causal) row_limit = min(row_limit, pos + causal_off);
            float m_mine = -INFINITY;  // reference max my partial sum is relative to
            float l = 0.f;
            int b = wg;  // S buffer of block j = j % 3
            for (int j = wg; j < n; j += 2) {
                const uint32_t tmem_s = tmem_base + lane_base + b * 128;
                progress<kTrace>(params, sched.last_idx, j, 400);
                WAITN(bar(L::kBarSFull + 2 * b + wg), sf_phase[b], 7);
                sf_phase[b] ^= 1u;
                tc_fence_after_sync();
                tr.ev(1);
                uint32_t s[128];
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    tmem_ld32(tmem_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
                tmem_wait_ld();
                tr.ev(2);
                const int col0 = j * kBlockN;
                const bool need_mask = (col0 + kBlockN > params.seqlen_k) ||
                                       (params.causal && (col0 + kBlockN - 1 > w.pos0 + causal_off));
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
                const float bmax = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));

                // reference max left by the owner of block j-1 (the other warpgroup)
                float m_old = -INFINITY;
                if (j > 0) {
                    progress<kTrace>(params, sched.last_idx, j, 401);
                    WAITN(bar(L::kBarMax + (1 - wg)), mx_phase, 16);
                    mx_phase ^= 1u;
                    m_old = ld_shared_f32(max_addr);
                }
                if (m_old != m_mine) {  // the partner raised the reference since my last block
                    l *= ex2_approx((m_mine - m_old) * scale_log2);  // m_mine == -inf -> l is 0 anyway
                    m_mine = m_old;
                }
                // lazy rescale: move the reference only when this block exceeds it by more than 2^8
                float alpha = 1.f;
                if ((bmax - m_old) * scale_log2 > kRescaleThreshold) {  // false for NaN (-inf - -inf)
                    alpha = ex2_approx((m_old - bmax) * scale_log2);
                    m_mine = bmax;
                    l *= alpha;
                }
                st_shared_f32(max_addr, m_mine);
                st_shared_f32(smem_base + L::kScale + (b * 128 + r) * 4, alpha);
                __syncwarp();
                if (lane == 0) {
                    if (j + 1 < n) mbar_arrive(bar(L::kBarMax + wg));
                    mbar_arrive(bar(L::kBarScale + b));
                }
                tr.ev(3);

                const float msc = (m_mine == -INFINITY) ? 0.f : m_mine * scale_log2;
                const uint64_t nm2 = pack_f32x2(-msc, -msc);
                uint64_t la = pack_f32x2(0.f, 0.f), lb = pack_f32x2(0.f, 0.f);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t pk[16];
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int i = c * 32 + 2 * q;
                        const uint64_t x2 = fma_f32x2(
                            pack_f32x2(__uint_as_float(s[i]), __uint_as_float(s[i + 1])), sc2, nm2);
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
                if (lane == 0) mbar_arrive(bar(L::kBarPFull + b));
                tr.ev(4);
                float l0, l1;
                unpack_f32x2(add_f32x2(la, lb), l0, l1);
                l += l0 + l1;
                b += 2;
                if (b >= kSBufs) b -= kSBufs;
            }
            // hand my partial sum (and the reference it is relative to) to the correction warpgroup
            const uint32_t fa = smem_base + L::kFinal + tp * (2 * 2 * 128 * 4) + wg * 1024 + r * 4;
            st_shared_f32(fa, l);
            st_shared_f32(fa + 512, m_mine);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(L::kBarFinal + tp * 2 + wg));
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
        __threadfence();
        if (atomicAdd(params.sched + 1, 1) == int(gridDim.x) - 1) {
            params.sched[0] = 0;
            params.sched[1] = 0;
            __threadfence();
        }
    }
}

}  // namespace v3
}  // namespace fa100
