// FlashAttention-2 forward for sm_100a: warp-specialised tcgen05 / TMEM / TMA kernel.
//
// Replaces the reference's flash_attention_v2 (csrc/flash_attention_template.cuh:138-564),
// Mask (csrc/mask.cuh:30-88) and launch_flash_attention (template.cuh:567-583).  Same math
// (fp32 S and O accumulators, exp2-domain online softmax, P rounded to T before P·V, row sums
// from unrounded P, bottom-right causal alignment, kv-head = q-head / G), entirely different
// machine mapping:
//
//   CTA = one work tile = 2 M-tiles of 128 query rows sharing one K/V stream (384 threads)
//     warps 0-3  softmax warpgroup for M-tile 0   (one thread per query row, no shuffles)
//     warps 4-7  softmax warpgroup for M-tile 1
//     warp  8    MMA issuer (one elected thread): S_t = Q_t·K_jᵀ (SS), O_t += P_t·V_j (TS)
//     warp  9    TMA producer: Q tiles once, K_j / V_j through an mbarrier ring
//     warps 10-11 idle (they exist so that setmaxnreg can move their registers to the softmax WGs)
//   TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,256+D) O1 [256+D,256+2D);
//     P_t (16-bit, packed two per column) aliases the first 64 columns of S_t.
//   SMEM: Q0, Q1, kStages x (K or V) tiles, all 128B-swizzled [rows][64] panels as written by
//     TMA; K-major UMMA descriptors for Q and K, MN-major for V (no transpose anywhere).
//
// An M-tile's 128 rows are (head-in-group g, position p) pairs: row r -> g = r / P, p = r % P,
// with P = rows_per_head.  Unpacked: P = 128, one q head.  GQA-packed: P = 128 / G and the G
// q heads of one kv head share the tile (the Q TMA box is (64, P, G) of a 5-D tensor map).
#pragma once
#include "fa_sm100_ptx.cuh"

namespace fa100 {

struct KernelParams {
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
    int32_t q_tile_bytes;   // bytes TMA delivers for one Q M-tile (all panels)
    int32_t causal;
    float scale_log2;       // softmax_scale * log2(e)
    unsigned int* status;   // [0] abort flag/code, [1..3] detail
};

constexpr int kBlockM = 128;
constexpr int kBlockN = 128;
constexpr int kPanelBytes = 128 * 128;  // 128 rows x 64 16-bit elements
constexpr int kNumThreads = 384;  // 3 warpgroups: softmax 0, softmax 1, {MMA, TMA, 2 idle}
constexpr int kMmaWarp = 8;
constexpr int kLoadWarp = 9;
// setmaxnreg can only redistribute what the CTA was launched with: ptxas gives 168 registers to
// each of the 384 threads (3 warps per SM sub-partition), so 2 x kRegsSoftmax + kRegsIssue must not
// exceed 3 x 168 = 504 — otherwise setmaxnreg.inc blocks forever (a hang no mbarrier watchdog sees).
constexpr int kRegsSoftmax = 216;
constexpr int kRegsIssue = 72;
static_assert(2 * kRegsSoftmax + kRegsIssue <= 3 * 168, "register budget exceeds launch allocation");
constexpr float kRescaleThreshold = 8.0f;  // lazy rescale: only when the max grows by > 2^8

template <int kD, int kStages>
struct SmemLayout {
    static constexpr int kPanels = kD / 64;
    static constexpr int kTileBytes = kPanels * kPanelBytes;  // a Q M-tile or one K / V block
    static constexpr int kQ0 = 0;
    static constexpr int kQ1 = kTileBytes;
    static constexpr int kKV = 2 * kTileBytes;
    static constexpr int kBars = kKV + kStages * kTileBytes;
    // barrier slots (8 B each)
    static constexpr int kBarQFull = 0;                     // [2]
    static constexpr int kBarKVFull = 2;                    // [kStages]
    static constexpr int kBarKVEmpty = 2 + kStages;         // [kStages]
    static constexpr int kBarSFull = 2 + 2 * kStages;       // [2]
    static constexpr int kBarPFull = 4 + 2 * kStages;       // [2]
    static constexpr int kBarOFull = 6 + 2 * kStages;       // [2]
    static constexpr int kNumBars = 8 + 2 * kStages;
    static constexpr int kTmemPtr = kBars + kNumBars * 8;
    static constexpr int kTotal = kTmemPtr + 16;
    static constexpr int kDynamicBytes = kTotal + 1024;  // slack to align the base to 1024 B
};

// Number of 128-wide KV blocks M-tile `t` of this CTA has to visit (0 = nothing to do).
__device__ __forceinline__ int tile_num_blocks(const KernelParams& p, int tile_pos0) {
    if (tile_pos0 >= p.seqlen_q) return 0;
    int kv_end = p.seqlen_k;
    if (p.causal) {
        int maxpos = min(tile_pos0 + p.rows_per_head, p.seqlen_q) - 1;
        kv_end = min(kv_end, maxpos + (p.seqlen_k - p.seqlen_q) + 1);
    }
    return kv_end <= 0 ? 0 : (kv_end + kBlockN - 1) / kBlockN;
}

template <int kD, bool kBf16, int kStages>
__global__ void __launch_bounds__(kNumThreads, 1)
fa_fwd_sm100_kernel(const __grid_constant__ KernelParams params) {
    using L = SmemLayout<kD, kStages>;
    constexpr int kPanels = L::kPanels;
    constexpr uint32_t kTileBytes = L::kTileBytes;

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    unsigned int* status = params.status;

    auto bar = [&](int slot) -> uint32_t { return smem_base + L::kBars + slot * 8; };

    // ---- work tile decode: heaviest causal tiles get the lowest block ids (dispatched first) ----
    const int hb = params.head_units * params.batch;
    const int mrank = blockIdx.x / hb;
    const int rem = blockIdx.x - mrank * hb;
    const int mblock = params.num_m_blocks - 1 - mrank;
    const int hu = rem % params.head_units;
    const int b = rem / params.head_units;
    const int P = params.rows_per_head;
    const bool packed = params.pack_g > 1;
    const int kv_head = packed ? hu : hu / params.group;
    const int g_coord = packed ? 0 : hu % params.group;
    const int pos0[2] = {mblock * 2 * P, mblock * 2 * P + P};
    const int nblk[2] = {tile_num_blocks(params, pos0[0]), tile_num_blocks(params, pos0[1])};
    const int nmax = max(nblk[0], nblk[1]);

    // ---- one-time setup ----
    if (warp == kLoadWarp && lane == 0) {
        prefetch_tensormap(&params.tm_q);
        prefetch_tensormap(&params.tm_k);
        prefetch_tensormap(&params.tm_v);
        mbar_init(bar(L::kBarQFull + 0), 1);
        mbar_init(bar(L::kBarQFull + 1), 1);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(L::kBarKVFull + s), 1);
            mbar_init(bar(L::kBarKVEmpty + s), 1);
        }
        for (int t = 0; t < 2; ++t) {
            mbar_init(bar(L::kBarSFull + t), 1);
            mbar_init(bar(L::kBarPFull + t), 128);
            mbar_init(bar(L::kBarOFull + t), 1);
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

    if (warp >= 8) {
      reg_dealloc<kRegsIssue>();
      if (warp == kLoadWarp) {
        // =========================== TMA producer ===========================
        if (lane == 0) {
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                if (nblk[t] > 0) {
                    const uint32_t qbar = bar(L::kBarQFull + t);
                    mbar_arrive_expect_tx(qbar, params.q_tile_bytes);
                    for (int pn = 0; pn < kPanels; ++pn)
                        tma_load_5d(smem_base + (t ? L::kQ1 : L::kQ0) + pn * kPanelBytes,
                                    &params.tm_q, qbar, pn * 64, pos0[t], g_coord, kv_head, b);
                }
            }
            int slot = 0;
            uint32_t phase = 0;
            for (int it = 0; it < 2 * nmax; ++it) {  // K0 V0 K1 V1 ...
                const int j = it >> 1;
                const CUtensorMap* tm = (it & 1) ? &params.tm_v : &params.tm_k;
                mbar_wait(bar(L::kBarKVEmpty + slot), phase ^ 1u, status, 1);
                const uint32_t fb = bar(L::kBarKVFull + slot);
                mbar_arrive_expect_tx(fb, kTileBytes);
                for (int pn = 0; pn < kPanels; ++pn)
                    tma_load_4d(smem_base + L::kKV + slot * kTileBytes + pn * kPanelBytes, tm, fb,
                                pn * 64, j * kBlockN, kv_head, b);
                if (++slot == kStages) { slot = 0; phase ^= 1u; }
            }
        }
      } else if (warp == kMmaWarp) {
        // =========================== MMA issuer ===========================
        if (lane == 0) {
            constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 128, false, false);
            constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
            const uint32_t tmem_s[2] = {tmem_base, tmem_base + 128};
            const uint32_t tmem_o[2] = {tmem_base + 256, tmem_base + 256 + kD};
            const uint32_t sq[2] = {smem_base + L::kQ0, smem_base + L::kQ1};

            auto issue_s = [&](int t, uint32_t kaddr) {
#pragma unroll
                for (int ks = 0; ks < kD / 16; ++ks) {
                    const uint32_t off = (ks >> 2) * kPanelBytes + (ks & 3) * 32;
                    umma_ss(tmem_s[t], make_smem_desc_sw128(sq[t] + off, 16, 1024),
                            make_smem_desc_sw128(kaddr + off, 16, 1024), idesc_qk, ks > 0);
                }
            };
            auto issue_pv = [&](int t, uint32_t vaddr, bool acc) {
#pragma unroll
                for (int ks = 0; ks < kBlockN / 16; ++ks) {
                    umma_ts(tmem_o[t], tmem_s[t] + ks * 8,
                            make_smem_desc_sw128(vaddr + ks * 2048, kPanelBytes, 1024), idesc_pv,
                            (acc || ks > 0) ? 1u : 0u);
                }
            };

            int slot = 0;
            uint32_t phase = 0;
            auto advance = [&]() { if (++slot == kStages) { slot = 0; phase ^= 1u; } };

            if (nmax > 0) {
                // prologue: S_t(0) = Q_t · K_0ᵀ
                int kslot = slot; uint32_t kphase = phase; advance();
                mbar_wait(bar(L::kBarKVFull + kslot), kphase, status, 2);
                tc_fence_after_sync();
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (nblk[t] > 0) {
                        mbar_wait(bar(L::kBarQFull + t), 0, status, 3);
                        issue_s(t, smem_base + L::kKV + kslot * kTileBytes);
                        tc_commit(bar(L::kBarSFull + t));
                    }
                }
                tc_commit(bar(L::kBarKVEmpty + kslot));

                for (int j = 0; j < nmax; ++j) {
                    const int vslot = slot; const uint32_t vphase = phase; advance();
                    const bool has_next = (j + 1 < nmax);
                    if (has_next) { kslot = slot; kphase = phase; advance(); }
                    mbar_wait(bar(L::kBarKVFull + vslot), vphase, status, 4);
                    bool k_ready = false;
                    const uint32_t vaddr = smem_base + L::kKV + vslot * kTileBytes;
                    const uint32_t kaddr = smem_base + L::kKV + kslot * kTileBytes;
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        if (j < nblk[t]) {
                            mbar_wait(bar(L::kBarPFull + t), j & 1, status, 5);
                            tc_fence_after_sync();
                            issue_pv(t, vaddr, j > 0);
                            if (j + 1 < nblk[t]) {
                                if (!k_ready) {
                                    mbar_wait(bar(L::kBarKVFull + kslot), kphase, status, 6);
                                    tc_fence_after_sync();
                                    k_ready = true;
                                }
                                issue_s(t, kaddr);
                                tc_commit(bar(L::kBarSFull + t));
                            } else {
                                tc_commit(bar(L::kBarOFull + t));
                            }
                        }
                    }
                    tc_commit(bar(L::kBarKVEmpty + vslot));
                    if (has_next) tc_commit(bar(L::kBarKVEmpty + kslot));
                }
            }
        }
      }
    } else {
        // =========================== softmax warpgroups ===========================
        reg_alloc<kRegsSoftmax>();
        const int t = warp >> 2;                 // M-tile handled by this warpgroup
        const int r = threadIdx.x & 127;         // query row inside the M-tile == TMEM lane
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t tmem_s = tmem_base + lane_base + t * 128;
        const uint32_t tmem_o = tmem_base + lane_base + 256 + t * kD;
        const int my_pos0 = mblock * 2 * P + t * P;
        const int n = tile_num_blocks(params, my_pos0);
        const int g = r / P;
        const int pos = my_pos0 + (r - g * P);
        const bool row_valid = (g < params.pack_g) && (pos < params.seqlen_q);
        const float scale_log2 = params.scale_log2;
        const int causal_off = params.seqlen_k - params.seqlen_q;
        int row_limit = params.seqlen_k - 1;                       // last visible column
        if (params.causal) row_limit = min(row_limit, pos + causal_off);

        float m = -INFINITY;  // running row max (raw score units)
        float l = 0.f;        // running row sum of exp2, relative to m

        for (int j = 0; j < n; ++j) {
            mbar_wait(bar(L::kBarSFull + t), j & 1, status, 7);
            tc_fence_after_sync();

            uint32_t s[128];
#pragma unroll
            for (int c = 0; c < 4; ++c)
                tmem_ld32(tmem_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
            tmem_wait_ld();

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

            float mx0 = __uint_as_float(s[0]), mx1 = __uint_as_float(s[1]);
            float mx2 = __uint_as_float(s[2]), mx3 = __uint_as_float(s[3]);
#pragma unroll
            for (int i = 4; i < 128; i += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(s[i]));
                mx1 = fmaxf(mx1, __uint_as_float(s[i + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(s[i + 2]));
                mx3 = fmaxf(mx3, __uint_as_float(s[i + 3]));
            }
            const float m_new = fmaxf(m, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)));

            if (j == 0) {
                m = m_new;
            } else {
                // lazy rescale: keep the stale max unless some row of this warp grew by > 2^8
                const bool grow = (m_new - m) * scale_log2 > kRescaleThreshold;
                if (__any_sync(0xffffffffu, grow)) {
                    const float alpha = (m_new == -INFINITY) ? 1.f
                                                             : ex2_approx((m - m_new) * scale_log2);
                    m = m_new;
                    l *= alpha;
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
                }
            }

            const float msc = (m == -INFINITY) ? 0.f : m * scale_log2;
            float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
            uint32_t pk[64];
#pragma unroll
            for (int i = 0; i < 128; i += 4) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(s[i]), scale_log2, -msc));
                const float p1 = ex2_approx(fmaf(__uint_as_float(s[i + 1]), scale_log2, -msc));
                const float p2 = ex2_approx(fmaf(__uint_as_float(s[i + 2]), scale_log2, -msc));
                const float p3 = ex2_approx(fmaf(__uint_as_float(s[i + 3]), scale_log2, -msc));
                l0 += p0; l1 += p1; l2 += p2; l3 += p3;
                pk[i / 2] = pack2<kBf16>(p0, p1);
                pk[i / 2 + 1] = pack2<kBf16>(p2, p3);
            }
            l += (l0 + l1) + (l2 + l3);
            tmem_st32(tmem_s, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
            tmem_st32(tmem_s + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
            tmem_wait_st();
            tc_fence_before_sync();
            mbar_arrive(bar(L::kBarPFull + t));
        }

        // ---- epilogue: O / l -> T -> global (row-contiguous 16 B stores) ----
        if (my_pos0 < params.seqlen_q) {
            const int head = packed ? hu * params.group + g : hu;
            uint8_t* orow = reinterpret_cast<uint8_t*>(params.o) +
                            2 * (static_cast<int64_t>(b) * params.o_stride_b +
                                 static_cast<int64_t>(head) * params.o_stride_h +
                                 static_cast<int64_t>(pos) * params.o_stride_s);
            if (n > 0) {
                mbar_wait(bar(L::kBarOFull + t), 0, status, 8);
                tc_fence_after_sync();
            }
            const float inv = (l > 0.f) ? __frcp_rn(l) : 0.f;
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
        }
    }

    // ---- teardown ----
    tc_fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace fa100
