// fa_sm100_decode.cuh — split-KV decode kernels (Sq == 1) for B200.
//
// SURVEY §8(f) rank 2.  The reference handles a single query position by packing the q heads of a
// kv group into the rows of one M tile and running the prefill kernel, one CTA per (batch, kv-head)
// (csrc/flash_attention_api.cpp:72-83, 128-133); "Flash Decoding" is a TODO there (README.md:20).
// With one CTA per unit a decode step uses B·Hkv SMs at most and each of them walks the whole K/V
// sequence alone, so the step is far from the bound that matters here: HBM bandwidth (K and V are
// read exactly once, 2·Sk·D·sizeof(T) bytes per unit; everything else is noise).
//
// Shape of this kernel (bandwidth first, the math is an afterthought at 8 flop/byte):
//   * grid = B·Hkv·nsplit CTAs of 4 warps, two CTAs per SM; the host picks nsplit so that one wave
//     covers the SMs (fa_sm100_host.cu: decode_plan).  A CTA owns `chunk` consecutive keys of one
//     (batch, kv-head) unit.
//   * every warp streams its own 16-key tiles (tile i of the chunk goes to warp i % 4) through a
//     private 3-stage cp.async ring — 16 B per lane per copy, rows of D·2 bytes are contiguous, so
//     every request is a full 32 B sector group; no block-wide barrier in the loop.  Up to
//     2 CTAs x 4 warps x 2 tiles x 8 KB = 128 KB are in flight per SM.
//   * the G <= 16 q heads of the group are the 16 rows of a warp-level m16n8k16 MMA: S = Q·Kᵀ
//     (K fragments by ldmatrix from XOR-swizzled shared memory), exp2-domain online softmax in
//     registers (fp32 state, row sum from unrounded P, P rounded to T before P·V — the same
//     numerics as the prefill kernel), O += P·V (V fragments by ldmatrix.trans).  The 5th-gen
//     tensor cores (tcgen05, M >= 64) have nothing to offer a 16-row problem that is HBM-bound.
//   * the four warps merge (m, l, O) through shared memory; with nsplit > 1 the splits of a unit
//     are merged inside the same launch: 2 .. 8 splits run as one thread-block cluster and rank 0
//     reads its peers' partials out of their shared memory (DSMEM); beyond that each CTA writes its
//     partial to the workspace and the last CTA of the unit to arrive (one atomic counter per
//     unit) merges them.
//   * 16 copies a lane and tile: everything about their addresses that does not change from tile
//     to tile is computed once (TileCopy) — the loop was issue-bound on address arithmetic.
//   * paged KV with per-sequence lengths (fa_decode_flat_sm100): the host cannot size a split per
//     sequence (the lengths are device data), so the CTAs of one wave share the FLATTENED key space
//     instead: every (batch, kv-head) unit owns ceil(len / 64) blocks of 64 keys, laid end to end,
//     and CTA i takes blocks [i·W, (i+1)·W) of that line, whichever units they belong to.  A CTA
//     walks its range unit by unit ("segments"); a segment that covers its unit writes O, any
//     other writes a partial, and the last CTA of a unit to arrive merges.  Every CTA streams the
//     same number of bytes however ragged the batch is.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include <cooperative_groups.h>

#include <type_traits>

namespace fa100 {
namespace decode {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTileKeys = 16;   // keys per warp tile
constexpr int kStages = 3;      // cp.async ring depth per warp
constexpr int kRows = 16;       // MMA M: q heads of one kv group (G <= 16), rest zero
constexpr float kNegBig = -1.0e30f;  // finite stand-in for -inf in the running max
constexpr int kFlatBlock = kWarps * kTileKeys;  // 64 keys: granularity of the flattened key space
constexpr int kFlatMaxBatch = 2048;  // sequences whose prefix sums a CTA keeps in shared memory

struct DecodeParams {
    const void* q;
    const void* k;
    const void* v;
    void* o;
    int batch, heads_kv, group, seqlen_k, headdim;
    int64_t q_sb, q_sh;        // element strides: batch, head
    int64_t k_sb, k_sh, k_ss;  // batch, head, sequence
    int64_t v_sb, v_sh, v_ss;
    int64_t o_sb, o_sh;
    float scale_log2;
    int nsplit;       // CTAs per (batch, kv-head) unit
    int chunk;        // keys per CTA, multiple of kWarps * kTileKeys
    float* part_o;    // [units][nsplit][kRows][kD]  un-normalised O relative to part_ml.m
    float* part_ml;   // [units][nsplit][kRows][2]   (m in the exp2 domain, l)
    int* counters;    // [units], zero on entry, zero again on exit (library-owned ring slot)
    // paged KV (kPaged): k / v are page pools, key j of batch b lives in page
    // block_table[b * bt_stride + j / page_size] at in-page token j % page_size
    const int32_t* block_table;
    const int32_t* seqlens;   // [batch] keys per sequence (seqlen_k is then only the split bound)
    int64_t bt_stride;
    int64_t k_sp, v_sp;       // element stride between pages
    int page_size;            // multiple of kTileKeys
    int page_shift;           // log2(page_size) when it is a power of two, else -1
    // flattened key space (fa_decode_flat_sm100): part_o / part_ml hold 2 slots per CTA (its
    // first and its last segment), counters one int per CTA
    int flat_min_blocks;      // smallest range of a CTA, in blocks of kFlatBlock keys
};

template <int kD>
struct Smem {
    static constexpr int kRowBytes = kD * 2;
    static constexpr int kTileBytes = kTileKeys * kRowBytes;        // one K or V tile
    static constexpr int kStageBytes = 2 * kTileBytes;              // K + V
    static constexpr int kWarpBytes = kStages * kStageBytes;
    static constexpr int kRingBytes = kWarps * kWarpBytes;          // 96 KB at D=128
    // the merge buffers reuse the ring once it has drained
    static constexpr int kMergeO = 0;                                // float [kWarps][kRows][kD]
    static constexpr int kMergeML = kWarps * kRows * kD * 4;         // float [kWarps][kRows][2]
    static constexpr int kFlag = kMergeML + kWarps * kRows * 2 * 4;  // int
    // cluster merge: this CTA's combined partial, read by rank 0 of the cluster through DSMEM
    static constexpr int kClusterO = (kFlag + 16 + 255) / 256 * 256;  // float [kRows][kD]
    static constexpr int kClusterML = kClusterO + kRows * kD * 4;     // float [kRows][2]
    static constexpr int kTotal = kRingBytes;
    static_assert(kClusterML + kRows * 2 * 4 <= kRingBytes, "merge buffers must fit in the ring");
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// 16-byte async copy global -> shared, zero-filled when !valid (src is then not dereferenced).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int bytes = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if constexpr (kBf16) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    } else {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
}
template <bool kBf16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    if constexpr (kBf16) {
        __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&t);
    } else {
        __half2 t = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&t);
    }
}
template <bool kBf16>
__device__ __forceinline__ void store_out(void* base, int64_t idx, float x) {
    if constexpr (kBf16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(x);
    else reinterpret_cast<__half*>(base)[idx] = __float2half_rn(x);
}
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}


// ---- one warp copies one 16-key K tile and V tile into a ring stage --------------------------------
// The kernels issue 48 % of the cycles they are resident, two warps to a scheduler: instructions
// count, and 64-bit address arithmetic per 16-byte copy was 40 % of them.  Everything that does
// not change from tile to tile is computed once per kernel here: lane l copies chunk c = l % kChunks
// of rows l / kChunks + j · (32 / kChunks); the shared-memory offsets are XOR-swizzled per row.
template <int kD>
struct TileCopy {
    static constexpr int kChunks = kD / 8;                      // 16-byte chunks per row
    static constexpr int kCopies = kTileKeys * kChunks / 32;    // cp.async per lane per K (or V) tile
    static constexpr int kRowStep = 32 / kChunks;               // rows between a lane's copies
    int64_t k_lane, v_lane;   // BYTE offset of the lane's first copy inside a tile
    int64_t k_step, v_step;   // byte offset between its copies
    uint32_t soff[kCopies];   // byte offsets inside the stage's K (or V) tile
    int row, chunk;           // first row and chunk of the lane

    __device__ __forceinline__ void init(const DecodeParams& p, int lane) {
        using L = Smem<kD>;
        row = lane / kChunks;
        chunk = lane - row * kChunks;
        k_lane = (row * p.k_ss + chunk * 8) * 2;
        v_lane = (row * p.v_ss + chunk * 8) * 2;
        k_step = kRowStep * p.k_ss * 2;
        v_step = kRowStep * p.v_ss * 2;
#pragma unroll
        for (int j = 0; j < kCopies; ++j) {
            const int r = row + j * kRowStep;
            soff[j] = r * L::kRowBytes + ((chunk ^ (r & 7)) << 4);
        }
    }
    // kt / vt: row 0 of the tile (head and page applied); rows >= valid_rows and chunks >= D / 8
    // are zero-filled.  `safe` is any mapped address (zero-size copies still carry one).
    template <typename T>
    __device__ __forceinline__ void copy(uint32_t dst, const T* kt, const T* vt, int valid_rows, int D,
                                         const T* safe) const {
        using L = Smem<kD>;
        const char* ks = reinterpret_cast<const char*>(kt) + k_lane;
        const char* vs = reinterpret_cast<const char*>(vt) + v_lane;
        if (valid_rows >= kTileKeys && D == kD) {
#pragma unroll
            for (int j = 0; j < kCopies; ++j) {
                cp_async16(dst + soff[j], ks, true);
                cp_async16(dst + L::kTileBytes + soff[j], vs, true);
                ks += k_step;
                vs += v_step;
            }
        } else {
#pragma unroll
            for (int j = 0; j < kCopies; ++j) {
                const bool ok = row + j * kRowStep < valid_rows && chunk * 8 < D;
                cp_async16(dst + soff[j], ok ? ks : reinterpret_cast<const char*>(safe), ok);
                cp_async16(dst + L::kTileBytes + soff[j], ok ? vs : reinterpret_cast<const char*>(safe), ok);
                ks += k_step;
                vs += v_step;
            }
        }
    }
};

// ---- one key range of one (batch, kv-head) unit ---------------------------------------------------
// All four warps stream keys [key_begin, key_end) of unit (b, kvh) — key_begin a multiple of 16 —
// and leave their (m, l, O) triples in the merge buffers of shared memory (which overlay the ring:
// the ring is idle on entry and idle again on return).  Ends with a __syncthreads().
template <int kD, bool kBf16, bool kPaged>
__device__ __forceinline__ void attend_range(const DecodeParams& p, uint8_t* smem, int b, int kvh,
                                             int key_begin, int key_end) {
    using L = Smem<kD>;
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    constexpr int kKSteps = kD / 16;       // MMA k-steps of Q·Kᵀ
    constexpr int kNBlocks = kD / 8;       // MMA n-blocks of P·V

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = p.group, D = p.headdim;
    const int ntiles = key_end > key_begin ? (key_end - key_begin + kTileKeys - 1) / kTileKeys : 0;

    const T* kbase = reinterpret_cast<const T*>(p.k) + (kPaged ? 0 : b * p.k_sb) + kvh * p.k_sh;
    const T* vbase = reinterpret_cast<const T*>(p.v) + (kPaged ? 0 : b * p.v_sb) + kvh * p.v_sh;
    const uint32_t ring = smem_addr(smem) + warp * L::kWarpBytes;
    // Paged: a 16-key tile never straddles a page (page_size % 16 == 0, tiles start at multiples
    // of 16), so one block-table entry serves the tile; the entry of the NEXT tile this warp will
    // issue is loaded one issue ahead so that its latency never sits in front of the copies.
    const int32_t* bt = kPaged ? p.block_table + b * p.bt_stride : nullptr;
    int phys_pref = 0;
    if constexpr (kPaged) {
        const int key0 = key_begin + warp * kTileKeys;
        if (warp < ntiles) phys_pref = __ldg(bt + (p.page_shift >= 0 ? key0 >> p.page_shift : key0 / p.page_size));
    }

    // ---- per-warp cp.async ring ----
    TileCopy<kD> tc;
    tc.init(p, lane);
    auto issue_tile = [&](int t) {  // t: tile index inside the range (this warp's), may be >= ntiles
        if (t < ntiles) {
            const int stage = (t / kWarps) % kStages;
            const uint32_t dst = ring + stage * L::kStageBytes;
            const int key0 = key_begin + t * kTileKeys;
            const T* kb = kbase;
            const T* vb = vbase;
            int row0 = key0;  // token index of the tile's first row relative to kb / vb
            if constexpr (kPaged) {
                row0 = p.page_shift >= 0 ? key0 & (p.page_size - 1) : key0 % p.page_size;
                kb += static_cast<int64_t>(phys_pref) * p.k_sp;
                vb += static_cast<int64_t>(phys_pref) * p.v_sp;
                if (t + kWarps < ntiles) {
                    const int keyn = key0 + kWarps * kTileKeys;
                    phys_pref = __ldg(bt + (p.page_shift >= 0 ? keyn >> p.page_shift : keyn / p.page_size));
                }
            }
            tc.copy(dst, kb + static_cast<int64_t>(row0) * p.k_ss, vb + static_cast<int64_t>(row0) * p.v_ss,
                    key_end - key0, D, kbase);
        }
        cp_async_commit();  // always commit: the wait below counts groups
    };

    float o[kNBlocks][4];
#pragma unroll
    for (int j = 0; j < kNBlocks; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m0 = kNegBig, m1 = kNegBig;  // running max (exp2 domain) of rows lane/4 and lane/4 + 8
    float l0 = 0.f, l1 = 0.f;          // this thread's share of the row sums

    issue_tile(warp);
    issue_tile(warp + kWarps);

    // ---- Q fragments (A operand, rows = heads of the group), straight from global, under the
    //      latency of the first two K/V tiles ----
    uint32_t qa[kKSteps][4];
    {
        const T* qbase = reinterpret_cast<const T*>(p.q) + b * p.q_sb + static_cast<int64_t>(kvh) * G * p.q_sh;
        const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
        for (int ks = 0; ks < kKSteps; ++ks) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int row = r0 + (i & 1) * 8, col = ks * 16 + c0 + (i >> 1) * 8;
                uint32_t x = 0;
                if (row < G && col < D) x = *reinterpret_cast<const uint32_t*>(qbase + row * p.q_sh + col);
                qa[ks][i] = x;
            }
        }
    }

    for (int t = warp; t < ntiles; t += kWarps) {
        issue_tile(t + 2 * kWarps);
        cp_async_wait<2>();
        __syncwarp();
        const int stage = (t / kWarps) % kStages;
        const uint32_t ks_addr = ring + stage * L::kStageBytes;
        const uint32_t vs_addr = ks_addr + L::kTileBytes;

        // S = Q · Kᵀ : 16 heads x 16 keys
        float s[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
        {
            const int krow = ((lane >> 4) << 3) + (lane & 7);  // key row this lane addresses
            const int khalf = (lane >> 3) & 1;                 // d-low / d-high 8x8 matrix
#pragma unroll
            for (int ks = 0; ks < kKSteps; ++ks) {
                uint32_t b0, b1, b2, b3;
                const int c = 2 * ks + khalf;
                ldmatrix_x4(ks_addr + krow * L::kRowBytes + ((c ^ (krow & 7)) << 4), b0, b1, b2, b3);
                mma_16816<kBf16>(s[0], qa[ks], b0, b1);
                mma_16816<kBf16>(s[1], qa[ks], b2, b3);
            }
        }
        // mask keys past the end of this range (only the last tile can have them)
        const int key0 = key_begin + t * kTileKeys;
        if (key0 + kTileKeys > key_end) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int key = key0 + j * 8 + (lane & 3) * 2 + (i & 1);
                    if (key >= key_end) s[j][i] = -INFINITY;
                }
        }
        // online softmax, exp2 domain
        float mx0 = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
        float mx1 = fmaxf(fmaxf(s[0][2], s[0][3]), fmaxf(s[1][2], s[1][3]));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0 * p.scale_log2), mn1 = fmaxf(m1, mx1 * p.scale_log2);
        const float a0 = fast_exp2(m0 - mn0), a1 = fast_exp2(m1 - mn1);
        m0 = mn0; m1 = mn1;
        float pr[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            pr[j][0] = fast_exp2(fmaf(s[j][0], p.scale_log2, -mn0));
            pr[j][1] = fast_exp2(fmaf(s[j][1], p.scale_log2, -mn0));
            pr[j][2] = fast_exp2(fmaf(s[j][2], p.scale_log2, -mn1));
            pr[j][3] = fast_exp2(fmaf(s[j][3], p.scale_log2, -mn1));
        }
        l0 = fmaf(l0, a0, (pr[0][0] + pr[0][1]) + (pr[1][0] + pr[1][1]));
        l1 = fmaf(l1, a1, (pr[0][2] + pr[0][3]) + (pr[1][2] + pr[1][3]));
        if (__any_sync(0xffffffffu, a0 != 1.f || a1 != 1.f)) {
#pragma unroll
            for (int j = 0; j < kNBlocks; ++j) { o[j][0] *= a0; o[j][1] *= a0; o[j][2] *= a1; o[j][3] *= a1; }
        }
        // P (rounded to T) as the A operand of P·V
        uint32_t pa[4];
        pa[0] = pack2<kBf16>(pr[0][0], pr[0][1]);
        pa[1] = pack2<kBf16>(pr[0][2], pr[0][3]);
        pa[2] = pack2<kBf16>(pr[1][0], pr[1][1]);
        pa[3] = pack2<kBf16>(pr[1][2], pr[1][3]);
        {
            const int vrow = (((lane >> 3) & 1) << 3) + (lane & 7);  // key row
            const int vsel = lane >> 4;                              // which of the two d-chunks
#pragma unroll
            for (int jj = 0; jj < kNBlocks / 2; ++jj) {
                uint32_t b0, b1, b2, b3;
                const int c = 2 * jj + vsel;
                ldmatrix_x4_trans(vs_addr + vrow * L::kRowBytes + ((c ^ (vrow & 7)) << 4), b0, b1, b2, b3);
                mma_16816<kBf16>(o[2 * jj], pa, b0, b1);
                mma_16816<kBf16>(o[2 * jj + 1], pa, b2, b3);
            }
        }
        __syncwarp();  // the stage is overwritten by the next iteration's copies
    }
    cp_async_wait<0>();
    // row sums: the four lanes of a quad hold disjoint keys
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    __syncthreads();  // every warp is done with its ring: reuse it for the merge

    float* mo = reinterpret_cast<float*>(smem + L::kMergeO) + warp * kRows * kD;
    float* mml = reinterpret_cast<float*>(smem + L::kMergeML) + warp * kRows * 2;
    {
        const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < kNBlocks; ++j) {
            *reinterpret_cast<float2*>(mo + r0 * kD + j * 8 + c0) = make_float2(o[j][0], o[j][1]);
            *reinterpret_cast<float2*>(mo + (r0 + 8) * kD + j * 8 + c0) = make_float2(o[j][2], o[j][3]);
        }
        if ((lane & 3) == 0) {
            mml[r0 * 2] = m0; mml[r0 * 2 + 1] = l0;
            mml[(r0 + 8) * 2] = m1; mml[(r0 + 8) * 2 + 1] = l1;
        }
    }
    __syncthreads();

}

// Combine the four warps' triples left by attend_range and either write O of unit (b, kvh)
// (final: the range was the whole unit) or the partial: O [kRows][kD] to dst_o, (m, l) pairs to
// dst_ml (a workspace slot, or this CTA's own shared memory for the cluster merge).
template <int kD, bool kBf16>
__device__ __forceinline__ void write_range(const DecodeParams& p, const uint8_t* smem, int b, int kvh,
                                            bool final, float* dst_o, float* dst_ml) {
    using L = Smem<kD>;
    const int G = p.group, D = p.headdim;
    const float* all_o = reinterpret_cast<const float*>(smem + L::kMergeO);
    const float* all_ml = reinterpret_cast<const float*>(smem + L::kMergeML);
    for (int idx = threadIdx.x; idx < G * kD; idx += kThreads) {
        const int row = idx / kD, d = idx - row * kD;
        float m = kNegBig;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) m = fmaxf(m, all_ml[(w * kRows + row) * 2]);
        float acc = 0.f, l = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const float f = fast_exp2(all_ml[(w * kRows + row) * 2] - m);
            acc = fmaf(all_o[(w * kRows + row) * kD + d], f, acc);
            l = fmaf(all_ml[(w * kRows + row) * 2 + 1], f, l);
        }
        if (final) {
            if (d < D) {
                const float inv = l > 0.f ? __frcp_rn(l) : 0.f;
                store_out<kBf16>(p.o, b * p.o_sb + (static_cast<int64_t>(kvh) * G + row) * p.o_sh + d, acc * inv);
            }
        } else {
            dst_o[row * kD + d] = acc;
            if (d == 0) {
                dst_ml[row * 2] = m;
                dst_ml[row * 2 + 1] = l;
            }
        }
    }
}

// After write_range(final = false): count this CTA in on `counter`; true for the last of the
// unit's `n` pieces to arrive (block-uniform; every thread must call it).
template <int kD>
__device__ __forceinline__ bool arrive_last(uint8_t* smem, int* counter, int n) {
    using L = Smem<kD>;
    int* flag = reinterpret_cast<int*>(smem + L::kFlag);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *flag = (atomicAdd(counter, 1) == n - 1);
    __syncthreads();
    const bool last = *flag != 0;
    __syncthreads();  // the merge reuses the shared memory the flag lives in
    if (last) __threadfence();
    return last;
}

// Merge the `n` pieces of unit (b, kvh) and write its O.  ml_of(i) / o_of(i): the 32 (m, l)
// floats and the [kRows][kD] partial O of piece i — workspace slots in L2 (kL2: ld.global.cg) or
// the shared memory of the CTAs of a cluster (plain generic loads).  Uses the first n·128 bytes of
// the (idle) ring as scratch; every thread of the CTA must call it.
template <int kD, bool kBf16, bool kL2, typename MlOf, typename OOf>
__device__ __forceinline__ void merge_pieces(const DecodeParams& p, uint8_t* smem, int b, int kvh, int n,
                                             MlOf ml_of, OOf o_of) {
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    const int G = p.group, D = p.headdim;
    auto ld4 = [](const float* ptr) {
        if constexpr (kL2) return __ldcg(reinterpret_cast<const float4*>(ptr));
        else return *reinterpret_cast<const float4*>(ptr);
    };
    // Every load below is independent of the previous one (the partials come from L2 at some
    // 600 ns a round trip: a serial loop over the pieces would cost more than the whole K/V stream).
    // 1. all (m, l) pairs of the unit -> shared memory, float4 loads (a piece's 32 floats are contiguous)
    float* sm_ml = reinterpret_cast<float*>(smem);  // [n][kRows][2], later (f, -) in place
    {
        constexpr int kPer = kRows * 2 / 4;  // float4s per piece
        const int n4 = n * kPer;
#pragma unroll 4
        for (int i = threadIdx.x; i < n4; i += kThreads) {
            const int sp = i / kPer, j = i - sp * kPer;
            reinterpret_cast<float4*>(sm_ml)[i] = ld4(ml_of(sp) + j * 4);
        }
    }
    __syncthreads();
    // 2. per row: global max, total sum, and the weight of every piece (8 lanes per row)
    {
        const int row = threadIdx.x >> 3, part = threadIdx.x & 7;
        float m = kNegBig;
        for (int sp = part; sp < n; sp += 8) m = fmaxf(m, sm_ml[(sp * kRows + row) * 2]);
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
        float l = 0.f;
        for (int sp = part; sp < n; sp += 8)
            l = fmaf(sm_ml[(sp * kRows + row) * 2 + 1], fast_exp2(sm_ml[(sp * kRows + row) * 2] - m), l);
        l += __shfl_xor_sync(0xffffffffu, l, 1);
        l += __shfl_xor_sync(0xffffffffu, l, 2);
        l += __shfl_xor_sync(0xffffffffu, l, 4);
        const float inv = l > 0.f ? __frcp_rn(l) : 0.f;
        for (int sp = part; sp < n; sp += 8)
            sm_ml[(sp * kRows + row) * 2] = fast_exp2(sm_ml[(sp * kRows + row) * 2] - m) * inv;
    }
    __syncthreads();
    // 3. O[row][d..d+3] = sum over pieces of weight * partial O, float4 per thread, 8 loads in flight
    for (int idx = threadIdx.x; idx < G * (kD / 4); idx += kThreads) {
        const int row = idx / (kD / 4), d = (idx - row * (kD / 4)) * 4;
        if (d >= D) continue;
        const int in_piece = row * kD + d;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int sp = 0; sp < n; sp += 8) {
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                x[u] = sp + u < n ? ld4(o_of(sp + u) + in_piece) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float f = sp + u < n ? sm_ml[((sp + u) * kRows + row) * 2] : 0.f;
                acc.x = fmaf(x[u].x, f, acc.x); acc.y = fmaf(x[u].y, f, acc.y);
                acc.z = fmaf(x[u].z, f, acc.z); acc.w = fmaf(x[u].w, f, acc.w);
            }
        }
        const int64_t o_idx = b * p.o_sb + (static_cast<int64_t>(kvh) * G + row) * p.o_sh + d;
        uint2 packed;
        packed.x = pack2<kBf16>(acc.x, acc.y);
        packed.y = pack2<kBf16>(acc.z, acc.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.o) + o_idx) = packed;  // D % 8 == 0, 16 B-aligned rows
    }
    __syncthreads();
}

// Keys of sequence b of a paged batch.  The contract is 0 <= seqlens[b] <= max_seqlen_k (the block
// table is only that wide); a value outside it is clamped rather than followed out of the table.
__device__ __forceinline__ int paged_len(const DecodeParams& p, int b) {
    return min(max(__ldg(p.seqlens + b), 0), p.seqlen_k);
}

// ---- one CTA per (unit, split): contiguous K/V, or paged K/V with the per-sequence split ---------
// kCluster: the nsplit (2 .. 8) CTAs of a unit are one thread-block cluster.  Every CTA leaves its
// partial in its own shared memory, the cluster synchronises, and rank 0 merges straight out of its
// peers' shared memory (DSMEM): no partials through L2, no fence, no atomic, no second round trip —
// the fixed cost that dominates decode calls of a few microseconds.  Same arithmetic in the same
// order as the workspace merge: bit-equal results.
template <int kD, bool kBf16, bool kPaged, bool kCluster = false>
__global__ void __launch_bounds__(kThreads, 2)
fa_decode_sm100(const DecodeParams p) {
    using L = Smem<kD>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int unit = blockIdx.x / p.nsplit, split = blockIdx.x - unit * p.nsplit;
    const int b = unit / p.heads_kv, kvh = unit - b * p.heads_kv;
    const int key_begin = split * p.chunk;
    const int seqlen = kPaged ? paged_len(p, b) : p.seqlen_k;
    const int key_end = min(seqlen, key_begin + p.chunk);

    attend_range<kD, kBf16, kPaged>(p, smem, b, kvh, key_begin, key_end);
    if constexpr (kCluster) {
        namespace cg = cooperative_groups;
        cg::cluster_group cluster = cg::this_cluster();
        float* own_o = reinterpret_cast<float*>(smem + L::kClusterO);
        float* own_ml = reinterpret_cast<float*>(smem + L::kClusterML);
        write_range<kD, kBf16>(p, smem, b, kvh, false, own_o, own_ml);
        cluster.sync();  // every partial of the unit is in place (and visible cluster-wide)
        if (split == 0)
            merge_pieces<kD, kBf16, false>(
                p, smem, b, kvh, p.nsplit,
                [&](int i) { return static_cast<const float*>(cluster.map_shared_rank(own_ml, i)); },
                [&](int i) { return static_cast<const float*>(cluster.map_shared_rank(own_o, i)); });
        cluster.sync();  // a CTA's shared memory must outlive the reads of rank 0
    } else {
        const int64_t ubase = static_cast<int64_t>(unit) * p.nsplit;
        write_range<kD, kBf16>(p, smem, b, kvh, p.nsplit == 1, p.part_o + (ubase + split) * kRows * kD,
                               p.part_ml + (ubase + split) * kRows * 2);
        if (p.nsplit == 1) return;
        if (!arrive_last<kD>(smem, p.counters + unit, p.nsplit)) return;
        const float* ml0 = p.part_ml + ubase * kRows * 2;
        const float* o0 = p.part_o + ubase * kRows * kD;
        merge_pieces<kD, kBf16, true>(
            p, smem, b, kvh, p.nsplit, [ml0](int i) { return ml0 + i * kRows * 2; },
            [o0](int i) { return o0 + static_cast<int64_t>(i) * kRows * kD; });
        if (threadIdx.x == 0) p.counters[unit] = 0;  // re-arm for the next launch on this slot
    }
}

// ---- paged K/V, the CTAs share the flattened key space -------------------------------------------
// Unit u = (b, kvh) owns n(b) = max(1, ceil(len(b) / 64)) blocks of the line, units in (b, kvh)
// order: offset(u) = Hkv · pre[b] + kvh · n(b) with pre the exclusive prefix sum of n.  CTA c takes
// blocks [c·W, (c+1)·W), W = max(flat_min_blocks, ceil(total / gridDim.x)).  Its range is a run of
// segments, one per unit it touches; only the first and the last can be partial, so a CTA owns two
// workspace slots (2c: first segment, 2c + 1: any later one).  A unit cut into pieces spans the
// CTAs c0 .. c1 and is the LAST segment of c0, so counters[c0] is its own; piece i of it sits in
// slot 2(c0 + i) for i > 0 and in 2·c0 or 2·c0 + 1 for i = 0 (the former iff the unit starts
// exactly at c0's range).
//
// A block is 64 keys = one 16-key tile for each of the four warps, so all warps change segment at
// the same block index and the copy pipeline never drains inside a CTA: the tiles of blocks i + 1
// and i + 2 — of the next unit, if block i ends a segment — are in flight while block i is computed
// and while the four warps merge, through the ring stage block i just vacated.
template <int kD>
struct FlatSmem {
    using L = Smem<kD>;
    static constexpr int kPre = L::kRingBytes;                       // int [kFlatMaxBatch + 1]
    static constexpr int kWarpTotals = kPre + (kFlatMaxBatch + 4) * 4;  // int [kWarps]
    static constexpr int kML = kWarpTotals + kWarps * 4;             // float [kWarps][kRows][2]
    static constexpr int kRowScale = kML + kWarps * kRows * 2 * 4;   // float [kRows][2]: (max, 1 / sum)
    static constexpr int kFlag = kRowScale + kRows * 2 * 4;          // int
    static constexpr int kTotal = kFlag + 16;
    static_assert(kRows * kD * 4 <= L::kStageBytes, "a warp's O must fit in one ring stage");
    static_assert(kML % 16 == 0, "alignment");
};

struct FlatCursor {
    int b, kvh, blk;  // block `blk` of unit (b, kvh)
    int n, len;       // blocks and keys of sequence b
};

template <int kD, bool kBf16>
__global__ void __launch_bounds__(kThreads, 2)
fa_decode_flat_sm100(const DecodeParams p) {
    using L = Smem<kD>;
    using F = FlatSmem<kD>;
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    constexpr int kKSteps = kD / 16;
    constexpr int kNBlocks = kD / 8;
    extern __shared__ __align__(128) uint8_t smem[];
    int* pre = reinterpret_cast<int*>(smem + F::kPre);
    const int B = p.batch, Hkv = p.heads_kv, G = p.group, D = p.headdim;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- exclusive prefix sum of the block counts: thread i owns a run of `per` sequences ----
    {
        const int per = (B + kThreads - 1) / kThreads;
        const int first = threadIdx.x * per, last = min(B, first + per);
        int sum = 0;
        for (int i = first; i < last; ++i) sum += max(1, (paged_len(p, i) + kFlatBlock - 1) / kFlatBlock);
        int incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += y;
        }
        int* wtot = reinterpret_cast<int*>(smem + F::kWarpTotals);
        if (lane == 31) wtot[warp] = incl;
        __syncthreads();
        int base = incl - sum;
        for (int w = 0; w < warp; ++w) base += wtot[w];
        for (int i = first; i < last; ++i) {
            pre[i] = base;
            base += max(1, (paged_len(p, i) + kFlatBlock - 1) / kFlatBlock);
        }
        if (threadIdx.x == kThreads - 1) pre[B] = base;  // the last thread's run ends the line (or is empty)
        __syncthreads();
    }
    const int total = Hkv * pre[B];
    const int W = max(p.flat_min_blocks, (total + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x));
    const int cta = blockIdx.x;
    const int start = cta * W;
    if (start >= total) return;
    const int nblk = min(total, start + W) - start;

    // ---- the unit `start` falls into: largest b with Hkv · pre[b] <= start ----
    FlatCursor cc;
    {
        int lo = 0, hi = B;  // invariant: Hkv · pre[lo] <= start < Hkv · pre[hi]
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (Hkv * pre[mid] <= start) lo = mid; else hi = mid;
        }
        cc.b = lo;
        cc.n = pre[lo + 1] - pre[lo];
        cc.kvh = (start - Hkv * pre[lo]) / cc.n;
        cc.blk = start - Hkv * pre[lo] - cc.kvh * cc.n;
        cc.len = paged_len(p, lo);
    }
    FlatCursor ic = cc;  // the copies run two blocks ahead of the math
    auto advance = [&](FlatCursor& c) {
        if (++c.blk < c.n) return;
        c.blk = 0;
        if (++c.kvh < Hkv) return;
        c.kvh = 0;
        if (++c.b < B) {
            c.n = pre[c.b + 1] - pre[c.b];
            c.len = paged_len(p, c.b);
        } else {
            c.n = 1;  // past the line: never issued, never computed
            c.len = 0;
        }
    };
    // This warp's tile of a block starts at key blk·64 + warp·16; it never straddles a page.
    auto table_entry = [&](const FlatCursor& c) {
        const int key0 = c.blk * kFlatBlock + warp * kTileKeys;
        const int pg = p.page_shift >= 0 ? key0 >> p.page_shift : key0 / p.page_size;
        return key0 < c.len ? __ldg(p.block_table + c.b * p.bt_stride + pg) : 0;
    };


    // ---- per-warp cp.async ring, continuous over the CTA's whole range ----
    const uint32_t ring = smem_addr(smem) + warp * L::kWarpBytes;
    const T* kpool = reinterpret_cast<const T*>(p.k);
    const T* vpool = reinterpret_cast<const T*>(p.v);
    TileCopy<kD> tc;
    tc.init(p, lane);
    int phys = table_entry(ic);  // page of the tile the next issue copies
    int issued = 0, istage = 0;
    // Copies of block `issued` (at ic).  kSameUnit: the caller knows that the block after it is
    // in the same unit and inside the CTA's range, which is all the cursor logic there is to skip.
    auto issue_block = [&](auto same_unit) {
        constexpr bool kSameUnit = decltype(same_unit)::value;
        if (kSameUnit || issued < nblk) {
            const int key0 = ic.blk * kFlatBlock + warp * kTileKeys;
            if (key0 < ic.len) {
                const int row0 = p.page_shift >= 0 ? key0 & (p.page_size - 1) : key0 % p.page_size;
                tc.copy(ring + istage * L::kStageBytes,
                        kpool + static_cast<int64_t>(phys) * p.k_sp + ic.kvh * p.k_sh + static_cast<int64_t>(row0) * p.k_ss,
                        vpool + static_cast<int64_t>(phys) * p.v_sp + ic.kvh * p.v_sh + static_cast<int64_t>(row0) * p.v_ss,
                        ic.len - key0, D, kpool);
            }
            ++issued;
            istage = istage == kStages - 1 ? 0 : istage + 1;
            if constexpr (kSameUnit) {
                ++ic.blk;
                phys = table_entry(ic);  // one issue ahead: its latency is off the path
            } else {
                advance(ic);
                if (issued < nblk) {
                    phys = table_entry(ic);
                    if (ic.blk == 0 && lane < G * 2) {  // a new unit two blocks ahead: its q rows -> L2
                        const T* qrow = reinterpret_cast<const T*>(p.q) + ic.b * p.q_sb +
                                        (static_cast<int64_t>(ic.kvh) * G + (lane >> 1)) * p.q_sh + (lane & 1) * 64;
                        if ((lane & 1) * 64 < D) asm volatile("prefetch.global.L2 [%0];" ::"l"(qrow));
                    }
                }
            }
        }
        cp_async_commit();  // always commit: the waits count groups
    };
    issue_block(std::false_type{});
    issue_block(std::false_type{});

    float o[kNBlocks][4];
    float m0, m1, l0, l1;
    uint32_t qa[kKSteps][4];
    int cstage = 0;
    float* ml_all = reinterpret_cast<float*>(smem + F::kML);

    // The warp's tile of block cc.blk: S = Q·Kᵀ, online softmax, O += P·V.  kFull: all 16 keys
    // exist (every block but the last of a unit).
    auto compute_block = [&](auto full) {
        constexpr bool kFull = decltype(full)::value;
        const uint32_t ks_addr = ring + cstage * L::kStageBytes;
        const uint32_t vs_addr = ks_addr + L::kTileBytes;
        const int key0 = cc.blk * kFlatBlock + warp * kTileKeys;
        if (kFull || key0 < cc.len) {
            float s[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            {
                const int krow = ((lane >> 4) << 3) + (lane & 7);  // key row this lane addresses
                const int khalf = (lane >> 3) & 1;                 // d-low / d-high 8x8 matrix
#pragma unroll
                for (int ks = 0; ks < kKSteps; ++ks) {
                    uint32_t b0, b1, b2, b3;
                    const int c = 2 * ks + khalf;
                    ldmatrix_x4(ks_addr + krow * L::kRowBytes + ((c ^ (krow & 7)) << 4), b0, b1, b2, b3);
                    mma_16816<kBf16>(s[0], qa[ks], b0, b1);
                    mma_16816<kBf16>(s[1], qa[ks], b2, b3);
                }
            }
            if (!kFull && key0 + kTileKeys > cc.len) {  // keys past the end of the sequence
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int key = key0 + j * 8 + (lane & 3) * 2 + (u & 1);
                        if (key >= cc.len) s[j][u] = -INFINITY;
                    }
            }
            // online softmax, exp2 domain
            float mx0 = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
            float mx1 = fmaxf(fmaxf(s[0][2], s[0][3]), fmaxf(s[1][2], s[1][3]));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0 * p.scale_log2), mn1 = fmaxf(m1, mx1 * p.scale_log2);
            const float a0 = fast_exp2(m0 - mn0), a1 = fast_exp2(m1 - mn1);
            m0 = mn0; m1 = mn1;
            float pr[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                pr[j][0] = fast_exp2(fmaf(s[j][0], p.scale_log2, -mn0));
                pr[j][1] = fast_exp2(fmaf(s[j][1], p.scale_log2, -mn0));
                pr[j][2] = fast_exp2(fmaf(s[j][2], p.scale_log2, -mn1));
                pr[j][3] = fast_exp2(fmaf(s[j][3], p.scale_log2, -mn1));
            }
            l0 = fmaf(l0, a0, (pr[0][0] + pr[0][1]) + (pr[1][0] + pr[1][1]));
            l1 = fmaf(l1, a1, (pr[0][2] + pr[0][3]) + (pr[1][2] + pr[1][3]));
            if (__any_sync(0xffffffffu, a0 != 1.f || a1 != 1.f)) {
#pragma unroll
                for (int j = 0; j < kNBlocks; ++j) { o[j][0] *= a0; o[j][1] *= a0; o[j][2] *= a1; o[j][3] *= a1; }
            }
            uint32_t pa[4];
            pa[0] = pack2<kBf16>(pr[0][0], pr[0][1]);
            pa[1] = pack2<kBf16>(pr[0][2], pr[0][3]);
            pa[2] = pack2<kBf16>(pr[1][0], pr[1][1]);
            pa[3] = pack2<kBf16>(pr[1][2], pr[1][3]);
            {
                const int vrow = (((lane >> 3) & 1) << 3) + (lane & 7);  // key row
                const int vsel = lane >> 4;                              // which of the two d-chunks
#pragma unroll
                for (int jj = 0; jj < kNBlocks / 2; ++jj) {
                    uint32_t b0, b1, b2, b3;
                    const int c = 2 * jj + vsel;
                    ldmatrix_x4_trans(vs_addr + vrow * L::kRowBytes + ((c ^ (vrow & 7)) << 4), b0, b1, b2, b3);
                    mma_16816<kBf16>(o[2 * jj], pa, b0, b1);
                    mma_16816<kBf16>(o[2 * jj + 1], pa, b2, b3);
                }
            }
        }
        __syncwarp();  // the stage is free: the next issue (or the merge) overwrites it
        cstage = cstage == kStages - 1 ? 0 : cstage + 1;
    };

    // q fragments of unit (b, kvh): rows = the G heads of the group, straight from global
    auto load_q = [&](int b, int kvh) {
        const T* qbase = reinterpret_cast<const T*>(p.q) + b * p.q_sb + static_cast<int64_t>(kvh) * G * p.q_sh;
        const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
        for (int ks = 0; ks < kKSteps; ++ks) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int row = r0 + (u & 1) * 8, col = ks * 16 + c0 + (u >> 1) * 8;
                uint32_t x = 0;
                if (row < G && col < D) x = __ldg(reinterpret_cast<const uint32_t*>(qbase + row * p.q_sh + col));
                qa[ks][u] = x;
            }
        }
    };
    load_q(cc.b, cc.kvh);  // goes out before the first wait: one latency, not two

    for (int i = 0; i < nblk;) {
        // ---- one segment: blocks cc.blk .. of unit (cc.b, cc.kvh), as far as the range goes ----
        const int seg_blk0 = cc.blk;
        const int seg_end = i + min(nblk - i, cc.n - cc.blk);
        const bool first_segment = i == 0;
        const bool unit_ends = seg_end - i == cc.n - cc.blk;
#pragma unroll
        for (int j = 0; j < kNBlocks; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
        m0 = m1 = kNegBig;
        l0 = l1 = 0.f;
        // lean part: blocks i + 2 and i + 3 are in this segment, block i is not the unit's last
        for (; i + 3 < seg_end; ++i) {
            issue_block(std::true_type{});  // block i + 2, into the stage block i - 1 vacated
            cp_async_wait<2>();             // blocks i .. i + 2 are pending: i has landed
            __syncwarp();
            compute_block(std::true_type{});
            ++cc.blk;
        }
        // the segment's last three blocks: the copies may run into the next unit
        for (; i < seg_end; ++i) {
            issue_block(std::false_type{});
            cp_async_wait<2>();
            __syncwarp();
            compute_block(std::false_type{});
            ++cc.blk;
        }
        --cc.blk;  // the segment's last block
        const FlatCursor done = cc;
        advance(cc);
        // q of the next segment: its registers are free now, and the loads fly during the merge
        if (i < nblk) load_q(cc.b, cc.kvh);

        // ---- merge the four warps through the stage just vacated ----
        const int mstage = cstage == 0 ? kStages - 1 : cstage - 1;
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        {
            float* mo = reinterpret_cast<float*>(smem + warp * L::kWarpBytes + mstage * L::kStageBytes);
            float* mml = ml_all + warp * kRows * 2;
            const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
            for (int j = 0; j < kNBlocks; ++j) {
                *reinterpret_cast<float2*>(mo + r0 * kD + j * 8 + c0) = make_float2(o[j][0], o[j][1]);
                *reinterpret_cast<float2*>(mo + (r0 + 8) * kD + j * 8 + c0) = make_float2(o[j][2], o[j][3]);
            }
            if ((lane & 3) == 0) {
                mml[r0 * 2] = m0; mml[r0 * 2 + 1] = l0;
                mml[(r0 + 8) * 2] = m1; mml[(r0 + 8) * 2 + 1] = l1;
            }
        }
        __syncthreads();
        const bool whole = seg_blk0 == 0 && unit_ends;
        const int64_t slot = 2 * cta + (first_segment ? 0 : 1);
        const int64_t o_row0 = done.b * p.o_sb + static_cast<int64_t>(done.kvh) * G * p.o_sh;
        for (int idx = threadIdx.x; idx < G * kD; idx += kThreads) {
            const int row = idx / kD, d = idx - row * kD;
            float m = kNegBig;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) m = fmaxf(m, ml_all[(w * kRows + row) * 2]);
            float acc = 0.f, l = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const float f = fast_exp2(ml_all[(w * kRows + row) * 2] - m);
                const float* wo = reinterpret_cast<const float*>(smem + w * L::kWarpBytes + mstage * L::kStageBytes);
                acc = fmaf(wo[row * kD + d], f, acc);
                l = fmaf(ml_all[(w * kRows + row) * 2 + 1], f, l);
            }
            if (whole) {
                if (d < D) store_out<kBf16>(p.o, o_row0 + row * p.o_sh + d, acc * (l > 0.f ? __frcp_rn(l) : 0.f));
            } else {
                p.part_o[(slot * kRows + row) * kD + d] = acc;
                if (d == 0) {
                    p.part_ml[(slot * kRows + row) * 2] = m;
                    p.part_ml[(slot * kRows + row) * 2 + 1] = l;
                }
            }
        }
        if (!whole) {
            // ---- one piece of a unit that spans several CTAs: the last to arrive merges ----
            const int off = start + i - 1 - done.blk;  // offset(u) on the line (block i - 1 was the segment's last)
            const int c0 = off / W, npieces = (off + done.n - 1) / W - c0 + 1;
            const int slot0 = 2 * c0 + (off == c0 * W ? 0 : 1);
            int* flag = reinterpret_cast<int*>(smem + F::kFlag);
            // Release by ONE thread after the barrier (the fence is cumulative over what the
            // barrier made visible to it): a fence in every thread would also wait for every
            // thread's copies in flight, i.e. drain the pipeline this kernel keeps full.
            __syncthreads();
            if (threadIdx.x == 0) {
                __threadfence();
                *flag = (atomicAdd(p.counters + c0, 1) == npieces - 1);
            }
            __syncthreads();
            if (*flag != 0) {
                __threadfence();
                auto slot_of = [c0, slot0](int k) { return static_cast<int64_t>(k == 0 ? slot0 : 2 * (c0 + k)); };
                // 1. per row: max and 1 / sum over the pieces (8 lanes a row, (m, l) from L2)
                float* rowscale = reinterpret_cast<float*>(smem + F::kRowScale);
                {
                    const int row = threadIdx.x >> 3, part = threadIdx.x & 7;
                    float m = kNegBig, l = 0.f;
                    for (int sp = part; sp < npieces; sp += 8) {
                        const float2 ml = __ldcg(reinterpret_cast<const float2*>(p.part_ml + (slot_of(sp) * kRows + row) * 2));
                        const float mn = fmaxf(m, ml.x);
                        l = fmaf(l, fast_exp2(m - mn), ml.y * fast_exp2(ml.x - mn));
                        m = mn;
                    }
#pragma unroll
                    for (int d = 1; d < 8; d <<= 1) {
                        const float mo = __shfl_xor_sync(0xffffffffu, m, d), lo = __shfl_xor_sync(0xffffffffu, l, d);
                        const float mn = fmaxf(m, mo);
                        l = l * fast_exp2(m - mn) + lo * fast_exp2(mo - mn);
                        m = mn;
                    }
                    if (part == 0) {
                        rowscale[row * 2] = m;
                        rowscale[row * 2 + 1] = l > 0.f ? __frcp_rn(l) : 0.f;
                    }
                }
                __syncthreads();
                // 2. O[row][d..d+3] = sum over pieces of weight · partial O; 8 independent loads in flight
                for (int idx = threadIdx.x; idx < G * (kD / 4); idx += kThreads) {
                    const int row = idx / (kD / 4), d = (idx - row * (kD / 4)) * 4;
                    if (d >= D) continue;
                    const float rm = rowscale[row * 2], rinv = rowscale[row * 2 + 1];
                    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int sp = 0; sp < npieces; sp += 8) {
                        float4 x[8];
                        float pm[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const bool ok = sp + u < npieces;
                            const int64_t sl = ok ? slot_of(sp + u) : 0;
                            x[u] = ok ? __ldcg(reinterpret_cast<const float4*>(p.part_o + (sl * kRows + row) * kD + d))
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
                            pm[u] = ok ? __ldcg(p.part_ml + (sl * kRows + row) * 2) : kNegBig;
                        }
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const float f = sp + u < npieces ? fast_exp2(pm[u] - rm) * rinv : 0.f;
                            acc.x = fmaf(x[u].x, f, acc.x); acc.y = fmaf(x[u].y, f, acc.y);
                            acc.z = fmaf(x[u].z, f, acc.z); acc.w = fmaf(x[u].w, f, acc.w);
                        }
                    }
                    uint2 packed;
                    packed.x = pack2<kBf16>(acc.x, acc.y);
                    packed.y = pack2<kBf16>(acc.z, acc.w);
                    *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.o) + o_row0 + row * p.o_sh + d) = packed;
                }
                if (threadIdx.x == 0) p.counters[c0] = 0;  // re-arm for the next launch on this slot
            }
        }
        __syncthreads();  // the merge stage and the flag are free again
    }
    cp_async_wait<0>();
}

}  // namespace decode
}  // namespace fa100
