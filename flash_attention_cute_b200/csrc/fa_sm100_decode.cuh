// fa_sm100_decode.cuh — split-KV decode kernel (Sq == 1) for B200.
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
//   * the four warps merge (m, l, O) through shared memory; with nsplit > 1 the CTA writes its
//     partial to the workspace and the last CTA of a unit to arrive (one atomic counter per unit)
//     merges the splits and writes O — one launch, no second kernel.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include <type_traits>

namespace fa100 {
namespace decode {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTileKeys = 16;   // keys per warp tile
constexpr int kStages = 3;      // cp.async ring depth per warp
constexpr int kRows = 16;       // MMA M: q heads of one kv group (G <= 16), rest zero
constexpr float kNegBig = -1.0e30f;  // finite stand-in for -inf in the running max

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
    static constexpr int kTotal = kRingBytes > kFlag + 16 ? kRingBytes : kFlag + 16;
    static_assert(kFlag + 16 <= kRingBytes, "merge buffers must fit in the ring");
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

template <int kD, bool kBf16, bool kPaged>
__global__ void __launch_bounds__(kThreads, 2)
fa_decode_sm100(const DecodeParams p) {
    using L = Smem<kD>;
    constexpr int kChunks = kD / 8;        // 16-byte chunks per row
    constexpr int kCopies = kTileKeys * kChunks / 32;  // cp.async per lane per K (or V) tile
    constexpr int kKSteps = kD / 16;       // MMA k-steps of Q·Kᵀ
    constexpr int kNBlocks = kD / 8;       // MMA n-blocks of P·V
    extern __shared__ __align__(128) uint8_t smem[];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int unit = blockIdx.x / p.nsplit, split = blockIdx.x - unit * p.nsplit;
    const int b = unit / p.heads_kv, kvh = unit - b * p.heads_kv;
    const int G = p.group, D = p.headdim;
    const int key_begin = split * p.chunk;
    const int seqlen = kPaged ? p.seqlens[b] : p.seqlen_k;
    const int key_end = min(seqlen, key_begin + p.chunk);
    const int ntiles = key_end > key_begin ? (key_end - key_begin + kTileKeys - 1) / kTileKeys : 0;

    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    const T* kbase = reinterpret_cast<const T*>(p.k) + (kPaged ? 0 : b * p.k_sb) + kvh * p.k_sh;
    const T* vbase = reinterpret_cast<const T*>(p.v) + (kPaged ? 0 : b * p.v_sb) + kvh * p.v_sh;
    const uint32_t ring = smem_addr(smem) + warp * L::kWarpBytes;
    // Paged: a 16-key tile never straddles a page (page_size % 16 == 0, tiles start at multiples
    // of 16), so one block-table entry serves the tile; the entry of the NEXT tile this warp will
    // issue is loaded one issue ahead so that its latency never sits in front of the copies.
    const int32_t* bt = kPaged ? p.block_table + b * p.bt_stride : nullptr;
    int phys_pref = 0;
    if constexpr (kPaged) {
        if (warp < ntiles) phys_pref = __ldg(bt + (key_begin + warp * kTileKeys) / p.page_size);
    }

    // ---- per-warp cp.async ring ----
    auto issue_tile = [&](int t) {  // t: tile index inside the chunk (this warp's), may be >= ntiles
        if (t < ntiles) {
            const int stage = (t / kWarps) % kStages;
            const uint32_t dst = ring + stage * L::kStageBytes;
            const int key0 = key_begin + t * kTileKeys;
            const T* kb = kbase;
            const T* vb = vbase;
            int row0 = key0;  // token index of the tile's first row relative to kb / vb
            if constexpr (kPaged) {
                const int pg = key0 / p.page_size;
                row0 = key0 - pg * p.page_size;
                kb += static_cast<int64_t>(phys_pref) * p.k_sp;
                vb += static_cast<int64_t>(phys_pref) * p.v_sp;
                if (t + kWarps < ntiles) phys_pref = __ldg(bt + (key0 + kWarps * kTileKeys) / p.page_size);
            }
#pragma unroll
            for (int j = 0; j < kCopies; ++j) {
                const int i = lane + 32 * j;
                const int row = i / kChunks, c = i - row * kChunks;
                const bool ok = key0 + row < key_end && c * 8 < D;
                const uint32_t off = row * L::kRowBytes + ((c ^ (row & 7)) << 4);
                cp_async16(dst + off, ok ? kb + static_cast<int64_t>(row0 + row) * p.k_ss + c * 8 : kbase, ok);
                cp_async16(dst + L::kTileBytes + off, ok ? vb + static_cast<int64_t>(row0 + row) * p.v_ss + c * 8 : vbase, ok);
            }
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
        // mask keys past the end of this chunk (only the last tile can have them)
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

    const float* all_o = reinterpret_cast<const float*>(smem + L::kMergeO);
    const float* all_ml = reinterpret_cast<const float*>(smem + L::kMergeML);
    const int64_t part_base = (static_cast<int64_t>(unit) * p.nsplit + split) * kRows;
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
        if (p.nsplit == 1) {
            if (d < D) {
                const float inv = l > 0.f ? __frcp_rn(l) : 0.f;
                store_out<kBf16>(p.o, b * p.o_sb + (static_cast<int64_t>(kvh) * G + row) * p.o_sh + d, acc * inv);
            }
        } else {
            p.part_o[(part_base + row) * kD + d] = acc;
            if (d == 0) {
                p.part_ml[(part_base + row) * 2] = m;
                p.part_ml[(part_base + row) * 2 + 1] = l;
            }
        }
    }
    if (p.nsplit == 1) return;

    // ---- last CTA of the unit merges the splits ----
    int* flag = reinterpret_cast<int*>(smem + L::kFlag);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *flag = (atomicAdd(p.counters + unit, 1) == p.nsplit - 1);
    __syncthreads();
    const bool last = *flag != 0;
    __syncthreads();  // the merge below reuses the shared memory the flag lives in
    if (!last) return;
    __threadfence();
    // Every load below is independent of the previous one (the partials come from L2 at some
    // 600 ns a round trip: a serial loop over the splits would cost more than the whole K/V stream).
    // 1. all (m, l) pairs of the unit -> shared memory, coalesced float4 loads
    const int64_t ubase = static_cast<int64_t>(unit) * p.nsplit * kRows;
    float* sm_ml = reinterpret_cast<float*>(smem);  // [nsplit][kRows][2], later (f, -) in place
    {
        const float4* src = reinterpret_cast<const float4*>(p.part_ml + ubase * 2);
        const int n4 = p.nsplit * kRows * 2 / 4;
#pragma unroll 4
        for (int i = threadIdx.x; i < n4; i += kThreads)
            reinterpret_cast<float4*>(sm_ml)[i] = __ldcg(src + i);
    }
    __syncthreads();
    // 2. per row: global max, total sum, and the weight of every split (8 lanes per row)
    {
        const int row = threadIdx.x >> 3, part = threadIdx.x & 7;
        float m = kNegBig;
        for (int sp = part; sp < p.nsplit; sp += 8) m = fmaxf(m, sm_ml[(sp * kRows + row) * 2]);
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
        float l = 0.f;
        for (int sp = part; sp < p.nsplit; sp += 8)
            l = fmaf(sm_ml[(sp * kRows + row) * 2 + 1], fast_exp2(sm_ml[(sp * kRows + row) * 2] - m), l);
        l += __shfl_xor_sync(0xffffffffu, l, 1);
        l += __shfl_xor_sync(0xffffffffu, l, 2);
        l += __shfl_xor_sync(0xffffffffu, l, 4);
        const float inv = l > 0.f ? __frcp_rn(l) : 0.f;
        for (int sp = part; sp < p.nsplit; sp += 8)
            sm_ml[(sp * kRows + row) * 2] = fast_exp2(sm_ml[(sp * kRows + row) * 2] - m) * inv;
    }
    __syncthreads();
    // 3. O[row][d..d+3] = sum over splits of weight * partial O, float4 per thread, 8 loads in flight
    for (int idx = threadIdx.x; idx < G * (kD / 4); idx += kThreads) {
        const int row = idx / (kD / 4), d = (idx - row * (kD / 4)) * 4;
        if (d >= D) continue;
        const float4* src = reinterpret_cast<const float4*>(p.part_o + (ubase + row) * kD + d);
        const int64_t step = static_cast<int64_t>(kRows) * kD / 4;  // float4s between splits
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int sp = 0;
        for (; sp + 8 <= p.nsplit; sp += 8) {
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = __ldcg(src + (sp + u) * step);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float f = sm_ml[((sp + u) * kRows + row) * 2];
                acc.x = fmaf(x[u].x, f, acc.x); acc.y = fmaf(x[u].y, f, acc.y);
                acc.z = fmaf(x[u].z, f, acc.z); acc.w = fmaf(x[u].w, f, acc.w);
            }
        }
        {
            float4 x[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                x[u] = sp + u < p.nsplit ? __ldcg(src + (sp + u) * step) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float f = sp + u < p.nsplit ? sm_ml[((sp + u) * kRows + row) * 2] : 0.f;
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
    if (threadIdx.x == 0) p.counters[unit] = 0;  // re-arm for the next launch on this workspace
}

}  // namespace decode
}  // namespace fa100
