// fa_sm100_rope.cuh — rotary position embedding on q and k in one pass.
//
// The step in front of the attention op in the reference's model path is
//     q_embed = q * cos + rotate_half(q) * sin        (models/rope_attn_fwd.py:8-12, 34-38, 87-88)
// for q and k, evaluated by PyTorch as two slices, a negation, a cat, two multiplies and an add per
// tensor: six full passes over q and k and five temporaries.  It is pure HBM traffic; this kernel
// does it in one pass (read q, k once, write q', k' once; cos/sin rows are shared by all heads of a
// position and stay in L1/L2), on the [B, S, H, D]-memory views the projections produce, without
// touching their strides.
//
// Arithmetic is the reference's, rounding for rounding: PyTorch evaluates each 16-bit op in fp32
// and rounds its result to the tensor dtype, so  out = rn(rn(x·cos) + rn(x_rot·sin))  with
// x_rot = (-x2, x1).  The kernel reproduces exactly these three roundings — results are bit-equal
// to the eager expression (tests/test_rope.py).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace fa100 {
namespace rope {

constexpr int kThreads = 256;

struct RopeParams {
    const void* q;
    const void* k;
    void* q_out;
    void* k_out;
    const void* cos;
    const void* sin;
    int batch, heads_q, heads_kv, seqlen, headdim;
    int64_t q_s[3], k_s[3], qo_s[3], ko_s[3];  // element strides: batch, head, sequence
    int64_t cs_sb, cs_ss;                      // cos / sin: batch (0 = broadcast), sequence
    long long total;                           // B * S * (Hq + Hkv) * (D / 16) work items
};

// out = rn(rn(x * c) + rn(r * s)) on a pair of packed 16-bit values.  The packed hardware ops
// round the exact product / sum once; PyTorch's "compute in fp32, round to the dtype" gives the
// same bits (the fp32 product of two 16-bit values is exact, and an fp32 sum that is inexact has
// an addend too small to reach a rounding midpoint of the larger one).  The _rn forms keep ptxas
// from fusing multiply and add into one FMA, which would drop a rounding.
template <bool kBf16>
__device__ __forceinline__ uint32_t rot2(uint32_t x, uint32_t r, uint32_t rsign, uint32_t c, uint32_t s) {
    r ^= rsign;  // negate both lanes (0x80008000) or neither (0)
    if constexpr (kBf16) {
        using V = __nv_bfloat162;
        const V a = __hmul2_rn(*reinterpret_cast<const V*>(&x), *reinterpret_cast<const V*>(&c));
        const V b = __hmul2_rn(*reinterpret_cast<const V*>(&r), *reinterpret_cast<const V*>(&s));
        const V o = __hadd2_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&o);
    } else {
        using V = __half2;
        const V a = __hmul2_rn(*reinterpret_cast<const V*>(&x), *reinterpret_cast<const V*>(&c));
        const V b = __hmul2_rn(*reinterpret_cast<const V*>(&r), *reinterpret_cast<const V*>(&s));
        const V o = __hadd2_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&o);
    }
}

// Four (batch, position) rows — all heads of q and k, contiguous in the [B, S, H, D] memory the
// projections write — per loop turn of a CTA (their loads are issued together); a thread owns one 16-byte chunk of the first half
// of a head and its partner in the second half.  32-bit index arithmetic only: at 64 B of traffic
// per item the budget is ~350 thread instructions per item per SM (HBM 6.5 TB/s), which 64-bit
// divisions and an fp32 emulation of the three roundings had used up (4.5 TB/s measured).
template <bool kBf16>
__global__ void __launch_bounds__(kThreads)
fa_rope_sm100(const RopeParams p) {
    const int half = p.headdim / 2;
    const unsigned chunks = p.headdim / 16;  // 16-byte chunks in one half of a head
    const unsigned items = (p.heads_q + p.heads_kv) * chunks;
    const unsigned rows = static_cast<unsigned>(p.batch) * p.seqlen;
    constexpr uint32_t kNeg = 0x80008000u;
    constexpr int kRowsPerTurn = 4;  // independent rows per loop turn: 8 data loads in flight per thread
    // Few heads or a small head dimension (Qwen2-0.5B: 16 heads x 4 chunks = 64 items a row): the
    // block covers rpp rows side by side instead of leaving three quarters of its threads idle.
    const unsigned rpp = items >= kThreads ? 1u : kThreads / items;
    const unsigned sub = rpp > 1 ? threadIdx.x / items : 0u;
    const unsigned j0 = threadIdx.x - sub * items;
    const unsigned jstep = rpp > 1 ? items : static_cast<unsigned>(kThreads);
    const unsigned rstep = gridDim.x * rpp;
    if (sub >= rpp) return;
    for (unsigned row0 = blockIdx.x * rpp + sub; row0 < rows; row0 += kRowsPerTurn * rstep) {
        for (unsigned j = j0; j < items; j += jstep) {
            const unsigned h = j / chunks, c = j - h * chunks;
            const bool is_q = h < static_cast<unsigned>(p.heads_q);
            const unsigned hh = is_q ? h : h - p.heads_q;
            const int64_t* ss = is_q ? p.q_s : p.k_s;
            const int64_t* os = is_q ? p.qo_s : p.ko_s;
            const uint16_t* sbase = reinterpret_cast<const uint16_t*>(is_q ? p.q : p.k) + hh * ss[1] + c * 8;
            uint16_t* dbase = reinterpret_cast<uint16_t*>(is_q ? p.q_out : p.k_out) + hh * os[1] + c * 8;
            uint4 x1[kRowsPerTurn], x2[kRowsPerTurn], c1[kRowsPerTurn], c2[kRowsPerTurn], s1[kRowsPerTurn], s2[kRowsPerTurn];
            uint16_t* dst[kRowsPerTurn];
#pragma unroll
            for (int u = 0; u < kRowsPerTurn; ++u) {
                const unsigned row = row0 + u * rstep;
                dst[u] = nullptr;
                if (row < rows) {
                    const unsigned b = row / static_cast<unsigned>(p.seqlen), s = row - b * p.seqlen;
                    const uint16_t* src = sbase + b * ss[0] + s * ss[2];
                    dst[u] = dbase + b * os[0] + s * os[2];
                    const int64_t cs = b * p.cs_sb + s * p.cs_ss + c * 8;
                    const uint16_t* cosr = reinterpret_cast<const uint16_t*>(p.cos) + cs;
                    const uint16_t* sinr = reinterpret_cast<const uint16_t*>(p.sin) + cs;
                    x1[u] = *reinterpret_cast<const uint4*>(src);
                    x2[u] = *reinterpret_cast<const uint4*>(src + half);
                    c1[u] = __ldg(reinterpret_cast<const uint4*>(cosr));
                    c2[u] = __ldg(reinterpret_cast<const uint4*>(cosr + half));
                    s1[u] = __ldg(reinterpret_cast<const uint4*>(sinr));
                    s2[u] = __ldg(reinterpret_cast<const uint4*>(sinr + half));
                }
            }
#pragma unroll
            for (int u = 0; u < kRowsPerTurn; ++u) {
                if (dst[u] == nullptr) continue;
                uint4 o1, o2;
                // first half: x1·cos1 + (−x2)·sin1 ; second half: x2·cos2 + x1·sin2
                o1.x = rot2<kBf16>(x1[u].x, x2[u].x, kNeg, c1[u].x, s1[u].x);
                o1.y = rot2<kBf16>(x1[u].y, x2[u].y, kNeg, c1[u].y, s1[u].y);
                o1.z = rot2<kBf16>(x1[u].z, x2[u].z, kNeg, c1[u].z, s1[u].z);
                o1.w = rot2<kBf16>(x1[u].w, x2[u].w, kNeg, c1[u].w, s1[u].w);
                o2.x = rot2<kBf16>(x2[u].x, x1[u].x, 0u, c2[u].x, s2[u].x);
                o2.y = rot2<kBf16>(x2[u].y, x1[u].y, 0u, c2[u].y, s2[u].y);
                o2.z = rot2<kBf16>(x2[u].z, x1[u].z, 0u, c2[u].z, s2[u].z);
                o2.w = rot2<kBf16>(x2[u].w, x1[u].w, 0u, c2[u].w, s2[u].w);
                *reinterpret_cast<uint4*>(dst[u]) = o1;
                *reinterpret_cast<uint4*>(dst[u] + half) = o2;
            }
        }
    }
}

}  // namespace rope
}  // namespace fa100
