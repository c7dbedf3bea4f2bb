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

template <bool kBf16>
struct Cvt;
template <>
struct Cvt<true> {
    static __device__ __forceinline__ float2 up(uint32_t x) {
        return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&x));
    }
    static __device__ __forceinline__ float rn(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
    static __device__ __forceinline__ uint32_t down(float a, float b) {
        __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&t);
    }
};
template <>
struct Cvt<false> {
    static __device__ __forceinline__ float2 up(uint32_t x) {
        return __half22float2(*reinterpret_cast<const __half2*>(&x));
    }
    static __device__ __forceinline__ float rn(float x) { return __half2float(__float2half_rn(x)); }
    static __device__ __forceinline__ uint32_t down(float a, float b) {
        __half2 t = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&t);
    }
};

// out = rn(rn(x * c) + rn(r * s)) on a pair of packed 16-bit values
template <bool kBf16>
__device__ __forceinline__ uint32_t rot2(uint32_t x, uint32_t r, float rsign, uint32_t c, uint32_t s) {
    using C = Cvt<kBf16>;
    const float2 xf = C::up(x), rf = C::up(r), cf = C::up(c), sf = C::up(s);
    const float a0 = C::rn(xf.x * cf.x), a1 = C::rn(xf.y * cf.y);
    const float b0 = C::rn(rsign * rf.x * sf.x), b1 = C::rn(rsign * rf.y * sf.y);
    return C::down(a0 + b0, a1 + b1);
}

template <bool kBf16>
__global__ void __launch_bounds__(kThreads)
fa_rope_sm100(const RopeParams p) {
    const int half = p.headdim / 2;
    const int chunks = p.headdim / 16;  // 16-byte chunks in one half of a head
    const int htot = p.heads_q + p.heads_kv;
    const long long stride = static_cast<long long>(gridDim.x) * kThreads;
    for (long long i = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x; i < p.total; i += stride) {
        const int c = static_cast<int>(i % chunks);
        long long t = i / chunks;
        const int h = static_cast<int>(t % htot);
        t /= htot;
        const int s = static_cast<int>(t % p.seqlen);
        const int b = static_cast<int>(t / p.seqlen);
        const bool is_q = h < p.heads_q;
        const int hh = is_q ? h : h - p.heads_q;
        const int64_t* ss = is_q ? p.q_s : p.k_s;
        const int64_t* os = is_q ? p.qo_s : p.ko_s;
        const uint16_t* src = reinterpret_cast<const uint16_t*>(is_q ? p.q : p.k) + b * ss[0] + hh * ss[1] + s * ss[2] + c * 8;
        uint16_t* dst = reinterpret_cast<uint16_t*>(is_q ? p.q_out : p.k_out) + b * os[0] + hh * os[1] + s * os[2] + c * 8;
        const int64_t cs = b * p.cs_sb + s * p.cs_ss + c * 8;
        const uint16_t* cosp = reinterpret_cast<const uint16_t*>(p.cos) + cs;
        const uint16_t* sinp = reinterpret_cast<const uint16_t*>(p.sin) + cs;
        const uint4 x1 = *reinterpret_cast<const uint4*>(src);
        const uint4 x2 = *reinterpret_cast<const uint4*>(src + half);
        const uint4 c1 = __ldg(reinterpret_cast<const uint4*>(cosp));
        const uint4 c2 = __ldg(reinterpret_cast<const uint4*>(cosp + half));
        const uint4 s1 = __ldg(reinterpret_cast<const uint4*>(sinp));
        const uint4 s2 = __ldg(reinterpret_cast<const uint4*>(sinp + half));
        uint4 o1, o2;
        // first half: x1·cos1 + (−x2)·sin1 ; second half: x2·cos2 + x1·sin2
        o1.x = rot2<kBf16>(x1.x, x2.x, -1.f, c1.x, s1.x);
        o1.y = rot2<kBf16>(x1.y, x2.y, -1.f, c1.y, s1.y);
        o1.z = rot2<kBf16>(x1.z, x2.z, -1.f, c1.z, s1.z);
        o1.w = rot2<kBf16>(x1.w, x2.w, -1.f, c1.w, s1.w);
        o2.x = rot2<kBf16>(x2.x, x1.x, 1.f, c2.x, s2.x);
        o2.y = rot2<kBf16>(x2.y, x1.y, 1.f, c2.y, s2.y);
        o2.z = rot2<kBf16>(x2.z, x1.z, 1.f, c2.z, s2.z);
        o2.w = rot2<kBf16>(x2.w, x1.w, 1.f, c2.w, s2.w);
        *reinterpret_cast<uint4*>(dst) = o1;
        *reinterpret_cast<uint4*>(dst + half) = o2;
    }
}

}  // namespace rope
}  // namespace fa100
