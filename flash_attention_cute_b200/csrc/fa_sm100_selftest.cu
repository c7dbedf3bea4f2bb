// UMMA / TMA / TMEM primitive self-test (C ABI: fa_sm100_selftest_umma, include/fa_sm100.h).
// One CTA runs the exact descriptor recipes of fa_sm100_fwd.cuh in isolation:
//   S = Q·Kᵀ      SS-mode tcgen05.mma, K-major 128B-swizzled Q and K panels written by TMA
//   P = T(S * p_scale) packed two-per-column into TMEM by tcgen05.st
//   O = P·V       TS-mode tcgen05.mma, A from TMEM, V MN-major 128B-swizzled panels
// and dumps S and O as fp32 so tests/test_primitives.py can compare against torch matmuls.
#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/fa_sm100.h"
#include "fa_sm100_ptx.cuh"

extern "C" int fa_sm100_internal_wait_ctx(int device, fa100::WaitCtx* out);
extern "C" int fa_sm100_internal_encode_2d(CUtensorMap* map, int is_bf16, const void* base,
                                           uint64_t rows, uint64_t cols);
extern "C" void fa_sm100_internal_count_launch(void);
extern "C" int fa_sm100_internal_fail(int code, const char* msg);

namespace fa100 {

struct SelftestParams {
    CUtensorMap tm_q, tm_k, tm_v;  // 2-D (d, row) maps, box (64, 128)
    float* s_out;
    float* o_out;
    float p_scale;
    uint32_t v_lbo, v_sbo;
    int32_t headdim;
    WaitCtx wait;
};

template <int kD, bool kBf16>
__global__ void __launch_bounds__(128, 1) selftest_kernel(const __grid_constant__ SelftestParams p) {
    constexpr int kPanels = kD / 64;
    constexpr int kPanelBytes = 128 * 128;
    constexpr uint32_t kTile = kPanels * kPanelBytes;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t sq = base, sk = base + kTile, sv = base + 2 * kTile;
    const uint32_t bars = base + 3 * kTile;  // [0] tma full, [1] S done, [2] O done
    const uint32_t tptr = bars + 64;
    const int warp = threadIdx.x >> 5;

    if (threadIdx.x == 0) {
        mbar_init(bars + 0, 1);
        mbar_init(bars + 8, 1);
        mbar_init(bars + 16, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        tmem_alloc(tptr, 256);
        tmem_relinquish_alloc_permit();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (tptr - base));

    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bars, 3 * kTile);
        for (int pn = 0; pn < kPanels; ++pn) {
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                ::"r"(sq + pn * kPanelBytes), "l"(reinterpret_cast<uint64_t>(&p.tm_q)), "r"(bars), "r"(pn * 64), "r"(0) : "memory");
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                ::"r"(sk + pn * kPanelBytes), "l"(reinterpret_cast<uint64_t>(&p.tm_k)), "r"(bars), "r"(pn * 64), "r"(0) : "memory");
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                ::"r"(sv + pn * kPanelBytes), "l"(reinterpret_cast<uint64_t>(&p.tm_v)), "r"(bars), "r"(pn * 64), "r"(0) : "memory");
        }
        (void)mbar_wait(bars, 0, p.wait, 101);  // a timeout is reported through the status word
        tc_fence_after_sync();
        constexpr uint32_t idesc_qk = make_idesc_f16(kBf16, 128, 128, false, false);
#pragma unroll
        for (int ks = 0; ks < kD / 16; ++ks) {
            const uint32_t off = (ks >> 2) * kPanelBytes + (ks & 3) * 32;
            umma_ss(tmem, make_smem_desc_sw128(sq + off, 16, 1024),
                    make_smem_desc_sw128(sk + off, 16, 1024), idesc_qk, ks > 0);
        }
        tc_commit(bars + 8);
    }

    // every thread owns one row
    const int r = threadIdx.x;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    (void)mbar_wait(bars + 8, 0, p.wait, 102);
    tc_fence_after_sync();
    uint32_t pk[64];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint32_t s[32];
        tmem_ld32(tmem + lane_base + c * 32, s);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) p.s_out[r * 128 + c * 32 + i] = __uint_as_float(s[i]);
#pragma unroll
        for (int i = 0; i < 32; i += 2)
            pk[c * 16 + i / 2] =
                pack2<kBf16>(__uint_as_float(s[i]) * p.p_scale, __uint_as_float(s[i + 1]) * p.p_scale);
    }
    tmem_st32(tmem + lane_base, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
    tmem_st32(tmem + lane_base + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
    tmem_wait_st();
    tc_fence_before_sync();
    __syncthreads();

    if (threadIdx.x == 0) {
        tc_fence_after_sync();
        constexpr uint32_t idesc_pv = make_idesc_f16(kBf16, 128, kD, false, true);
        const uint32_t lbo = p.v_lbo ? p.v_lbo : kPanelBytes;
        const uint32_t sbo = p.v_sbo ? p.v_sbo : 1024;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
            umma_ts(tmem + 128, tmem + ks * 8, make_smem_desc_sw128(sv + ks * 2048, lbo, sbo),
                    idesc_pv, ks > 0);
        tc_commit(bars + 16);
    }
    (void)mbar_wait(bars + 16, 0, p.wait, 103);
    tc_fence_after_sync();
#pragma unroll
    for (int c = 0; c < kD / 32; ++c) {
        uint32_t o[32];
        tmem_ld32(tmem + lane_base + 128 + c * 32, o);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (c * 32 + i < p.headdim) p.o_out[r * p.headdim + c * 32 + i] = __uint_as_float(o[i]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after_sync();
        tmem_dealloc(tmem, 256);
    }
}

}  // namespace fa100

extern "C" int fa_sm100_selftest_umma(const void* q, const void* k, const void* v, float* s_out,
                                      float* o_out, int32_t headdim, int32_t dtype, float p_scale,
                                      uint32_t v_lbo, uint32_t v_sbo, int32_t device, void* stream) {
    if (headdim <= 0 || headdim > 128 || headdim % 8)
        return fa_sm100_internal_fail(FA_SM100_ERR_INVALID_ARGUMENT, "selftest: bad headdim");
    fa100::SelftestParams sp;
    int rc0 = fa_sm100_internal_wait_ctx(device, &sp.wait);
    if (rc0) return rc0;
    cudaSetDevice(device);
    const int bf16 = dtype == FA_SM100_BF16;
    int rc;
    if ((rc = fa_sm100_internal_encode_2d(&sp.tm_q, bf16, q, 128, headdim))) return rc;
    if ((rc = fa_sm100_internal_encode_2d(&sp.tm_k, bf16, k, 128, headdim))) return rc;
    if ((rc = fa_sm100_internal_encode_2d(&sp.tm_v, bf16, v, 128, headdim))) return rc;
    sp.s_out = s_out;
    sp.o_out = o_out;
    sp.p_scale = p_scale;
    sp.v_lbo = v_lbo;
    sp.v_sbo = v_sbo;
    sp.headdim = headdim;
    const int kd = headdim <= 64 ? 64 : 128;
    const int smem = 3 * (kd / 64) * 128 * 128 + 128 + 1024;
    using Fn = void (*)(const fa100::SelftestParams);
    Fn fn = kd == 128 ? (bf16 ? fa100::selftest_kernel<128, true> : fa100::selftest_kernel<128, false>)
                      : (bf16 ? fa100::selftest_kernel<64, true> : fa100::selftest_kernel<64, false>);
    cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(fn),
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return fa_sm100_internal_fail(FA_SM100_ERR_CUDA, cudaGetErrorString(e));
    fn<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(sp);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fa_sm100_internal_fail(FA_SM100_ERR_CUDA, cudaGetErrorString(e));
    fa_sm100_internal_count_launch();
    return FA_SM100_OK;
}
