// tcgen05.mma issue-rate microbenchmark (debug tool): cycles per MMA for the shapes the forward
// kernel could use.  Operands are whatever is in shared memory / TMEM (values do not matter).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/ubench_mma tools/ubench_mma.cu
#include <cstdio>
#include <cuda_runtime.h>

#include "../flash_attention_cute_b200/csrc/fa_sm100_ptx.cuh"

using namespace fa100;

// mode 0: SS (A, B from smem), K-major both.  mode 1: TS (A from TMEM), B MN-major (like P·V).
template <int kN, int kMode>
__global__ void __launch_bounds__(128, 1) bench(long long* out, int iters, int distinct) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t base = smem_u32(smem);
    const uint32_t bar = base + 196608;
    const uint32_t tptr = bar + 16;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    if (threadIdx.x < 32) {
        tmem_alloc(tptr, 512);
        tmem_relinquish_alloc_permit();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + 196608 + 16);
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = make_idesc_f16(true, 128, kN, false, kMode == 1);
        long long t0 = clock64_now();
        // descriptors precomputed: the loop body is 8 back-to-back tcgen05.mma
        uint64_t da[8], db[8];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
            da[ks] = make_smem_desc_sw128(base + off, 16, 1024);
            db[ks] = kMode == 0 ? make_smem_desc_sw128(base + 32768 + (ks & 3) * 32768 + off, 16, 1024)
                                : make_smem_desc_sw128(base + 32768 + (ks & 3) * 32768 + ks * 2048, 16384, 1024);
        }
        (void)distinct;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                if constexpr (kMode == 0) umma_ss(tmem + (ks & 1) * 128, da[ks], db[ks], idesc, 1);
                else umma_ts(tmem + 256, tmem + ks * 8, db[ks], idesc, 1);
            }
        }
        tc_commit(bar);
        long long t1 = clock64_now();
        unsigned int dummy_status = 0;
        while (!mbar_try_wait(bar, 0)) {}
        long long t2 = clock64_now();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
        (void)dummy_status;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) {
        tc_fence_after_sync();
        tmem_dealloc(tmem, 512);
    }
}

template <int kN, int kMode>
void run(const char* name, int ctas) {
    long long* out;
    cudaMalloc(&out, 16);
    const int smem = 196608 + 64;
    cudaFuncSetAttribute(bench<kN, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 512;
    for (int rep = 0; rep < 2; ++rep) bench<kN, kMode><<<ctas, 128, smem>>>(out, iters, 4);
    long long h[2] = {0, 0};
    cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaDeviceSynchronize();
    const double per = double(h[1]) / (iters * 8.0);
    printf("%-28s ctas=%3d  %7.2f cycles/MMA (issue loop %7.2f)  ideal %d  -> %5.1f%% of peak  [%s]\n", name, ctas, per,
           double(h[0]) / (iters * 8.0), kN / 2, 100.0 * (kN / 2) / per, cudaGetErrorString(e));
    cudaFree(out);
}

int main() {
    for (int ctas : {1, 148}) {
        run<128, 0>("SS M128 N128 K16 (QK^T)", ctas);
        run<64, 0>("SS M128 N64  K16 (QK^T/64)", ctas);
        run<256, 0>("SS M128 N256 K16", ctas);
        run<128, 1>("TS M128 N128 K16 (PV)", ctas);
        run<64, 1>("TS M128 N64  K16 (PV D=64)", ctas);
    }
    return 0;
}
