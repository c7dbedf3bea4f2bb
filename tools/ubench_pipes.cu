// Pipe-rate microbenchmark for the softmax inner loop (debug tool, not part of the product).
// For each op: ILP independent dependency chains per thread, W warps per SM sub-partition,
// one CTA on one SM; prints cycles per warp-instruction per sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes tools/ubench_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 512

template <int OP>
__device__ __forceinline__ void step(float& a, float& b, uint64_t& p, float k0, float k1) {
    if constexpr (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
    if constexpr (OP == 1) asm volatile("fma.rn.ftz.f32 %0, %0, %1, %2;" : "+f"(a) : "f"(k0), "f"(k1));
    if constexpr (OP == 2) asm volatile("fma.rn.ftz.f32 %0, %0, %1, 0f3F317218;" : "+f"(a) : "f"(k0));
    if constexpr (OP == 3) {
        uint64_t kk0, kk1;
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(kk0) : "f"(k0));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(kk1) : "f"(k1));
        asm volatile("fma.rn.ftz.f32x2 %0, %0, %1, %2;" : "+l"(p) : "l"(kk0), "l"(kk1));
    }
    if constexpr (OP == 4) {
        uint64_t kk0;
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(kk0) : "f"(k0));
        asm volatile("add.rn.ftz.f32x2 %0, %0, %1;" : "+l"(p) : "l"(kk0));
    }
    if constexpr (OP == 5) asm volatile("add.rn.ftz.f32 %0, %0, %1;" : "+f"(a) : "f"(k0));
    if constexpr (OP == 6) asm volatile("max.f32 %0, %0, %1;" : "+f"(a) : "f"(k0));
    if constexpr (OP == 7) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a) : "f"(k0), "f"(k1));
    if constexpr (OP == 8) {  // shl+add (LEA)
        uint32_t x = __float_as_uint(a), y = __float_as_uint(b);
        asm volatile("{.reg .b32 t; shl.b32 t, %1, 23; add.s32 %0, %0, t;}" : "+r"(x) : "r"(y));
        a = __uint_as_float(x);
    }
    if constexpr (OP == 9) {  // pack two f32 -> bf16x2
        uint32_t r;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a), "f"(b));
        a = __uint_as_float(r);
    }
    if constexpr (OP == 10) asm volatile("add.rm.ftz.f32 %0, %0, 0f4B400000;" : "+f"(a));
    if constexpr (OP == 11) {  // f32x2 add with RM
        uint64_t kk0;
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(kk0) : "f"(k0));
        asm volatile("add.rm.ftz.f32x2 %0, %0, %1;" : "+l"(p) : "l"(kk0));
    }
    if constexpr (OP == 12) {  // scalar emulated exp2 (imm forms), 1 element
        float x = fmaxf(a, -127.f);
        float fl, r, fr, q;
        asm volatile("add.rm.ftz.f32 %0, %1, 0f4B400000;" : "=f"(fl) : "f"(x));
        asm volatile("add.rn.ftz.f32 %0, %1, 0fCB400000;" : "=f"(r) : "f"(fl));
        asm volatile("sub.rn.ftz.f32 %0, %1, %2;" : "=f"(fr) : "f"(x), "f"(r));
        asm volatile("fma.rn.ftz.f32 %0, %1, 0f3D9DD553, 0f3E691BC5;" : "=f"(q) : "f"(fr));
        asm volatile("fma.rn.ftz.f32 %0, %0, %1, 0f3F31F32C;" : "+f"(q) : "f"(fr));
        asm volatile("fma.rn.ftz.f32 %0, %0, %1, 0f3F800000;" : "+f"(q) : "f"(fr));
        uint32_t xi = __float_as_uint(q), fi = __float_as_uint(fl);
        asm volatile("{.reg .b32 t; shl.b32 t, %1, 23; add.s32 %0, %0, t;}" : "+r"(xi) : "r"(fi));
        a = -__uint_as_float(xi);
    }
    if constexpr (OP == 13) {  // packed emulated exp2, 2 elements (as in the kernel)
        float x0, x1;
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(x0), "=f"(x1) : "l"(p));
        x0 = fmaxf(x0, -127.f);
        x1 = fmaxf(x1, -127.f);
        uint64_t xc, magic, fl, r, fr, q, c3, c2, c1, one;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(xc) : "f"(x0), "f"(x1));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(magic) : "f"(12582912.f));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(c3) : "f"(0.07706704f));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(0.22764499f));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(c1) : "f"(0.69511676f));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(one) : "f"(1.f));
        asm volatile("add.rm.ftz.f32x2 %0, %1, %2;" : "=l"(fl) : "l"(xc), "l"(magic));
        asm volatile("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(fl), "l"(magic));
        asm volatile("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(fr) : "l"(xc), "l"(r));
        asm volatile("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(q) : "l"(fr), "l"(c3), "l"(c2));
        asm volatile("fma.rn.ftz.f32x2 %0, %0, %1, %2;" : "+l"(q) : "l"(fr), "l"(c1));
        asm volatile("fma.rn.ftz.f32x2 %0, %0, %1, %2;" : "+l"(q) : "l"(fr), "l"(one));
        float q0, q1, f0, f1;
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(q0), "=f"(q1) : "l"(q));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f0), "=f"(f1) : "l"(fl));
        uint32_t a0 = __float_as_uint(q0) + (__float_as_uint(f0) << 23);
        uint32_t a1 = __float_as_uint(q1) + (__float_as_uint(f1) << 23);
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(-__uint_as_float(a0)), "f"(-__uint_as_float(a1)));
    }
    if constexpr (OP == 14) {  // MUFU + independent scalar emulation side by side (1 + 1 elements)
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b));
        step<12>(a, b, p, k0, k1);
    }
    if constexpr (OP == 15) asm volatile("mul.rn.ftz.f32 %0, %0, %1;" : "+f"(a) : "f"(k0));
}

template <int OP, int ILP>
__global__ void bench(float* out, long long* cycles, float k0, float k1, int warps) {
    if ((threadIdx.x >> 5) >= warps) return;
    float a[ILP], b[ILP];
    uint64_t p[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        a[i] = -0.001f * (threadIdx.x + i);
        b[i] = 0.5f + i;
        p[i] = (uint64_t(__float_as_uint(a[i])) << 32) | __float_as_uint(a[i]);
    }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) step<OP>(a[i], b[i], p[i], k0, k1);
    }
    long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc += a[i] + b[i] + float(p[i] & 0xff);
    out[threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[0] = t1 - t0;
}

template <int OP>
void run(const char* name, int instr_per_step) {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 4096);
    cudaMalloc(&cyc, 8);
    printf("%-34s", name);
    const int wps[] = {1, 2, 4};
    for (int wi = 0; wi < 3; ++wi) {
        int warps = wps[wi] * 4;
        bench<OP, 8><<<1, 512>>>(out, cyc, 0.999f, -0.0001f, warps);
        bench<OP, 8><<<1, 512>>>(out, cyc, 0.999f, -0.0001f, warps);
        long long c = 0;
        cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        // per sub-partition: wps[wi] warps each issued ITERS*8 steps
        double per_step = double(c) / (double(ITERS) * 8 * wps[wi]);
        printf("  w/smsp=%d: %6.2f cyc/step (%5.2f/instr)", wps[wi], per_step, per_step / instr_per_step);
    }
    printf("\n");
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    run<0>("MUFU.EX2", 1);
    run<1>("FFMA 3-reg", 1);
    run<2>("FFMA imm", 1);
    run<15>("FMUL", 1);
    run<3>("FFMA2 (f32x2)", 1);
    run<4>("FADD2 (f32x2)", 1);
    run<11>("FADD2.RM (f32x2)", 1);
    run<5>("FADD", 1);
    run<10>("FADD.RM imm", 1);
    run<6>("FMNMX", 1);
    run<7>("FMNMX3", 1);
    run<8>("SHL+IADD (LEA)", 1);
    run<9>("F2FP.BF16 pack", 1);
    run<12>("emu exp2 scalar (1 elem)", 1);
    run<13>("emu exp2 packed (2 elem)", 1);
    run<14>("MUFU + scalar emu (2 elem)", 1);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
