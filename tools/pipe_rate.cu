// pipe_rate.cu -- per-SMSP issue interval and dependent latency of the instructions the softmax
// warps are made of (FFMA, FFMA2, FADD2, MUFU.EX2, FMNMX, FMNMX3, F2FP pack, LEA), one warp per SMSP.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) { uint64_t d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float fmx(float a, float b) { float d; asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float fmx3(float a, float b, float c) { float d; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ uint32_t pk(float a, float b) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ uint32_t ex2h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t ex2b2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t pkh(float a, float b) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ uint32_t hfma2(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint32_t lea(uint32_t a, uint32_t b) { uint32_t r; asm volatile("{ .reg .u32 t; shl.b32 t, %1, 23; add.u32 %0, t, %2; }" : "=r"(r) : "r"(a), "r"(b)); return r; }

// OP: which instruction; CH: number of independent chains (ILP)
template <int OP, int CH>
__global__ void k(float* out, long long* cyc, float seed)
{
    float v[CH]; uint64_t w[CH]; uint32_t u[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { v[i] = seed + i + threadIdx.x; w[i] = ((uint64_t)__float_as_uint(v[i]) << 32) | __float_as_uint(v[i] * 0.5f); u[i] = i + threadIdx.x; }
    const float c1 = seed * 0.999f, c2 = seed * 0.001f;
    const uint64_t C1 = ((uint64_t)__float_as_uint(c1) << 32) | __float_as_uint(c1), C2 = ((uint64_t)__float_as_uint(c2) << 32) | __float_as_uint(c2);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 256; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (OP == 0) v[i] = ffma(v[i], c1, c2);
                if (OP == 1) w[i] = f2_fma(w[i], C1, C2);
                if (OP == 2) w[i] = f2_add(w[i], C2);
                if (OP == 3) v[i] = ex2(v[i]);
                if (OP == 4) v[i] = fmx(v[i], c1);
                if (OP == 5) v[i] = fmx3(v[i], c1, c2);
                if (OP == 6) u[i] = pk(__uint_as_float(u[i]), c1);
                if (OP == 7) u[i] = lea(u[i], u[i]);
                if (OP == 8) { v[i] = ex2(v[i]); w[i] = f2_fma(w[i], C1, C2); }                      // MUFU + FFMA2 co-issue
                if (OP == 9) { v[i] = ex2(v[i]); w[i] = f2_fma(w[i], C1, C2); w[i] = f2_add(w[i], C2); u[i] = pk(__uint_as_float(u[i]), c1); }
                if (OP == 10) { w[i] = f2_fma(w[i], C1, C2); u[i] = lea(u[i], u[i]); }               // FMA pipe + ALU pipe
                if (OP == 11) { v[i] = ffma(v[i], c1, c2); u[i] = lea(u[i], u[i]); }
                if (OP == 12) u[i] = ex2h2(u[i]);                                                    // 2 x MUFU.EX2.F16 (one per half)
                if (OP == 13) u[i] = ex2b2(u[i]);                                                    // 2 x MUFU.EX2.BF16
                if (OP == 14) u[i] = pkh(__uint_as_float(u[i]), c1);                                 // F2FP.F16.F32.PACK_AB
                if (OP == 15) u[i] = hfma2(u[i], u[i], u[i]);                                        // HFMA2
            }
    }
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += v[i] + __uint_as_float((uint32_t)w[i]) + __uint_as_float((uint32_t)(w[i] >> 32)) + __uint_as_float(u[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int OP, int CH> void run(const char* name, int warps, float* out, long long* cyc)
{
    k<OP, CH><<<1, 32 * warps>>>(out, cyc, 0.5f);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-28s chains %2d warps/SMSP %d: %.2f clk per warp-instr-group\n", name, CH, warps / 4, (double)h / (256.0 * 4 * CH));
}
#define RUN(OP, NAME) run<OP, 1>(NAME, 4, out, cyc); run<OP, 8>(NAME, 4, out, cyc); run<OP, 8>(NAME, 8, out, cyc);

int main()
{
    float* out; long long* cyc;
    cudaMalloc(&out, 4096 * 4); cudaMalloc(&cyc, 8);
    RUN(0, "FFMA") RUN(1, "FFMA2") RUN(2, "FADD2") RUN(3, "MUFU.EX2") RUN(4, "FMNMX") RUN(5, "FMNMX3") RUN(6, "F2FP.BF16.PACK") RUN(7, "SHL+IADD (LEA)")
    RUN(12, "ex2.f16x2 (2 MUFU.EX2.F16)") RUN(13, "ex2.bf16x2 (2 MUFU.EX2.BF16)") RUN(14, "F2FP.F16.PACK") RUN(15, "HFMA2")
    RUN(8, "MUFU + FFMA2") RUN(9, "MUFU+FFMA2+FADD2+F2FP") RUN(10, "FFMA2 + LEA") RUN(11, "FFMA + LEA")
    return 0;
}
