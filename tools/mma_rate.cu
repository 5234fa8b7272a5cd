// mma_rate.cu -- tcgen05.mma issue/execute rate microbenchmark (one CTA per SM, garbage operands):
// how many SM clocks does one 128xNx16 bf16 MMA take when A/B come from shared memory (SS) or A from
// TMEM (TS)?  Used to separate "tensor pipe is busy" from "operand fetch is the limit" in the
// attention kernels.   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I cuda_flashattention_b200/csrc
#include "ptx.cuh"
#include <stdlib.h>
using namespace b200;

// mode 0: SS, A K-major [128 x 128], B K-major [N x 128]        (S = Q K^T)
// mode 1: TS, A in TMEM, B MN-major [128(k) x N]                (O += P V)
// mode 2: alternate 8 SS + 8 TS                                 (the forward's steady state)
// mode 3: mode 0, and after every group of 8 a wait on an mbarrier that completed long ago  (cost of a ready wait)
// mode 4: mode 3 + tcgen05.fence::after_thread_sync behind the wait  (does the fence hold the issuing thread until its
//         earlier MMAs have drained?  r1(d): the forward traces fit that; this prices it)
// mode 5: mode 0 + the fence alone after every group
__global__ void __launch_bounds__(128, 1) rate_kernel(int mode, int N, int groups, long long* out)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 192 * 1024);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
    for (int i = threadIdx.x; i < 192 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); fence_mbar_init(); mbar_arrive(bar + 1); }   // bar[1]: phase 0 complete
    if (threadIdx.x < 32) { tmem_alloc<512>(slot); tmem_relinquish(); }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x < 32) {
        if (elect_one()) {
            const uint32_t idesc_ss = make_idesc(kFmtBF16, 128, N, false, false);
            const uint32_t idesc_ts = make_idesc(kFmtBF16, 128, N, false, true);
            const uint32_t a_base = smem_u32(smem), b_base = a_base + 32 * 1024, v_base = a_base + 64 * 1024;
            const uint64_t ad = make_sdesc(a_base, 16, 1024, 2), bd = make_sdesc(b_base, 16, 1024, 2);
            const uint64_t vd = make_sdesc(v_base, 128 * 128, 1024, 2);
            const long long t0 = clock64();
            for (int g = 0; g < groups; ++g) {
                if (mode == 0 || mode >= 2) {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)
                        umma_f16_ss(tmem, sdesc_advance(ad, (kk >> 2) * 16384 + (kk & 3) * 32),
                                    sdesc_advance(bd, (kk >> 2) * 16384 + (kk & 3) * 32), idesc_ss, 1);
                }
                if (mode == 1 || mode == 2) {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)
                        umma_f16_ts(tmem + 256, tmem + 128 + 8 * kk, sdesc_advance(vd, kk * 2048), idesc_ts, 1);
                }
                if (mode == 3 || mode == 4) mbar_wait(bar + 1, 0);
                if (mode == 4 || mode == 5) tc_fence_after();
            }
            umma_commit(bar);
            const long long t1 = clock64();
            mbar_wait(bar, 0);
            const long long t2 = clock64();
            if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}

int main()
{
    long long* d; cudaMalloc(&d, 16);
    const int smem = 1024 + 192 * 1024 + 64;
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const char* names[6] = {"SS (A,B smem K-major)", "TS (A tmem, B smem MN-major)", "8 SS + 8 TS alternating",
                            "SS x8 + ready wait", "SS x8 + ready wait + fence", "SS x8 + fence"};
    for (int grid : {1, 148})
        for (int mode = 0; mode < 6; ++mode)
            for (int N : {64, 128, 256}) {
                if (N == 256 && mode != 0) continue;
                const int groups = 64;
                rate_kernel<<<grid, 128, smem>>>(mode, N, groups, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
                const int n_mma = groups * (mode == 2 ? 16 : 8);
                printf("grid %3d  %-30s N=%3d: issue %.1f clk/MMA, complete %.1f clk/MMA (ideal %.0f)\n", grid, names[mode], N,
                       (double)h[0] / n_mma, (double)h[1] / n_mma, 128.0 * N * 16 * 2 / 8192);
            }
    return 0;
}
