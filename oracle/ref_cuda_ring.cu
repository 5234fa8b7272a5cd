// ref_cuda_ring.cu -- C-linkage shim around the reference's UNMODIFIED CUDA ring-attention kernel.
//
// TEST / MEASUREMENT INFRASTRUCTURE ONLY (see naive_oracle.c header).  No attention code here: the reference kernel file is
// #included where it lies under REF_ROOT (src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:13-140,
// ring_attention_forward_kernel with its (m, l, unnormalised O) carry and the last_step normalisation) together with the
// reference's own util/nccl_utils.h and the mpi.h it vendors under headers/openmpi.  The ranks of a P-way ring are run ONE
// AFTER THE OTHER on one GPU: for every rank the kernel is launched P times with the launch configuration and the K/V
// schedule of the reference's host loop (ring_attention_forward, :159-214: Br = Bc = 64, 4 warps, block (rank - step) mod P,
// last_step on the final launch); the K/V block a real ring would have RECEIVED is addressed in place in the full K / V.
// That is what tests/test_gpu_refcuda.py compares our ring path with (north_star: "checked ... against ... its CUDA
// FA2/ring kernels").  Built by `make -C oracle refcuda` into oracle/_ref/ (git-ignored, travels to the GPU box).
#include <assert.h>
#include <vector>
#include <iostream>
#include <cmath>
#include <cstdlib>
#include <cuda_runtime.h>
#include "03_flash_attention_v2_ring/common/ring_attention_kernel.cu"

// Q, K, V: [N, d] fp32 on the device, N = P * n;  O: [N, d];  L, M: [N + 64] floats (the kernel reads L / M of the rows
// that pad the last 64-row block, :68-71).  Returns cudaGetLastError() after a device synchronisation, -1 for bad shapes.
extern "C" __attribute__((visibility("default")))
int ref_cuda_ring_forward_emulated(const float* Q, const float* K, const float* V, float* O, float* L, float* M,
                                   int N, int d, int P, float scale)
{
    if (d > 64 || P < 1 || N < P || N % P != 0) return -1;
    const int n = N / P;
    constexpr int Br = 64, Bc = 64, num_warps = 4;
    const int smem_size = 2 * Bc * d * sizeof(float);
    const dim3 grid(div_up(n, Br), 1, 1), block(num_warps * WARP_SIZE, 1, 1);
    cudaMemset(O, 0, (size_t)N * d * sizeof(float));
    cudaMemset(L, 0, (size_t)(N + 64) * sizeof(float));
    init_array<<<div_up(N + 64, 256), 256>>>(M, N + 64, -INFINITY);
    for (int rank = 0; rank < P; ++rank)
        for (int step = 0; step < P; ++step) {
            const int kv_block_idx = (rank - step + P) % P;
            ring_attention_forward_kernel<Br, Bc, 64, num_warps><<<grid, block, smem_size>>>(
                Q + (size_t)rank * n * d, K + (size_t)kv_block_idx * n * d, V + (size_t)kv_block_idx * n * d,
                O + (size_t)rank * n * d, L + (size_t)rank * n, M + (size_t)rank * n, n, n, d, scale, step == P - 1);
        }
    cudaDeviceSynchronize();
    return (int)cudaGetLastError();
}
