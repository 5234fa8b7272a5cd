// ref_cuda_fwd.cu -- C-linkage shim around the reference's UNMODIFIED CUDA FlashAttention-2 forward.
//
// TEST / MEASUREMENT INFRASTRUCTURE ONLY (see naive_oracle.c header).  No attention code here: the reference kernel file
// is #included where it lies under REF_ROOT (src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:13-172, the
// Br = Bc = 32, 64-thread configuration its ring and 1-GPU drivers use) and its host wrapper re-exported with C linkage,
// so that tests/ and tools/ref_cuda_compare.py can run the reference's own GPU kernel on the same inputs as ours
// (north_star: "checked ... against the repo's own CPU naive attention and its CUDA FA2/ring kernels").  Built by
// `make -C oracle refcuda` into oracle/_ref/ (git-ignored, travels to the GPU box).  No reference source is copied.
// (the reference header relies on its includer for these, as its drivers do: 03_attention_1GPU.cu:1-12)
#include <assert.h>
#include <vector>
#include <iostream>
#include <cmath>
#include <cstdlib>
#include <cuda_runtime.h>
#include "03_flash_attention_v2_ring/common/flash_attention_kernel.cu"

extern "C" __attribute__((visibility("default")))
int ref_cuda_fa2_forward(const float* Q, const float* K, const float* V, float* O, float* L, int seq_len, int head_dim, float scale)
{
    if (head_dim > 64) return -1;              // the reference asserts head_dim <= 64 (:152)
    flash_attention_2_forward(Q, K, V, O, L, seq_len, head_dim, scale);
    return (int)cudaGetLastError();
}
