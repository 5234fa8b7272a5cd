// ref_cuda_bwd.cu -- C-linkage shim around the reference's UNMODIFIED CUDA FlashAttention-2 backward
// (src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:47-299).  Same rules as ref_cuda_fwd.cu: test /
// measurement infrastructure only, the reference file is #included where it lies, nothing is copied.
// Note the reference wrapper needs 4 * 32 * head_dim * 4 bytes of dynamic shared memory and never opts in above 48 KB:
// head_dim = 128 fails to launch; the shim reports that instead of dropping it.
#include "02_flash_attention_v2_backward/flash_attention_backward_kernel.cu"

extern "C" __attribute__((visibility("default")))
int ref_cuda_fa2_backward(const float* Q, const float* K, const float* V, const float* O, const float* L, const float* dO,
                          float* dQ, float* dK, float* dV, int seq_len, int head_dim, float scale)
{
    flash_attention_2_backward(Q, K, V, O, L, dO, dQ, dK, dV, seq_len, head_dim, scale);
    return (int)cudaGetLastError();
}
