// static_scratch.cu -- b200_fa2_bwd_static: the batched backward for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The backward's scratch (D = rowsum(dO o O), the padded statistics and the fp32 dQ accumulator, see
// tc_bwd_workspace_bytes / tc_bwd_tf32_workspace_bytes) lives in ONE file-scope device array of
// B200_STATIC_SCRATCH_BYTES.  Every GPU that loads the library holds that array until the process exits.  Every call
// on a device shares it and finds the previous call's contents in it: the kernels rewrite D and the statistics and
// zero the dQ accumulator (cudaMemsetAsync on the caller's stream) before they use them, so calls ordered on the device
// are independent -- calls that OVERLAP on the GPU (different streams, no event between them) race on it.
//
// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address; the address the kernels get
// is looked up once per device with cudaGetSymbolAddress while that device is current.
#include "common.cuh"
#include <mutex>

namespace b200 {

static __device__ __align__(256) unsigned char g_static_scratch[B200_STATIC_SCRATCH_BYTES];

static void* g_static_ptr[64];          // device address of g_static_scratch on device i (0 = not looked up yet)
static std::mutex g_static_mu;

static int static_scratch_address(void** out)
{
    B200_TRY(device_check());
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        B200_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));      // the current device's copy of the array
        if (!p) return B200_ERR_CUDA;
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_fa2_bwd_static(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                                   void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype,
                                   cudaStream_t st)
{
    if (B < 0 || H < 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    const size_t need = b200_fa2_bwd_workspace_bytes(B, H, Nq, Nkv, d, dtype);
    if (need > (size_t)B200_STATIC_SCRATCH_BYTES) return B200_ERR_WORKSPACE;
    void* ws = nullptr;
    // need == 0: empty or invalid arguments -- b200_fa2_bwd returns its status for them before it looks at the workspace
    if (need > 0) B200_TRY(static_scratch_address(&ws));
    return b200_fa2_bwd(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nq, Nkv, d, scale, dtype, ws, need > 0 ? (size_t)B200_STATIC_SCRATCH_BYTES : 0, st);
}
