// common.cuh -- internal declarations shared by the translation units of libb200attn.so.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/b200attn.h"

namespace b200 {

void set_last_cuda_error(int code);
void count_launch(int n = 1);       // bookkeeping for b200_debug_launch_count()

#define B200_CUDA_TRY(call)                                   \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) {                             \
            ::b200::set_last_cuda_error((int)e__);            \
            (void)cudaGetLastError();                         \
            return B200_ERR_CUDA;                             \
        }                                                     \
    } while (0)

#define B200_TRY(call)                                        \
    do {                                                      \
        int s__ = (call);                                     \
        if (s__ != B200_OK) return s__;                       \
    } while (0)

// Causal masking (extension; the reference is non-causal): key j is visible to query i iff j - i <= coff, with
// coff = Nkv - Nq (bottom-right aligned) when causal and kNoCausal otherwise.
constexpr int kNoCausal = 0x3fffffff;

// ---- tensor layouts of the extended entry points ---------------------------------------------------
// Element strides of a [B, H, N, d] tensor whose last dimension is contiguous: the reference's contiguous slabs,
// token-major [B, N, H, d] views, row sub-blocks (ring steps) and grouped K/V heads are all instances.
struct Lay { long long sb, sh, sn; };
inline Lay lay_contiguous(int H, int N, int d) { return Lay{(long long)H * N * d, (long long)N * d, (long long)d}; }
struct FwdArgs {
    const void *Q, *K, *V; void* O; float* L;      // L: [B*H rows of ldL floats], row (b*H + h) starts at L + (b*H + h) * ldL
    Lay lq, lk, lv, lo; long long ldL;
    int B, H, Hkv, Nq, Nkv, d; float scale; int dtype; bool causal;
};
struct BwdArgs {
    const void *Q, *K, *V, *dO; const float *L, *D;   // D contiguous [B*H, Nq]; L rows ldL apart
    void *dQ, *dK, *dV;
    Lay lq, lk, lv, ldo, ldq, ldk, ldv;
    int B, H, Hkv, Nq, Nkv, d; float scale; int dtype; bool causal;
    long long ldL = 0;                                 // floats between consecutive (b,h) rows of L; 0 = Nq
};

// ---- generic CUDA-core kernels (generic_simt.cu) ------------------------------------------------
// mode 0: O + LSE (L may be NULL);  mode 1: O + l (into L) + m (into M)   [FA1 API]
int simt_fwd(const void* Q, const void* K, const void* V, void* O, float* L, float* M, int BH, int Nq, int Nkv, int d,
             float scale, int dtype, int mode, cudaStream_t st, int coff = kNoCausal);
int rowdot(const void* dO, const void* O, float* D, size_t rows, int d, int dtype, cudaStream_t st);
// dQ may be NULL (skip), or dK and dV both NULL (skip)
int simt_bwd(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* D,
             void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, int dtype, cudaStream_t st,
             int coff = kNoCausal);
int rowdot_strided(const void* dO, const void* O, const Lay& ldo, const Lay& lo, float* D, int B, int H, int N, int d, int dtype,
                   cudaStream_t st);
int lse_merge(void* Oa, float* La, const void* Ob, const float* Lb, size_t rows, int d, int dtype, cudaStream_t st);
// (O, L) = merge of n partials at Op + s * o_stride (elements) / Lp + s * l_stride, plus (O, L) itself when `running`
// upper_mask (causal zigzag ring): partial s with bit s set only covers rows >= c of every Nl-row slab
int lse_merge_multi(void* O, float* L, const void* Op, const float* Lp, size_t o_stride, size_t l_stride, int n, bool running,
                    size_t rows, int d, int dtype, cudaStream_t st, int Nl = 0, int c = 0, unsigned upper_mask = 0u);

// ---- tcgen05 tensor-core kernels (fwd_sm100.cu / bwd_sm100.cu) ------------------------------------
bool tc_fwd_supported(int dtype, int d, int Nq, int Nkv);
int tc_fwd(const void* Q, const void* K, const void* V, void* O, float* L, int BH, int Nq, int Nkv, int d,
           float scale, int dtype, cudaStream_t st, bool causal = false);
// strided / grouped-head form (tensor-core shapes only: tc_fwd_supported, 16-byte aligned bases and strides)
int tc_fwd_strided(const FwdArgs& a, cudaStream_t st);
int launch_fwd_split(const FwdArgs& a, int BH_valid, cudaStream_t st);   // experimental variant (fwd_split_sm100.cu), bf16 d = 128
bool fwd_pair_eligible(const FwdArgs& a);                                // experimental variant (fwd_pair_sm100.cu): CTA pairs,
int launch_fwd_pair(const FwdArgs& a, int BH_valid, cudaStream_t st);    //   bf16 d = 128 non-causal, Nq % 512 == 0, Nkv % 128 == 0
int launch_fwd_shs(const FwdArgs& a, int BH_valid, cudaStream_t st);     // experimental variant (fwd_shs_sm100.cu), bf16 d = 128
int launch_fwd_dbuf(const FwdArgs& a, int BH_valid, cudaStream_t st);    // the shipped bf16 d = 128 forward (fwd_dbuf_sm100.cu)
// persistent ring forward (fwd_dbuf_sm100.cu, RING = true): one launch walks all P K/V blocks, waiting on arrival counters
bool fwd_ring_persistent_supported(int dtype, int d, int BH, int Nl, int nranks);
int launch_fwd_ring_persistent(const void* Q, const void* K, const void* V, void* O, float* L, const void* Kslots, const void* Vslots,
                               const uint32_t* flags, uint32_t epoch, int BH, int Nl, int nranks, float scale, cudaStream_t st);
bool tc_bwd_supported(int dtype, int d, int Nq, int Nkv);
size_t tc_bwd_workspace_bytes(int BH, int Nq, int Nkv, int d, int dtype);
int tc_bwd(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* D,
           void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, int dtype,
           void* ws, size_t ws_bytes, cudaStream_t st, bool causal = false);
int tc_bwd_strided(const BwdArgs& a, void* ws, size_t ws_bytes, cudaStream_t st);

// fp32 I/O backward on tf32 tensor cores (bwd_tf32_sm100.cu): d = 64, contiguous
bool tc_bwd_tf32_supported(int dtype, int d, int Nq, int Nkv);
size_t tc_bwd_tf32_workspace_bytes(int BH, int Nq, int Nkv);
int tc_bwd_tf32(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* D,
                void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, float scale, void* ws, size_t ws_bytes, cudaStream_t st,
                bool causal = false);
int pad_stats(const float* L, const float* D, float* L2p, float* Dp, int BH, int Nq, cudaStream_t st);

// ring backward building blocks (bf16, d = 128): padded statistics once, then accumulate-only steps
bool tc_bwd_ring_supported(int dtype, int d, int BH, const void* Q, const void* K, const void* V, const void* dO);
size_t tc_bwd_ring_stats_bytes(int BH, int Nq);
int tc_bwd_ring_prepare(const float* L, const float* D, void* stats, int BH, int Nq, cudaStream_t st);
int tc_bwd_ring_step(const void* Q, const void* K, const void* V, const void* dO, const void* stats,
                     float* dQacc, void* dKp, void* dVp, int BH, int Nq, int Nkv, float scale, cudaStream_t st);
int tc_bwd_ring_step_sub(const void* Q, const void* K, const void* V, const void* dO, const void* stats,
                         float* dQacc, void* dKp, void* dVp, int BH, int Nl, int q_row0, int Nq, int Nkv, bool causal,
                         float scale, cudaStream_t st);

int tc_bwd_ring_step_heads(const void* Q, const void* K, const void* V, const void* dO, const float* L2p, const float* Dp,
                           float* dQacc, void* dKp, void* dVp, int h0, int nh, int Nl, int q_row0, int Nq, int Nkv, bool causal,
                           float scale, cudaStream_t st);

// ---- per-device scratch arena (api.cu) ------------------------------------------------------------
// Grows monotonically; reuse across calls is safe because all users are stream-ordered on the
// caller's stream and growth synchronises the device before freeing the old block.
int arena_get(size_t bytes, void** out);
int device_check();          // B200_OK iff the current device is compute capability 10.x
int num_sms();

}  // namespace b200
