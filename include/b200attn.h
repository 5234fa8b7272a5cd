/*
 * b200attn.h -- C-ABI of the B200-native (sm_100a) exact-attention library.
 *
 * This is the drop-in boundary for the reference's hot path.  The reference
 * (terryye/cuda_FlashAttention) has no FFI: its boundary is a set of C++ free functions defined
 * in .cu files that the drivers #include textually.  Each entry point below names the reference
 * function it replaces (paths relative to the reference root); compat/ holds source-compat
 * shims that re-create the reference's C++ signatures on top of this ABI, and INTEGRATION.md
 * shows the binding a maintainer would add.
 *
 * Conventions (same as the reference unless noted):
 *   - All tensor pointers are DEVICE pointers owned by the caller; nothing is retained.
 *   - Layout: one head is row-major [N, d] contiguous (row stride d); batched tensors are
 *     [B, H, N, d] contiguous so every (b,h) slab is one reference call; L / l / m are [B, H, N] fp32.
 *   - Non-causal, no mask, no dropout, no bias.  `scale` multiplies QK^T; scale == 0 selects
 *     1/sqrt(d) (reference src/util/naive_attention.h:9).  LSE is the natural log of the sum of
 *     exp of the SCALED scores.
 *   - Calls enqueue work on `stream` and return without synchronising (reference wrappers launch
 *     on the legacy default stream and return; pass stream = 0 for identical behaviour).
 *   - Every function returns B200_OK (0) or a negative b200_status; nothing aborts or asserts.
 *     (The reference returns void and drops errors; see b200_strerror.)
 *   - There is no CPU fallback: without a CUDA device / on a non-sm_100 device the calls fail with
 *     B200_ERR_NO_DEVICE / B200_ERR_ARCH.
 *   - Threads and streams: the calls may be made from several host threads.  The internal scratch
 *     arena is ONE buffer per device, reused by every call that is given workspace = NULL, so such
 *     calls must be ordered on the device (same stream, or event dependencies); calls that are meant
 *     to overlap on different streams of one device pass their own workspace
 *     (b200_fa2_bwd_workspace_bytes).  A call that has to grow the arena synchronises the device.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif
#ifndef NCCL_H_
typedef struct ncclComm* ncclComm_t;
#endif

typedef enum {
    B200_OK = 0,
    B200_ERR_INVALID_ARGUMENT = -1,
    B200_ERR_UNSUPPORTED_HEAD_DIM = -2,
    B200_ERR_CUDA = -3,          /* a CUDA runtime/driver call failed; b200_last_cuda_error() has the code */
    B200_ERR_NO_DEVICE = -4,
    B200_ERR_ARCH = -5,          /* device is not compute capability 10.x */
    B200_ERR_WORKSPACE = -6,     /* caller-provided workspace too small */
    B200_ERR_NCCL = -7,
    B200_ERR_UNSUPPORTED = -8
} b200_status;

typedef enum {
    B200_DTYPE_BF16 = 0,         /* bf16 I/O, bf16 tensor-core operands, fp32 accumulation */
    B200_DTYPE_F32 = 1,          /* fp32 I/O (the reference's type), tf32 tensor-core operands, fp32 accumulation */
    B200_DTYPE_FP8_E4M3 = 2      /* FORWARD ONLY, d = 128: Q, K, V are fp8 e4m3 bytes, O is bf16, L fp32; kind::f8f6f4
                                  * tensor-core operands, P rounded to e4m3, fp32 accumulation.  Unit scales: fold any
                                  * per-tensor dequantisation of Q / K into `scale` and apply V's to O.  Calls that the fp8
                                  * kernel cannot take (other d, unaligned pointers, backward, ring) return
                                  * B200_ERR_UNSUPPORTED -- there is no generic fp8 path. */
} b200_dtype;

B200_API const char* b200_strerror(int status);
B200_API int b200_last_cuda_error(void);
B200_API int b200_version(void);

/* Which kernel family a (dtype, d, Nq, Nkv) problem is routed to: 1 = tcgen05 tensor-core path,
 * 0 = generic CUDA-core path (tiny / odd shapes; still on the GPU). */
B200_API int b200_fwd_uses_tensor_cores(int dtype, int d, int Nq, int Nkv);
B200_API int b200_bwd_uses_tensor_cores(int dtype, int d, int Nq, int Nkv);

/* The same question for ONE CONCRETE CALL: which family b200_fa2_fwd(_ex) / b200_fa2_bwd(_ex) will run with these
 * pointers, shape, scale and flags.  Beyond (dtype, d) the tensor-core kernels need 16-byte aligned base pointers (TMA),
 * scale > 0 after the "0 means 1/sqrt(d)" rule (forward), and the fp32 backward is d = 64 only; a call that
 * misses one of these still runs, on the generic CUDA-core kernels (~15 TFLOP/s), and this is how a caller finds out.
 * Returns B200_ROUTE_TENSOR_CORE, B200_ROUTE_GENERIC or a negative b200_status for arguments the call would reject.
 * Pure host functions (no device needed); the calls themselves use exactly these predicates. */
#define B200_ROUTE_GENERIC 0
#define B200_ROUTE_TENSOR_CORE 1
B200_API int b200_fa2_fwd_route(const void* Q, const void* K, const void* V, const void* O, int Nq, int Nkv, int d,
                       float scale, int dtype, int flags);
B200_API int b200_fa2_bwd_route(const void* Q, const void* K, const void* V, const void* dO, const void* dQ, const void* dK,
                       const void* dV, int Nq, int Nkv, int d, float scale, int dtype, int flags);

/* Test hook: non-zero routes every call to the generic CUDA-core kernels (used by the GPU tests to
 * cross-check the tensor-core kernels on the device at sizes the CPU oracle cannot reach).
 * Returns the previous setting.  Not a fallback: both families are GPU code in this library. */
B200_API int b200_debug_force_generic(int on);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches claim). */
B200_API unsigned long long b200_debug_launch_count(void);

/* ---- reference-signature entry points (one head, fp32) -------------------------------------- */

/* Replaces flash_attention_2_forward(Q,K,V,O,L,seq_len,head_dim,softmax_scale):
 *   src/02_flash_attention_v2_forward/flash_attention_kernel.cu:300-343 and
 *   src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:133-172.   L may be NULL. */
B200_API int b200_fa2_fwd_f32(const float* Q, const float* K, const float* V, float* O, float* L,
                     int seq_len, int head_dim, float softmax_scale, cudaStream_t stream);

/* Replaces flash_attention_2_backward(Q,K,V,O,L,dO,dQ,dK,dV,seq_len,head_dim,softmax_scale):
 *   src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:249-299.
 * dQ, dK, dV are fully overwritten (the reference memsets dK/dV then accumulates). Scratch for
 * D = rowsum(dO o O) is taken from an internal per-device arena (stream-ordered). */
B200_API int b200_fa2_bwd_f32(const float* Q, const float* K, const float* V, const float* O, const float* L,
                     const float* dO, float* dQ, float* dK, float* dV,
                     int seq_len, int head_dim, float softmax_scale, cudaStream_t stream);

/* Replaces flash_attention(d_Q,d_K,d_V,d_O,d_l,d_m,N,d,Bc,M): src/01_flash_attention_v1/main.cu:7-70.
 * Outputs normalised O, l = sum_j exp(s_j - m) and m = rowmax (NOT the LSE); scale fixed at
 * 1/sqrt(d) (main.cu:60).  The reference synchronises before returning; so does this call. */
B200_API int b200_fa1_fwd_f32(const float* Q, const float* K, const float* V, float* O, float* l, float* m,
                     int N, int d, cudaStream_t stream);

/* ---- batched entry points ([B,H,N,d] contiguous) ---------------------------------------------- */

B200_API int b200_fa2_fwd(const void* Q, const void* K, const void* V, void* O, float* L,
                 int B, int H, int Nq, int Nkv, int d, float scale, int dtype, cudaStream_t stream);

/* ---- extended entry points (flags) ----------------------------------------------------------------
 * B200_FLAG_CAUSAL: query row i attends to keys 0 .. i + (Nkv - Nq) only (bottom-right aligned, needs
 * Nkv >= Nq).  Not in the reference, which is non-causal; its own docs list it as the next step
 * (src/02_flash_attention_v2_backward/__info__/README.md:173).  d in {64,128} runs on the tensor-core kernels for any
 * Nq <= Nkv (masked diagonal tiles; tiles above the diagonal are skipped), the fp32 (tf32) backward at d = 64 included;
 * every other d runs the generic CUDA-core kernels.  flags = 0 is b200_fa2_fwd / b200_fa2_bwd. */
#define B200_FLAG_CAUSAL 1
B200_API int b200_fa2_fwd_ex(const void* Q, const void* K, const void* V, void* O, float* L,
                    int B, int H, int Nq, int Nkv, int d, float scale, int dtype, int flags, cudaStream_t stream);
B200_API int b200_fa2_bwd_ex(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                    void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype,
                    int flags, void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* ---- descriptor entry points: strided layouts and grouped K/V heads (the callers' side) ----
 * The reference stops at contiguous single-head [N,d] slabs (src/02_flash_attention_v2_forward/flash_attention_kernel.cu:300); models hold
 * token-major [B,N,H,d] activations and share K/V heads between query heads (GQA / MQA).  A descriptor gives every
 * tensor its batch / head / row stride in ELEMENTS (last dimension contiguous; all three zero = contiguous
 * [B,H,N,d]), Hkv <= H with H % Hkv == 0 (query head h reads K/V head h / (H/Hkv); dK/dV are summed over the group),
 * and the same flags as the _ex calls.  L (and the internal D) stay contiguous [B,H,Nq] fp32.
 * Non-contiguous or grouped descriptors need the tensor-core shapes (d in {64,128}; backward bf16) and 16-byte
 * aligned bases and strides; anything else returns B200_ERR_UNSUPPORTED (no silent slow path). */
typedef struct b200_layout { long long batch_stride, head_stride, row_stride; } b200_layout;
typedef struct b200_attn_desc {
    int B, H, Hkv, Nq, Nkv, d, dtype, flags;
    float scale;                                   /* 0 => 1/sqrt(d) */
    b200_layout q, k, v, o, d_o, dq, dk, dv;       /* o is also the layout of O in the backward */
} b200_attn_desc;
B200_API int b200_fa2_fwd_desc(const b200_attn_desc* desc, const void* Q, const void* K, const void* V, void* O, float* L,
                      cudaStream_t stream);
B200_API int b200_fa2_bwd_desc(const b200_attn_desc* desc, const void* Q, const void* K, const void* V, const void* O,
                      const float* L, const void* dO, void* dQ, void* dK, void* dV, void* workspace, size_t workspace_bytes,
                      cudaStream_t stream);

/* ---- packed variable-length batches ------------------------------------------------
 * Q/O/dO/dQ are [total_q, H, d], K/V/dK/dV [total_k, Hkv, d] (head / row strides from the descriptor's layouts, all
 * zero = contiguous; batch strides, B, Nq, Nkv are ignored); sequence i is rows cu_seqlens[i] .. cu_seqlens[i+1]-1
 * (HOST arrays of n_seqs + 1 ints).  L is [H, total_q] fp32.  One launch sequence per batch entry on the tensor-core
 * kernels (d in {64,128}; backward bf16; causal needs q length <= k length per sequence, bottom-right aligned); the
 * sequences are independent launches and are spread over a few internal streams that fork from and join back into
 * `stream` (the call stays stream-ordered for the caller); the backward takes its scratch from the library's arena. */
B200_API int b200_fa2_fwd_varlen(const b200_attn_desc* desc, int n_seqs, const int* cu_seqlens_q, const int* cu_seqlens_k,
                        const void* Q, const void* K, const void* V, void* O, float* L, cudaStream_t stream);
B200_API int b200_fa2_bwd_varlen(const b200_attn_desc* desc, int n_seqs, const int* cu_seqlens_q, const int* cu_seqlens_k,
                        const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                        void* dQ, void* dK, void* dV, cudaStream_t stream);

/* ---- host-buffer entry points --------------------------------------------------------------------
 * The reference's drivers own HOST arrays and copy them to the device around every call
 * (src/02_flash_attention_v2_forward/main.cu:42-64, src/02_flash_attention_v2_backward/main.cu:261-289).
 * These do the same job for [B,H,N,d] host tensors (pinned memory for full speed), but pipelined:
 * the (b,h) slabs are cut into chunks and host->device copies, attention and device->host copies
 * of consecutive chunks overlap on three internal streams, so the call runs at PCIe speed instead
 * of copy + compute + copy.  Stream-ordered on `stream` (not host-synchronous); the host buffers
 * must stay valid until `stream` has passed the call.
 *   b200_fa2_fwd_host     : O, L              <- Q, K, V
 *   b200_fa2_fwd_bwd_host : O, L, dQ, dK, dV  <- Q, K, V, dO   (forward + backward of every chunk) */
B200_API int b200_fa2_fwd_host(const void* hQ, const void* hK, const void* hV, void* hO, float* hL,
                      int B, int H, int N, int d, float scale, int dtype, cudaStream_t stream);
B200_API int b200_fa2_fwd_bwd_host(const void* hQ, const void* hK, const void* hV, const void* hdO,
                          void* hO, float* hL, void* hdQ, void* hdK, void* hdV,
                          int B, int H, int N, int d, float scale, int dtype, cudaStream_t stream);

/* Workspace the batched backward needs (D vector and, on the tensor-core path, an fp32 dQ
 * accumulator).  Pass workspace = NULL to let the library use its internal arena. */
B200_API size_t b200_fa2_bwd_workspace_bytes(int B, int H, int Nq, int Nkv, int d, int dtype);

B200_API int b200_fa2_bwd(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                 void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype,
                 void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* b200_fa2_bwd for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments and
 * results as b200_fa2_bwd without the workspace pair: the scratch is a file-scope device array of
 * B200_STATIC_SCRATCH_BYTES inside the library (csrc/static_scratch.cu), held by every GPU that loads the library until
 * the process exits.  The call runs no cudaMalloc / cudaFree / allocating library routine; its first use on a device
 * looks the array's address up (cudaGetSymbolAddress), so make one call per device before starting a capture.
 * Problems whose b200_fa2_bwd_workspace_bytes() exceeds B200_STATIC_SCRATCH_BYTES are rejected with
 * B200_ERR_WORKSPACE and nothing is launched (32 MiB covers e.g. bf16 B*H = 8, N = 4096, d = 128).
 * All calls on a device share the one array, which keeps the previous call's contents; the call itself rewrites or
 * zeroes what it reads, so calls ordered on the device are independent, but calls that OVERLAP on the GPU (different
 * streams without an event between them) race on it -- use b200_fa2_bwd with one workspace per stream for those. */
#define B200_STATIC_SCRATCH_BYTES (32u << 20)
B200_API int b200_fa2_bwd_static(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                        void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype,
                        cudaStream_t stream);

/* D[b,h,i] = sum_c dO[b,h,i,c] * O[b,h,i,c]   (reference flash_attention_backward_kernel.cu:97-120) */
B200_API int b200_rowdot(const void* dO, const void* O, float* D, int B, int H, int N, int d, int dtype, cudaStream_t stream);

/* In place on (Oa, La): merge two normalised partial results by log-sum-exp:
 *   L = log(e^La + e^Lb),  O = e^(La-L) Oa + e^(Lb-L) Ob.
 * Equivalent to the reference ring kernel's (m, l, O_unnorm) carry,
 * src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:67-79,109-139. */
B200_API int b200_lse_merge(void* Oa, float* La, const void* Ob, const float* Lb,
                   int B, int H, int N, int d, int dtype, cudaStream_t stream);

/* ---- ring attention (sequence sharded over nranks GPUs, one process per GPU) ------------------ */

/* KV block that `rank` attends to at ring step `step`: (rank - step) mod nranks
 * (reference common/ring_attention_kernel.cu:198).  Pure host function. */
B200_API int b200_ring_kv_block(int rank, int step, int nranks);

/* Replaces ring_attention_forward(Q_local,K_local,V_local,O_local,L_local,total_seq_len,
 * local_seq_len,head_dim,softmax_scale,comm,rank,nranks):
 *   src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:143-239, with
 *   ring_exchange_kv (src/util/nccl_utils.h:133-142) on an internal side stream.
 * Batched: local shards are [B,H,N_local,d]; L_local is [B,H,N_local] (LSE).  K_local/V_local are
 * NOT modified (the reference overwrites them for nranks >= 3).  nranks == 1 needs no communicator.
 * The ring calls are COLLECTIVE: every rank of `comm` makes the same sequence of ring calls with the same shapes (the
 * copy-engine transport counts on it -- its hand-off counters are derived from the call sequence, and its scratch is
 * re-exported collectively when a call needs a larger one).  If a ring call returns an error on any rank, the
 * transport state of that communicator is undefined: destroy it (b200_ring_comm_destroy) and create a new one before
 * the next ring call, as one would after a failed NCCL collective. */
B200_API int b200_ring_fwd(const void* Q_local, const void* K_local, const void* V_local, void* O_local, float* L_local,
                  int B, int H, int N_local, int d, float scale, int dtype,
                  ncclComm_t comm, int rank, int nranks, cudaStream_t stream);

/* Ring backward (not in the reference): Q,dO,O,L,dQ stay, (K,V,dK,dV) rotate so that
 * dK/dV arrive home fully reduced. */
B200_API int b200_ring_bwd(const void* Q_local, const void* K_local, const void* V_local, const void* O_local,
                  const float* L_local, const void* dO_local, void* dQ_local, void* dK_local, void* dV_local,
                  int B, int H, int N_local, int d, float scale, int dtype,
                  ncclComm_t comm, int rank, int nranks, cudaStream_t stream);

/* Causal ring forward (flags = B200_FLAG_CAUSAL; flags = 0 is b200_ring_fwd).  ZIGZAG sharding: the sequence of
 * P*N_local rows is cut into 2P chunks of N_local/2 rows and rank r holds chunk r followed by chunk 2P-1-r
 * (b200_ring_zigzag_chunk gives the two chunk indices), which gives every rank the same work under the causal mask --
 * the load-balanced placement the reference's notes list as future work (__info__/README.md:173).  Every step after
 * the first is an unmasked rectangle (all local queries x the first half of the visiting block, or the second half of
 * the local queries x the whole block).  Needs the tensor-core shapes (d in {64,128}), N_local even, P <= 32. */
B200_API int b200_ring_fwd_ex(const void* Q_local, const void* K_local, const void* V_local, void* O_local, float* L_local,
                     int B, int H, int N_local, int d, float scale, int dtype, int flags,
                     ncclComm_t comm, int rank, int nranks, cudaStream_t stream);
/* Causal ring backward on the same zigzag sharding (bf16, d = 128, N_local a multiple of 128). */
B200_API int b200_ring_bwd_ex(const void* Q_local, const void* K_local, const void* V_local, const void* O_local,
                     const float* L_local, const void* dO_local, void* dQ_local, void* dK_local, void* dV_local,
                     int B, int H, int N_local, int d, float scale, int dtype, int flags,
                     ncclComm_t comm, int rank, int nranks, cudaStream_t stream);
/* chunk index (0 .. 2P-1) held in half `half` (0 / 1) of rank `rank`'s local rows under the zigzag sharding */
B200_API int b200_ring_zigzag_chunk(int rank, int half, int nranks);

/* Host-buffer ring call: the reference's ring driver owns HOST arrays and copies its shard to the device around the call
 * (src/03_flash_attention_v2_ring/04_ring_attention.cu:97-112).  This does the same for [B,H,N_local,d] host shards
 * (pinned memory for full speed), pipelined like b200_fa2_fwd_bwd_host: the (b,h) slabs are cut into >= 2 chunks and the
 * host->device copy, ring forward + backward and device->host copy of consecutive chunks overlap.  Collective (every
 * rank calls it with the same shapes); stream-ordered on `stream`. */
B200_API int b200_ring_fwd_bwd_host(const void* hQ, const void* hK, const void* hV, const void* hdO,
                           void* hO, float* hL, void* hdQ, void* hdK, void* hdV,
                           int B, int H, int N_local, int d, float scale, int dtype, int flags,
                           ncclComm_t comm, int rank, int nranks, cudaStream_t stream);

/* NCCL bootstrap without MPI (replaces init_nccl_comm / cleanup_mpi_nccl, src/util/nccl_utils.h:29-103):
 * rank 0 calls b200_ring_unique_id and ships the 128 bytes to the other ranks by any means
 * (file, env, torch.distributed store), then every rank calls b200_ring_comm_init. */
B200_API int b200_ring_unique_id(void* out128);
B200_API int b200_ring_comm_init(ncclComm_t* comm, const void* id128, int rank, int nranks);
B200_API int b200_ring_comm_destroy(ncclComm_t comm);
/* ring_exchange_kv (src/util/nccl_utils.h:133-142): grouped send(next)/recv(prev) of two buffers. */
B200_API int b200_ring_exchange_kv(const void* send_k, void* recv_k, const void* send_v, void* recv_v, size_t bytes_each,
                          int next_rank, int prev_rank, ncclComm_t comm, cudaStream_t stream);

/* Data plane of the ring calls (K/V blocks and, in the backward, the travelling dK/dV accumulators):
 *   B200_RING_TRANSPORT_AUTO (-1, default) : copy-engine pushes over NVLink through CUDA-IPC mappings when every rank can
 *       map its neighbours, else NCCL; the environment variable B200_RING_TRANSPORT=nccl selects NCCL;
 *   B200_RING_TRANSPORT_COPY_ENGINE (0)    : as AUTO, ignoring the environment variable;
 *   B200_RING_TRANSPORT_NCCL (1)           : grouped ncclSend/ncclRecv -- exactly what the reference's ring_exchange_kv
 *       does (src/util/nccl_utils.h:133-142).
 * Collective like the ring calls: every rank sets the same mode between the same two ring calls. */
#define B200_RING_TRANSPORT_AUTO (-1)
#define B200_RING_TRANSPORT_COPY_ENGINE 0
#define B200_RING_TRANSPORT_NCCL 1
B200_API int b200_ring_set_transport(int mode);

/* Schedule of the ring FORWARD:
 *   B200_RING_FWD_AUTO (0, default)  : STEPWISE (the faster one on every shape measured at 2 and 8 GPUs);
 *   B200_RING_FWD_STEPWISE (1)       : one attention launch per ring step into its own partial (O_s, LSE_s), one n-way
 *       log-sum-exp merge at the end -- the reference's schedule (common/ring_attention_kernel.cu:197-226) without its
 *       per-step cudaDeviceSynchronize;
 *   B200_RING_FWD_PERSISTENT (2)     : ONE launch per call.  Every K/V block has its own receive slot, the copy engines
 *       forward the blocks around the ring independently of the compute, and the kernel's TMA producer waits on the
 *       arrival counters they write, carrying (m, l, O) across the blocks in registers / TMEM (device-side
 *       waiting instead of host-sequenced steps; no partial buffers, no merge).  Needs bf16, d = 128, N_local a multiple
 *       of 64, non-causal and the copy-engine transport; other calls run STEPWISE.
 * Collective like b200_ring_set_transport. */
#define B200_RING_FWD_AUTO 0
#define B200_RING_FWD_STEPWISE 1
#define B200_RING_FWD_PERSISTENT 2
B200_API int b200_ring_set_forward_mode(int mode);

/* Schedule of the ring BACKWARD's gradient reduction:
 *   B200_RING_BWD_AUTO (0, default)     : HOME when usable (copy-engine transport, bf16 d = 128 tensor-core steps,
 *       non-causal, <= 64 ranks): the step that processes block b pushes its bf16 (dK, dV) partial straight to rank b,
 *       which sums its own and the P - 1 received partials in one pass -- no travelling accumulators, no per-step
 *       accumulate kernels, half the NVLink bytes;
 *   B200_RING_BWD_TRAVELLING (1)        : fp32 (dK, dV) accumulators hop around the ring one step behind the K/V blocks
 *       (also what the NCCL transport, fp32 and the causal zigzag ring use).
 * Collective like b200_ring_set_transport. */
#define B200_RING_BWD_AUTO 0
#define B200_RING_BWD_TRAVELLING 1
B200_API int b200_ring_set_backward_mode(int mode);

/* Frees the library's per-device scratch arena (optional; also released at process exit). */
B200_API int b200_release_workspace(void);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
