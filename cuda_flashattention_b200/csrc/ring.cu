// ring.cu -- ring attention over NCCL (sequence sharded over nranks GPUs, one process per GPU).
//
// Replaces the reference host loop ring_attention_forward
// (src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:143-239) and ring_exchange_kv
// (src/util/nccl_utils.h:133-142).  Differences by design:
//   * the K/V rotation runs on an internal side stream ordered by events -- no
//     cudaDeviceSynchronize per step (reference :220), no per-call cudaMalloc/stream creation;
//   * K_local / V_local are never written (the reference scribbles on them for nranks >= 3);
//   * every step produces a normalised partial (O_s, LSE_s) with the single-GPU forward kernel
//     and the partials are merged by log-sum-exp, which is algebraically the reference's
//     (m, l, O_unnorm) carry (:67-79, :109-139);
//   * a ring BACKWARD exists (the reference has none): (K, V, dK, dV) rotate, Q/dO/O/L/dQ stay.
#include "common.cuh"
#include <nccl.h>
#include <mutex>

namespace b200 {

#define B200_NCCL_TRY(call)                                  \
    do {                                                     \
        ncclResult_t r__ = (call);                           \
        if (r__ != ncclSuccess) return B200_ERR_NCCL;        \
    } while (0)

struct RingState {
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t ev_main = nullptr;          // recorded on the caller's stream
    cudaEvent_t ev_comm[2] = {nullptr, nullptr};
    cudaEvent_t ev_compute[2] = {nullptr, nullptr};
    cudaEvent_t ev_acc = nullptr;           // ring backward: accumulator hop finished
    void* buf = nullptr;                    // cached scratch
    size_t bytes = 0;
};
static RingState g_ring[64];
static std::mutex g_ring_mu;

static int ring_state(RingState** out)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lk(g_ring_mu);
    RingState& s = g_ring[dev];
    if (!s.comm_stream) {
        B200_CUDA_TRY(cudaStreamCreateWithFlags(&s.comm_stream, cudaStreamNonBlocking));
        B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_main, cudaEventDisableTiming));
        B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_acc, cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i) {
            B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_comm[i], cudaEventDisableTiming));
            B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_compute[i], cudaEventDisableTiming));
        }
    }
    *out = &s;
    return B200_OK;
}
static int ring_scratch(RingState* s, size_t bytes, void** out)
{
    if (s->bytes < bytes) {
        if (s->buf) { B200_CUDA_TRY(cudaDeviceSynchronize()); B200_CUDA_TRY(cudaFree(s->buf)); s->buf = nullptr; s->bytes = 0; }
        B200_CUDA_TRY(cudaMalloc(&s->buf, bytes));
        s->bytes = bytes;
    }
    *out = s->buf;
    return B200_OK;
}
static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// acc[i] += (float)part[i]
template <typename T>
__global__ void __launch_bounds__(256) accum_kernel(float* __restrict__ acc, const T* __restrict__ part, size_t n)
{
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        float4 a = *reinterpret_cast<float4*>(acc + i);
        float p0, p1, p2, p3;
        if constexpr (sizeof(T) == 4) {
            float4 p = *reinterpret_cast<const float4*>(part + i); p0 = p.x; p1 = p.y; p2 = p.z; p3 = p.w;
        } else {
            uint2 raw = *reinterpret_cast<const uint2*>(part + i);
            __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&raw.x), hi = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
            p0 = __low2float(lo); p1 = __high2float(lo); p2 = __low2float(hi); p3 = __high2float(hi);
        }
        a.x += p0; a.y += p1; a.z += p2; a.w += p3;
        *reinterpret_cast<float4*>(acc + i) = a;
    } else {
        for (; i < n; ++i) {
            if constexpr (sizeof(T) == 4) acc[i] += part[i]; else acc[i] += __bfloat162float(part[i]);
        }
    }
}
// out[i] = (T)acc[i]
template <typename T>
__global__ void __launch_bounds__(256) cast_kernel(T* __restrict__ out, const float* __restrict__ acc, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        if constexpr (sizeof(T) == 4) out[i] = acc[i]; else out[i] = __float2bfloat16_rn(acc[i]);
    }
}
static int accum(float* acc, const void* part, size_t n, int dtype, cudaStream_t st)
{
    unsigned grid = (unsigned)((n / 4 + 255) / 256 + 1);
    if (dtype == B200_DTYPE_F32) accum_kernel<float><<<grid, 256, 0, st>>>(acc, (const float*)part, n);
    else accum_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(acc, (const __nv_bfloat16*)part, n);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}
static int cast_out(void* out, const float* acc, size_t n, int dtype, cudaStream_t st)
{
    unsigned grid = (unsigned)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    if (dtype == B200_DTYPE_F32) cast_kernel<float><<<grid, 256, 0, st>>>((float*)out, acc, n);
    else cast_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)out, acc, n);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

static int exchange(const void* const* send, void* const* recv, const size_t* bytes, int n, int next, int prev,
                    ncclComm_t comm, cudaStream_t st)
{
    B200_NCCL_TRY(ncclGroupStart());
    for (int i = 0; i < n; ++i) {
        B200_NCCL_TRY(ncclSend(send[i], bytes[i], ncclInt8, next, comm, st));
        B200_NCCL_TRY(ncclRecv(recv[i], bytes[i], ncclInt8, prev, comm, st));
    }
    B200_NCCL_TRY(ncclGroupEnd());
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_ring_kv_block(int rank, int step, int nranks)
{
    if (nranks <= 0) return -1;
    int b = (rank - step) % nranks;
    return b < 0 ? b + nranks : b;
}

int b200_ring_unique_id(void* out128)
{
    if (!out128) return B200_ERR_INVALID_ARGUMENT;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    B200_NCCL_TRY(ncclGetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return B200_OK;
}

int b200_ring_comm_init(ncclComm_t* comm, const void* id128, int rank, int nranks)
{
    if (!comm || !id128 || nranks <= 0 || rank < 0 || rank >= nranks) return B200_ERR_INVALID_ARGUMENT;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    B200_NCCL_TRY(ncclCommInitRank(comm, nranks, id, rank));
    return B200_OK;
}

int b200_ring_comm_destroy(ncclComm_t comm)
{
    if (!comm) return B200_OK;
    B200_NCCL_TRY(ncclCommDestroy(comm));
    return B200_OK;
}

int b200_ring_exchange_kv(const void* send_k, void* recv_k, const void* send_v, void* recv_v, size_t bytes_each,
                          int next_rank, int prev_rank, ncclComm_t comm, cudaStream_t st)
{
    if (!comm || !send_k || !recv_k || !send_v || !recv_v) return B200_ERR_INVALID_ARGUMENT;
    const void* s[2] = {send_k, send_v};
    void* r[2] = {recv_k, recv_v};
    size_t b[2] = {bytes_each, bytes_each};
    return exchange(s, r, b, 2, next_rank, prev_rank, comm, st);
}

int b200_ring_fwd(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nl, int d,
                  float scale, int dtype, ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks || B < 0 || H < 0 || Nl < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F32) return B200_ERR_INVALID_ARGUMENT;
    if ((size_t)B * H * Nl == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L) return B200_ERR_INVALID_ARGUMENT;     // the merge needs the LSE
    if (nranks == 1) return b200_fa2_fwd(Q, K, V, O, L, B, H, Nl, Nl, d, scale, dtype, st);
    if (!comm) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());

    RingState* rs;
    B200_TRY(ring_state(&rs));
    const size_t es = dtype == B200_DTYPE_F32 ? 4 : 2;
    const size_t kv_bytes = (size_t)B * H * Nl * d * es;
    const size_t l_bytes = (size_t)B * H * Nl * sizeof(float);
    // scratch: 2 x (K,V) receive buffers + partial O + partial LSE
    const size_t total = 4 * align256(kv_bytes) + align256(kv_bytes) + align256(l_bytes);
    void* base;
    B200_TRY(ring_scratch(rs, total, &base));
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    void* Kb[2]; void* Vb[2];
    Kb[0] = p; p += align256(kv_bytes); Vb[0] = p; p += align256(kv_bytes);
    Kb[1] = p; p += align256(kv_bytes); Vb[1] = p; p += align256(kv_bytes);
    void* Op = p; p += align256(kv_bytes);
    float* Lp = reinterpret_cast<float*>(p);

    const int next = (rank + 1) % nranks, prev = (rank - 1 + nranks) % nranks;
    // the side stream may start once everything already enqueued on the caller's stream is done
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));

    const void* Kc = K; const void* Vc = V;
    for (int step = 0; step < nranks; ++step) {
        const int slot = step & 1;            // receive slot for the block used at step+1
        if (step < nranks - 1) {
            // receive slot `slot` held the block that step-1 attended to; wait until that read is done
            if (step >= 2) B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_compute[slot], 0));
            const void* s[2] = {Kc, Vc};
            void* r[2] = {Kb[slot], Vb[slot]};
            size_t b[2] = {kv_bytes, kv_bytes};
            B200_TRY(exchange(s, r, b, 2, next, prev, comm, rs->comm_stream));
            B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[slot], rs->comm_stream));
        }
        // attention of the local queries against the current block (overlaps the exchange above)
        if (step == 0) {
            B200_TRY(b200_fa2_fwd(Q, Kc, Vc, O, L, B, H, Nl, Nl, d, scale, dtype, st));
        } else {
            B200_TRY(b200_fa2_fwd(Q, Kc, Vc, Op, Lp, B, H, Nl, Nl, d, scale, dtype, st));
            B200_TRY(b200_lse_merge(O, L, Op, Lp, B, H, Nl, d, dtype, st));
        }
        if (step < nranks - 1) {
            // block of step (in slot^1 for step >= 1) is now free for the receive of step+1
            B200_CUDA_TRY(cudaEventRecord(rs->ev_compute[slot ^ 1], st));
            B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[slot], 0));
            Kc = Kb[slot]; Vc = Vb[slot];
        }
    }
    return B200_OK;
}

int b200_ring_bwd(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                  void* dQ, void* dK, void* dV, int B, int H, int Nl, int d, float scale, int dtype,
                  ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks || B < 0 || H < 0 || Nl < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F32) return B200_ERR_INVALID_ARGUMENT;
    if ((size_t)B * H * Nl == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L || !dO || !dQ || !dK || !dV) return B200_ERR_INVALID_ARGUMENT;
    if (nranks == 1) return b200_fa2_bwd(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nl, Nl, d, scale, dtype, nullptr, 0, st);
    if (!comm) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());

    RingState* rs;
    B200_TRY(ring_state(&rs));
    const size_t es = dtype == B200_DTYPE_F32 ? 4 : 2;
    const size_t n_el = (size_t)B * H * Nl * d;
    const size_t kv_bytes = n_el * es, acc_bytes = n_el * sizeof(float);
    // scratch: 2 x (K,V) recv, 3 partial grads (I/O dtype), dQ fp32 acc, 2 x (dK,dV) fp32 acc
    const size_t total = 4 * align256(kv_bytes) + 3 * align256(kv_bytes) + align256(acc_bytes) + 4 * align256(acc_bytes);
    void* base;
    B200_TRY(ring_scratch(rs, total, &base));
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    void* Kb[2]; void* Vb[2]; float* dKa[2]; float* dVa[2];
    Kb[0] = p; p += align256(kv_bytes); Vb[0] = p; p += align256(kv_bytes);
    Kb[1] = p; p += align256(kv_bytes); Vb[1] = p; p += align256(kv_bytes);
    void* dQp = p; p += align256(kv_bytes);
    void* dKp = p; p += align256(kv_bytes);
    void* dVp = p; p += align256(kv_bytes);
    float* dQa = reinterpret_cast<float*>(p); p += align256(acc_bytes);
    for (int i = 0; i < 2; ++i) {
        dKa[i] = reinterpret_cast<float*>(p); p += align256(acc_bytes);
        dVa[i] = reinterpret_cast<float*>(p); p += align256(acc_bytes);
    }
    const int next = (rank + 1) % nranks, prev = (rank - 1 + nranks) % nranks;

    B200_CUDA_TRY(cudaMemsetAsync(dQa, 0, acc_bytes, st));
    B200_CUDA_TRY(cudaMemsetAsync(dKa[0], 0, acc_bytes, st));
    B200_CUDA_TRY(cudaMemsetAsync(dVa[0], 0, acc_bytes, st));
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));

    // Step s: block b = (rank - s) mod P is resident (K/V in Kc/Vc, its running fp32 dK/dV in acc[cur]).
    // Compute partial grads, add them, then pass (dK,dV) accumulators on; K/V for step s+1 are
    // prefetched on the side stream while step s computes.  After the last step one more hop
    // brings every block's dK/dV home.
    const void* Kc = K; const void* Vc = V;
    int cur = 0;                                       // which fp32 accumulator pair is resident
    for (int step = 0; step < nranks; ++step) {
        const int slot = step & 1;
        if (step < nranks - 1) {
            if (step >= 2) B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_compute[slot], 0));
            const void* s[2] = {Kc, Vc};
            void* r[2] = {Kb[slot], Vb[slot]};
            size_t b[2] = {kv_bytes, kv_bytes};
            B200_TRY(exchange(s, r, b, 2, next, prev, comm, rs->comm_stream));
            B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[slot], rs->comm_stream));
        }
        B200_TRY(b200_fa2_bwd(Q, Kc, Vc, O, L, dO, dQp, dKp, dVp, B, H, Nl, Nl, d, scale, dtype, nullptr, 0, st));
        B200_TRY(accum(dQa, dQp, n_el, dtype, st));
        // the running dK/dV of this block arrive from the previous rank while we compute
        if (step >= 1) B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_acc, 0));
        B200_TRY(accum(dKa[cur], dKp, n_el, dtype, st));
        B200_TRY(accum(dVa[cur], dVp, n_el, dtype, st));
        // hand the accumulators to the next rank (queued behind this step's K/V prefetch)
        B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
        B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
        {
            const void* s[2] = {dKa[cur], dVa[cur]};
            void* r[2] = {dKa[cur ^ 1], dVa[cur ^ 1]};
            size_t b[2] = {acc_bytes, acc_bytes};
            B200_TRY(exchange(s, r, b, 2, next, prev, comm, rs->comm_stream));
        }
        B200_CUDA_TRY(cudaEventRecord(rs->ev_acc, rs->comm_stream));
        cur ^= 1;
        if (step < nranks - 1) {
            B200_CUDA_TRY(cudaEventRecord(rs->ev_compute[slot ^ 1], st));
            B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[slot], 0));
            Kc = Kb[slot]; Vc = Vb[slot];
        }
    }
    B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_acc, 0));
    // after P hops the accumulators of block `rank` are back home
    B200_TRY(cast_out(dQ, dQa, n_el, dtype, st));
    B200_TRY(cast_out(dK, dKa[cur], n_el, dtype, st));
    B200_TRY(cast_out(dV, dVa[cur], n_el, dtype, st));
    return B200_OK;
}

}  // extern "C"
