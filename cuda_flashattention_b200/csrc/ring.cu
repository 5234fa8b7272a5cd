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
#include <cuda.h>
#include <algorithm>
#include <mutex>
#include <math.h>
#include <stdlib.h>
#include <string.h>

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
    cudaStream_t attn_stream[2] = {nullptr, nullptr};   // ring forward: steps alternate between two streams so the
    cudaEvent_t ev_attn[2] = {nullptr, nullptr}, ev_merge[2] = {nullptr, nullptr};        // next step's CTAs fill the SMs the current step's last wave leaves idle
    void* buf = nullptr;                    // cached scratch
    size_t bytes = 0;
    bool ready = false;                     // every stream / event above exists
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
    if (!s.ready) {
        // (a failed creation leaves `ready` false: the next call retries what is still missing instead of running
        // with null handles)
        if (!s.comm_stream) B200_CUDA_TRY(cudaStreamCreateWithFlags(&s.comm_stream, cudaStreamNonBlocking));
        if (!s.ev_main) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_main, cudaEventDisableTiming));
        if (!s.ev_acc) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_acc, cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i) {
            if (!s.ev_comm[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_comm[i], cudaEventDisableTiming));
            if (!s.ev_compute[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_compute[i], cudaEventDisableTiming));
            if (!s.attn_stream[i]) B200_CUDA_TRY(cudaStreamCreateWithFlags(&s.attn_stream[i], cudaStreamNonBlocking));
            if (!s.ev_attn[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_attn[i], cudaEventDisableTiming));
            if (!s.ev_merge[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&s.ev_merge[i], cudaEventDisableTiming));
        }
        s.ready = true;
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

// ---------------------------------------------------------------------------------------------
// Copy-engine transport.  NCCL's send/recv kernels occupy SMs for the whole transfer, and an SM
// that hosts an NCCL CTA cannot host an attention CTA (ours take the whole register file), which
// costs ~20 % of a ring step.  So the K/V (and dK/dV accumulator) blocks are PUSHED into the next
// rank's receive buffer with cudaMemcpyAsync over NVLink -- DMA engines, zero SMs -- through a
// CUDA-IPC mapping of the peer's scratch, and ordered with stream memory operations on small
// counters (cuStreamWriteValue32 into the peer's flags, cuStreamWaitValue32 on our own).  NCCL is
// still the bootstrap (all-gather of the IPC handles on the caller's communicator) and the
// fallback when IPC / peer access is unavailable or B200_RING_TRANSPORT=nccl.
// Every rank runs the same call sequence, so "how many times has my neighbour filled / released
// buffer x" is known locally: the counters only ever grow and are never reset between calls.
// ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_memop32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static PFN_memop32 g_wait32 = nullptr, g_write32 = nullptr;
static bool load_memops()
{
    if (g_wait32 && g_write32) return true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return false;
    g_wait32 = reinterpret_cast<PFN_memop32>(p);
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return false;
    g_write32 = reinterpret_cast<PFN_memop32>(p);
    return true;
}

// F_PRDY + s: persistent forward -- block s of the current call has landed in receive slot s - 1 (value = call epoch)
// F_HOME + s: "home" backward -- the dK/dV partial computed at ring step s + 1 has landed in home slot s (value = call epoch)
// F_DONE_FROM + r: rank r has finished (read everything of) as many ring calls as the value says
enum { F_RDY_KV = 0, F_FREE_KV = 2, F_RDY_ACC = 4, F_FREE_ACC = 6, F_PRDY = 16, F_HOME = 16 + 256, F_DONE_FROM = 16 + 512,
       F_COUNT = 16 + 768, kMaxPeers = 64 };

// transport preference (b200_ring_set_transport): -1 = B200_RING_TRANSPORT decides (default: copy engines), 0 = copy engines,
// 1 = NCCL send/recv.  Collective like the ring calls themselves: every rank holds the same value.
static int g_transport_pref = -1;
// forward schedule (b200_ring_set_forward_mode): 0 = automatic, 1 = one launch per ring step + n-way LSE merge,
// 2 = one persistent launch that walks all blocks (needs the copy-engine transport; else falls back to 1)
static int g_fwd_mode = 0;
// backward schedule (b200_ring_set_backward_mode): 0 = automatic (partials pushed to the block's home rank when the
// copy-engine transport and the tensor-core step kernel are usable), 1 = travelling fp32 accumulators
static int g_bwd_mode = 0;

struct PeerLink {
    ncclComm_t comm = nullptr;
    int rank = -1, nranks = 0;
    bool ok = false;                        // copy-engine transport usable with `comm`
    uint32_t* flags = nullptr;              // my counters (written by my neighbours)
    uint8_t* xchg = nullptr;                // device staging for the handle all-gather
    uint8_t* next_buf = nullptr;            // IPC mapping of the next rank's scratch
    uint32_t* next_flags = nullptr;
    uint32_t* prev_flags = nullptr;
    // every peer's scratch and counters are mapped (the home backward pushes partial gradients straight to the rank that
    // owns the block; the call-end handshake tells every peer); next_* / prev_* are aliases into these
    uint8_t* peer_buf[kMaxPeers] = {};
    uint32_t* peer_flags[kMaxPeers] = {};
    bool all_peers = false;                 // nranks <= kMaxPeers: peer_buf / peer_flags are valid for every rank != rank
    uint32_t bwd_epoch = 0;                 // home backward calls made on this link
    size_t exported = 0;                    // scratch size the neighbours have mapped
    void* exported_ptr = nullptr;           // ... and its address (an NCCL-transport call may have re-allocated it since)
    uint32_t calls_done = 0;                // ring calls finished on this link (F_CALL_DONE handshake)
    uint32_t pers_epoch = 0;                // persistent forward calls made on this link (value of their arrival flags)
    uint32_t push_kv[2] = {0, 0}, in_kv[2] = {0, 0}, rel_kv[2] = {0, 0};
    uint32_t push_acc[2] = {0, 0}, in_acc[2] = {0, 0}, out_acc[2] = {0, 0};
};
static PeerLink g_link[64];

struct LinkMsg {                            // what every rank publishes
    cudaIpcMemHandle_t buf, flags;
    unsigned long long bytes;
    int ok;
    int pad;
};

static void link_close(PeerLink& L)
{
    for (int i = 0; i < kMaxPeers; ++i) {
        if (L.peer_buf[i]) { cudaIpcCloseMemHandle(L.peer_buf[i]); L.peer_buf[i] = nullptr; }
        if (L.peer_flags[i]) { cudaIpcCloseMemHandle(L.peer_flags[i]); L.peer_flags[i] = nullptr; }
    }
    L.next_buf = nullptr; L.next_flags = nullptr; L.prev_flags = nullptr; L.all_peers = false;
    (void)cudaGetLastError();
}

static int link_allgather(PeerLink& L, RingState* rs, const LinkMsg& mine, LinkMsg* all)
{
    const size_t sz = sizeof(LinkMsg);
    B200_CUDA_TRY(cudaMemcpyAsync(L.xchg, &mine, sz, cudaMemcpyHostToDevice, rs->comm_stream));
    B200_NCCL_TRY(ncclAllGather(L.xchg, L.xchg + 256, sz, ncclInt8, L.comm, rs->comm_stream));
    B200_CUDA_TRY(cudaMemcpyAsync(all, L.xchg + 256, sz * L.nranks, cudaMemcpyDeviceToHost, rs->comm_stream));
    B200_CUDA_TRY(cudaStreamSynchronize(rs->comm_stream));
    return B200_OK;
}

// Makes sure the scratch is at least `need` bytes and that the copy-engine link (if usable at
// all) points at the neighbours' current scratch.  Collective: every rank of `comm` gets here in
// the same call with the same `need`.  Returns L.ok through *use_ce.
static int link_prepare(RingState* rs, ncclComm_t comm, int rank, int nranks, size_t need, void** scratch, bool* use_ce)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    PeerLink& L = g_link[dev];
    static int env_ce = -1;
    if (env_ce < 0) { const char* e = getenv("B200_RING_TRANSPORT"); env_ce = (e && !strcmp(e, "nccl")) ? 0 : 1; }
    const int want_ce = g_transport_pref < 0 ? env_ce : (g_transport_pref == 0 ? 1 : 0);
    const bool new_comm = (L.comm != comm || L.rank != rank || L.nranks != nranks);
    // (re-export also when a call on the NCCL transport re-allocated the scratch the neighbours still have mapped)
    const bool grow = rs->bytes < need || (L.ok && (rs->buf != L.exported_ptr || rs->bytes != L.exported));
    if (!want_ce || nranks > 255) {
        B200_TRY(ring_scratch(rs, need, scratch));
        *use_ce = false;
        return B200_OK;
    }
    if (!new_comm && !grow) { *scratch = rs->buf; *use_ce = L.ok; return B200_OK; }

    // ---- (re)build the link: quiesce, drop the old mappings, agree, re-map
    B200_CUDA_TRY(cudaDeviceSynchronize());
    link_close(L);
    if (new_comm) {
        uint32_t* keep_flags = L.flags; uint8_t* keep_xchg = L.xchg;
        L = PeerLink();
        L.flags = keep_flags; L.xchg = keep_xchg;
        L.comm = comm; L.rank = rank; L.nranks = nranks;
    }
    if (!L.flags) {
        B200_CUDA_TRY(cudaMalloc(&L.flags, F_COUNT * sizeof(uint32_t)));
        B200_CUDA_TRY(cudaMalloc(&L.xchg, 256 + 256 * 256));
    }
    if (new_comm) B200_CUDA_TRY(cudaMemset(L.flags, 0, F_COUNT * sizeof(uint32_t)));
    LinkMsg mine, all[256];
    memset(&mine, 0, sizeof(mine));
    static_assert(sizeof(LinkMsg) <= 256, "LinkMsg must fit its all-gather slot");
    // round 1 doubles as a barrier: after it nobody maps my old scratch any more and every rank has zeroed its flags
    B200_TRY(link_allgather(L, rs, mine, all));
    B200_TRY(ring_scratch(rs, need, scratch));
    bool ok = load_memops();
    if (ok) ok = cudaIpcGetMemHandle(&mine.buf, rs->buf) == cudaSuccess && cudaIpcGetMemHandle(&mine.flags, L.flags) == cudaSuccess;
    mine.bytes = rs->bytes; mine.ok = ok ? 1 : 0;
    (void)cudaGetLastError();
    B200_TRY(link_allgather(L, rs, mine, all));
    const int next = (rank + 1) % nranks, prev = (rank - 1 + nranks) % nranks;
    for (int r = 0; r < nranks; ++r) ok = ok && all[r].ok && all[r].bytes == rs->bytes;
    if (ok) {
        // map the neighbours (always) and, for rings of up to kMaxPeers ranks, every other rank as well
        const bool everyone = nranks <= kMaxPeers;
        for (int r = 0; r < nranks && ok; ++r) {
            if (r == rank || !(everyone || r == next || r == prev)) continue;
            const int slot = everyone ? r : (r == next ? 0 : 1);
            void* pb = nullptr; void* pf = nullptr;
            ok = cudaIpcOpenMemHandle(&pb, all[r].buf, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            L.peer_buf[slot] = reinterpret_cast<uint8_t*>(pb);
            if (ok) ok = cudaIpcOpenMemHandle(&pf, all[r].flags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            L.peer_flags[slot] = reinterpret_cast<uint32_t*>(pf);
        }
        if (ok) {
            L.next_buf = L.peer_buf[everyone ? next : 0];
            L.next_flags = L.peer_flags[everyone ? next : 0];
            L.prev_flags = L.peer_flags[everyone ? prev : (prev == next ? 0 : 1)];
            L.all_peers = everyone;
        }
        (void)cudaGetLastError();
    }
    // round 3: the link is used only if EVERY rank managed to map its neighbours
    memset(&mine, 0, sizeof(mine));
    mine.ok = ok ? 1 : 0;
    B200_TRY(link_allgather(L, rs, mine, all));
    for (int r = 0; r < nranks; ++r) ok = ok && all[r].ok;
    if (!ok) link_close(L);
    L.ok = ok; L.exported = rs->bytes; L.exported_ptr = rs->buf;
    *use_ce = ok;
    return B200_OK;
}

#define B200_CU_TRY(call) do { if ((call) != CUDA_SUCCESS) return B200_ERR_CUDA; } while (0)
static inline int wait_ge(cudaStream_t st, uint32_t* addr, uint32_t v)
{
    B200_CU_TRY(g_wait32(st, (CUdeviceptr)(uintptr_t)addr, v, CU_STREAM_WAIT_VALUE_GEQ));
    return B200_OK;
}
static inline int write_val(cudaStream_t st, uint32_t* addr, uint32_t v)
{
    B200_CU_TRY(g_write32(st, (CUdeviceptr)(uintptr_t)addr, v, CU_STREAM_WRITE_VALUE_DEFAULT));
    return B200_OK;
}
// Cross-call handshake of the copy-engine transport.  The per-buffer counters only order the hand-offs INSIDE a call;
// the scratch regions of consecutive calls alias (forward partials / backward accumulators / receive slots), and a
// fast rank could push the first block of call n+1 into a neighbour that is still reading call n's results.  So every
// call ends by telling the PREVIOUS rank "I have read everything call n left in my scratch" (call_end, ordered behind
// the call's last kernel on the caller's stream), and begins by waiting for that message from the NEXT rank -- the
// only rank it writes into -- before its first push (call_begin).
static int call_begin(PeerLink& L, RingState* rs)
{
    const int next = (L.rank + 1) % L.nranks;
    if (L.calls_done > 0) B200_TRY(wait_ge(rs->comm_stream, L.flags + F_DONE_FROM + next, L.calls_done));
    return B200_OK;
}
// ... and before the first write of this call into ANY other rank `peer` (the home backward pushes to every rank)
static int call_begin_peer(PeerLink& L, RingState* rs, int peer)
{
    if (L.calls_done > 0) B200_TRY(wait_ge(rs->comm_stream, L.flags + F_DONE_FROM + peer, L.calls_done));
    return B200_OK;
}
static int call_end(PeerLink& L, RingState* rs, cudaStream_t st)
{
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
    ++L.calls_done;
    if (L.all_peers) {
        for (int r = 0; r < L.nranks; ++r)
            if (r != L.rank) B200_TRY(write_val(rs->comm_stream, L.peer_flags[r] + F_DONE_FROM + L.rank, L.calls_done));
        return B200_OK;
    }
    return write_val(rs->comm_stream, L.prev_flags + F_DONE_FROM + L.rank, L.calls_done);   // only the previous rank writes into us
}

// push two buffers into the next rank's K/V receive slot
static int ce_push_kv(PeerLink& L, RingState* rs, int slot, const void* k, const void* v, size_t off_k, size_t off_v, size_t bytes)
{
    const uint32_t n = ++L.push_kv[slot];
    if (n > 1) B200_TRY(wait_ge(rs->comm_stream, L.flags + F_FREE_KV + slot, n - 1));   // neighbour done with the previous fill
    B200_CUDA_TRY(cudaMemcpyAsync(L.next_buf + off_k, k, bytes, cudaMemcpyDeviceToDevice, rs->comm_stream));
    B200_CUDA_TRY(cudaMemcpyAsync(L.next_buf + off_v, v, bytes, cudaMemcpyDeviceToDevice, rs->comm_stream));
    return write_val(rs->comm_stream, L.next_flags + F_RDY_KV + slot, n);
}

// acc[i] (+)= (float)part[i] for the PAIR of tensors selected by blockIdx.y (dK and dV in one launch).  INIT: acc = part
// (step 0 of the ring backward: the travelling accumulators start from this rank's contribution -- no memset, no read).
template <typename T, bool INIT>
__global__ void __launch_bounds__(256) accum_kernel(float* __restrict__ acc0, const T* __restrict__ part0,
                                                    float* __restrict__ acc1, const T* __restrict__ part1, size_t n)
{
    float* __restrict__ acc = blockIdx.y ? acc1 : acc0;
    const T* __restrict__ part = blockIdx.y ? part1 : part0;
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (!INIT) a = *reinterpret_cast<float4*>(acc + i);
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
            const float pv = sizeof(T) == 4 ? (float)part[i] : __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(&part[i]));
            acc[i] = INIT ? pv : acc[i] + pv;
        }
    }
}
// the same for the first `nrows` rows of every slab of `slab_rows` rows of d elements (d % 4 == 0): causal ring steps
// that only produced gradients for the first half of the visiting K/V block
template <typename T>
__global__ void __launch_bounds__(256) accum_rows_kernel(float* __restrict__ acc, const T* __restrict__ part, size_t slabs,
                                                         int slab_rows, int nrows, int d)
{
    const size_t per = (size_t)nrows * d / 4;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= slabs * per) return;
    const size_t slab = idx / per, i = slab * (size_t)slab_rows * d + (idx - slab * per) * 4;
    float4 a = *reinterpret_cast<float4*>(acc + i);
    if constexpr (sizeof(T) == 4) {
        const float4 p = *reinterpret_cast<const float4*>(part + i);
        a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
    } else {
        uint2 raw = *reinterpret_cast<const uint2*>(part + i);
        __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&raw.x), hi = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
        a.x += __low2float(lo); a.y += __high2float(lo); a.z += __low2float(hi); a.w += __high2float(hi);
    }
    *reinterpret_cast<float4*>(acc + i) = a;
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
// one launch for up to two (accumulator, partial) pairs of n elements each; acc1 == nullptr: a single pair
static int accum2(float* acc0, const void* part0, float* acc1, const void* part1, size_t n, int dtype, bool init, cudaStream_t st)
{
    if (n == 0) return B200_OK;
    dim3 grid((unsigned)((n / 4 + 255) / 256 + 1), acc1 ? 2 : 1);
    if (dtype == B200_DTYPE_F32) {
        if (init) accum_kernel<float, true><<<grid, 256, 0, st>>>(acc0, (const float*)part0, acc1, (const float*)part1, n);
        else accum_kernel<float, false><<<grid, 256, 0, st>>>(acc0, (const float*)part0, acc1, (const float*)part1, n);
    } else {
        if (init) accum_kernel<__nv_bfloat16, true><<<grid, 256, 0, st>>>(acc0, (const __nv_bfloat16*)part0, acc1, (const __nv_bfloat16*)part1, n);
        else accum_kernel<__nv_bfloat16, false><<<grid, 256, 0, st>>>(acc0, (const __nv_bfloat16*)part0, acc1, (const __nv_bfloat16*)part1, n);
    }
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}
static int accum(float* acc, const void* part, size_t n, int dtype, cudaStream_t st)
{
    return accum2(acc, part, nullptr, nullptr, n, dtype, false, st);
}
static int accum_rows(float* acc, const void* part, size_t slabs, int slab_rows, int nrows, int d, int dtype, cudaStream_t st)
{
    const size_t n4 = slabs * (size_t)nrows * d / 4;
    const unsigned grid = (unsigned)((n4 + 255) / 256);
    if (dtype == B200_DTYPE_F32) accum_rows_kernel<float><<<grid, 256, 0, st>>>(acc, (const float*)part, slabs, slab_rows, nrows, d);
    else accum_rows_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(acc, (const __nv_bfloat16*)part, slabs, slab_rows, nrows, d);
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

// out = (T)(own + sum of `nslots` partials `slot_stride` elements apart) for the PAIR of tensors selected by blockIdx.y
__global__ void __launch_bounds__(256) reduce_partials_kernel(__nv_bfloat16* __restrict__ out0, const __nv_bfloat16* __restrict__ own0,
                                                              const __nv_bfloat16* __restrict__ slots0, __nv_bfloat16* __restrict__ out1,
                                                              const __nv_bfloat16* __restrict__ own1, const __nv_bfloat16* __restrict__ slots1,
                                                              size_t n8, int nslots, size_t slot_stride)
{
    __nv_bfloat16* __restrict__ out = blockIdx.y ? out1 : out0;
    const __nv_bfloat16* __restrict__ own = blockIdx.y ? own1 : own0;
    const __nv_bfloat16* __restrict__ slots = blockIdx.y ? slots1 : slots0;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n8; i += stride) {
        float acc[8];
        auto add = [&](const __nv_bfloat16* src, bool first) {
            const uint4 raw = reinterpret_cast<const uint4*>(src)[i];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(h[k]);
                acc[2 * k] = first ? f.x : acc[2 * k] + f.x;
                acc[2 * k + 1] = first ? f.y : acc[2 * k + 1] + f.y;
            }
        };
        add(own, true);
        for (int s = 0; s < nslots; ++s) add(slots + (size_t)s * slot_stride, false);
        uint4 o;
        __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
        for (int k = 0; k < 4; ++k) oh[k] = __floats2bfloat162_rn(acc[2 * k], acc[2 * k + 1]);
        reinterpret_cast<uint4*>(out)[i] = o;
    }
}

}  // namespace b200
using namespace b200;

// "Home" ring backward (copy-engine transport, bf16 d = 128 tensor-core steps, non-causal): K/V rotate as in the forward,
// but the gradients do not travel with them.  The step that processes block b writes its bf16 (dK, dV) partial into a
// rotating pair of local buffers and the copy engine pushes it STRAIGHT TO RANK b -- the block's home -- into the slot
// of that step; after the last step every rank sums its own partial and the P - 1 that arrived (one pass, fp32 in
// registers, one rounding) into dK / dV.  Against the travelling fp32 accumulators of ring_bwd_impl this halves the
// NVLink bytes (bf16 partials instead of fp32 accumulators), removes the per-step accumulate kernels and the serial hop
// chain, and leaves only the last step's push exposed (cut into head groups like before).
static int ring_bwd_home(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                         void* dQ, void* dK, void* dV, int B, int H, int Nl, int d, float scale, ncclComm_t comm, int rank, int nranks,
                         cudaStream_t st, bool* done)
{
    *done = false;
    if (nranks > kMaxPeers) return B200_OK;
    RingState* rs;
    B200_TRY(ring_state(&rs));
    const int BH = B * H, dtype = B200_DTYPE_BF16;
    const size_t n_el = (size_t)BH * Nl * d, kv_bytes = n_el * 2, acc_bytes = n_el * sizeof(float);
    const size_t kvb = align256(kv_bytes);
    const size_t d_bytes = align256((size_t)BH * Nl * sizeof(float));
    const size_t stats_bytes = tc_bwd_ring_stats_bytes(BH, Nl);
    // scratch: 2 x (K,V) recv | own partial + 2 rotating partials (dK,dV each) | (P-1) home slots (dK,dV each) | dQ fp32 | D | stats
    const size_t total = 4 * kvb + 6 * kvb + 2 * (size_t)(nranks - 1) * kvb + align256(acc_bytes) + d_bytes + align256(stats_bytes);
    void* base;
    bool ce = false;
    B200_TRY(link_prepare(rs, comm, rank, nranks, total, &base, &ce));
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    PeerLink& lk = g_link[dev];
    if (!ce || !lk.all_peers) return B200_OK;                       // (the caller falls back to the travelling accumulators)
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    auto off = [&](const void* q) { return (size_t)(reinterpret_cast<const uint8_t*>(q) - reinterpret_cast<uint8_t*>(base)); };
    void* Kb[2]; void* Vb[2];
    Kb[0] = p; p += kvb; Vb[0] = p; p += kvb; Kb[1] = p; p += kvb; Vb[1] = p; p += kvb;
    uint8_t* dKp[3]; uint8_t* dVp[3];                                // [2] = this rank's own block (step 0), [0/1] rotate
    for (int i = 0; i < 3; ++i) { dKp[i] = p; p += kvb; dVp[i] = p; p += kvb; }
    uint8_t* homeK = p; p += (size_t)(nranks - 1) * kvb;
    uint8_t* homeV = p; p += (size_t)(nranks - 1) * kvb;
    float* dQa = reinterpret_cast<float*>(p); p += align256(acc_bytes);
    float* Dvec = reinterpret_cast<float*>(p); p += d_bytes;
    void* stats = p;
    const int Nl_pad64 = (Nl + 63) / 64 * 64;
    const float* L2p = reinterpret_cast<const float*>(stats);
    const float* Dp = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(stats) + align256((size_t)BH * Nl_pad64 * sizeof(float)));
    const float sc = scale == 0.0f ? 1.0f / sqrtf((float)d) : scale;
    const uint32_t epoch = ++lk.bwd_epoch;

    B200_TRY(rowdot(dO, O, Dvec, (size_t)BH * Nl, d, dtype, st));
    B200_TRY(tc_bwd_ring_prepare(L, Dvec, stats, BH, Nl, st));
    B200_CUDA_TRY(cudaMemsetAsync(dQa, 0, acc_bytes, st));
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
    B200_TRY(call_begin(lk, rs));

    const void* Kc = K; const void* Vc = V;
    for (int step = 0; step < nranks; ++step) {
        const int slot = step & 1, cur_slot = slot ^ 1;
        if (step >= 1) {
            const uint32_t k = ++lk.in_kv[cur_slot];
            B200_TRY(wait_ge(st, lk.flags + F_RDY_KV + cur_slot, k));
            B200_TRY(wait_ge(rs->comm_stream, lk.flags + F_RDY_KV + cur_slot, k));
        }
        if (step < nranks - 1) B200_TRY(ce_push_kv(lk, rs, slot, Kc, Vc, off(Kb[slot]), off(Vb[slot]), kv_bytes));
        const int pb = step == 0 ? 2 : (step & 1);
        // the push that last read this rotating pair (step - 2) must be over before the kernel overwrites it
        if (step >= 3) B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_compute[pb], 0));
        const int home = (rank - step + nranks) % nranks;            // the block of this step lives on rank `home`
        int groups = 1;
        if (step == nranks - 1) {
            const long long ctas = (long long)BH * ((Nl + 127) / 128);
            groups = (int)std::max<long long>(1, std::min<long long>(std::min(4, BH), ctas / (6 * num_sms())));
        }
        for (int g = 0; g < groups; ++g) {
            const int h0 = (int)((long long)BH * g / groups), h1 = (int)((long long)BH * (g + 1) / groups);
            const size_t b0 = (size_t)h0 * Nl * d * 2, nb = (size_t)(h1 - h0) * Nl * d * 2;      // byte range of the head group
            B200_TRY(tc_bwd_ring_step_heads(Q, Kc, Vc, dO, L2p, Dp, dQa, dKp[pb], dVp[pb], h0, h1 - h0, Nl, 0, Nl, Nl, false, sc, st));
            if (step == 0) continue;
            B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
            B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
            if (g == 0) B200_TRY(call_begin_peer(lk, rs, home));
            B200_CUDA_TRY(cudaMemcpyAsync(lk.peer_buf[home] + off(homeK) + (size_t)(step - 1) * kvb + b0, dKp[pb] + b0, nb, cudaMemcpyDeviceToDevice, rs->comm_stream));
            B200_CUDA_TRY(cudaMemcpyAsync(lk.peer_buf[home] + off(homeV) + (size_t)(step - 1) * kvb + b0, dVp[pb] + b0, nb, cudaMemcpyDeviceToDevice, rs->comm_stream));
            if (g == groups - 1) {
                B200_TRY(write_val(rs->comm_stream, lk.peer_flags[home] + F_HOME + step - 1, epoch));
                B200_CUDA_TRY(cudaEventRecord(rs->ev_compute[pb], rs->comm_stream));
                // this step's K/V block (its compute and onward push are both ordered before this point) goes back to the sender
                B200_TRY(write_val(rs->comm_stream, lk.prev_flags + F_FREE_KV + cur_slot, ++lk.rel_kv[cur_slot]));
            }
        }
        if (step < nranks - 1) { Kc = Kb[slot]; Vc = Vb[slot]; }
    }
    // dQ never left: its cast runs under the last pushes
    B200_TRY(cast_out(dQ, dQa, n_el, dtype, st));
    for (int s = 0; s + 1 < nranks; ++s) B200_TRY(wait_ge(st, lk.flags + F_HOME + s, epoch));
    {
        const size_t n8 = n_el / 8;                                  // (d = 128: a multiple of 8)
        dim3 grid((unsigned)std::min<size_t>((n8 + 255) / 256, (size_t)148 * 16), 2);
        reduce_partials_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<__nv_bfloat16*>(dK), reinterpret_cast<const __nv_bfloat16*>(dKp[2]),
                                                     reinterpret_cast<const __nv_bfloat16*>(homeK), reinterpret_cast<__nv_bfloat16*>(dV),
                                                     reinterpret_cast<const __nv_bfloat16*>(dVp[2]), reinterpret_cast<const __nv_bfloat16*>(homeV),
                                                     n8, nranks - 1, kvb / 2);
        B200_CUDA_TRY(cudaGetLastError());
        count_launch();
    }
    // the caller may reuse K/V once its stream has passed this call (their last reader is the first K/V push)
    B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[0], rs->comm_stream));
    B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[0], 0));
    B200_TRY(call_end(lk, rs, st));
    *done = true;
    return B200_OK;
}

namespace b200 {

static int exchange(const void* const* send, void* const* recv, const size_t* bytes, int n, int next, int prev,
                    ncclComm_t comm, cudaStream_t st)
{
#ifdef B200_EXPERIMENTS
    static int dbg = -1;                 // side libraries only: B200_RING_DBG=1 skips the transfers (timing only, wrong results)
    if (dbg < 0) { const char* e = getenv("B200_RING_DBG"); dbg = e ? atoi(e) : 0; }
    if (dbg & 1) return B200_OK;
#endif
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

int b200_ring_set_transport(int mode)
{
    if (mode < -1 || mode > 1) return B200_ERR_INVALID_ARGUMENT;
    g_transport_pref = mode;
    return B200_OK;
}

int b200_ring_set_forward_mode(int mode)
{
    if (mode < 0 || mode > 2) return B200_ERR_INVALID_ARGUMENT;
    g_fwd_mode = mode;
    return B200_OK;
}

int b200_ring_set_backward_mode(int mode)
{
    if (mode < 0 || mode > 1) return B200_ERR_INVALID_ARGUMENT;
    g_bwd_mode = mode;
    return B200_OK;
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
    for (int i = 0; i < 64; ++i)
        if (g_link[i].comm == comm) { link_close(g_link[i]); g_link[i].comm = nullptr; g_link[i].ok = false; }
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

}  // extern "C"

// Causal ring (B200_FLAG_CAUSAL): ZIGZAG sharding.  The sequence is cut into 2P chunks of c = Nl/2 rows and rank r
// holds chunk r (local rows [0,c)) followed by chunk 2P-1-r (rows [c,2c)) -- the load-balanced placement the
// reference's notes ask for (__info__/README.md:173).  With that placement every ring step is a plain rectangle:
//   step 0 (own block)            : causal attention on the local 2c x 2c problem;
//   block from a rank b < r       : ALL local queries against the FIRST half of the block (chunk b precedes both
//                                   local chunks, chunk 2P-1-b follows both);
//   block from a rank b > r       : the SECOND half of the local queries against ALL of the block.
// Each step is 2c^2 score pairs on every rank -- no rank idles -- and needs no mask beyond step 0; the halves are
// row sub-blocks of the same buffers, addressed through the strided forward.
static int ring_fwd_impl(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nl, int d,
                         float scale, int dtype, int flags, ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks || B < 0 || H < 0 || Nl < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F32) return B200_ERR_INVALID_ARGUMENT;
    if ((flags & ~B200_FLAG_CAUSAL) != 0) return B200_ERR_INVALID_ARGUMENT;
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if ((size_t)B * H * Nl == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L) return B200_ERR_INVALID_ARGUMENT;     // the merge needs the LSE
    if (nranks == 1) return b200_fa2_fwd_ex(Q, K, V, O, L, B, H, Nl, Nl, d, scale, dtype, flags, st);
    if (!comm) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    const int c = Nl / 2;
    if (causal) {
        if ((Nl & 1) || nranks > 32) return B200_ERR_INVALID_ARGUMENT;
        const uintptr_t al = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)O;
        if (!b200_fwd_uses_tensor_cores(dtype, d, c, c) || (al & 15) || B > 65535 || H > 65535) return B200_ERR_UNSUPPORTED;
        if (!(scale > 0.0f)) scale = 1.0f / sqrtf((float)d);
    }

    RingState* rs;
    B200_TRY(ring_state(&rs));
    const size_t es = dtype == B200_DTYPE_F32 ? 4 : 2;
    const size_t kv_bytes = (size_t)B * H * Nl * d * es;
    const size_t l_bytes = (size_t)B * H * Nl * sizeof(float);

    // ---- persistent forward: every block gets its OWN receive slot, the copy engines forward block after
    // block around the ring with no reference to the compute (push s waits only for arrival s), and ONE attention
    // launch walks the blocks in arrival order, its TMA producer polling the arrival counters the pushes write.  No
    // per-step launches, no partial outputs, no merge kernel; the kernel never waits on a kernel of another rank, only
    // on copy-engine traffic that is independent of every SM.
    {
        const uintptr_t al = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)O;
        // AUTO = stepwise.  The persistent schedule is built, correct (RING_OK at 2 / 4 / 8 ranks) and measured -- and does not
        // pay on this machine: a CTA of the persistent kernel needs block s+1 one CTA-block-time (~0.12 ms at Nl = 8192) after
        // block s, the stepwise schedule only a whole launch later, so the first wave stalls on the hop chain (P - 1 sequential
        // hops of transfer + two stream mem-ops) while holding every SM, and the wave quantisation that costs the stepwise
        // schedule its efficiency (512 CTAs on 148 SMs) is the same for a persistent launch of the same CTAs.
        // tools/ring_modes.py: 2 GPUs x0.64 ... x1.04 (profiles/r2c_ring_modes_2gpu.txt), 8 GPUs x0.86 ... x0.98 at 128-512
        // CTAs per rank (profiles/r2d_ring_modes_8gpu.txt); C4's shape in the fwd+bwd loop 7.2 ms against 4.2 ms.
        // b200_ring_set_forward_mode(B200_RING_FWD_PERSISTENT) selects it.
        const bool pays = g_fwd_mode == 2;
        const bool want = g_fwd_mode != 1 && pays && !causal && (long long)B * H <= 65535 && (al & 15) == 0 &&
                          fwd_ring_persistent_supported(dtype, d, B * H, Nl, nranks) && nranks <= 256;
        if (want) {
            const size_t need = 2 * (size_t)(nranks - 1) * kv_bytes;      // (kv_bytes is a multiple of 16 KB here)
            void* base;
            bool ce = false;
            B200_TRY(link_prepare(rs, comm, rank, nranks, need, &base, &ce));
            if (ce) {
                int dev = 0;
                B200_CUDA_TRY(cudaGetDevice(&dev));
                PeerLink& lk = g_link[dev];
                uint8_t* Ks = reinterpret_cast<uint8_t*>(base);
                uint8_t* Vs = Ks + (size_t)(nranks - 1) * kv_bytes;
                const uint32_t epoch = ++lk.pers_epoch;
                const float sc = scale > 0.0f ? scale : 1.0f / sqrtf((float)d);
                B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
                B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
                B200_TRY(call_begin(lk, rs));
                for (int s = 0; s + 1 < nranks; ++s) {
                    // block s (the local shard for s = 0, else what arrived in slot s - 1) goes into the next rank's slot s
                    if (s >= 1) B200_TRY(wait_ge(rs->comm_stream, lk.flags + F_PRDY + s, epoch));
                    const void* ks = s == 0 ? K : (const void*)(Ks + (size_t)(s - 1) * kv_bytes);
                    const void* vs = s == 0 ? V : (const void*)(Vs + (size_t)(s - 1) * kv_bytes);
                    B200_CUDA_TRY(cudaMemcpyAsync(lk.next_buf + (size_t)s * kv_bytes, ks, kv_bytes, cudaMemcpyDeviceToDevice, rs->comm_stream));
                    B200_CUDA_TRY(cudaMemcpyAsync(lk.next_buf + (size_t)(nranks - 1 + s) * kv_bytes, vs, kv_bytes, cudaMemcpyDeviceToDevice,
                                                  rs->comm_stream));
                    B200_TRY(write_val(rs->comm_stream, lk.next_flags + F_PRDY + s + 1, epoch));
                    // the caller may reuse K/V (the source of the FIRST push only) once its stream has passed this call
                    if (s == 0) B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[0], rs->comm_stream));
                }
                B200_TRY(launch_fwd_ring_persistent(Q, K, V, O, L, Ks, Vs, lk.flags + F_PRDY, epoch, B * H, Nl, nranks, sc, st));
                B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[0], 0));
                B200_TRY(call_end(lk, rs, st));
                return B200_OK;
            }
        }
    }
    // scratch: 2 x (K,V) receive buffers + NB x (partial O + partial LSE).  One partial per ring step when that fits
    // in 4 GB (it does for every shape in BASELINE.json): no step then waits for a merge -- they are all folded by one
    // n-way merge at the end -- and each attention kernel can start in the tail of the previous one.
    const size_t part_bytes = align256(kv_bytes) + align256(l_bytes);
    int NB = (int)std::min<size_t>((size_t)std::min(nranks, 32), std::max<size_t>(2, ((size_t)4 << 30) / part_bytes));   // (merge: <= 32 partials)
    if (const char* e = getenv("B200_RING_MAX_PARTIALS")) NB = std::max(2, std::min(NB, atoi(e)));   // test hook (same on all ranks)
    const size_t total = 4 * align256(kv_bytes) + (size_t)NB * part_bytes;
    void* base;
    bool ce = false;
    B200_TRY(link_prepare(rs, comm, rank, nranks, total, &base, &ce));
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    PeerLink& lk = g_link[dev];
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    void* Kb[2]; void* Vb[2]; size_t offK[2], offV[2];
    for (int i = 0; i < 2; ++i) {
        offK[i] = (size_t)(p - reinterpret_cast<uint8_t*>(base)); Kb[i] = p; p += align256(kv_bytes);
        offV[i] = (size_t)(p - reinterpret_cast<uint8_t*>(base)); Vb[i] = p; p += align256(kv_bytes);
    }
    // partial s: O at Opart + s * part_bytes, LSE right behind it
    uint8_t* Opart = p;
    float* Lpart = reinterpret_cast<float*>(p + align256(kv_bytes));
    auto Opb = [&](int i) { return (void*)(Opart + (size_t)i * part_bytes); };
    auto Lpb = [&](int i) { return reinterpret_cast<float*>(Opart + (size_t)i * part_bytes + align256(kv_bytes)); };
    const bool deferred = ce || causal;       // one partial per step + n-way merge (the NCCL non-causal path keeps pairwise merges)
    if (!deferred) NB = 2;
    unsigned upper_mask = 0u;                 // causal: partials of the current group that only cover the rows >= c
    const Lay lay = lay_contiguous(H, Nl, d);

    const int next = (rank + 1) % nranks, prev = (rank - 1 + nranks) % nranks;
    // the side streams may start once everything already enqueued on the caller's stream is done
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
    if (ce) {
        B200_CUDA_TRY(cudaStreamWaitEvent(rs->attn_stream[0], rs->ev_main, 0));
        B200_CUDA_TRY(cudaStreamWaitEvent(rs->attn_stream[1], rs->ev_main, 0));
        B200_TRY(call_begin(lk, rs));
    }

    const void* Kc = K; const void* Vc = V;
    for (int step = 0; step < nranks; ++step) {
        const int slot = step & 1;            // receive slot for the block used at step+1
        const int cur_slot = slot ^ 1;        // slot holding the block of this step (step >= 1)
        // copy-engine mode: the attention of consecutive steps runs on two alternating streams (the merges stay
        // on the caller's stream), so a step's last, partly filled wave overlaps the next step's first one
        cudaStream_t as = ce ? rs->attn_stream[step & 1] : st;
        const int pb = deferred ? step % NB : (step & 1);
        void* Op = Opb(pb); float* Lp = Lpb(pb);
        if (ce && step >= 1) {
            // the block of this step was pushed into cur_slot by the previous rank: both streams wait for it
            const uint32_t k = ++lk.in_kv[cur_slot];
            B200_TRY(wait_ge(as, lk.flags + F_RDY_KV + cur_slot, k));
            B200_TRY(wait_ge(rs->comm_stream, lk.flags + F_RDY_KV + cur_slot, k));
        }
        if (step < nranks - 1) {
            if (ce) {
                B200_TRY(ce_push_kv(lk, rs, slot, Kc, Vc, offK[slot], offV[slot], kv_bytes));
            } else {
                // receive slot `slot` held the block that step-1 attended to; wait until that read is done
                if (step >= 2) B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_compute[slot], 0));
                const void* s[2] = {Kc, Vc};
                void* r[2] = {Kb[slot], Vb[slot]};
                size_t b[2] = {kv_bytes, kv_bytes};
                B200_TRY(exchange(s, r, b, 2, next, prev, comm, rs->comm_stream));
                B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[slot], rs->comm_stream));
            }
        }
        // attention of the local queries against the current block (overlaps the exchange above)
        if (deferred) {
            // (only when the partials did not all fit) the group merge that consumed this buffer must be done
            if (ce && step >= NB) B200_CUDA_TRY(cudaStreamWaitEvent(as, rs->ev_merge[0], 0));
            const int blk = (rank - step + nranks) % nranks;
            if (!causal) {
                B200_TRY(b200_fa2_fwd(Q, Kc, Vc, Op, Lp, B, H, Nl, Nl, d, scale, dtype, as));
            } else if (step == 0) {
                B200_TRY(b200_fa2_fwd_ex(Q, Kc, Vc, Op, Lp, B, H, Nl, Nl, d, scale, dtype, B200_FLAG_CAUSAL, as));
            } else if (blk < rank) {
                FwdArgs a{Q, Kc, Vc, Op, Lp, lay, lay, lay, lay, (long long)Nl, B, H, H, Nl, c, d, scale, dtype, false};
                B200_TRY(tc_fwd_strided(a, as));
            } else {
                const size_t ro = (size_t)c * d * es;              // byte offset of local row c
                FwdArgs a{(const uint8_t*)Q + ro, Kc, Vc, (uint8_t*)Op + ro, Lp + c, lay, lay, lay, lay, (long long)Nl,
                          B, H, H, c, Nl, d, scale, dtype, false};
                B200_TRY(tc_fwd_strided(a, as));
                upper_mask |= 1u << pb;
            }
            if (ce) {
                B200_CUDA_TRY(cudaEventRecord(rs->ev_attn[step & 1], as));
                B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_attn[step & 1], 0));
            }
            if (pb == NB - 1 || step == nranks - 1) {
                // fold partials 0 .. pb (and, after the first group, the running result) into (O, L)
                B200_TRY(lse_merge_multi(O, L, Opart, Lpart, part_bytes / es, part_bytes / sizeof(float), pb + 1, step >= NB,
                                         (size_t)B * H * Nl, d, dtype, st, Nl, c, upper_mask));
                upper_mask = 0u;
                if (ce) B200_CUDA_TRY(cudaEventRecord(rs->ev_merge[0], st));
            }
        } else {
            if (step == 0) B200_TRY(b200_fa2_fwd(Q, Kc, Vc, O, L, B, H, Nl, Nl, d, scale, dtype, as));
            else B200_TRY(b200_fa2_fwd(Q, Kc, Vc, Op, Lp, B, H, Nl, Nl, d, scale, dtype, as));
            if (step >= 1) B200_TRY(b200_lse_merge(O, L, Op, Lp, B, H, Nl, d, dtype, st));
        }
        if (ce) {
            if (step >= 1) {
                // cur_slot is free again once this step's attention AND its onward push are done: tell the sender
                B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_attn[step & 1], 0));
                B200_TRY(write_val(rs->comm_stream, lk.prev_flags + F_FREE_KV + cur_slot, ++lk.rel_kv[cur_slot]));
            }
        } else if (step < nranks - 1) {
            // block of step (in slot^1 for step >= 1) is now free for the receive of step+1
            B200_CUDA_TRY(cudaEventRecord(rs->ev_compute[slot ^ 1], st));
            B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[slot], 0));
        }
        if (step < nranks - 1) { Kc = Kb[slot]; Vc = Vb[slot]; }
    }
    if (ce) {
        // the caller may reuse K/V (the source of the first push) once its stream has passed this call
        B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[0], rs->comm_stream));
        B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[0], 0));
        B200_TRY(call_end(lk, rs, st));           // (behind the final merge: the partials have been read)
    }
    return B200_OK;
}

extern "C" {

int b200_ring_fwd(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nl, int d,
                  float scale, int dtype, ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    return ring_fwd_impl(Q, K, V, O, L, B, H, Nl, d, scale, dtype, 0, comm, rank, nranks, st);
}

int b200_ring_zigzag_chunk(int rank, int half, int nranks)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks || (half != 0 && half != 1)) return -1;
    return half == 0 ? rank : 2 * nranks - 1 - rank;
}

int b200_ring_fwd_ex(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nl, int d,
                     float scale, int dtype, int flags, ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    return ring_fwd_impl(Q, K, V, O, L, B, H, Nl, d, scale, dtype, flags, comm, rank, nranks, st);
}

}  // extern "C"

// Causal (zigzag) ring backward: the rectangles of ring_fwd_impl, with the dK/dV partial of a "first half of the
// block" step added into the first half of the travelling accumulators only.
static int ring_bwd_impl(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                         void* dQ, void* dK, void* dV, int B, int H, int Nl, int d, float scale, int dtype, int flags,
                         ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks || B < 0 || H < 0 || Nl < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F32) return B200_ERR_INVALID_ARGUMENT;
    if ((flags & ~B200_FLAG_CAUSAL) != 0) return B200_ERR_INVALID_ARGUMENT;
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if ((size_t)B * H * Nl == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L || !dO || !dQ || !dK || !dV) return B200_ERR_INVALID_ARGUMENT;
    if (nranks == 1) return b200_fa2_bwd_ex(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nl, Nl, d, scale, dtype, flags, nullptr, 0, st);
    if (!comm) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    const int c = Nl / 2;
    if (causal) {
        // needs the tensor-core step kernel (bf16, d = 128) and half-blocks that are whole 64-row statistics tiles
        if (!(tc_bwd_ring_supported(dtype, d, B * H, Q, K, V, dO) && b200_bwd_uses_tensor_cores(dtype, d, Nl, Nl)) || Nl % 128 != 0)
            return B200_ERR_UNSUPPORTED;
    }

    RingState* rs;
    B200_TRY(ring_state(&rs));
    const size_t es = dtype == B200_DTYPE_F32 ? 4 : 2;
    const size_t n_el = (size_t)B * H * Nl * d;
    const size_t kv_bytes = n_el * es, acc_bytes = n_el * sizeof(float);
    // tensor-core steps (bf16, d = 128): D and the padded statistics are computed once per call and each step adds
    // its dQ straight into the fp32 accumulator; otherwise every step is a full b200_fa2_bwd + accumulate
    const bool fused = tc_bwd_ring_supported(dtype, d, B * H, Q, K, V, dO) && b200_bwd_uses_tensor_cores(dtype, d, Nl, Nl);
    if (fused && !causal && g_bwd_mode != 1) {
        bool done = false;
        B200_TRY(ring_bwd_home(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nl, d, scale, comm, rank, nranks, st, &done));
        if (done) return B200_OK;
    }
    const size_t d_bytes = align256((size_t)B * H * Nl * sizeof(float));
    const size_t stats_bytes = fused ? tc_bwd_ring_stats_bytes(B * H, Nl) : 0;
    // scratch: 2 x (K,V) recv, 3 partial grads (I/O dtype), dQ fp32 acc, 2 x (dK,dV) fp32 acc, D, padded stats
    const size_t total = 4 * align256(kv_bytes) + 3 * align256(kv_bytes) + align256(acc_bytes) + 4 * align256(acc_bytes) +
                         d_bytes + align256(stats_bytes);
    void* base;
    bool ce = false;
    B200_TRY(link_prepare(rs, comm, rank, nranks, total, &base, &ce));
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    PeerLink& lk = g_link[dev];
    uint8_t* p = reinterpret_cast<uint8_t*>(base);
    auto off = [&](const void* q) { return (size_t)(reinterpret_cast<const uint8_t*>(q) - reinterpret_cast<uint8_t*>(base)); };
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
    float* Dvec = reinterpret_cast<float*>(p); p += d_bytes;
    void* stats = p;
    const int next = (rank + 1) % nranks, prev = (rank - 1 + nranks) % nranks;

    if (fused) {
        B200_TRY(rowdot(dO, O, Dvec, (size_t)B * H * Nl, d, dtype, st));
        B200_TRY(tc_bwd_ring_prepare(L, Dvec, stats, B * H, Nl, st));
    }
    B200_CUDA_TRY(cudaMemsetAsync(dQa, 0, acc_bytes, st));
    // (dKa[0] / dVa[0] are written, not accumulated into, by step 0: no memset)
    const int BH = B * H;
    const int Nl_pad64 = (Nl + 63) / 64 * 64;
    const float* L2p = reinterpret_cast<const float*>(stats);
    const float* Dp = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(stats) + align256((size_t)BH * Nl_pad64 * sizeof(float)));
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
    if (ce) B200_TRY(call_begin(lk, rs));

    // Step s: block b = (rank - s) mod P is resident (K/V in Kc/Vc, its running fp32 dK/dV in acc[cur]).
    // Compute partial grads, add them, then pass (dK,dV) accumulators on; K/V for step s+1 are
    // prefetched on the side stream while step s computes.  After the last step one more hop
    // brings every block's dK/dV home.
    const void* Kc = K; const void* Vc = V;
    int cur = 0;                                       // which fp32 accumulator pair is resident
    for (int step = 0; step < nranks; ++step) {
        const int slot = step & 1, cur_slot = slot ^ 1;
        if (ce && step >= 1) {
            const uint32_t k = ++lk.in_kv[cur_slot];
            B200_TRY(wait_ge(st, lk.flags + F_RDY_KV + cur_slot, k));
            B200_TRY(wait_ge(rs->comm_stream, lk.flags + F_RDY_KV + cur_slot, k));
        }
        if (step < nranks - 1) {
            if (ce) {
                B200_TRY(ce_push_kv(lk, rs, slot, Kc, Vc, off(Kb[slot]), off(Vb[slot]), kv_bytes));
            } else {
                if (step >= 2) B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_compute[slot], 0));
                const void* s[2] = {Kc, Vc};
                void* r[2] = {Kb[slot], Vb[slot]};
                size_t b[2] = {kv_bytes, kv_bytes};
                B200_TRY(exchange(s, r, b, 2, next, prev, comm, rs->comm_stream));
                B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[slot], rs->comm_stream));
            }
        }
        const int blk = (rank - step + nranks) % nranks;
        const bool first_half_only = causal && step > 0 && blk < rank;   // dK/dV of this step exist for rows [0, c) only
        // The LAST step is cut into head groups: a group's accumulators start their final hop (home) while the next group
        // computes, so only the last group's transfer is exposed (r1: the whole 2 x acc_bytes hop was -- 0.75 ms of an
        // 11.9 ms call at 2 GPUs).  Groups keep >= ~6 waves of CTAs each, so small per-rank head counts are not split.
        int groups = 1;
        if (fused && !causal && step == nranks - 1) {
            const long long ctas = (long long)BH * ((Nl + 127) / 128);
            groups = (int)std::max<long long>(1, std::min<long long>(std::min(4, BH), ctas / (6 * num_sms())));
        }
        const size_t es_part = dtype == B200_DTYPE_F32 ? 4 : 2;
        for (int g = 0; g < groups; ++g) {
            const int h0 = (int)((long long)BH * g / groups), h1 = (int)((long long)BH * (g + 1) / groups);
            const size_t e0 = (size_t)h0 * Nl * d, ne = (size_t)(h1 - h0) * Nl * d;     // element range of this head group
            if (fused) {
                const float sc = scale == 0.0f ? 1.0f / sqrtf((float)d) : scale;
                if (!causal) B200_TRY(tc_bwd_ring_step_heads(Q, Kc, Vc, dO, L2p, Dp, dQa, dKp, dVp, h0, h1 - h0, Nl, 0, Nl, Nl, false, sc, st));
                else if (step == 0) B200_TRY(tc_bwd_ring_step_sub(Q, Kc, Vc, dO, stats, dQa, dKp, dVp, B * H, Nl, 0, Nl, Nl, true, sc, st));
                else if (blk < rank) B200_TRY(tc_bwd_ring_step_sub(Q, Kc, Vc, dO, stats, dQa, dKp, dVp, B * H, Nl, 0, Nl, c, false, sc, st));
                else B200_TRY(tc_bwd_ring_step_sub(Q, Kc, Vc, dO, stats, dQa, dKp, dVp, B * H, Nl, c, c, Nl, false, sc, st));
            } else {
                B200_TRY(b200_fa2_bwd(Q, Kc, Vc, O, L, dO, dQp, dKp, dVp, B, H, Nl, Nl, d, scale, dtype, nullptr, 0, st));
                B200_TRY(accum(dQa, dQp, n_el, dtype, st));
            }
            // the running dK/dV of this block arrive from the previous rank while we compute
            if (step >= 1 && g == 0) {
                if (ce) B200_TRY(wait_ge(st, lk.flags + F_RDY_ACC + cur, ++lk.in_acc[cur]));
                else B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_acc, 0));
            }
            if (first_half_only) {
                B200_TRY(accum_rows(dKa[cur], dKp, (size_t)B * H, Nl, c, d, dtype, st));
                B200_TRY(accum_rows(dVa[cur], dVp, (size_t)B * H, Nl, c, d, dtype, st));
            } else {
                // (step 0: the accumulators START as this rank's contribution -- written, not added to a zeroed buffer)
                B200_TRY(accum2(dKa[cur] + e0, (const uint8_t*)dKp + e0 * es_part, dVa[cur] + e0, (const uint8_t*)dVp + e0 * es_part, ne, dtype,
                                step == 0, st));
            }
            // hand the accumulators to the next rank (queued behind this step's K/V prefetch)
            B200_CUDA_TRY(cudaEventRecord(rs->ev_main, st));
            B200_CUDA_TRY(cudaStreamWaitEvent(rs->comm_stream, rs->ev_main, 0));
            const bool last_group = g == groups - 1;
            if (ce) {
                const int dst = cur ^ 1;
                // the neighbour's accumulator pair `dst` must have been passed on as often as ours was
                if (g == 0 && lk.out_acc[dst] > 0) B200_TRY(wait_ge(rs->comm_stream, lk.flags + F_FREE_ACC + dst, lk.out_acc[dst]));
                B200_CUDA_TRY(cudaMemcpyAsync(lk.next_buf + off(dKa[dst] + e0), dKa[cur] + e0, ne * sizeof(float), cudaMemcpyDeviceToDevice, rs->comm_stream));
                B200_CUDA_TRY(cudaMemcpyAsync(lk.next_buf + off(dVa[dst] + e0), dVa[cur] + e0, ne * sizeof(float), cudaMemcpyDeviceToDevice, rs->comm_stream));
                if (last_group) {
                    B200_TRY(write_val(rs->comm_stream, lk.next_flags + F_RDY_ACC + dst, ++lk.push_acc[dst]));
                    B200_TRY(write_val(rs->comm_stream, lk.prev_flags + F_FREE_ACC + cur, ++lk.out_acc[cur]));
                    if (step >= 1)      // this step's K/V block (compute + onward push both ordered before this point) is released too
                        B200_TRY(write_val(rs->comm_stream, lk.prev_flags + F_FREE_KV + cur_slot, ++lk.rel_kv[cur_slot]));
                }
            } else {
                const void* sx[2] = {dKa[cur] + e0, dVa[cur] + e0};
                void* rx[2] = {dKa[cur ^ 1] + e0, dVa[cur ^ 1] + e0};
                size_t bx[2] = {ne * sizeof(float), ne * sizeof(float)};
                B200_TRY(exchange(sx, rx, bx, 2, next, prev, comm, rs->comm_stream));
                if (last_group) B200_CUDA_TRY(cudaEventRecord(rs->ev_acc, rs->comm_stream));
            }
        }
        cur ^= 1;
        if (!ce && step < nranks - 1) {
            B200_CUDA_TRY(cudaEventRecord(rs->ev_compute[slot ^ 1], st));
            B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[slot], 0));
        }
        if (step < nranks - 1) { Kc = Kb[slot]; Vc = Vb[slot]; }
    }
    // dQ never travelled: its cast runs under the accumulators' last hop
    B200_TRY(cast_out(dQ, dQa, n_el, dtype, st));
    // after P hops the accumulators of block `rank` are back home (in pair `cur`)
    if (ce) {
        B200_TRY(wait_ge(st, lk.flags + F_RDY_ACC + cur, ++lk.in_acc[cur]));
        B200_CUDA_TRY(cudaEventRecord(rs->ev_comm[0], rs->comm_stream));
        B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_comm[0], 0));
    } else {
        B200_CUDA_TRY(cudaStreamWaitEvent(st, rs->ev_acc, 0));
    }
    B200_TRY(cast_out(dK, dKa[cur], n_el, dtype, st));
    B200_TRY(cast_out(dV, dVa[cur], n_el, dtype, st));
    if (ce) B200_TRY(call_end(lk, rs, st));          // (behind the casts: the home accumulators have been read)
    return B200_OK;
}

extern "C" {

int b200_ring_bwd(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                  void* dQ, void* dK, void* dV, int B, int H, int Nl, int d, float scale, int dtype,
                  ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    return ring_bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nl, d, scale, dtype, 0, comm, rank, nranks, st);
}

int b200_ring_bwd_ex(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                     void* dQ, void* dK, void* dV, int B, int H, int Nl, int d, float scale, int dtype, int flags,
                     ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    return ring_bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, B, H, Nl, d, scale, dtype, flags, comm, rank, nranks, st);
}

}  // extern "C"
