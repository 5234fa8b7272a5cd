// api.cu -- the extern "C" boundary declared in include/b200attn.h: argument validation, routing
// between the tcgen05 tensor-core kernels and the generic CUDA-core kernels, scratch arena.
#include "common.cuh"
#include <math.h>
#include <atomic>
#include <mutex>

namespace b200 {

static thread_local int g_last_cuda_error = 0;
void set_last_cuda_error(int code) { g_last_cuda_error = code; }
static std::atomic<unsigned long long> g_launches{0};     // callers may launch from several host threads
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

struct Arena { void* ptr = nullptr; size_t bytes = 0; };
static Arena g_arena[64];
static std::mutex g_arena_mu;

int arena_get(size_t bytes, void** out)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lk(g_arena_mu);
    Arena& a = g_arena[dev];
    if (a.bytes < bytes) {
        if (a.ptr) { B200_CUDA_TRY(cudaDeviceSynchronize()); B200_CUDA_TRY(cudaFree(a.ptr)); a.ptr = nullptr; a.bytes = 0; }
        size_t want = (bytes + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
        B200_CUDA_TRY(cudaMalloc(&a.ptr, want));
        a.bytes = want;
    }
    *out = a.ptr;
    return B200_OK;
}

static int g_dev_state[64];   // 0 unknown, 1 ok, -1 wrong arch
static int g_dev_sms[64];
int device_check()
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_last_cuda_error((int)e); (void)cudaGetLastError(); return B200_ERR_NO_DEVICE; }
    if (dev < 0 || dev >= 64) return B200_ERR_INVALID_ARGUMENT;
    if (g_dev_state[dev] == 0) {
        int major = 0, sms = 0;
        e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess) { set_last_cuda_error((int)e); (void)cudaGetLastError(); return B200_ERR_NO_DEVICE; }
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        g_dev_sms[dev] = sms;
        g_dev_state[dev] = (major == 10) ? 1 : -1;
    }
    return g_dev_state[dev] == 1 ? B200_OK : B200_ERR_ARCH;
}
int num_sms()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    return g_dev_sms[dev] > 0 ? g_dev_sms[dev] : 148;
}

static int g_force_generic = 0;   // test hook: route everything to the CUDA-core kernels (GPU cross-check)

static inline bool valid_dtype(int dt) { return dt == B200_DTYPE_BF16 || dt == B200_DTYPE_F32; }
static inline float resolve_scale(float s, int d) { return s == 0.0f ? 1.0f / sqrtf((float)d) : s; }

// The routing predicates, shared by the calls and by b200_fa2_fwd_route / b200_fa2_bwd_route so that what a caller is
// told is what runs.  Tensor-core path: supported (dtype, d), 16-byte aligned bases (TMA), scale > 0 (the forward folds
// scale into the running max); the fp32 (tf32) backward is d = 64.  Everything else: generic CUDA-core kernels.
static bool fwd_routes_tc(const void* Q, const void* K, const void* V, const void* O, int dtype, int d, int Nq, int Nkv,
                          float scale, bool causal)
{
    const uintptr_t al = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)O;
    (void)causal;                   // (causal runs on the tensor cores for any Nq <= Nkv since r2e)
    return !g_force_generic && tc_fwd_supported(dtype, d, Nq, Nkv) && (al & 15) == 0 && scale > 0.0f;
}
static bool bwd_routes_tc(const void* Q, const void* K, const void* V, const void* dO, const void* dQ, const void* dK,
                          const void* dV, int dtype, int d, int Nq, int Nkv, bool causal)
{
    const uintptr_t al = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)dO | (uintptr_t)dQ | (uintptr_t)dK | (uintptr_t)dV;
    if (g_force_generic || (al & 15) != 0) return false;
    (void)causal;                                                            // (causal runs on both tensor-core backwards)
    if (tc_bwd_tf32_supported(dtype, d, Nq, Nkv)) return true;               // fp32 I/O, d = 64: tf32 kernel
    return tc_bwd_supported(dtype, d, Nq, Nkv);
}

static int fwd_impl(const void* Q, const void* K, const void* V, void* O, float* L, int BH, int Nq, int Nkv, int d,
                    float scale, int dtype, cudaStream_t st, int flags = 0)
{
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if ((flags & ~B200_FLAG_CAUSAL) != 0 || (causal && Nq > Nkv)) return B200_ERR_INVALID_ARGUMENT;
    if (!(valid_dtype(dtype) || dtype == B200_DTYPE_FP8_E4M3) || BH < 0 || Nq < 0 || Nkv < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (BH == 0 || Nq == 0) return B200_OK;
    if (Nkv == 0) return B200_ERR_INVALID_ARGUMENT;            // softmax over an empty set is undefined
    if (!Q || !K || !V || !O) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    scale = resolve_scale(scale, d);
    // tensor-core causal path: square problems (the usual self-attention case); everything else runs on the generic kernels
    if (fwd_routes_tc(Q, K, V, O, dtype, d, Nq, Nkv, scale, causal))
        return tc_fwd(Q, K, V, O, L, BH, Nq, Nkv, d, scale, dtype, st, causal);
    if (dtype == B200_DTYPE_FP8_E4M3) return B200_ERR_UNSUPPORTED;        // fp8 exists on the tensor-core kernel only
    return simt_fwd(Q, K, V, O, L, nullptr, BH, Nq, Nkv, d, scale, dtype, 0, st, causal ? Nkv - Nq : kNoCausal);
}

static size_t bwd_ws_bytes(int BH, int Nq, int Nkv, int d, int dtype)
{
    size_t dbytes = (((size_t)BH * Nq * sizeof(float)) + 255) & ~(size_t)255;
    size_t tc = tc_bwd_supported(dtype, d, Nq, Nkv) ? tc_bwd_workspace_bytes(BH, Nq, Nkv, d, dtype) : 0;
    if (tc_bwd_tf32_supported(dtype, d, Nq, Nkv)) tc = tc_bwd_tf32_workspace_bytes(BH, Nq, Nkv);
    return dbytes + tc;
}

static int bwd_impl(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                    void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, int dtype,
                    void* ws, size_t ws_bytes, cudaStream_t st, int flags = 0)
{
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if ((flags & ~B200_FLAG_CAUSAL) != 0 || (causal && Nq > Nkv)) return B200_ERR_INVALID_ARGUMENT;
    if (dtype == B200_DTYPE_FP8_E4M3) return B200_ERR_UNSUPPORTED;         // forward-only dtype
    if (!valid_dtype(dtype) || BH < 0 || Nq < 0 || Nkv < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (BH == 0 || (Nq == 0 && Nkv == 0)) return B200_OK;
    if (Nq == 0 || Nkv == 0) return B200_ERR_INVALID_ARGUMENT;
    if (!Q || !K || !V || !O || !L || !dO || !dQ || !dK || !dV) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    scale = resolve_scale(scale, d);
    const size_t need = bwd_ws_bytes(BH, Nq, Nkv, d, dtype);
    if (ws == nullptr) B200_TRY(arena_get(need, &ws));
    else if (ws_bytes < need) return B200_ERR_WORKSPACE;
    float* D = reinterpret_cast<float*>(ws);
    const size_t dbytes = (((size_t)BH * Nq * sizeof(float)) + 255) & ~(size_t)255;
    B200_TRY(rowdot(dO, O, D, (size_t)BH * Nq, d, dtype, st));
    if (bwd_routes_tc(Q, K, V, dO, dQ, dK, dV, dtype, d, Nq, Nkv, causal)) {
        if (dtype == B200_DTYPE_F32)
            return tc_bwd_tf32(Q, K, V, dO, L, D, dQ, dK, dV, BH, Nq, Nkv, scale, reinterpret_cast<uint8_t*>(ws) + dbytes, need - dbytes, st,
                               causal);
        return tc_bwd(Q, K, V, dO, L, D, dQ, dK, dV, BH, Nq, Nkv, d, scale, dtype,
                      reinterpret_cast<uint8_t*>(ws) + dbytes, need - dbytes, st, causal);
    }
    return simt_bwd(Q, K, V, dO, L, D, dQ, dK, dV, BH, Nq, Nkv, d, scale, dtype, st, causal ? Nkv - Nq : kNoCausal);
}

// ---- host-buffer pipeline (b200_fa2_fwd_host / b200_fa2_fwd_bwd_host) ------------------------------
struct HostPipe {
    cudaStream_t h2d = nullptr, comp = nullptr, d2h = nullptr;
    cudaEvent_t ev_start = nullptr, ev_end = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
    void* buf = nullptr;
    size_t bytes = 0;
    bool ready = false;                          // every stream / event above exists
};
static HostPipe g_pipe[64];
static std::mutex g_pipe_mu;

// `compute(din, dout, dL, nh, stream)` enqueues the attention of one chunk of nh (b,h) slabs on `stream`;
// min_chunks > 1 asks for at least that many chunks (ring: few slabs per rank, but they must still pipeline)
template <class Compute>
static int host_pipeline(const void* hQ, const void* hK, const void* hV, const void* hdO, void* hO, float* hL,
                         void* hdQ, void* hdK, void* hdV, bool do_bwd, int B, int H, int N, int d,
                         int dtype, cudaStream_t st, int min_chunks, Compute&& compute)
{
    if (!valid_dtype(dtype) || B < 0 || H < 0 || N < 0 || d <= 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    const int BH = B * H;
    if (BH == 0 || N == 0) return B200_OK;
    if (!hQ || !hK || !hV || !hO || !hL || (do_bwd && (!hdO || !hdQ || !hdK || !hdV))) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_pipe_mu);
    HostPipe& P = g_pipe[dev];
    if (!P.ready) {                              // (a failed creation is retried by the next call, never run with null handles)
        if (!P.h2d) B200_CUDA_TRY(cudaStreamCreateWithFlags(&P.h2d, cudaStreamNonBlocking));
        if (!P.comp) B200_CUDA_TRY(cudaStreamCreateWithFlags(&P.comp, cudaStreamNonBlocking));
        if (!P.d2h) B200_CUDA_TRY(cudaStreamCreateWithFlags(&P.d2h, cudaStreamNonBlocking));
        if (!P.ev_start) B200_CUDA_TRY(cudaEventCreateWithFlags(&P.ev_start, cudaEventDisableTiming));
        if (!P.ev_end) B200_CUDA_TRY(cudaEventCreateWithFlags(&P.ev_end, cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i) {
            if (!P.ev_in[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&P.ev_in[i], cudaEventDisableTiming));
            if (!P.ev_done[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&P.ev_done[i], cudaEventDisableTiming));
            if (!P.ev_out[i]) B200_CUDA_TRY(cudaEventCreateWithFlags(&P.ev_out[i], cudaEventDisableTiming));
        }
        P.ready = true;
    }
    const size_t es = dtype == B200_DTYPE_F32 ? 4 : 2;
    const size_t slab = (size_t)N * d * es, lslab = (size_t)N * sizeof(float);
    // chunk = as many (b,h) slabs as make ONE WAVE of forward CTAs (256 query rows each), between 4 and 32 MB per tensor:
    // the call is PCIe-bound, so what counts is how soon the first copy-out starts and how late the last copy-in ends
    // (C3, r2h: 4 heads = 8 MB chunks 22.8 ms per pipelined step / 23.4 ms for one call alone; 16 heads = 32 MB: 23.5 / 26.4).
    // The ring call (min_chunks > 1: every chunk is a whole ring pass) keeps the 32 MB rule.
    static size_t chunk_bytes = 0;                 // tuning hook: B200_HOST_CHUNK_MB fixes the chunk size
    static bool chunk_fixed = false;
    if (!chunk_bytes) {
        const char* e = getenv("B200_HOST_CHUNK_MB");
        chunk_fixed = e && atoi(e) > 0;
        chunk_bytes = (size_t)(chunk_fixed ? atoi(e) : 32) << 20;
    }
    int hc = (int)(chunk_bytes / slab);
    if (!chunk_fixed && min_chunks <= 1) {
        const int ctas_per_head = (N + 255) / 256;
        const int h_wave = num_sms() / ctas_per_head, h_min = (int)((((size_t)4 << 20) + slab - 1) / slab);
        if (h_wave < hc) hc = h_wave;
        if (hc < h_min) hc = h_min;
    }
    if (min_chunks > 1 && hc > (BH + min_chunks - 1) / min_chunks) hc = (BH + min_chunks - 1) / min_chunks;
    if (hc < 1) hc = 1;
    if (hc > BH) hc = BH;
    const int nchunks = (BH + hc - 1) / hc;
    const int n_in = do_bwd ? 4 : 3, n_out = do_bwd ? 4 : 1;
    const size_t tb = (((size_t)hc * slab) + 255) & ~(size_t)255, lb = (((size_t)hc * lslab) + 255) & ~(size_t)255;
    const size_t per_set = (size_t)(n_in + n_out) * tb + lb;
    const size_t need = 2 * per_set;
    if (P.bytes < need) {
        if (P.buf) { B200_CUDA_TRY(cudaDeviceSynchronize()); B200_CUDA_TRY(cudaFree(P.buf)); P.buf = nullptr; P.bytes = 0; }
        B200_CUDA_TRY(cudaMalloc(&P.buf, need));
        P.bytes = need;
    }
    B200_CUDA_TRY(cudaEventRecord(P.ev_start, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(P.h2d, P.ev_start, 0));
    B200_CUDA_TRY(cudaStreamWaitEvent(P.d2h, P.ev_start, 0));
    const uint8_t* hin[4] = {(const uint8_t*)hQ, (const uint8_t*)hK, (const uint8_t*)hV, (const uint8_t*)hdO};
    uint8_t* hout[4] = {(uint8_t*)hO, (uint8_t*)hdQ, (uint8_t*)hdK, (uint8_t*)hdV};
    for (int c = 0; c < nchunks; ++c) {
        const int b = c & 1, h0 = c * hc, nh = (BH - h0) < hc ? (BH - h0) : hc;
        uint8_t* base = reinterpret_cast<uint8_t*>(P.buf) + (size_t)b * per_set;
        uint8_t* din[4]; uint8_t* dout[4];
        for (int i = 0; i < 4; ++i) { din[i] = base + (size_t)i * tb; dout[i] = base + (size_t)(n_in + i) * tb; }
        float* dL = reinterpret_cast<float*>(base + (size_t)(n_in + n_out) * tb);
        // host -> device (the input set b is free once the compute of the chunk that used it last -- chunk c-2, or a
        // chunk of the previous call, which may still be in flight on these streams -- is done)
        B200_CUDA_TRY(cudaStreamWaitEvent(P.h2d, P.ev_done[b], 0));
        for (int i = 0; i < n_in; ++i)
            B200_CUDA_TRY(cudaMemcpyAsync(din[i], hin[i] + (size_t)h0 * slab, (size_t)nh * slab, cudaMemcpyHostToDevice, P.h2d));
        B200_CUDA_TRY(cudaEventRecord(P.ev_in[b], P.h2d));
        // attention (the output set b is free once chunk c-2 has been copied out)
        B200_CUDA_TRY(cudaStreamWaitEvent(P.comp, P.ev_in[b], 0));
        B200_CUDA_TRY(cudaStreamWaitEvent(P.comp, P.ev_out[b], 0));
        B200_TRY(compute(din, dout, dL, nh, P.comp));
        B200_CUDA_TRY(cudaEventRecord(P.ev_done[b], P.comp));
        // device -> host
        B200_CUDA_TRY(cudaStreamWaitEvent(P.d2h, P.ev_done[b], 0));
        for (int i = 0; i < n_out; ++i)
            B200_CUDA_TRY(cudaMemcpyAsync(hout[i] + (size_t)h0 * slab, dout[i], (size_t)nh * slab, cudaMemcpyDeviceToHost, P.d2h));
        B200_CUDA_TRY(cudaMemcpyAsync(hL + (size_t)h0 * N, dL, (size_t)nh * lslab, cudaMemcpyDeviceToHost, P.d2h));
        B200_CUDA_TRY(cudaEventRecord(P.ev_out[b], P.d2h));
    }
    B200_CUDA_TRY(cudaEventRecord(P.ev_end, P.d2h));
    B200_CUDA_TRY(cudaStreamWaitEvent(st, P.ev_end, 0));
    return B200_OK;
}

// ---- side streams of the packed variable-length calls -------------------------------------------------
// The sequences of a packed batch are independent launches; on ONE stream a batch of short sequences leaves most SMs idle
// (a 512-token sequence of 16 heads is 32 forward CTAs).  They are spread round-robin over kVarlenStreams internal streams
// that fork from and join back into the caller's stream with events, so the call stays stream-ordered for the caller.
#ifndef B200_VARLEN_STREAMS
#define B200_VARLEN_STREAMS 4                    // (a side build with 1 is the single-stream behaviour, for A/B timing)
#endif
constexpr int kVarlenStreams = B200_VARLEN_STREAMS;
struct VarlenStreams {
    cudaStream_t s[kVarlenStreams] = {};
    cudaEvent_t fork = nullptr, join[kVarlenStreams] = {};
    bool ready = false;
};
static VarlenStreams g_vs[64];
static std::mutex g_vs_mu;                       // held while a call enqueues: the events are shared by the device's calls

static int varlen_streams(VarlenStreams** out)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_INVALID_ARGUMENT;
    VarlenStreams& v = g_vs[dev];
    if (!v.ready) {                              // (a failed creation is retried by the next call)
        if (!v.fork) B200_CUDA_TRY(cudaEventCreateWithFlags(&v.fork, cudaEventDisableTiming));
        for (int k = 0; k < kVarlenStreams; ++k) {
            if (!v.s[k]) B200_CUDA_TRY(cudaStreamCreateWithFlags(&v.s[k], cudaStreamNonBlocking));
            if (!v.join[k]) B200_CUDA_TRY(cudaEventCreateWithFlags(&v.join[k], cudaEventDisableTiming));
        }
        v.ready = true;
    }
    *out = &v;
    return B200_OK;
}
static int varlen_fork(VarlenStreams* v, int used, cudaStream_t st)
{
    B200_CUDA_TRY(cudaEventRecord(v->fork, st));
    for (int k = 0; k < used; ++k) B200_CUDA_TRY(cudaStreamWaitEvent(v->s[k], v->fork, 0));
    return B200_OK;
}
static int varlen_join(VarlenStreams* v, int used, cudaStream_t st)
{
    for (int k = 0; k < used; ++k) {
        B200_CUDA_TRY(cudaEventRecord(v->join[k], v->s[k]));
        B200_CUDA_TRY(cudaStreamWaitEvent(st, v->join[k], 0));
    }
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" {

const char* b200_strerror(int s)
{
    switch (s) {
        case B200_OK: return "ok";
        case B200_ERR_INVALID_ARGUMENT: return "invalid argument";
        case B200_ERR_UNSUPPORTED_HEAD_DIM: return "unsupported head dimension";
        case B200_ERR_CUDA: return "CUDA call failed (see b200_last_cuda_error)";
        case B200_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
        case B200_ERR_ARCH: return "device is not compute capability 10.x (sm_100a kernels only)";
        case B200_ERR_WORKSPACE: return "workspace too small";
        case B200_ERR_NCCL: return "NCCL call failed";
        case B200_ERR_UNSUPPORTED: return "unsupported configuration";
        default: return "unknown status";
    }
}
int b200_last_cuda_error(void) { return g_last_cuda_error; }
int b200_version(void) { return 100; }
unsigned long long b200_debug_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
#ifdef B200_TRACE
}
namespace b200 { int fwd_trace_read(long long* out, int n); int bwd_trace_read(long long* out, int n); int dbuf_trace_read(long long* out, int n); }
extern "C" {
__attribute__((visibility("default"))) int b200_debug_trace_read(long long* out, int n)
{
    const char* e = getenv("B200_FWD_VARIANT");           // the double-buffered forward (the default) keeps its own trace
    return (!e || atoi(e) == 4) ? b200::dbuf_trace_read(out, n) : b200::fwd_trace_read(out, n);
}
__attribute__((visibility("default"))) int b200_debug_trace_read_bwd(long long* out, int n) { return b200::bwd_trace_read(out, n); }
#endif
int b200_debug_force_generic(int on) { int old = g_force_generic; g_force_generic = on ? 1 : 0; return old; }

int b200_fwd_uses_tensor_cores(int dtype, int d, int Nq, int Nkv) { return tc_fwd_supported(dtype, d, Nq, Nkv) ? 1 : 0; }
int b200_bwd_uses_tensor_cores(int dtype, int d, int Nq, int Nkv)
{ return (tc_bwd_supported(dtype, d, Nq, Nkv) || tc_bwd_tf32_supported(dtype, d, Nq, Nkv)) ? 1 : 0; }

int b200_fa2_fwd_route(const void* Q, const void* K, const void* V, const void* O, int Nq, int Nkv, int d, float scale,
                       int dtype, int flags)
{
    if (!(valid_dtype(dtype) || dtype == B200_DTYPE_FP8_E4M3) || Nq < 1 || Nkv < 1 || d <= 0 || (flags & ~B200_FLAG_CAUSAL) != 0)
        return B200_ERR_INVALID_ARGUMENT;
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if (causal && Nq > Nkv) return B200_ERR_INVALID_ARGUMENT;
    if (dtype == B200_DTYPE_FP8_E4M3 && !fwd_routes_tc(Q, K, V, O, dtype, d, Nq, Nkv, resolve_scale(scale, d), causal)) return B200_ERR_UNSUPPORTED;
    return fwd_routes_tc(Q, K, V, O, dtype, d, Nq, Nkv, resolve_scale(scale, d), causal) ? B200_ROUTE_TENSOR_CORE : B200_ROUTE_GENERIC;
}
int b200_fa2_bwd_route(const void* Q, const void* K, const void* V, const void* dO, const void* dQ, const void* dK,
                       const void* dV, int Nq, int Nkv, int d, float scale, int dtype, int flags)
{
    (void)scale;
    if (!valid_dtype(dtype) || Nq < 1 || Nkv < 1 || d <= 0 || (flags & ~B200_FLAG_CAUSAL) != 0) return B200_ERR_INVALID_ARGUMENT;
    const bool causal = (flags & B200_FLAG_CAUSAL) != 0;
    if (causal && Nq > Nkv) return B200_ERR_INVALID_ARGUMENT;
    return bwd_routes_tc(Q, K, V, dO, dQ, dK, dV, dtype, d, Nq, Nkv, causal) ? B200_ROUTE_TENSOR_CORE : B200_ROUTE_GENERIC;
}

int b200_fa2_fwd_f32(const float* Q, const float* K, const float* V, float* O, float* L, int N, int d, float scale, cudaStream_t st)
{ return fwd_impl(Q, K, V, O, L, 1, N, N, d, scale, B200_DTYPE_F32, st); }

int b200_fa2_bwd_f32(const float* Q, const float* K, const float* V, const float* O, const float* L, const float* dO,
                     float* dQ, float* dK, float* dV, int N, int d, float scale, cudaStream_t st)
{ return bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, 1, N, N, d, scale, B200_DTYPE_F32, nullptr, 0, st); }

int b200_fa1_fwd_f32(const float* Q, const float* K, const float* V, float* O, float* l, float* m, int N, int d, cudaStream_t st)
{
    if (N < 0 || d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (N == 0) return B200_OK;
    if (!Q || !K || !V || !O || !l || !m) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    // (l, m) are the un-logged statistics; the generic kernel emits them directly.
    B200_TRY(simt_fwd(Q, K, V, O, l, m, 1, N, N, d, 1.0f / sqrtf((float)d), B200_DTYPE_F32, 1, st));
    B200_CUDA_TRY(cudaStreamSynchronize(st));      // reference flash_attention() synchronises (01/main.cu:64)
    return B200_OK;
}

int b200_fa2_fwd(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nq, int Nkv, int d,
                 float scale, int dtype, cudaStream_t st)
{
    if (B < 0 || H < 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    return fwd_impl(Q, K, V, O, L, B * H, Nq, Nkv, d, scale, dtype, st);
}

int b200_fa2_fwd_host(const void* hQ, const void* hK, const void* hV, void* hO, float* hL,
                      int B, int H, int N, int d, float scale, int dtype, cudaStream_t st)
{
    return host_pipeline(hQ, hK, hV, nullptr, hO, hL, nullptr, nullptr, nullptr, false, B, H, N, d, dtype, st, 1,
                         [&](uint8_t** din, uint8_t** dout, float* dL, int nh, cudaStream_t cs) {
                             return fwd_impl(din[0], din[1], din[2], dout[0], dL, nh, N, N, d, scale, dtype, cs);
                         });
}

int b200_fa2_fwd_bwd_host(const void* hQ, const void* hK, const void* hV, const void* hdO, void* hO, float* hL,
                          void* hdQ, void* hdK, void* hdV, int B, int H, int N, int d, float scale, int dtype, cudaStream_t st)
{
    return host_pipeline(hQ, hK, hV, hdO, hO, hL, hdQ, hdK, hdV, true, B, H, N, d, dtype, st, 1,
                         [&](uint8_t** din, uint8_t** dout, float* dL, int nh, cudaStream_t cs) {
                             B200_TRY(fwd_impl(din[0], din[1], din[2], dout[0], dL, nh, N, N, d, scale, dtype, cs));
                             return bwd_impl(din[0], din[1], din[2], dout[0], dL, din[3], dout[1], dout[2], dout[3], nh, N, N, d,
                                             scale, dtype, nullptr, 0, cs);
                         });
}

int b200_ring_fwd_bwd_host(const void* hQ, const void* hK, const void* hV, const void* hdO, void* hO, float* hL,
                           void* hdQ, void* hdK, void* hdV, int B, int H, int N_local, int d, float scale, int dtype, int flags,
                           ncclComm_t comm, int rank, int nranks, cudaStream_t st)
{
    // every chunk is one ring forward + backward over its heads (collective: all ranks cut B*H the same way)
    return host_pipeline(hQ, hK, hV, hdO, hO, hL, hdQ, hdK, hdV, true, B, H, N_local, d, dtype, st, 2,
                         [&](uint8_t** din, uint8_t** dout, float* dL, int nh, cudaStream_t cs) {
                             B200_TRY(b200_ring_fwd_ex(din[0], din[1], din[2], dout[0], dL, 1, nh, N_local, d, scale, dtype, flags,
                                                       comm, rank, nranks, cs));
                             return b200_ring_bwd_ex(din[0], din[1], din[2], dout[0], dL, din[3], dout[1], dout[2], dout[3], 1, nh,
                                                     N_local, d, scale, dtype, flags, comm, rank, nranks, cs);
                         });
}

int b200_fa2_fwd_ex(const void* Q, const void* K, const void* V, void* O, float* L, int B, int H, int Nq, int Nkv, int d,
                    float scale, int dtype, int flags, cudaStream_t st)
{
    if (B < 0 || H < 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    return fwd_impl(Q, K, V, O, L, B * H, Nq, Nkv, d, scale, dtype, st, flags);
}

int b200_fa2_bwd_ex(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                    void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype, int flags,
                    void* ws, size_t ws_bytes, cudaStream_t st)
{
    if (B < 0 || H < 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    return bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, B * H, Nq, Nkv, d, scale, dtype, ws, ws_bytes, st, flags);
}

static bool lay_is_zero(const b200_layout& l) { return l.batch_stride == 0 && l.head_stride == 0 && l.row_stride == 0; }
static Lay lay_of(const b200_layout& l, int H, int N, int d)
{
    return lay_is_zero(l) ? lay_contiguous(H, N, d) : Lay{l.batch_stride, l.head_stride, l.row_stride};
}
static bool lay_is_contiguous(const b200_layout& l, int H, int N, int d)
{
    const Lay c = lay_contiguous(H, N, d);
    return lay_is_zero(l) || (l.batch_stride == c.sb && l.head_stride == c.sh && l.row_stride == c.sn);
}
static int desc_check(const b200_attn_desc* p)
{
    if (!p || !valid_dtype(p->dtype) || p->B < 0 || p->H < 0 || p->Nq < 0 || p->Nkv < 0 || p->d <= 0) return B200_ERR_INVALID_ARGUMENT;
    if (p->H > 0 && (p->Hkv < 1 || p->Hkv > p->H || p->H % p->Hkv != 0)) return B200_ERR_INVALID_ARGUMENT;
    if ((p->flags & ~B200_FLAG_CAUSAL) != 0 || ((p->flags & B200_FLAG_CAUSAL) && p->Nq > p->Nkv)) return B200_ERR_INVALID_ARGUMENT;
    if ((long long)p->B * p->H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    return B200_OK;
}

int b200_fa2_fwd_desc(const b200_attn_desc* p, const void* Q, const void* K, const void* V, void* O, float* L, cudaStream_t st)
{
    B200_TRY(desc_check(p));
    const bool plain = p->Hkv == p->H && lay_is_contiguous(p->q, p->H, p->Nq, p->d) && lay_is_contiguous(p->o, p->H, p->Nq, p->d) &&
                       lay_is_contiguous(p->k, p->H, p->Nkv, p->d) && lay_is_contiguous(p->v, p->H, p->Nkv, p->d);
    if (plain) return fwd_impl(Q, K, V, O, L, p->B * p->H, p->Nq, p->Nkv, p->d, p->scale, p->dtype, st, p->flags);
    if ((size_t)p->B * p->H == 0 || (p->Nq == 0 && p->Nkv == 0)) return B200_OK;
    if (p->Nq == 0 || p->Nkv == 0 || !Q || !K || !V || !O || !L) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    const bool causal = (p->flags & B200_FLAG_CAUSAL) != 0;
    FwdArgs a{Q, K, V, O, L, lay_of(p->q, p->H, p->Nq, p->d), lay_of(p->k, p->Hkv, p->Nkv, p->d), lay_of(p->v, p->Hkv, p->Nkv, p->d),
              lay_of(p->o, p->H, p->Nq, p->d), (long long)p->Nq, p->B, p->H, p->Hkv, p->Nq, p->Nkv, p->d,
              resolve_scale(p->scale, p->d), p->dtype, causal};
    return tc_fwd_strided(a, st);
}

int b200_fa2_bwd_desc(const b200_attn_desc* p, const void* Q, const void* K, const void* V, const void* O, const float* L,
                      const void* dO, void* dQ, void* dK, void* dV, void* ws, size_t ws_bytes, cudaStream_t st)
{
    B200_TRY(desc_check(p));
    const int H = p->H, d = p->d;
    const bool plain = p->Hkv == H && lay_is_contiguous(p->q, H, p->Nq, d) && lay_is_contiguous(p->o, H, p->Nq, d) &&
                       lay_is_contiguous(p->d_o, H, p->Nq, d) && lay_is_contiguous(p->dq, H, p->Nq, d) &&
                       lay_is_contiguous(p->k, H, p->Nkv, d) && lay_is_contiguous(p->v, H, p->Nkv, d) &&
                       lay_is_contiguous(p->dk, H, p->Nkv, d) && lay_is_contiguous(p->dv, H, p->Nkv, d);
    if (plain) return bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, p->B * H, p->Nq, p->Nkv, d, p->scale, p->dtype, ws, ws_bytes, st, p->flags);
    if ((size_t)p->B * H == 0 || (p->Nq == 0 && p->Nkv == 0)) return B200_OK;
    if (p->Nq == 0 || p->Nkv == 0 || !Q || !K || !V || !O || !L || !dO || !dQ || !dK || !dV) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    if (!tc_bwd_supported(p->dtype, d, p->Nq, p->Nkv)) return B200_ERR_UNSUPPORTED;
    const int BH = p->B * H;
    const size_t need = bwd_ws_bytes(BH, p->Nq, p->Nkv, d, p->dtype);
    if (ws == nullptr) B200_TRY(arena_get(need, &ws));
    else if (ws_bytes < need) return B200_ERR_WORKSPACE;
    float* D = reinterpret_cast<float*>(ws);
    const size_t dbytes = (((size_t)BH * p->Nq * sizeof(float)) + 255) & ~(size_t)255;
    const Lay lo = lay_of(p->o, H, p->Nq, d), ldo = lay_of(p->d_o, H, p->Nq, d);
    B200_TRY(rowdot_strided(dO, O, ldo, lo, D, p->B, H, p->Nq, d, p->dtype, st));
    BwdArgs a{Q, K, V, dO, L, D, dQ, dK, dV, lay_of(p->q, H, p->Nq, d), lay_of(p->k, p->Hkv, p->Nkv, d), lay_of(p->v, p->Hkv, p->Nkv, d),
              ldo, lay_of(p->dq, H, p->Nq, d), lay_of(p->dk, p->Hkv, p->Nkv, d), lay_of(p->dv, p->Hkv, p->Nkv, d),
              p->B, H, p->Hkv, p->Nq, p->Nkv, d, resolve_scale(p->scale, d), p->dtype, (p->flags & B200_FLAG_CAUSAL) != 0};
    return tc_bwd_strided(a, reinterpret_cast<uint8_t*>(ws) + dbytes, need - dbytes, st);
}

// ---- packed variable-length batches ------------------------------------------------------------------
static int varlen_check(const b200_attn_desc* p, int n, const int* cq, const int* ck, int* max_q, int* max_k)
{
    if (!p || n < 0 || (n > 0 && (!cq || !ck)) || !valid_dtype(p->dtype) || p->H < 1 || p->Hkv < 1 || p->H % p->Hkv != 0 || p->d <= 0)
        return B200_ERR_INVALID_ARGUMENT;
    if ((p->flags & ~B200_FLAG_CAUSAL) != 0) return B200_ERR_INVALID_ARGUMENT;
    *max_q = *max_k = 0;
    for (int i = 0; i < n; ++i) {
        const int nq = cq[i + 1] - cq[i], nk = ck[i + 1] - ck[i];
        if (nq < 0 || nk < 0 || ((nq == 0) != (nk == 0))) return B200_ERR_INVALID_ARGUMENT;
        if ((p->flags & B200_FLAG_CAUSAL) && nq > nk) return B200_ERR_INVALID_ARGUMENT;   // bottom-right aligned mask needs Nq <= Nkv
        if (nq > *max_q) *max_q = nq;
        if (nk > *max_k) *max_k = nk;
    }
    return B200_OK;
}
// head / row strides of a packed [total, heads, d] tensor (all-zero layout = exactly that, contiguous)
static Lay packed_lay(const b200_layout& l, int heads, int d)
{
    return lay_is_zero(l) ? Lay{0, (long long)d, (long long)heads * d} : Lay{0, l.head_stride, l.row_stride};
}

int b200_fa2_fwd_varlen(const b200_attn_desc* p, int n_seqs, const int* cu_q, const int* cu_k, const void* Q, const void* K,
                        const void* V, void* O, float* L, cudaStream_t st)
{
    int mq, mk;
    B200_TRY(varlen_check(p, n_seqs, cu_q, cu_k, &mq, &mk));
    if (n_seqs == 0 || mq == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    const size_t es = p->dtype == B200_DTYPE_F32 ? 4 : 2;
    const Lay lq = packed_lay(p->q, p->H, p->d), lk = packed_lay(p->k, p->Hkv, p->d), lv = packed_lay(p->v, p->Hkv, p->d),
              lo = packed_lay(p->o, p->H, p->d);
    const long long total_q = cu_q[n_seqs];
    int n_active = 0;
    for (int i = 0; i < n_seqs; ++i) n_active += cu_q[i + 1] > cu_q[i];
    const int used = n_active < kVarlenStreams ? n_active : kVarlenStreams;      // 1: no side streams
    std::unique_lock<std::mutex> vs_lock(g_vs_mu, std::defer_lock);
    VarlenStreams* vs = nullptr;
    if (used > 1) { vs_lock.lock(); B200_TRY(varlen_streams(&vs)); B200_TRY(varlen_fork(vs, used, st)); }
    int rc = B200_OK;
    for (int i = 0, j = 0; i < n_seqs; ++i) {
        const int nq = cu_q[i + 1] - cu_q[i], nk = cu_k[i + 1] - cu_k[i];
        if (nq == 0) continue;
        const cudaStream_t si = used > 1 ? vs->s[j++ % used] : st;
        FwdArgs a{(const uint8_t*)Q + (size_t)cu_q[i] * lq.sn * es, (const uint8_t*)K + (size_t)cu_k[i] * lk.sn * es,
                  (const uint8_t*)V + (size_t)cu_k[i] * lv.sn * es, (uint8_t*)O + (size_t)cu_q[i] * lo.sn * es, L + cu_q[i],
                  lq, lk, lv, lo, total_q, 1, p->H, p->Hkv, nq, nk, p->d, resolve_scale(p->scale, p->d), p->dtype,
                  (p->flags & B200_FLAG_CAUSAL) != 0};
        rc = tc_fwd_strided(a, si);
        if (rc != B200_OK) break;
    }
    // (joined also after a failed launch: whatever was enqueued on the side streams stays ordered before the caller's next work)
    if (used > 1) { const int jr = varlen_join(vs, used, st); if (rc == B200_OK) rc = jr; }
    return rc;
}

int b200_fa2_bwd_varlen(const b200_attn_desc* p, int n_seqs, const int* cu_q, const int* cu_k, const void* Q, const void* K,
                        const void* V, const void* O, const float* L, const void* dO, void* dQ, void* dK, void* dV,
                        cudaStream_t st)
{
    int mq, mk;
    B200_TRY(varlen_check(p, n_seqs, cu_q, cu_k, &mq, &mk));
    if (n_seqs == 0 || mq == 0) return B200_OK;
    if (!Q || !K || !V || !O || !L || !dO || !dQ || !dK || !dV) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    if (!tc_bwd_supported(p->dtype, p->d, mq, mk)) return B200_ERR_UNSUPPORTED;
    const int H = p->H, d = p->d;
    const size_t es = 2;
    const Lay lq = packed_lay(p->q, H, d), lk = packed_lay(p->k, p->Hkv, d), lv = packed_lay(p->v, p->Hkv, d),
              lo = packed_lay(p->o, H, d), ldo = packed_lay(p->d_o, H, d), ldq = packed_lay(p->dq, H, d),
              ldk = packed_lay(p->dk, p->Hkv, d), ldv = packed_lay(p->dv, p->Hkv, d);
    // one workspace per side stream, each sized for the longest sequence; a stream's sequences run back to back on it
    int n_active = 0;
    for (int i = 0; i < n_seqs; ++i) n_active += cu_q[i + 1] > cu_q[i];
    const int used = n_active < kVarlenStreams ? n_active : kVarlenStreams;
    void* ws0;
    const size_t need = (bwd_ws_bytes(H, mq, mk, d, p->dtype) + 255) & ~(size_t)255;
    B200_TRY(arena_get(need * (size_t)(used > 1 ? used : 1), &ws0));
    std::unique_lock<std::mutex> vs_lock(g_vs_mu, std::defer_lock);
    VarlenStreams* vs = nullptr;
    if (used > 1) { vs_lock.lock(); B200_TRY(varlen_streams(&vs)); B200_TRY(varlen_fork(vs, used, st)); }
    const long long total_q = cu_q[n_seqs];
    int rc = B200_OK;
    for (int i = 0, j = 0; i < n_seqs; ++i) {
        const int nq = cu_q[i + 1] - cu_q[i], nk = cu_k[i + 1] - cu_k[i];
        if (nq == 0) continue;
        const int slot = used > 1 ? j++ % used : 0;
        const cudaStream_t si = used > 1 ? vs->s[slot] : st;
        uint8_t* ws = reinterpret_cast<uint8_t*>(ws0) + (size_t)slot * need;
        const size_t q0 = (size_t)cu_q[i], k0 = (size_t)cu_k[i];
        float* D = reinterpret_cast<float*>(ws);
        const size_t dbytes = (((size_t)H * nq * sizeof(float)) + 255) & ~(size_t)255;
        const void* dOi = (const uint8_t*)dO + q0 * ldo.sn * es;
        const void* Oi = (const uint8_t*)O + q0 * lo.sn * es;
        rc = rowdot_strided(dOi, Oi, ldo, lo, D, 1, H, nq, d, p->dtype, si);
        if (rc != B200_OK) break;
        BwdArgs a{(const uint8_t*)Q + q0 * lq.sn * es, (const uint8_t*)K + k0 * lk.sn * es, (const uint8_t*)V + k0 * lv.sn * es, dOi,
                  L + q0, D, (uint8_t*)dQ + q0 * ldq.sn * es, (uint8_t*)dK + k0 * ldk.sn * es, (uint8_t*)dV + k0 * ldv.sn * es,
                  lq, lk, lv, ldo, ldq, ldk, ldv, 1, H, p->Hkv, nq, nk, d, resolve_scale(p->scale, d), p->dtype,
                  (p->flags & B200_FLAG_CAUSAL) != 0, total_q};
        rc = tc_bwd_strided(a, ws + dbytes, need - dbytes, si);
        if (rc != B200_OK) break;
    }
    if (used > 1) { const int jr = varlen_join(vs, used, st); if (rc == B200_OK) rc = jr; }
    return rc;
}

size_t b200_fa2_bwd_workspace_bytes(int B, int H, int Nq, int Nkv, int d, int dtype)
{
    if (B <= 0 || H <= 0 || Nq <= 0 || Nkv <= 0 || d <= 0 || !valid_dtype(dtype)) return 0;
    return bwd_ws_bytes(B * H, Nq, Nkv, d, dtype);
}

int b200_fa2_bwd(const void* Q, const void* K, const void* V, const void* O, const float* L, const void* dO,
                 void* dQ, void* dK, void* dV, int B, int H, int Nq, int Nkv, int d, float scale, int dtype,
                 void* ws, size_t ws_bytes, cudaStream_t st)
{
    if (B < 0 || H < 0 || (long long)B * H > 0x7fffffffLL) return B200_ERR_INVALID_ARGUMENT;
    return bwd_impl(Q, K, V, O, L, dO, dQ, dK, dV, B * H, Nq, Nkv, d, scale, dtype, ws, ws_bytes, st);
}

int b200_rowdot(const void* dO, const void* O, float* D, int B, int H, int N, int d, int dtype, cudaStream_t st)
{
    if (B < 0 || H < 0 || N < 0 || d <= 0 || !valid_dtype(dtype)) return B200_ERR_INVALID_ARGUMENT;
    if ((size_t)B * H * N == 0) return B200_OK;
    if (!dO || !O || !D) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    return rowdot(dO, O, D, (size_t)B * H * N, d, dtype, st);
}

int b200_lse_merge(void* Oa, float* La, const void* Ob, const float* Lb, int B, int H, int N, int d, int dtype, cudaStream_t st)
{
    if (B < 0 || H < 0 || N < 0 || d <= 0 || !valid_dtype(dtype)) return B200_ERR_INVALID_ARGUMENT;
    if ((size_t)B * H * N == 0) return B200_OK;
    if (!Oa || !La || !Ob || !Lb) return B200_ERR_INVALID_ARGUMENT;
    B200_TRY(device_check());
    return lse_merge(Oa, La, Ob, Lb, (size_t)B * H * N, d, dtype, st);
}

int b200_release_workspace(void)
{
    std::lock_guard<std::mutex> lk(g_arena_mu);
    for (int i = 0; i < 64; ++i)
        if (g_arena[i].ptr) {
            int cur = 0;
            cudaGetDevice(&cur);
            cudaSetDevice(i);
            cudaFree(g_arena[i].ptr);
            cudaSetDevice(cur);
            g_arena[i].ptr = nullptr; g_arena[i].bytes = 0;
        }
    return B200_OK;
}

}  // extern "C"
