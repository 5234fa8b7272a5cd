// generic_simt.cu -- shape-agnostic CUDA-core attention kernels (fp32 math, fp32 or bf16 I/O).
//
// Role in the library: (1) every shape the tensor-core kernels do not cover -- the reference's
// tiny fixtures with d in {1,2,4,32}, N in {1..64}, odd head dims -- still runs on the GPU (there
// is no CPU fallback anywhere in the product); (2) an independent on-device cross-check for the
// tcgen05 kernels at sizes where the CPU oracle is too slow.
//
// Math follows the reference's FA2 formulation: online softmax with running (m, l)
// (reference src/util/attention_helper.h:78-130), natural-log LSE = m + log l
// (src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:112-121), backward
// P = exp(S - L), dS = P o (dP - D), D = rowsum(dO o O)
// (src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:97-120,157-193).
// Unlike the reference backward (dQ-stationary + global atomics) both gradient kernels here are
// owner-computes: one warp owns a dQ row or a dK/dV row, so results are deterministic.
#include "common.cuh"
#include <type_traits>

namespace b200 {

constexpr int kSimtWarps = 8;          // warps (= rows) per CTA
constexpr int kSimtTile = 32;          // keys (or queries) per shared-memory tile = one per lane
constexpr int kSimtMaxD = 256;

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Cooperative load of `rows` x d elements (row-major, stride d) into smem with row stride ld = d+1
// (odd stride -> lane j reading row j is bank-conflict free).  Rows >= valid are zero-filled.
template <typename T>
__device__ __forceinline__ void load_tile(float* dst, const T* src, int valid, int d, int ld) {
    for (int idx = threadIdx.x; idx < kSimtTile * d; idx += blockDim.x) {
        int r = idx / d, c = idx - r * d;
        dst[r * ld + c] = r < valid ? ldf(src + (size_t)r * d + c) : 0.0f;
    }
}

// ---------------------------------------------------------------------------------------------
// Forward.  grid = (ceil(Nq / kSimtWarps), B*H); one warp per query row.
// mode 0: write normalised O and LSE.   mode 1 (FA1 API): write normalised O, l and m.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kSimtWarps * 32)
simt_fwd_kernel(const T* __restrict__ Q, const T* __restrict__ K, const T* __restrict__ V, T* __restrict__ O,
                float* __restrict__ L, float* __restrict__ Mout, int Nq, int Nkv, int d, float scale, int mode, int coff)
{
    extern __shared__ float sm[];
    const int ld = d + 1;
    float* sK = sm;
    float* sV = sK + kSimtTile * ld;
    float* sQ = sV + kSimtTile * ld;                // [kSimtWarps][d]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t bh = blockIdx.y;
    Q += bh * (size_t)Nq * d; O += bh * (size_t)Nq * d;
    K += bh * (size_t)Nkv * d; V += bh * (size_t)Nkv * d;
    const int row = blockIdx.x * kSimtWarps + warp;
    const bool active = row < Nq;
    float* q = sQ + warp * d;
    if (active) for (int c = lane; c < d; c += 32) q[c] = ldf(Q + (size_t)row * d + c);

    float m = -INFINITY, l = 0.0f;
    float acc[kSimtMaxD / 32];
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) acc[i] = 0.0f;

    for (int t0 = 0; t0 < Nkv; t0 += kSimtTile) {
        const int valid = min(kSimtTile, Nkv - t0);
        __syncthreads();
        load_tile(sK, K + (size_t)t0 * d, valid, d, ld);
        load_tile(sV, V + (size_t)t0 * d, valid, d, ld);
        __syncthreads();
        if (!active) continue;
        float s = 0.0f;
        const float* kr = sK + lane * ld;
        for (int c = 0; c < d; ++c) s = fmaf(q[c], kr[c], s);
        const bool vis = lane < valid && (t0 + lane) - row <= coff;          // causal: key j visible iff j <= row + coff
        s = vis ? s * scale : -INFINITY;
        const float m_new = fmaxf(m, warp_max(s));
        if (m_new == -INFINITY) continue;                                    // whole tile masked, nothing seen yet
        const float p = vis ? expf(s - m_new) : 0.0f;
        const float alpha = expf(m - m_new);          // m = -inf on the first tile -> 0
        l = alpha * l + warp_sum(p);
        m = m_new;
#pragma unroll
        for (int i = 0; i < kSimtMaxD / 32; ++i) acc[i] *= alpha;
        for (int j = 0; j < valid; ++j) {
            const float pj = __shfl_sync(0xffffffffu, p, j);
            const float* vr = sV + j * ld;
#pragma unroll
            for (int i = 0; i < kSimtMaxD / 32; ++i) {
                const int c = lane + 32 * i;
                if (c < d) acc[i] = fmaf(pj, vr[c], acc[i]);
            }
        }
    }
    if (!active) return;
    const float inv = 1.0f / l;
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) {
        const int c = lane + 32 * i;
        if (c < d) stf(O + (size_t)row * d + c, acc[i] * inv);
    }
    if (lane == 0) {
        if (mode == 0) { if (L) L[bh * Nq + row] = m + logf(l); }
        else { L[bh * Nq + row] = l; Mout[bh * Nq + row] = m; }
    }
}

// ---------------------------------------------------------------------------------------------
// D = rowsum(dO o O)  (reference flash_attention_backward_kernel.cu:97-120).  One warp per row.
// Also used by the tensor-core backward.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
rowdot_kernel(const T* __restrict__ dO, const T* __restrict__ O, float* __restrict__ D, size_t rows, int d)
{
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float s = 0.0f;
    for (int c = lane; c < d; c += 32) s = fmaf(ldf(dO + row * d + c), ldf(O + row * d + c), s);
    s = warp_sum(s);
    if (lane == 0) D[row] = s;
}

// The tensor-core shapes (bf16, d = 32 * EPL with EPL = 2 or 4 elements per lane, 8-byte aligned rows): one 4- or 8-byte load
// per lane, tensor and row instead of EPL two-byte loads, and two rows per warp so that four loads are in flight per lane.
// At C3 (2 x 268 MB read) the one-element-per-load kernel above runs at 3.9 TB/s; this is the HBM-bound helper of every backward.
template <int EPL>
__global__ void __launch_bounds__(256)
rowdot_bf16_vec_kernel(const __nv_bfloat16* __restrict__ dO, const __nv_bfloat16* __restrict__ O, float* __restrict__ D, size_t rows)
{
    using V = typename std::conditional<EPL == 4, uint2, uint32_t>::type;
    const size_t r0 = 2 * ((size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
    if (r0 >= rows) return;
    const int lane = threadIdx.x & 31;
    const bool two = r0 + 1 < rows;
    const V* a = reinterpret_cast<const V*>(dO) + r0 * 32 + lane;
    const V* b = reinterpret_cast<const V*>(O) + r0 * 32 + lane;
    const V a0 = a[0], b0 = b[0];
    V a1 = a0, b1 = b0;
    if (two) { a1 = a[32]; b1 = b[32]; }
    auto dot = [](const V& x, const V& y) {
        const __nv_bfloat162* xp = reinterpret_cast<const __nv_bfloat162*>(&x);
        const __nv_bfloat162* yp = reinterpret_cast<const __nv_bfloat162*>(&y);
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < EPL / 2; ++i) {
            const float2 xf = __bfloat1622float2(xp[i]), yf = __bfloat1622float2(yp[i]);
            s = fmaf(xf.x, yf.x, s);
            s = fmaf(xf.y, yf.y, s);
        }
        return s;
    };
    const float s0 = warp_sum(dot(a0, b0)), s1 = warp_sum(dot(a1, b1));
    if (lane == 0) { D[r0] = s0; if (two) D[r0 + 1] = s1; }
}

// D for strided [B, H, N, d] views (extended entry points); D itself is contiguous [B*H, N]
template <typename T>
__global__ void __launch_bounds__(256)
rowdot_strided_kernel(const T* __restrict__ dO, const T* __restrict__ O, Lay ldo, Lay lo, float* __restrict__ D, int H, int N,
                      int d, size_t rows)
{
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const size_t bh = row / N;
    const int n = (int)(row - bh * N), b = (int)(bh / H), h = (int)(bh - (size_t)b * H);
    const T* pa = dO + (size_t)b * ldo.sb + (size_t)h * ldo.sh + (size_t)n * ldo.sn;
    const T* pb = O + (size_t)b * lo.sb + (size_t)h * lo.sh + (size_t)n * lo.sn;
    float s = 0.0f;
    for (int c = lane; c < d; c += 32) s = fmaf(ldf(pa + c), ldf(pb + c), s);
    s = warp_sum(s);
    if (lane == 0) D[row] = s;
}

// ---------------------------------------------------------------------------------------------
// dQ.  grid = (ceil(Nq / kSimtWarps), B*H); one warp per query row, lanes = keys of the tile.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kSimtWarps * 32)
simt_bwd_dq_kernel(const T* __restrict__ Q, const T* __restrict__ K, const T* __restrict__ V, const T* __restrict__ dO,
                   const float* __restrict__ L, const float* __restrict__ D, T* __restrict__ dQ,
                   int Nq, int Nkv, int d, float scale, int coff)
{
    extern __shared__ float sm[];
    const int ld = d + 1;
    float* sK = sm;
    float* sV = sK + kSimtTile * ld;
    float* sQ = sV + kSimtTile * ld;                // [kSimtWarps][d]
    float* sG = sQ + kSimtWarps * d;                // [kSimtWarps][d]  (dO rows)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t bh = blockIdx.y;
    Q += bh * (size_t)Nq * d; dO += bh * (size_t)Nq * d; dQ += bh * (size_t)Nq * d;
    K += bh * (size_t)Nkv * d; V += bh * (size_t)Nkv * d;
    L += bh * (size_t)Nq; D += bh * (size_t)Nq;
    const int row = blockIdx.x * kSimtWarps + warp;
    const bool active = row < Nq;
    float* q = sQ + warp * d;
    float* g = sG + warp * d;
    float Li = 0.0f, Di = 0.0f;
    if (active) {
        for (int c = lane; c < d; c += 32) { q[c] = ldf(Q + (size_t)row * d + c); g[c] = ldf(dO + (size_t)row * d + c); }
        Li = L[row]; Di = D[row];
    }
    float acc[kSimtMaxD / 32];
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) acc[i] = 0.0f;

    for (int t0 = 0; t0 < Nkv; t0 += kSimtTile) {
        const int valid = min(kSimtTile, Nkv - t0);
        __syncthreads();
        load_tile(sK, K + (size_t)t0 * d, valid, d, ld);
        load_tile(sV, V + (size_t)t0 * d, valid, d, ld);
        __syncthreads();
        if (!active) continue;
        float s = 0.0f, dp = 0.0f;
        const float* kr = sK + lane * ld;
        const float* vr = sV + lane * ld;
        for (int c = 0; c < d; ++c) { s = fmaf(q[c], kr[c], s); dp = fmaf(g[c], vr[c], dp); }
        const float p = (lane < valid && (t0 + lane) - row <= coff) ? expf(s * scale - Li) : 0.0f;
        const float ds = p * (dp - Di);
        for (int j = 0; j < valid; ++j) {
            const float dsj = __shfl_sync(0xffffffffu, ds, j);
            const float* kj = sK + j * ld;
#pragma unroll
            for (int i = 0; i < kSimtMaxD / 32; ++i) {
                const int c = lane + 32 * i;
                if (c < d) acc[i] = fmaf(dsj, kj[c], acc[i]);
            }
        }
    }
    if (!active) return;
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) {
        const int c = lane + 32 * i;
        if (c < d) stf(dQ + (size_t)row * d + c, acc[i] * scale);
    }
}

// ---------------------------------------------------------------------------------------------
// dK, dV.  grid = (ceil(Nkv / kSimtWarps), B*H); one warp per key row, lanes = queries of the tile.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kSimtWarps * 32)
simt_bwd_dkv_kernel(const T* __restrict__ Q, const T* __restrict__ K, const T* __restrict__ V, const T* __restrict__ dO,
                    const float* __restrict__ L, const float* __restrict__ D, T* __restrict__ dK, T* __restrict__ dV,
                    int Nq, int Nkv, int d, float scale, int coff)
{
    extern __shared__ float sm[];
    const int ld = d + 1;
    float* sQ = sm;                                 // [32][ld]
    float* sG = sQ + kSimtTile * ld;                // [32][ld]
    float* sKr = sG + kSimtTile * ld;               // [kSimtWarps][d]
    float* sVr = sKr + kSimtWarps * d;              // [kSimtWarps][d]
    float* sL = sVr + kSimtWarps * d;               // [32]
    float* sD = sL + kSimtTile;                     // [32]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t bh = blockIdx.y;
    Q += bh * (size_t)Nq * d; dO += bh * (size_t)Nq * d;
    K += bh * (size_t)Nkv * d; V += bh * (size_t)Nkv * d; dK += bh * (size_t)Nkv * d; dV += bh * (size_t)Nkv * d;
    L += bh * (size_t)Nq; D += bh * (size_t)Nq;
    const int row = blockIdx.x * kSimtWarps + warp;
    const bool active = row < Nkv;
    float* kr = sKr + warp * d;
    float* vr = sVr + warp * d;
    if (active)
        for (int c = lane; c < d; c += 32) { kr[c] = ldf(K + (size_t)row * d + c); vr[c] = ldf(V + (size_t)row * d + c); }
    float accK[kSimtMaxD / 32], accV[kSimtMaxD / 32];
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) { accK[i] = 0.0f; accV[i] = 0.0f; }

    for (int t0 = 0; t0 < Nq; t0 += kSimtTile) {
        const int valid = min(kSimtTile, Nq - t0);
        __syncthreads();
        load_tile(sQ, Q + (size_t)t0 * d, valid, d, ld);
        load_tile(sG, dO + (size_t)t0 * d, valid, d, ld);
        if (threadIdx.x < kSimtTile) {
            sL[threadIdx.x] = threadIdx.x < valid ? L[t0 + threadIdx.x] : 0.0f;
            sD[threadIdx.x] = threadIdx.x < valid ? D[t0 + threadIdx.x] : 0.0f;
        }
        __syncthreads();
        if (!active) continue;
        float s = 0.0f, dp = 0.0f;
        const float* qi = sQ + lane * ld;
        const float* gi = sG + lane * ld;
        for (int c = 0; c < d; ++c) { s = fmaf(qi[c], kr[c], s); dp = fmaf(gi[c], vr[c], dp); }
        const float p = (lane < valid && row - (t0 + lane) <= coff) ? expf(s * scale - sL[lane]) : 0.0f;
        const float ds = p * (dp - sD[lane]);
        for (int i2 = 0; i2 < valid; ++i2) {
            const float pi = __shfl_sync(0xffffffffu, p, i2);
            const float dsi = __shfl_sync(0xffffffffu, ds, i2);
            const float* qr = sQ + i2 * ld;
            const float* gr = sG + i2 * ld;
#pragma unroll
            for (int i = 0; i < kSimtMaxD / 32; ++i) {
                const int c = lane + 32 * i;
                if (c < d) { accV[i] = fmaf(pi, gr[c], accV[i]); accK[i] = fmaf(dsi, qr[c], accK[i]); }
            }
        }
    }
    if (!active) return;
#pragma unroll
    for (int i = 0; i < kSimtMaxD / 32; ++i) {
        const int c = lane + 32 * i;
        if (c < d) { stf(dK + (size_t)row * d + c, accK[i] * scale); stf(dV + (size_t)row * d + c, accV[i]); }
    }
}

// ---------------------------------------------------------------------------------------------
// Merge of two normalised partial results by log-sum-exp (ring attention, north_star):
//   L = log(e^La + e^Lb), O = e^(La-L) Oa + e^(Lb-L) Ob     in place on (Oa, La).
// One thread per 4 output elements; HBM-bound.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
lse_merge_kernel(T* __restrict__ Oa, float* __restrict__ La, const T* __restrict__ Ob, const float* __restrict__ Lb,
                 size_t rows, int d)
{
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const float la = La[row], lb = Lb[row];
    const float mx = fmaxf(la, lb);
    float wa, wb, lnew;
    if (mx == -INFINITY) { wa = 0.f; wb = 0.f; lnew = -INFINITY; }
    else {
        const float ea = expf(la - mx), eb = expf(lb - mx);
        const float tot = ea + eb;
        wa = ea / tot; wb = eb / tot; lnew = mx + logf(tot);
    }
    for (int c = lane; c < d; c += 32)
        stf(Oa + row * d + c, wa * ldf(Oa + row * d + c) + wb * ldf(Ob + row * d + c));
    __syncwarp();
    if (lane == 0) La[row] = lnew;
}

// ---------------------------------------------------------------------------------------------
// n-way merge (ring forward, deferred): partial s lives at Op + s * o_stride elements / Lp + s * l_stride;
// with `running` the current (O, L) is one more input.  L = log sum_s e^(L_s), O = sum_s e^(L_s - L) O_s,
// accumulated in fp32 and rounded once.  One warp per row, 4 elements per lane and trip; HBM-bound.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
lse_merge_multi_kernel(T* __restrict__ O, float* __restrict__ L, const T* __restrict__ Op, const float* __restrict__ Lp,
                       size_t o_stride, size_t l_stride, int n, int running, size_t rows, int d, int Nl, int c,
                       unsigned upper_mask)
{
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    // causal zigzag ring: the partials in upper_mask only cover the rows >= c of every Nl-row slab
    const unsigned skip = (upper_mask != 0u && (int)(row % (size_t)Nl) < c) ? upper_mask : 0u;
    auto lse_of = [&](int s) { return ((skip >> s) & 1u) ? -INFINITY : Lp[s * l_stride + row]; };
    float mx = running ? L[row] : -INFINITY;
    for (int s = 0; s < n; ++s) mx = fmaxf(mx, lse_of(s));
    float tot = 0.f;
    if (mx != -INFINITY) {
        if (running) tot += expf(L[row] - mx);
        for (int s = 0; s < n; ++s) tot += expf(lse_of(s) - mx);
    }
    const float inv = tot > 0.f ? 1.0f / tot : 0.f;
    for (int c = lane * 4; c < d; c += 128) {
        const int w = min(4, d - c);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        if (running) {
            const float wr = mx == -INFINITY ? 0.f : expf(L[row] - mx) * inv;
            for (int e = 0; e < w; ++e) acc[e] = wr * ldf(O + row * d + c + e);
        }
        for (int s = 0; s < n; ++s) {
            const float ws = mx == -INFINITY ? 0.f : expf(lse_of(s) - mx) * inv;
            if (ws == 0.f) continue;                                    // (also keeps stale bits of uncovered rows out)
            const T* src = Op + s * o_stride + row * d + c;
            if (w == 4 && (d & 3) == 0) {
                if constexpr (sizeof(T) == 2) {
                    const uint2 v = *reinterpret_cast<const uint2*>(src);
                    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&v.x), b = *reinterpret_cast<const __nv_bfloat162*>(&v.y);
                    acc[0] = fmaf(ws, __low2float(a), acc[0]); acc[1] = fmaf(ws, __high2float(a), acc[1]);
                    acc[2] = fmaf(ws, __low2float(b), acc[2]); acc[3] = fmaf(ws, __high2float(b), acc[3]);
                } else {
                    const float4 v = *reinterpret_cast<const float4*>(src);
                    acc[0] = fmaf(ws, v.x, acc[0]); acc[1] = fmaf(ws, v.y, acc[1]);
                    acc[2] = fmaf(ws, v.z, acc[2]); acc[3] = fmaf(ws, v.w, acc[3]);
                }
            } else {
                for (int e = 0; e < w; ++e) acc[e] = fmaf(ws, ldf(src + e), acc[e]);
            }
        }
        for (int e = 0; e < w; ++e) stf(O + row * d + c + e, acc[e]);
    }
    __syncwarp();
    if (lane == 0) L[row] = mx == -INFINITY ? -INFINITY : mx + logf(tot);
}

// ------------------------------------------------------------------------------------ launchers
static size_t simt_fwd_smem(int d) { return sizeof(float) * ((size_t)2 * kSimtTile * (d + 1) + (size_t)kSimtWarps * d); }
static size_t simt_dq_smem(int d) { return sizeof(float) * ((size_t)2 * kSimtTile * (d + 1) + (size_t)2 * kSimtWarps * d); }
static size_t simt_dkv_smem(int d) { return sizeof(float) * ((size_t)2 * kSimtTile * (d + 1) + (size_t)2 * kSimtWarps * d + 2 * kSimtTile); }

template <typename T>
static int simt_fwd_t(const T* Q, const T* K, const T* V, T* O, float* L, float* M, int BH, int Nq, int Nkv, int d,
                      float scale, int mode, int coff, cudaStream_t st)
{
    size_t smem = simt_fwd_smem(d);
    if (smem > 48 * 1024) B200_CUDA_TRY(cudaFuncSetAttribute(simt_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((Nq + kSimtWarps - 1) / kSimtWarps, BH);
    simt_fwd_kernel<T><<<grid, kSimtWarps * 32, smem, st>>>(Q, K, V, O, L, M, Nq, Nkv, d, scale, mode, coff);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

int simt_fwd(const void* Q, const void* K, const void* V, void* O, float* L, float* M, int BH, int Nq, int Nkv, int d,
             float scale, int dtype, int mode, cudaStream_t st, int coff)
{
    if (d > kSimtMaxD) return B200_ERR_UNSUPPORTED_HEAD_DIM;
    if (BH > 65535) return B200_ERR_INVALID_ARGUMENT;
    if (dtype == B200_DTYPE_F32)
        return simt_fwd_t((const float*)Q, (const float*)K, (const float*)V, (float*)O, L, M, BH, Nq, Nkv, d, scale, mode, coff, st);
    return simt_fwd_t((const __nv_bfloat16*)Q, (const __nv_bfloat16*)K, (const __nv_bfloat16*)V, (__nv_bfloat16*)O, L, M,
                      BH, Nq, Nkv, d, scale, mode, coff, st);
}

template <typename T>
static int rowdot_t(const T* dO, const T* O, float* D, size_t rows, int d, cudaStream_t st)
{
    if (rows == 0) return B200_OK;
    if (std::is_same<T, __nv_bfloat16>::value && (d == 128 || d == 64) && (((uintptr_t)dO | (uintptr_t)O) & 7) == 0) {
        const unsigned grid = (unsigned)((rows + 15) / 16);                  // 8 warps x 2 rows per block
        if (d == 128) rowdot_bf16_vec_kernel<4><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, D, rows);
        else rowdot_bf16_vec_kernel<2><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, D, rows);
        B200_CUDA_TRY(cudaGetLastError());
        count_launch();
        return B200_OK;
    }
    rowdot_kernel<T><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(dO, O, D, rows, d);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}
int rowdot(const void* dO, const void* O, float* D, size_t rows, int d, int dtype, cudaStream_t st)
{
    if (dtype == B200_DTYPE_F32) return rowdot_t((const float*)dO, (const float*)O, D, rows, d, st);
    return rowdot_t((const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, D, rows, d, st);
}

int rowdot_strided(const void* dO, const void* O, const Lay& ldo, const Lay& lo, float* D, int B, int H, int N, int d, int dtype,
                   cudaStream_t st)
{
    const size_t rows = (size_t)B * H * N;
    if (rows == 0) return B200_OK;
    const unsigned grid = (unsigned)((rows + 7) / 8);
    if (dtype == B200_DTYPE_F32) rowdot_strided_kernel<float><<<grid, 256, 0, st>>>((const float*)dO, (const float*)O, ldo, lo, D, H, N, d, rows);
    else rowdot_strided_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, ldo, lo, D, H, N, d, rows);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

template <typename T>
static int simt_bwd_t(const T* Q, const T* K, const T* V, const T* dO, const float* L, const float* D, T* dQ, T* dK, T* dV,
                      int BH, int Nq, int Nkv, int d, float scale, int coff, cudaStream_t st)
{
    size_t s1 = simt_dq_smem(d), s2 = simt_dkv_smem(d);
    if (s1 > 48 * 1024) B200_CUDA_TRY(cudaFuncSetAttribute(simt_bwd_dq_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s1));
    if (s2 > 48 * 1024) B200_CUDA_TRY(cudaFuncSetAttribute(simt_bwd_dkv_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s2));
    if (dQ) {
        dim3 g1((Nq + kSimtWarps - 1) / kSimtWarps, BH);
        simt_bwd_dq_kernel<T><<<g1, kSimtWarps * 32, s1, st>>>(Q, K, V, dO, L, D, dQ, Nq, Nkv, d, scale, coff);
        B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    }
    if (dK && dV) {
        dim3 g2((Nkv + kSimtWarps - 1) / kSimtWarps, BH);
        simt_bwd_dkv_kernel<T><<<g2, kSimtWarps * 32, s2, st>>>(Q, K, V, dO, L, D, dK, dV, Nq, Nkv, d, scale, coff);
        B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    }
    return B200_OK;
}

int simt_bwd(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* D,
             void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, int dtype, cudaStream_t st, int coff)
{
    if (d > kSimtMaxD) return B200_ERR_UNSUPPORTED_HEAD_DIM;
    if (BH > 65535) return B200_ERR_INVALID_ARGUMENT;
    if (dtype == B200_DTYPE_F32)
        return simt_bwd_t((const float*)Q, (const float*)K, (const float*)V, (const float*)dO, L, D, (float*)dQ, (float*)dK,
                          (float*)dV, BH, Nq, Nkv, d, scale, coff, st);
    return simt_bwd_t((const __nv_bfloat16*)Q, (const __nv_bfloat16*)K, (const __nv_bfloat16*)V, (const __nv_bfloat16*)dO, L, D,
                      (__nv_bfloat16*)dQ, (__nv_bfloat16*)dK, (__nv_bfloat16*)dV, BH, Nq, Nkv, d, scale, coff, st);
}

int lse_merge(void* Oa, float* La, const void* Ob, const float* Lb, size_t rows, int d, int dtype, cudaStream_t st)
{
    if (rows == 0) return B200_OK;
    unsigned grid = (unsigned)((rows + 7) / 8);
    if (dtype == B200_DTYPE_F32) lse_merge_kernel<float><<<grid, 256, 0, st>>>((float*)Oa, La, (const float*)Ob, Lb, rows, d);
    else lse_merge_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)Oa, La, (const __nv_bfloat16*)Ob, Lb, rows, d);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

int lse_merge_multi(void* O, float* L, const void* Op, const float* Lp, size_t o_stride, size_t l_stride, int n, bool running,
                    size_t rows, int d, int dtype, cudaStream_t st, int Nl, int c, unsigned upper_mask)
{
    if (n > 32) return B200_ERR_UNSUPPORTED;
    if (Nl <= 0) { Nl = 1; upper_mask = 0u; }
    if (rows == 0 || (n == 0 && running)) return B200_OK;
    unsigned grid = (unsigned)((rows + 7) / 8);
    if (dtype == B200_DTYPE_F32)
        lse_merge_multi_kernel<float><<<grid, 256, 0, st>>>((float*)O, L, (const float*)Op, Lp, o_stride, l_stride, n, running ? 1 : 0, rows, d,
                                                                Nl, c, upper_mask);
    else
        lse_merge_multi_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)O, L, (const __nv_bfloat16*)Op, Lp, o_stride, l_stride, n,
                                                                     running ? 1 : 0, rows, d, Nl, c, upper_mask);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

}  // namespace b200
