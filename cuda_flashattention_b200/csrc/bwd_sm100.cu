// bwd_sm100.cu -- FlashAttention-2 backward for sm_100a (bf16, d = 128 or 64): dK/dV-stationary CTAs,
// tcgen05 MMAs with every accumulator in TMEM, dQ reduced into an fp32 buffer by bulk add-reductions.
//
// Replaces flash_attention_2_backward_kernel (reference
// src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:47-246).  Same math --
// P = exp(S - L), dP = dO V^T, dS = P o (dP - D), dV += P^T dO, dK += scale dS^T Q,
// dQ += scale dS K with D = rowsum(dO o O) -- but the loop order is the one north_star asks for:
// a CTA owns one 128-row K/V tile (dK, dV accumulate in TMEM for the whole kernel, no atomics on
// them) and streams 64-row Q/dO tiles; the reference is dQ-stationary with smem + global atomics
// on dK/dV and 32x redundant dQ math (:196-231).
//
// Everything is computed TRANSPOSED so that a thread owns a kv row (TMEM lane):
//   M1  S^T  = K Q_i^T              SS   [128 kv x 64 q]   -> TMEM S[b] (double buffered)
//   M2  dP^T = V dO_i^T             SS   [128 kv x 64 q]   -> TMEM dP
//   compute warpgroups: P^T = exp2(S^T c - L_i log2e) -> TMEM (bf16, aliasing S[b]);
//                       dS^T = scale P^T o (dP^T - D_i) -> TMEM (bf16, aliasing dP) and smem
//   M3  dV  += P^T dO_i             TS   A = P^T from TMEM, B = dO_i as an MN-major operand
//   M4  dK  += dS^T Q_i             TS   A = dS^T from TMEM, B = Q_i as an MN-major operand
//   M5  dQ_i^T = K^T dS             SS   A = K (MN-major view), B = dS^T smem (MN-major) -> TMEM dQ
//   drain warpgroup: dQ_i^T TMEM -> registers -> two alternating 16 KB staging buffers ([q][d] fp32)
//                    -> cp.reduce.async.bulk (add.f32) into the fp32 dQ accumulator
// Extensions on the same kernel: causal mask (diagonal q tiles), grouped K/V heads (the CTA of a K/V head walks
// the q tiles of all its query heads), strided layouts (4-D tensor maps), d = 64 (half-width tiles).
// TMEM columns: S[0] 0..63, S[1] 64..127, dP 128..191, dQ^T 192..255, dV 256..383, dK 384..511.
// Issue order per iteration i:  M1(i+1) | M3(i) M4(i) | M2(i+1) | M5(i)   (see DESIGN.md for the slack analysis)
//
// Warps: 0 TMA producer, 1 MMA issuer, 2 TMEM allocator, 4-7 and 8-11 compute warpgroups (q columns
// 0-31 / 32-63 of the same kv rows), 12-15 dQ drain.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"
#include <stdlib.h>
#include <type_traits>

namespace b200 {

namespace bwd {
constexpr int D = 128, BN = 128, BM = 64, NS = 3;
constexpr int KV_BOX = BN * 128, KV_TILE = 2 * KV_BOX;        // 32 KB: [128 rows x 128 d] bf16 as 2 boxes of 64 d
constexpr int Q_BOX = BM * 128, Q_TILE = 2 * Q_BOX;           // 16 KB: [64 rows x 128 d]
constexpr int DS_TILE = BN * 128;                             // 16 KB: dS^T [128 kv rows x 64 q] bf16
constexpr int STAT_BYTES = BM * 4;                            // 256 B of L (or D) per Q tile
constexpr int THREADS = 512;
// smem carve-up (offsets from the 1024-aligned base)
constexpr int OFF_K = 0, OFF_V = OFF_K + KV_TILE, OFF_Q = OFF_V + KV_TILE, OFF_DO = OFF_Q + NS * Q_TILE;
constexpr int OFF_DS = OFF_DO + NS * Q_TILE, OFF_STAT = OFF_DS + 2 * DS_TILE;
#ifndef B200_DQ_RED
// dQ leaves through staging buffers and bulk reduce-adds (cp.reduce.async.bulk): STG_NBUF buffers of STG_ROWS q rows,
// used round-robin.  (-DB200_DQ_RED builds the r1(b) drain instead: per-warp red.global.add from registers.)
#ifndef B200_STG_ROWS
#define B200_STG_ROWS 32
#endif
constexpr int STG_ROWS = B200_STG_ROWS, STG_NBUF = BM / STG_ROWS, STG_BYTES = STG_ROWS * D * 4;
constexpr int OFF_STG = OFF_STAT + NS * 2 * STAT_BYTES;
constexpr int OFF_BAR = OFF_STG + STG_NBUF * STG_BYTES;
#else
constexpr int OFF_BAR = OFF_STAT + NS * 2 * STAT_BYTES;
#endif
constexpr int NBARS = 1 + 2 * NS + 2 + 1 + 2 + 1 + 2 + 1;
constexpr int SMEM = 1024 + OFF_BAR + NBARS * 8 + 16;
static_assert(SMEM <= 227 * 1024, "backward shared memory budget");
constexpr uint32_t TM_S = 0, TM_DP = 128, TM_DQ = 192, TM_DV = 256, TM_DK = 384;
}  // namespace bwd

#ifdef B200_TRACE
__device__ long long g_trace_bwd[3 * 8 * 8];
#endif
// Timing-only experiment hooks (WRONG RESULTS): compiled only into `make VARIANT=exp EXTRA=-DB200_EXPERIMENTS` side
// libraries, selected there by B200_BWD_DBG; the product build folds every B200_DBG(...) to false.
//   1 no dQ staging / reduction   2 plain dQ stores (B200_DQ_RED build)   4 no M5   8 no dS shared-memory copy
//   16 no M1   32 no M2   64 no M3 / M4   128 no exponentials
#ifdef B200_EXPERIMENTS
#define B200_DBG(bit) ((dbg & (bit)) != 0)
#else
#define B200_DBG(bit) false
#endif
// roles: 0 = compute warpgroup 0 (its warp 0), 1 = dQ drain warpgroup (its warp 0), 2 = MMA issuer
#define B200_TRB(role, i, ev) B200_TRACE_EVENT(g_trace_bwd, role, i, ev, (role) == 2 || (((threadIdx.x >> 5) & 3) == 0 && ((role) == 1 || threadIdx.x < 256)))

__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void lds_f32x4(float* dst, uint32_t saddr) {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(dst[0]), "=f"(dst[1]), "=f"(dst[2]), "=f"(dst[3]) : "r"(saddr));
}
__device__ __forceinline__ void bulk_reduce_add_f32(float* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                 ::"l"(reinterpret_cast<uint64_t>(gdst)), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void red_add_f32(float* gaddr, float v) {
    asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(gaddr), "f"(v) : "memory");
}

// One compute warpgroup (w = 0/1 owns q columns 32w..32w+31 of every Q tile; thread = kv row).
// The per-column statistics L_i, D_i are the same for every thread; they are fetched from shared
// memory into registers BEFORE the wait for the MMA they go with, because shared memory is the
// saturated resource of this kernel (the N = 64 SS MMAs alone ask for more than its 128 B/clk):
// a load issued on the critical path queues behind operand traffic for hundreds of cycles.
__device__ __forceinline__ void bwd_compute_role(int w, uint32_t tmem, uint8_t* smem, uint64_t* bar_in_full,
                                                 uint64_t* bar_S, uint64_t* bar_dP, uint64_t* bar_P, uint64_t* bar_dS,
                                                 int n_tot, int n_it, int qt_min, int i_off, int kv0, int causal,
                                                 float scale_log2, float scale, int dbg)
{
    (void)dbg;
    using namespace bwd;
    const int r = threadIdx.x & 127;                                   // kv row within the tile == TMEM lane
    const uint32_t lane_base = uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16;
    const bool lane0 = (threadIdx.x & 31) == 0;
    const uint32_t stat_base = smem_u32(smem + OFF_STAT) + 128 * w;    // + s * 2 * STAT_BYTES (+ STAT_BYTES for D)
    for (int i = 0, ii = 0; i < n_tot; ++i) {
        const int b = i & 1, s = i % NS;
        const uint32_t tS = tmem + TM_S + 64 * b + 32 * w + lane_base;
        const uint32_t tP = tmem + TM_DP + 32 * w + lane_base;
        // ---- P^T = exp2(S^T * scale*log2e - L*log2e)
        mbar_wait(&bar_in_full[s], (i / NS) & 1);                       // L_i, D_i landed (same barrier as Q_i, dO_i)
        float st[32];                                                   // L_i (log2 units), later D_i
#pragma unroll
        for (int c = 0; c < 8; ++c) lds_f32x4(st + 4 * c, stat_base + s * 2 * STAT_BYTES + 16 * c);
        mbar_wait(&bar_S[b], (i >> 1) & 1);
        tc_fence_after();
        B200_TRB(0, i, 0);
        uint32_t sv[32];
        B200_TMEM_LD_32(tS, sv);
        tmem_wait_ld();
        B200_TRB(0, i, 4);
        float p[32];
#pragma unroll
        for (int c = 0; c < 32; ++c)
            p[c] = B200_DBG(128) ? fmaf(__uint_as_float(sv[c]), scale_log2, -st[c]) : fast_exp2(fmaf(__uint_as_float(sv[c]), scale_log2, -st[c]));
        if (causal) {
            // key kv0 + r is hidden from the queries before it: only the two q tiles on the diagonal are affected
            int qt = ii + i_off;
            if (qt >= n_it) qt -= n_it;
            // (`causal` = 1 + (Nkv - Nq): key kv0 + r is visible to the queries q >= kv0 + r - (Nkv - Nq))
            const int first_visible = kv0 + r - (causal - 1) - (qt_min + qt) * BM - 32 * w;   // column (of this warpgroup's 32) of the first query that sees the key
            if (first_visible > 0) {
#pragma unroll
                for (int c = 0; c < 32; ++c) if (c < first_visible) p[c] = 0.0f;
            }
        }
        {
            uint32_t pk[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) pk[c] = pack_bf16x2(p[2 * c], p[2 * c + 1]);
            B200_TMEM_ST_16(tS, pk);
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) lds_f32x4(st + 4 * c, stat_base + s * 2 * STAT_BYTES + STAT_BYTES + 16 * c);   // D_i, ahead of dP
        B200_TRB(0, i, 5);
        tmem_wait_st();
        B200_TRB(0, i, 6);
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(&bar_P[b]);
        B200_TRB(0, i, 1);
        // ---- dS^T = scale * P^T o (dP^T - D)
        mbar_wait(bar_dP, i & 1);
        tc_fence_after();
        B200_TRB(0, i, 2);
        B200_TMEM_LD_32(tP, sv);
        tmem_wait_ld();
        uint32_t dk[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            const float a0 = p[2 * c] * (__uint_as_float(sv[2 * c]) - st[2 * c]) * scale;
            const float a1 = p[2 * c + 1] * (__uint_as_float(sv[2 * c + 1]) - st[2 * c + 1]) * scale;
            dk[c] = pack_bf16x2(a0, a1);
        }
        B200_TMEM_ST_16(tP, dk);
        // smem copy (B operand of M5): row r = 128 B, this warpgroup owns 16-byte chunks 4w..4w+3
        uint8_t* rowp = smem + OFF_DS + b * DS_TILE + r * 128;
        if (!B200_DBG(8)) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                *reinterpret_cast<uint4*>(rowp + (((4 * w + q) ^ (r & 7)) * 16)) = make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
        }
        fence_proxy_async_smem();
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(bar_dS);
        B200_TRB(0, i, 3);
        if (++ii == n_it) ii = 0;                                       // next q head of the group
    }
}

// dQ drain warpgroup: TMEM dQ_i^T [128 d lanes x 64 q] -> registers -> staging buffer ([q][d] fp32: a warp's 32
// lanes hold 32 consecutive d of one q row, so every st.shared is one conflict-free 128-byte row segment) -> one
// cp.reduce.async.bulk add-reduction of 32 contiguous accumulator rows per buffer.  The TMEM columns are released
// right after the load, so the next dQ MMA never waits for the reductions.  (-DB200_DQ_RED: per-warp red.global.add
// from registers instead -- the LSU store path carrying that stream cost 17 % of the kernel)
template <int DR>
__device__ __forceinline__ void bwd_drain_role(uint32_t tmem, uint64_t* bar_dQ_full, uint64_t* bar_dQ_empty,
                                               float* __restrict__ dQacc, int Nq, int dq_rows, int n_tot, int n_it, int qt_min,
                                               int i_off, size_t bhq0, int dbg, uint8_t* smem)
{
    using namespace bwd;
    (void)dbg;
    (void)dbg;
    const int r = threadIdx.x & 127;                                   // d index == TMEM lane
    const uint32_t tQ = tmem + TM_DQ + (uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16);
    const bool lane0 = (threadIdx.x & 31) == 0;
    for (int i = 0, ii = 0, g = 0; i < n_tot; ++i) {
        int qt = ii + i_off;                                           // the q tile this iteration works on
        if (qt >= n_it) qt -= n_it;
        qt += qt_min;
        const size_t bh = bhq0 + g;                                    // q head g of this K/V head's group
        if (++ii == n_it) { ii = 0; ++g; }
        mbar_wait(bar_dQ_full, i & 1);
        tc_fence_after();
        B200_TRB(1, i, 0);
        uint32_t v[64];
        B200_TMEM_LD_32(tQ, v);
        B200_TMEM_LD_32(tQ + 32, v + 32);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(bar_dQ_empty);                          // TMEM columns are free again
        B200_TRB(1, i, 1);
        float* dst = dQacc + (bh * dq_rows + (size_t)qt * BM) * DR + r;   // (dq_rows: rows per slab of the accumulator)
        const int q_valid = min(BM, Nq - qt * BM);                     // ragged last q tile: rows past Nq do not exist
#ifndef B200_DQ_RED
        if (!B200_DBG(1)) {
            const bool leader = (threadIdx.x & 127) == 0;
#pragma unroll
            for (int half = 0; half < BM / STG_ROWS; ++half) {
                float* stg = reinterpret_cast<float*>(smem + OFF_STG + half * STG_BYTES);
                if (leader) tma_store_wait_read<STG_NBUF - 1>();       // the reduce that last used this buffer has read it
                named_bar_sync(6, 128);
                if (r < DR) {
#pragma unroll
                    for (int q = 0; q < STG_ROWS; ++q) stg[q * DR + r] = __uint_as_float(v[STG_ROWS * half + q]);
                }
                fence_proxy_async_smem();
                named_bar_sync(6, 128);
                if (leader) {
                    const int rows = min(STG_ROWS, q_valid - STG_ROWS * half);
                    if (rows > 0) bulk_reduce_add_f32(dst - r + (size_t)STG_ROWS * half * DR, stg, (uint32_t)rows * DR * 4);
                    tma_store_commit();
                }
            }
        }
#else
        if (!B200_DBG(1) && r < DR) {                                    // (d = 64: lanes 64..127 hold the zero padding)
            if (B200_DBG(2)) {                                             // experiment: plain stores, same bytes, no L2 atomic
#pragma unroll
                for (int q = 0; q < BM; ++q) dst[q * DR] = __uint_as_float(v[q]);
            } else if (q_valid == BM) {
#pragma unroll
                for (int q = 0; q < BM; ++q) red_add_f32(dst + q * DR, __uint_as_float(v[q]));
            } else {
#pragma unroll
                for (int q = 0; q < BM; ++q) if (q < q_valid) red_add_f32(dst + q * DR, __uint_as_float(v[q]));
            }
        }
#endif
        B200_TRB(1, i, 2);
    }
#ifndef B200_DQ_RED
    if ((threadIdx.x & 127) == 0) tma_store_wait<0>();
#endif
}

// dV (w = 0) / dK (w = 1) epilogue: TMEM fp32 [128 kv x 128 d] -> bf16 -> swizzled smem -> TMA store.
template <int DR>
__device__ __forceinline__ void bwd_epilogue(int w, uint32_t tmem, uint8_t* smem, const CUtensorMap* tm, int kv0, int hk, int b)
{
    using namespace bwd;
    const int r = threadIdx.x & 127;
    const uint32_t tA = tmem + (w == 0 ? TM_DV : TM_DK) + (uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16);
    uint8_t* out = smem + (w == 0 ? OFF_V : OFF_K);                    // K / V tiles are dead after the last MMA
#pragma unroll
    for (int c = 0; c < DR / 32; ++c) {
        uint32_t o[32];
        B200_TMEM_LD_32(tA + 32 * c, o);
        tmem_wait_ld();
        uint8_t* rowp = out + (c >> 1) * KV_BOX + r * 128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o[8 * q]), __uint_as_float(o[8 * q + 1]));
            v.y = pack_bf16x2(__uint_as_float(o[8 * q + 2]), __uint_as_float(o[8 * q + 3]));
            v.z = pack_bf16x2(__uint_as_float(o[8 * q + 4]), __uint_as_float(o[8 * q + 5]));
            v.w = pack_bf16x2(__uint_as_float(o[8 * q + 6]), __uint_as_float(o[8 * q + 7]));
            *reinterpret_cast<uint4*>(rowp + ((((c & 1) * 4 + q) ^ (r & 7)) * 16)) = v;
        }
    }
    fence_proxy_async_smem();
    named_bar_sync(4 + w, 128);
    if (r == 0) {
        tma_store_4d(tm, out, 0, kv0, hk, b);
        if (DR == 128) tma_store_4d(tm, out + KV_BOX, 64, kv0, hk, b);
        tma_store_commit();
        tma_store_wait<0>();
    }
}

// DR = real head dimension (128 or 64).  The tile layout is the d = 128 one; for d = 64 only the first 64-wide box of
// every tile is loaded and the MMAs run over 64 (M1, M2: K extent; M3, M4: N extent).  M5 has d as its M extent, which
// stays 128: the second box of K is zero-filled once, so lanes 64..127 of dQ^T are zero and are not drained.
// r2 experiments on this kernel, all measured at C3 and all removed from the source again (DESIGN.md 4.2 has the reading):
//   * issue loop unrolled by six with merged waits: -4 %              (profiles/r2a_bwd_variant1_negative.txt)
//   * SS / TS MMAs interleaved as SS SS TS triples: correct, -8 %      (profiles/r2a_bwd_interleaved_negative.txt)
//   * K^T resident in TMEM (M5 as a TS MMA) with S single-buffered: correct, level (the S -> P -> M3 -> M1 chain, ~2300 clk,
//     replaces the pipe limit)
//   * mbarrier.try_wait suspend hints: no effect                       (profiles/r2a_try_suspend_no_effect.txt)
//   * dS hand-off split in two (TMEM copy ready -> M4 at once, shared-memory copy ready -> M5): level / -1 %
//                                                                      (profiles/r2e_bwd_split_ds_handoff_level.txt)
//   * M1 issued for PAIRS of q tiles (one N = 128 group into S[0] | S[1], Q slots box-major so that two tiles form one
//     128-row operand, dS single-buffered to pay for the fourth Q slot): correct on every tail, -6 % (11.6-11.7 against
//     10.9-11.1 ms): the scores of a pair can only be issued after P of its second tile is consumed, which puts the
//     compute warps on the critical path, and the N = 128 bursts starve their shared-memory loads / stores
//                                                                      (profiles/r2h_bwd_paired_m1_negative.txt, .patch)
// The subtraction profile (profiles/r2a_bwd_subtraction_profile.txt) and tools/mma_rate.cu explain why: an SS MMA with
// M = 128 costs >= 64 clk whatever N <= 128 is (the A tile is read from shared memory at 64 B/clk), so the three N = 64
// SS groups cost 3 x 8 x 64 = 1536 clk and the two TS groups 512: 2048 clk per iteration against the measured ~2090.
template <int DR>
__global__ void __launch_bounds__(bwd::THREADS, 1)
bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
           const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
           float* __restrict__ dQacc, const __grid_constant__ CUtensorMap tmdK,
           const __grid_constant__ CUtensorMap tmdV, const float* __restrict__ L2p, const float* __restrict__ Dp,
           int Nq, int stat_stride, int dq_rows, float scale_log2, float scale, int dbg, int causal, int G, int BHkv)
{
    using namespace bwd;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* bar_KV = bars;                     // [1]
    uint64_t* bar_in_full = bar_KV + 1;          // [NS]  Q_i, dO_i, L_i, D_i landed
    uint64_t* bar_in_empty = bar_in_full + NS;   // [NS]
    uint64_t* bar_S = bar_in_empty + NS;         // [2]   S^T(i) complete
    uint64_t* bar_dP = bar_S + 2;                // [1]   dP^T(i) complete
    uint64_t* bar_P = bar_dP + 1;                // [2]   P^T(i) written (one arrival per compute warp)
    uint64_t* bar_dS = bar_P + 2;                // [1]   dS^T(i) written to TMEM + smem (one arrival per compute warp)
    uint64_t* bar_dQ_full = bar_dS + 1;          // [1]
    uint64_t* bar_dQ_empty = bar_dQ_full + 1;    // [1]   (one arrival per drain warp)
    uint64_t* bar_done = bar_dQ_empty + 1;       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_done + 1);

    const int warp = threadIdx.x >> 5;
    // grid (kv tiles, K/V heads of the map, batches of the map).  One CTA serves its K/V tile for all G query heads of
    // the group (grouped-query attention; G = 1: MHA): dK/dV simply keep accumulating over the group's q tiles.
    const int hk = blockIdx.y, b = blockIdx.z, kv0 = blockIdx.x * BN;
    if (b * (int)gridDim.y + hk >= BHkv) return;                    // (ragged last batch of a re-split contiguous BH)
    const size_t bhq0 = ((size_t)b * gridDim.y + hk) * G;            // first q head of the group, in [B*H)
    const int n_q = (Nq + BM - 1) / BM;
    // Every CTA of a (b,h) slab walks the q tiles in the same cyclic order but starts at a different
    // one: at any moment the CTAs that run together reduce-add into DIFFERENT dQ tiles, so their
    // fp32 atomics do not queue up on the same L2 lines.
    // Causal (bottom-right aligned, Nq <= Nkv): the q tiles before the diagonal (rows < kv0 - (Nkv - Nq)) see none of this CTA's keys and are skipped;
    // low blockIdx.x = most work = scheduled first.
    const int qt_min = causal ? min(max(kv0 - (causal - 1), 0) / BM, n_q - 1) : 0;
    const int n_it = n_q - qt_min;
    const int i_off = (int)(((long long)blockIdx.x * n_q) / gridDim.x) % n_it;
    const int n_tot = G * n_it;

    if (threadIdx.x == 0) {
        mbar_init(bar_KV, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&bar_in_full[s], 1); mbar_init(&bar_in_empty[s], 1); }
        mbar_init(&bar_S[0], 1); mbar_init(&bar_S[1], 1); mbar_init(bar_dP, 1);
        mbar_init(&bar_P[0], 8); mbar_init(&bar_P[1], 8); mbar_init(bar_dS, 8);
        mbar_init(bar_dQ_full, 1); mbar_init(bar_dQ_empty, 4); mbar_init(bar_done, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmdO);
        tma_prefetch_desc(&tmdK); tma_prefetch_desc(&tmdV);
    }
    if (warp == 2) { tmem_alloc<512>(tmem_slot); tmem_relinquish(); }
    if (DR == 64) {
        for (int i = threadIdx.x; i < KV_BOX / 16; i += THREADS)
            reinterpret_cast<uint4*>(smem + OFF_K + KV_BOX)[i] = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    constexpr int NBOX = DR / 64;

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (elect_one()) {
            mbar_expect_tx(bar_KV, 2 * NBOX * KV_BOX);
            for (int bx = 0; bx < NBOX; ++bx) {
                tma_load_4d(smem + OFF_K + bx * KV_BOX, &tmK, bar_KV, bx * 64, kv0, hk, b, kEvictFirst);
                tma_load_4d(smem + OFF_V + bx * KV_BOX, &tmV, bar_KV, bx * 64, kv0, hk, b, kEvictFirst);
            }
            for (int i = 0, ii = 0, g = 0; i < n_tot; ++i) {
                const int s = i % NS;
                int qt = ii + i_off;
                if (qt >= n_it) qt -= n_it;
                qt += qt_min;
                const int h = hk * G + g;
                const float* Lrow = L2p + (bhq0 + g) * stat_stride;
                const float* Drow = Dp + (bhq0 + g) * stat_stride;
                if (++ii == n_it) { ii = 0; ++g; }
                mbar_wait(&bar_in_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&bar_in_full[s], 2 * NBOX * Q_BOX + 2 * STAT_BYTES);
                for (int bx = 0; bx < NBOX; ++bx) {
                    tma_load_4d(smem + OFF_Q + s * Q_TILE + bx * Q_BOX, &tmQ, &bar_in_full[s], bx * 64, qt * BM, h, b, kEvictLast);
                    tma_load_4d(smem + OFF_DO + s * Q_TILE + bx * Q_BOX, &tmdO, &bar_in_full[s], bx * 64, qt * BM, h, b, kEvictLast);
                }
                bulk_load_1d(smem + OFF_STAT + s * 2 * STAT_BYTES, Lrow + qt * BM, STAT_BYTES, &bar_in_full[s]);
                bulk_load_1d(smem + OFF_STAT + s * 2 * STAT_BYTES + STAT_BYTES, Drow + qt * BM, STAT_BYTES, &bar_in_full[s]);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        if (elect_one()) {
            constexpr uint32_t idesc_ss = make_idesc(kFmtBF16, 128, BM, false, false);   // M1, M2
            constexpr uint32_t idesc_ts = make_idesc(kFmtBF16, 128, DR, false, true);    // M3, M4
            constexpr uint32_t idesc_dq = make_idesc(kFmtBF16, 128, BM, true, true);     // M5
            const uint32_t sb = smem_u32(smem);
            auto mma_ss = [&](uint32_t a_tile, uint32_t b_tile, uint32_t d_col) {        // D = A[128 x d] B[64 x d]^T
                const uint64_t ad = make_sdesc(a_tile, 16, 1024, 2), bd = make_sdesc(b_tile, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < DR / 16; ++kk)
                    umma_f16_ss(tmem + d_col, sdesc_advance(ad, (kk >> 2) * KV_BOX + (kk & 3) * 32),
                                sdesc_advance(bd, (kk >> 2) * Q_BOX + (kk & 3) * 32), idesc_ss, kk > 0);
            };
            auto mma_ts = [&](uint32_t a_col, uint32_t b_tile, uint32_t d_col, bool acc) {   // D += A^T-in-TMEM[128 x 64] B[64 x d]
                const uint64_t bd = make_sdesc(b_tile, Q_BOX, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < BM / 16; ++kk)
                    umma_f16_ts(tmem + d_col, tmem + a_col + (kk < 2 ? 8 * kk : 32 + 8 * (kk - 2)),
                                sdesc_advance(bd, kk * 2048), idesc_ts, (acc || kk > 0) ? 1u : 0u);
            };
            auto M1 = [&](int i) { if (!B200_DBG(16)) mma_ss(sb + OFF_K, sb + OFF_Q + (i % NS) * Q_TILE, TM_S + 64 * (i & 1)); umma_commit(&bar_S[i & 1]); };
            auto M2 = [&](int i) { if (!B200_DBG(32)) mma_ss(sb + OFF_V, sb + OFF_DO + (i % NS) * Q_TILE, TM_DP); umma_commit(bar_dP); };

            mbar_wait(bar_KV, 0);
            mbar_wait(&bar_in_full[0], 0);
            tc_fence_after();
            M1(0); M2(0);
            auto gen = [&](const int i) {                       // one iteration, everything decided at run time
                const int b = i & 1, s = i % NS;
                if (i + 1 < n_tot) {
                    mbar_wait(&bar_in_full[(i + 1) % NS], ((i + 1) / NS) & 1);
                    tc_fence_after();
                    M1(i + 1);
                }
                B200_TRB(2, i, 0);
                mbar_wait(&bar_P[b], (i >> 1) & 1);
                tc_fence_after();
                B200_TRB(2, i, 1);
                if (!B200_DBG(64)) mma_ts(TM_S + 64 * b, sb + OFF_DO + s * Q_TILE, TM_DV, i > 0);          // M3: dV += P^T dO_i
                B200_TRB(2, i, 2);
                mbar_wait(bar_dS, i & 1);
                tc_fence_after();
                B200_TRB(2, i, 3);
                if (!B200_DBG(64)) mma_ts(TM_DP, sb + OFF_Q + s * Q_TILE, TM_DK, i > 0);                   // M4: dK += dS^T Q_i
                umma_commit(&bar_in_empty[s]);
                if (i + 1 < n_tot) M2(i + 1);
                B200_TRB(2, i, 4);
                mbar_wait(bar_dQ_empty, (i & 1) ^ 1);
                tc_fence_after();
                B200_TRB(2, i, 5);
                if (!B200_DBG(4)) {                                                     // M5: dQ_i^T = K^T dS
                    const uint64_t ad = make_sdesc(sb + OFF_K, KV_BOX, 1024, 2);
                    const uint64_t bd = make_sdesc(sb + OFF_DS + b * DS_TILE, 16, 1024, 2);
#pragma unroll
                    for (int kk = 0; kk < BN / 16; ++kk)
                        umma_f16_ss(tmem + TM_DQ, sdesc_advance(ad, kk * 2048), sdesc_advance(bd, kk * 2048), idesc_dq, kk > 0);
                }
                umma_commit(bar_dQ_full);
                B200_TRB(2, i, 6);
            };
            for (int i = 0; i < n_tot; ++i) gen(i);
            umma_commit(bar_done);
        }
    } else if (warp >= 4 && warp < 12) {
        const int w = (warp - 4) >> 2;
        bwd_compute_role(w, tmem, smem, bar_in_full, bar_S, bar_dP, bar_P, bar_dS, n_tot, n_it, qt_min, i_off, kv0, causal, scale_log2, scale, dbg);
        mbar_wait(bar_done, 0);
        tc_fence_after();
        bwd_epilogue<DR>(w, tmem, smem, w == 0 ? &tmdV : &tmdK, kv0, hk, b);
    } else if (warp >= 12) {
        bwd_drain_role<DR>(tmem, bar_dQ_full, bar_dQ_empty, dQacc, Nq, dq_rows, n_tot, n_it, qt_min, i_off, bhq0, dbg, smem);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------ host side
// L2p[bh][q] = L * log2(e), Dp[bh][q] = D, both padded to a multiple of 64 rows per slab with
// (+inf, 0): a padded query row then has P = exp2(0 - inf) = 0 and dS = 0, so ragged Nq needs no masks.
__global__ void __launch_bounds__(256)
pad_stats_kernel(const float* __restrict__ L, const float* __restrict__ D, float* __restrict__ L2p, float* __restrict__ Dp,
                 int Nq, int Nq_pad, size_t total, long long ldL)
{
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const size_t bh = idx / Nq_pad;
    const int q = (int)(idx - bh * Nq_pad);
    // ldL: floats between the L rows of consecutive (b,h) (Nq for the contiguous [B,H,Nq]; total_q for packed batches)
    if (q < Nq) { L2p[idx] = L[bh * ldL + q] * 1.4426950408889634f; Dp[idx] = D[bh * Nq + q]; }
    else { L2p[idx] = INFINITY; Dp[idx] = 0.0f; }
}

// (also used by the tf32 backward, bwd_tf32_sm100.cu)
int pad_stats(const float* L, const float* D, float* L2p, float* Dp, int BH, int Nq, cudaStream_t st)
{
    const int Nq_pad = (Nq + 63) / 64 * 64;
    const size_t total = (size_t)BH * Nq_pad;
    pad_stats_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, D, L2p, Dp, Nq, Nq_pad, total, Nq);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

__global__ void __launch_bounds__(256)
cast_dq_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out, size_t n4)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(acc)[i];
        uint2 o;
        o.x = pack_bf16x2(a.x, a.y);
        o.y = pack_bf16x2(a.z, a.w);
        reinterpret_cast<uint2*>(out)[i] = o;
    }
}

// fp32 accumulator [B*H, Nq, d] (contiguous) -> bf16 dQ with arbitrary batch / head / row strides; one warp per row
__global__ void __launch_bounds__(256)
cast_dq_strided_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out, Lay lay, int H, int Nq, int d, size_t rows)
{
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const size_t bh = row / Nq;
    const int n = (int)(row - bh * Nq), b = (int)(bh / H), h = (int)(bh - (size_t)b * H);
    const float* src = acc + row * d;
    __nv_bfloat16* dst = out + (size_t)b * lay.sb + (size_t)h * lay.sh + (size_t)n * lay.sn;
    for (int c = lane * 4; c < d; c += 128) {
        const float4 a = *reinterpret_cast<const float4*>(src + c);
        uint2 o;
        o.x = pack_bf16x2(a.x, a.y);
        o.y = pack_bf16x2(a.z, a.w);
        *reinterpret_cast<uint2*>(dst + c) = o;
    }
}

#ifdef B200_TRACE
int bwd_trace_read(long long* out, int n)
{
    if (n > 3 * 8 * 8) n = 3 * 8 * 8;
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_trace_bwd, n * sizeof(long long)));
    return n;
}
#endif

static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

bool tc_bwd_supported(int dtype, int d, int Nq, int Nkv)
{
    return dtype == B200_DTYPE_BF16 && (d == 128 || d == 64) && Nq >= 1 && Nkv >= 1;
}

size_t tc_bwd_workspace_bytes(int BH, int Nq, int Nkv, int d, int dtype)
{
    (void)Nkv; (void)dtype;
    const size_t Nq_pad = (size_t)(Nq + 63) / 64 * 64;
    return 2 * al256((size_t)BH * Nq_pad * sizeof(float)) + al256((size_t)BH * Nq * d * sizeof(float));
}

// One launch: (dK, dV) of every K/V head and the fp32 dQ contributions of its G query heads.
struct BwdLaunch {
    const void *Q, *K, *V, *dO; const float *L2p, *Dp; float* dQacc; void *dK, *dV;
    Lay lq, lk, lv, ldo, ldk, ldv;
    int Bm, Hkv, G, BHkv;            // grid / map batches and K/V heads, group size, valid K/V slabs (<= Bm * Hkv)
    int Nq, Nkv; float scale; bool causal;
    int stat_stride = 0, dq_rows = 0;    // floats per slab of L2p / Dp and rows per slab of dQacc; 0 = pad64(Nq) / Nq
};

template <int DR>
static int bwd_launch(const BwdLaunch& a, cudaStream_t st)
{
    using namespace bwd;
    // the opt-in shared-memory size is a per-device attribute of the function: set it once on every device used
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(bwd_kernel<DR>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int Nq_pad = (a.Nq + 63) / 64 * 64, Hq = a.Hkv * a.G;
    CUtensorMap tmQ, tmK, tmV, tmdO, tmdK, tmdV;
    const CUtensorMapDataType bf = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    int trc = 0;
    if ((trc = make_tmap_4d(&tmQ, a.Q, bf, 2, a.Bm, Hq, a.Nq, DR, a.lq.sb, a.lq.sh, a.lq.sn, BM, 64)) ||
        (trc = make_tmap_4d(&tmdO, a.dO, bf, 2, a.Bm, Hq, a.Nq, DR, a.ldo.sb, a.ldo.sh, a.ldo.sn, BM, 64)) ||
        (trc = make_tmap_4d(&tmK, a.K, bf, 2, a.Bm, a.Hkv, a.Nkv, DR, a.lk.sb, a.lk.sh, a.lk.sn, BN, 64)) ||
        (trc = make_tmap_4d(&tmV, a.V, bf, 2, a.Bm, a.Hkv, a.Nkv, DR, a.lv.sb, a.lv.sh, a.lv.sn, BN, 64)) ||
        (trc = make_tmap_4d(&tmdK, a.dK, bf, 2, a.Bm, a.Hkv, a.Nkv, DR, a.ldk.sb, a.ldk.sh, a.ldk.sn, BN, 64)) ||
        (trc = make_tmap_4d(&tmdV, a.dV, bf, 2, a.Bm, a.Hkv, a.Nkv, DR, a.ldv.sb, a.ldv.sh, a.ldv.sn, BN, 64)))
        return trc == -2 ? B200_ERR_UNSUPPORTED : B200_ERR_CUDA;   // -2: a stride is zero or not a multiple of 16 bytes
    dim3 grid((a.Nkv + BN - 1) / BN, a.Hkv, a.Bm);
#ifdef B200_EXPERIMENTS
    static int dbg = -1;                     // timing-only hooks, side libraries only (see B200_DBG above)
    if (dbg < 0) { const char* e = getenv("B200_BWD_DBG"); dbg = e ? atoi(e) : 0; }
#else
    const int dbg = 0;
#endif
    bwd_kernel<DR><<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmdO, a.dQacc, tmdK, tmdV, a.L2p, a.Dp, a.Nq,
                                                a.stat_stride ? a.stat_stride : Nq_pad, a.dq_rows ? a.dq_rows : a.Nq,
                                                a.scale * 1.4426950408889634f, a.scale, dbg, a.causal ? 1 + (a.Nkv - a.Nq) : 0, a.G, a.BHkv);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

static int bwd_launch_d128(const BwdLaunch& a, cudaStream_t st) { return bwd_launch<128>(a, st); }
static int bwd_launch_d64(const BwdLaunch& a, cudaStream_t st) { return bwd_launch<64>(a, st); }

// contiguous [BH, N, 128 or 64] slabs (the reference layout); BH is cut into (batches x heads) that fit the grid limits
static BwdLaunch contiguous_launch(const void* Q, const void* K, const void* V, const void* dO, const float* L2p, const float* Dp,
                                   float* dQacc, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, bool causal)
{
    const int Hm = BH <= 65535 ? BH : 32768, Bm = (BH + Hm - 1) / Hm;
    const Lay q = lay_contiguous(Hm, Nq, d), k = lay_contiguous(Hm, Nkv, d);
    return BwdLaunch{Q, K, V, dO, L2p, Dp, dQacc, dK, dV, q, k, k, q, k, k, Bm, Hm, 1, BH, Nq, Nkv, scale, causal};
}

// padded statistics + zeroed fp32 dQ accumulator in the workspace, one launch, then the cast (strided when dQ is)
static int bwd_run(const BwdLaunch& a, int BHq, int d, void* ws, void* dQ, const Lay* ldq, int B, int H,
                   const float* L, const float* Dvec, cudaStream_t st, long long ldL = 0)
{
    using namespace bwd;
    const int Nq = a.Nq, Nq_pad = (Nq + 63) / 64 * 64;
    uint8_t* p = reinterpret_cast<uint8_t*>(ws);
    float* L2p = reinterpret_cast<float*>(p); p += al256((size_t)BHq * Nq_pad * sizeof(float));
    float* Dp = reinterpret_cast<float*>(p);  p += al256((size_t)BHq * Nq_pad * sizeof(float));
    float* dQacc = reinterpret_cast<float*>(p);
    const size_t n_dq = (size_t)BHq * Nq * d;
    const size_t total = (size_t)BHq * Nq_pad;
    pad_stats_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, Dvec, L2p, Dp, Nq, Nq_pad, total, ldL > 0 ? ldL : Nq);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    B200_CUDA_TRY(cudaMemsetAsync(dQacc, 0, n_dq * sizeof(float), st));
    BwdLaunch b = a;
    b.L2p = L2p; b.Dp = Dp; b.dQacc = dQacc;
    if (d == 128) B200_TRY(bwd_launch_d128(b, st)); else B200_TRY(bwd_launch_d64(b, st));
    if (ldq == nullptr) {
        unsigned cgrid = (unsigned)((n_dq / 4 + 255) / 256);
        if (cgrid > 148u * 8) cgrid = 148u * 8;
        cast_dq_kernel<<<cgrid, 256, 0, st>>>(dQacc, reinterpret_cast<__nv_bfloat16*>(dQ), n_dq / 4);
    } else {
        const size_t rows = (size_t)BHq * Nq;
        cast_dq_strided_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(dQacc, reinterpret_cast<__nv_bfloat16*>(dQ), *ldq, H, Nq, d, rows);
        (void)B;
    }
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

int tc_bwd(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* Dvec,
           void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, int d, float scale, int dtype,
           void* ws, size_t ws_bytes, cudaStream_t st, bool causal)
{
    if (causal && Nq > Nkv) return B200_ERR_INVALID_ARGUMENT;
    const uintptr_t align = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)dO | (uintptr_t)dQ | (uintptr_t)dK | (uintptr_t)dV;
    if ((align & 15) != 0)
        return simt_bwd(Q, K, V, dO, L, Dvec, dQ, dK, dV, BH, Nq, Nkv, d, scale, dtype, st, causal ? Nkv - Nq : kNoCausal);
    if (ws_bytes < tc_bwd_workspace_bytes(BH, Nq, Nkv, d, dtype)) return B200_ERR_WORKSPACE;
    const BwdLaunch a = contiguous_launch(Q, K, V, dO, nullptr, nullptr, nullptr, dK, dV, BH, Nq, Nkv, d, scale, causal);
    return bwd_run(a, BH, d, ws, dQ, nullptr, 1, BH, L, Dvec, st);
}

// strided / grouped-head form: L, D contiguous [B*H, Nq]; workspace as tc_bwd_workspace_bytes(B*H, ...)
int tc_bwd_strided(const BwdArgs& g, void* ws, size_t ws_bytes, cudaStream_t st)
{
    if (!tc_bwd_supported(g.dtype, g.d, g.Nq, g.Nkv)) return B200_ERR_UNSUPPORTED;
    if (g.causal && g.Nq > g.Nkv) return B200_ERR_INVALID_ARGUMENT;
    if (g.B < 1 || g.H < 1 || g.Hkv < 1 || g.H % g.Hkv != 0 || g.B > 65535 || g.H > 65535) return B200_ERR_INVALID_ARGUMENT;
    const uintptr_t align = (uintptr_t)g.Q | (uintptr_t)g.K | (uintptr_t)g.V | (uintptr_t)g.dO | (uintptr_t)g.dQ | (uintptr_t)g.dK | (uintptr_t)g.dV;
    if ((align & 15) != 0 || (g.ldq.sn & 3) || (g.ldq.sh & 3) || (g.ldq.sb & 3)) return B200_ERR_UNSUPPORTED;
    if (ws_bytes < tc_bwd_workspace_bytes(g.B * g.H, g.Nq, g.Nkv, g.d, g.dtype)) return B200_ERR_WORKSPACE;
    const BwdLaunch a{g.Q, g.K, g.V, g.dO, nullptr, nullptr, nullptr, g.dK, g.dV, g.lq, g.lk, g.lv, g.ldo, g.ldk, g.ldv,
                      g.B, g.Hkv, g.H / g.Hkv, g.B * g.Hkv, g.Nq, g.Nkv, g.scale, g.causal};
    return bwd_run(a, g.B * g.H, g.d, ws, g.dQ, &g.ldq, g.B, g.H, g.L, g.D, st, g.ldL);
}

// ---- pieces for the ring backward: statistics once per call; per ring step one launch that adds
//      its dQ into the caller's fp32 accumulator (no zeroing, no cast) and writes bf16 dK/dV partials
bool tc_bwd_ring_supported(int dtype, int d, int BH, const void* Q, const void* K, const void* V, const void* dO)
{
    const uintptr_t align = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)dO;
    return dtype == B200_DTYPE_BF16 && d == 128 && (align & 15) == 0 && BH <= 65535;
}
size_t tc_bwd_ring_stats_bytes(int BH, int Nq)
{
    const size_t Nq_pad = (size_t)(Nq + 63) / 64 * 64;
    return 2 * al256((size_t)BH * Nq_pad * sizeof(float));
}
int tc_bwd_ring_prepare(const float* L, const float* Dvec, void* stats, int BH, int Nq, cudaStream_t st)
{
    const int Nq_pad = (Nq + 63) / 64 * 64;
    float* L2p = reinterpret_cast<float*>(stats);
    float* Dp = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(stats) + al256((size_t)BH * Nq_pad * sizeof(float)));
    const size_t total = (size_t)BH * Nq_pad;
    pad_stats_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(L, Dvec, L2p, Dp, Nq, Nq_pad, total, Nq);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}
// dQacc (fp32 [BH, Nq, 128]) += dQ(block);  dKp / dVp (bf16 [BH, Nkv, 128]) = dK(block), dV(block)
int tc_bwd_ring_step(const void* Q, const void* K, const void* V, const void* dO, const void* stats,
                     float* dQacc, void* dKp, void* dVp, int BH, int Nq, int Nkv, float scale, cudaStream_t st)
{
    return tc_bwd_ring_step_sub(Q, K, V, dO, stats, dQacc, dKp, dVp, BH, Nq, 0, Nq, Nkv, false, scale, st);
}

// The same on a row sub-block (causal zigzag ring): queries q_row0 .. q_row0 + Nq - 1 of every Nl-row slab against the
// first Nkv rows of the visiting block; all buffers keep their Nl-row slabs.  Only rows [0, Nkv) of dKp / dVp are written.
int tc_bwd_ring_step_sub(const void* Q, const void* K, const void* V, const void* dO, const void* stats,
                         float* dQacc, void* dKp, void* dVp, int BH, int Nl, int q_row0, int Nq, int Nkv, bool causal,
                         float scale, cudaStream_t st)
{
    const int Nl_pad = (Nl + 63) / 64 * 64;
    const float* L2p = reinterpret_cast<const float*>(stats);
    const float* Dp = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(stats) + al256((size_t)BH * Nl_pad * sizeof(float)));
    return tc_bwd_ring_step_heads(Q, K, V, dO, L2p, Dp, dQacc, dKp, dVp, 0, BH, Nl, q_row0, Nq, Nkv, causal, scale, st);
}

// ... and on a range of heads [h0, h0 + nh) of the call's B*H slabs (every pointer is the base of the WHOLE tensor; L2p / Dp
// are the two padded statistics arrays of tc_bwd_ring_prepare).  The ring backward cuts its last step into head groups so
// that a group's finished dK / dV accumulators travel home while the next group computes.
int tc_bwd_ring_step_heads(const void* Q, const void* K, const void* V, const void* dO, const float* L2p, const float* Dp,
                           float* dQacc, void* dKp, void* dVp, int h0, int nh, int Nl, int q_row0, int Nq, int Nkv, bool causal,
                           float scale, cudaStream_t st)
{
    const int Nl_pad = (Nl + 63) / 64 * 64;
    if (q_row0 != 0 && (q_row0 % 64 != 0)) return B200_ERR_UNSUPPORTED;     // statistics tiles must stay inside the slab's padding
    const size_t slab = (size_t)Nl * 128, qo = (size_t)h0 * slab + (size_t)q_row0 * 128, ko = (size_t)h0 * slab;
    const __nv_bfloat16 *Qb = reinterpret_cast<const __nv_bfloat16*>(Q), *Kb = reinterpret_cast<const __nv_bfloat16*>(K),
                        *Vb = reinterpret_cast<const __nv_bfloat16*>(V), *dOb = reinterpret_cast<const __nv_bfloat16*>(dO);
    BwdLaunch a = contiguous_launch(Qb + qo, Kb + ko, Vb + ko, dOb + qo, L2p + (size_t)h0 * Nl_pad + q_row0,
                                    Dp + (size_t)h0 * Nl_pad + q_row0, dQacc + qo, reinterpret_cast<__nv_bfloat16*>(dKp) + ko,
                                    reinterpret_cast<__nv_bfloat16*>(dVp) + ko, nh, Nq, Nkv, 128, scale, causal);
    // slabs are Nl rows apart whatever the sub-block covers
    const Lay full = lay_contiguous(a.Hkv, Nl, 128);
    a.lq = a.ldo = a.lk = a.lv = a.ldk = a.ldv = full;
    a.stat_stride = Nl_pad;
    a.dq_rows = Nl;
    return bwd_launch_d128(a, st);
}

}  // namespace b200
