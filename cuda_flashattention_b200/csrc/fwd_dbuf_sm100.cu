// fwd_dbuf_sm100.cu -- the shipped FlashAttention-2 forward for bf16, d = 128 (r2: promoted from experiment to default):
// 64-row K/V tiles with the score tile DOUBLE-BUFFERED in TMEM, so that no MMA ever waits for a softmax.
// Replaces flash_attention_2_forward_kernel (reference src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:13-131,
// src/util/attention_helper.h:40-132).
//
// Why: fwd_sm100.cu is bound by a dependency CHAIN, not by a unit (the CTA-pair experiment, fwd_pair_sm100.cu, halved
// the operand traffic and got slower by exactly the cross-CTA hand-off latency it added).  Per query tile the chain is
//     S(j) -> softmax(j) -> PV(j) -> S(j+1) -> softmax(j+1) ...
// because P(j) overwrites S(j) in place and S(j+1) goes into the same 128 TMEM columns: ~1850 clk of softmax +
// ~540 clk of hand-offs + 1024 clk of MMAs per 128 keys, and the two tiles of the CTA can only hide each other's MMA
// part (period ~3200 clk per 128 keys against 2048 clk of tensor work or of MUFU work).
// Here a K/V tile is 64 keys, S_t(j) is 64 fp32 columns and each query tile owns TWO score buffers:
//     S_t(j) lives in buffer j & 1;  S_t(j+2) is issued right after PV_t(j) (the MMA that frees the buffer),
// so S_t(j+1) is complete long before softmax_t(j) ends: the softmax warps run back to back (MUFU-bound), the MMA
// thread only ever waits for P, and the only remaining chain is softmax(j) -> PV(j) -> S(j+2).
// Cost: the N = 64 score MMAs are operand-fetch-bound (49 instead of 32 clk per K = 16 step, tools/mma_rate.cu), so the
// tensor work per 128 keys and CTA grows from 2048 to ~2590 clk -- still below fwd_sm100.cu's 3200-clk period.
// The O-rescale hand-off changes with it: "PV_t(j-1) is complete" can no longer be read off S_t(j)'s barrier (S_t(j)
// is now issued BEFORE PV_t(j-1)), so every PV commits to its own barrier bar_R[t], waited for only on the (rare)
// rescale path.
// Two more things turned out to matter (r1(d), profiles/README.md):
//   * ONE MMA-ISSUING THREAD PER QUERY TILE (warps 1 and 3).  With a single thread serving both tiles the kernel ran at
//     3.95 ms: in the trace the tensor pipe idles between the two tiles and the issuing thread falls ~1.5 steps behind
//     the softmax warps.  The MMA queue is shallow -- tools/mma_rate.cu prices a ready mbarrier wait between two MMA
//     groups at ~66 clk of pipe time (the fence behind it at ~8) -- so everything one thread does between two MMAs is a
//     pipe bubble; a second issuing thread fills it.  With one thread per tile: 3.30 ms.
//   * the FMA-pipe exp2 polynomial (B200_DBUF_POLY of every 16 pairs; ptx.cuh exp2_poly2).  It does not pay in
//     fwd_sm100.cu (chain-bound), but here the softmax warps run back to back and the MUFU is the co-limiter:
//     4 of 16 pairs -> 3.14 ms at C3 (fwd_sm100.cu: 3.28), 3.60 against 3.78 ms sustained under the power cap, causal
//     1.71 against 1.82 ms; ncu: tensor pipe 76 % active (fwd_sm100.cu 67 %), MUFU 58 %.
// STATUS: default since r2 -- the forward-facing -m gpu suite (plain / causal / layouts / host pipeline / ring emulation:
// 67 tests) passes on it and bench.py's forward went from 3.81 to 3.43 ms in the fwd+bwd loop (profiles/r2a_*).
//
// Everything else is fwd_sm100.cu: roles, thread = row softmax with speculation and lazy rescale, P published in two
// halves, causal / ragged masks, 4-D tensor maps (strided layouts, grouped K/V heads), epilogue through the Q tile.
// TMEM map (512 columns): S_t / P_t buffer u at 128 t + 64 u, O_t at 256 + 128 t.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"
#include <stdlib.h>
#include <type_traits>

namespace b200 {

#ifdef B200_TRACE
__device__ long long g_trace_dbuf[3 * 8 * 8];
#endif
#define FDB_TR(role, j, ev) B200_TRACE_EVENT(g_trace_dbuf, role, j, ev, (role) == 2 || ((threadIdx.x >> 5) & 3) == 0)

namespace fdb {
constexpr int D = 128, BC = 64, NT = 2;
constexpr int ST = 4;                                        // K and V stages (16 KB each)
constexpr int Q_BOX = 128 * 128, Q_TILE = 2 * Q_BOX;         // 32 KB: [128 rows x 128 d] bf16 as 2 boxes of 64 d
constexpr int KV_BOX = BC * 128, KV_TILE = 2 * KV_BOX;       // 16 KB: [64 keys x 128 d]
constexpr int THREADS = 128 + NT * 128;
constexpr int NBARS = NT * 9 + ST * 4;
constexpr int SMEM = 1024 + NT * Q_TILE + 2 * ST * KV_TILE + NBARS * 8 + 16;
constexpr int REGS_CTRL = 64, REGS_SOFTMAX = 216;            // setmaxnreg targets, as fwd_sm100.cu
constexpr float kThreshold = 8.0f;                           // lazy-rescale threshold (log2 units), as fwd_sm100.cu
static_assert(128 * REGS_CTRL + 256 * REGS_SOFTMAX <= 384 * 168, "register pool of the CTA");
static_assert(SMEM <= 227 * 1024, "shared memory budget");
}  // namespace fdb

// One softmax warpgroup (128 threads) serving query tile t.  Thread r owns query row r of the tile.
// bar_S / bar_Pa / bar_Pb point at the tile's PAIR of barriers (one per score buffer).
template <int POLY>
__device__ __forceinline__ void dbuf_softmax_role(int t, uint32_t tmem, uint8_t* sQ, uint64_t* bar_S, uint64_t* bar_Pa,
                                                  uint64_t* bar_Pb, uint64_t* bar_R, uint64_t* bar_O, const CUtensorMap* tmO,
                                                  float* __restrict__ Lout, long long ldL, int Nq, int Nkv, int nkv, int n_plain,
                                                  int causal, int q0, int h, int b, int bh, float scale_log2)
{
    using namespace fdb;
    const int wg_thread = threadIdx.x & 127;                       // == TMEM lane == row within the tile
    const int warp_in_wg = (threadIdx.x >> 5) & 3;
    const uint32_t lane_base = uint32_t(warp_in_wg * 32) << 16;
    const uint32_t tS0 = tmem + 128 * t + lane_base;               // score buffer 0 (buffer 1: + 64 columns)
    const uint32_t tO = tmem + 256 + 128 * t + lane_base;
    // first key column this row must NOT see: Nkv (ragged last tile), and row + 1 + (Nkv - Nq) under the causal mask
    // (`causal` carries the bottom-right offset: 0 = no mask, 1 + (Nkv - Nq) otherwise)
    const int col_end = causal ? min(Nkv, q0 + t * 128 + wg_thread + causal) : Nkv;
    const uint64_t scale2 = f2_pack(scale_log2, scale_log2);

    float m = -INFINITY, l = 0.0f;                                 // m in log2 units (already scaled)

    auto step = [&](auto masked_tag, const int j) {
        constexpr bool kMasked = decltype(masked_tag)::value;
        const int u = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const uint32_t tS = tS0 + 64 * u;
        mbar_wait(&bar_S[u], ph);
        tc_fence_after();
        FDB_TR(t, j, 0);
        uint32_t s[BC];
        B200_TMEM_LD_32(tS, s);
        B200_TMEM_LD_32(tS + 32, s + 32);
        tmem_wait_ld();
        FDB_TR(t, j, 1);
        if constexpr (kMasked) {
            const int lim = col_end - j * BC;
#pragma unroll
            for (int i = 0; i < BC; ++i) if (i >= lim) s[i] = 0xff800000u;       // -inf
        }
        // P for the 32 scores of chunk c with the running max m_use; partial row sums into acc
        uint64_t acc[4];
        auto exp_chunk = [&](const int c, const float m_use) {
            const uint64_t neg_m2 = f2_pack(-m_use, -m_use);
            uint32_t r[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const uint64_t x = f2_fma(f2_pack(__uint_as_float(s[32 * c + 2 * i]), __uint_as_float(s[32 * c + 2 * i + 1])), scale2, neg_m2);
                float p0, p1;
                // POLY of every 16 pairs take their exp2 from the FMA-pipe polynomial (ptx.cuh) instead of the MUFU,
                // spread evenly through the chunk so that both pipes stay busy
                const bool poly = !kMasked && POLY > 0 && ((i * POLY) % 16 < POLY);
                if (poly) {
                    exp2_poly2(x, p0, p1);
                } else {
                    float x0, x1;
                    f2_unpack(x, x0, x1);
                    p0 = fast_exp2(x0);
                    p1 = fast_exp2(x1);
                }
                if (c == 0 && i < 4) acc[i] = f2_pack(p0, p1); else acc[i & 3] = f2_add(acc[i & 3], f2_pack(p0, p1));
                r[i] = pack_bf16x2(p0, p1);
            }
            B200_TMEM_ST_16(tS + 16 * c, r);
        };
        // speculate the first half of P on the running max while the row max of this tile is reduced (fwd_sm100.cu)
        exp_chunk(0, m);
        float mq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            constexpr int QW = BC / 4;
            mq[q] = fmaxf(__uint_as_float(s[QW * q]), __uint_as_float(s[QW * q + 1]));
#pragma unroll
            for (int i = 2; i < QW; i += 2)
                mq[q] = fmaxf(mq[q], fmaxf(__uint_as_float(s[QW * q + i]), __uint_as_float(s[QW * q + i + 1])));
        }
        const float mx = fmaxf(fmaxf(mq[0], mq[1]), fmaxf(mq[2], mq[3]));
        const float m_new = fmaxf(m, mx * scale_log2);
        const bool need = (m_new - m) > kThreshold;
        float alpha = 1.0f;
        FDB_TR(t, j, 2);
        if (__any_sync(0xffffffffu, need)) {
            if (need) { alpha = fast_exp2(m - m_new); m = m_new; }
            if (j > 0) {
                // O_t is ours between the completion of PV_t(j-1) -- its own barrier: S_t(j) was issued before it -- and
                // the issue of PV_t(j), which waits for this warp's arrival below.  (S_t(j) complete implies PV_t(j-2)
                // complete, so bar_R has seen j-1 or j completions: the parity wait is unambiguous.)
                mbar_wait(bar_R, (j - 1) & 1);
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < D / 32; ++c) {
                    uint32_t o[32];
                    B200_TMEM_LD_32(tO + 32 * c, o);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    B200_TMEM_ST_32(tO + 32 * c, o);
                }
            }
            exp_chunk(0, m);                                       // redo the speculated half with the new max
        }
        FDB_TR(t, j, 3);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&bar_Pa[u]);     // first half of P is in TMEM: PV can start
        FDB_TR(t, j, 4);
        exp_chunk(1, m);
        float a0, a1, a2, a3;
        f2_unpack(f2_add(acc[0], acc[1]), a0, a1);
        f2_unpack(f2_add(acc[2], acc[3]), a2, a3);
        l = fmaf(l, alpha, (a0 + a1) + (a2 + a3));
        FDB_TR(t, j, 5);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&bar_Pb[u]);
        FDB_TR(t, j, 7);
    };

    for (int j = 0; j < n_plain; ++j) step(std::false_type{}, j);
    for (int j = n_plain; j < nkv; ++j) step(std::true_type{}, j); // (a full last tile masks nothing)

    // ---- epilogue: O_t / l -> smem (reusing the Q tile, 128B-swizzled) -> TMA store; LSE -> global
    mbar_wait(bar_O, 0);
    tc_fence_after();
    const float inv_l = 1.0f / l;
    uint8_t* sO = sQ + t * Q_TILE;
    const int r = wg_thread;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        B200_TMEM_LD_32(tO + 32 * c, o);
        tmem_wait_ld();
        // 32 bf16 = 64 B = 4 chunks of 16 B: box c/2, chunks (c&1)*4 .. +3
        uint8_t* rowp = sO + (c >> 1) * Q_BOX + r * 128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o[8 * q]) * inv_l, __uint_as_float(o[8 * q + 1]) * inv_l);
            v.y = pack_bf16x2(__uint_as_float(o[8 * q + 2]) * inv_l, __uint_as_float(o[8 * q + 3]) * inv_l);
            v.z = pack_bf16x2(__uint_as_float(o[8 * q + 4]) * inv_l, __uint_as_float(o[8 * q + 5]) * inv_l);
            v.w = pack_bf16x2(__uint_as_float(o[8 * q + 6]) * inv_l, __uint_as_float(o[8 * q + 7]) * inv_l);
            const int chunk = (c & 1) * 4 + q;
            *reinterpret_cast<uint4*>(rowp + ((chunk ^ (r & 7)) * 16)) = v;
        }
    }
    fence_proxy_async_smem();
    named_bar_sync(1 + t, 128);
    if (wg_thread == 0) {
#pragma unroll
        for (int bx = 0; bx < 2; ++bx) tma_store_4d(tmO, sO + bx * Q_BOX, bx * 64, q0 + t * 128, h, b);
        tma_store_commit();
        tma_store_wait<0>();
    }
    const int row = q0 + t * 128 + r;
    if (Lout != nullptr && row < Nq)
        Lout[(size_t)bh * ldL + row] = (m + log2f(l)) * 0.6931471805599453f;
}

// PEEL (B200_DBUF_PEEL=1, NOT yet run on a GPU): the issuers' steady-state iterations without any data-dependent branch
// between two runs of MMAs (see the note at the loop).  PEEL = false is the build all r1(d) measurements were taken on.
//
// RING = true: the PERSISTENT RING FORWARD ("one persistent kernel across all ring steps waiting on arrival
// flags"; replaces the reference's per-step relaunch + cudaDeviceSynchronize,
// src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:197-226).  The key sequence of this rank's queries is
// the concatenation of P = ring.blocks K/V blocks of ring.tiles_per_block tiles each: block 0 is the local shard (tmK /
// tmV), block s >= 1 sits in receive slot s - 1 of the rank's scratch (tmK2 / tmV2: the slot is the outermost map
// dimension) and is being pushed there by the previous rank's copy engine while this kernel runs.  Only the TMA
// producer knows: before the first tile of block s it polls the arrival counter ring.flags[s] (ld.acquire.sys) until it
// reaches ring.epoch.  MMA issuers and softmax warps see one long key loop, so (m, l, O) are carried in registers /
// TMEM across the blocks: no partial outputs, no LSE merge, one launch per call.
struct RingWalk { const uint32_t* flags; uint32_t epoch; int blocks, tiles_per_block; };

__device__ __forceinline__ void ring_wait_block(const uint32_t* flag, uint32_t epoch)
{
    uint32_t v;
    long long t0 = 0;
    for (uint32_t polls = 0;; ++polls) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if ((int32_t)(v - epoch) >= 0) break;
        __nanosleep(200);
        if ((polls & 0x3FFFu) == 0x3FFFu) {                 // watchdog (~20 s): a lost peer must not hang the GPU for good
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 40000000000ll) { printf("b200attn: ring block never arrived (flag %u, want %u)\n", v, epoch); __trap(); }
        }
    }
    asm volatile("fence.proxy.async;" ::: "memory");        // the block was written by a peer's DMA: order the TMA reads behind the flag
}

template <int POLY, bool PEEL, bool RING = false>
__global__ void __launch_bounds__(fdb::THREADS, 1)
fwd_dbuf_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                float* __restrict__ Lout, long long ldL, int Nq, int Nkv, float scale_log2, int causal, int G, int BH,
                const __grid_constant__ CUtensorMap tmK2, const __grid_constant__ CUtensorMap tmV2, RingWalk ring)
{
    using namespace fdb;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + NT * Q_TILE;
    uint8_t* sV = sK + ST * KV_TILE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ST * KV_TILE);
    uint64_t* bar_Q = bars;                 // [NT]     Q tile landed
    uint64_t* bar_S = bar_Q + NT;           // [NT][2]  S_t(j) complete in score buffer j & 1
    uint64_t* bar_Pa = bar_S + 2 * NT;      // [NT][2]  first half of P_t(j) written (one arrival per softmax warp)
    uint64_t* bar_Pb = bar_Pa + 2 * NT;     // [NT][2]  all of P_t(j) written
    uint64_t* bar_R = bar_Pb + 2 * NT;      // [NT]     PV_t(j) complete (O_t may be rescaled)
    uint64_t* bar_O = bar_R + NT;           // [NT]     last PV of tile t complete
    uint64_t* bar_Kf = bar_O + NT;          // [ST]
    uint64_t* bar_Ke = bar_Kf + ST;
    uint64_t* bar_Vf = bar_Ke + ST;
    uint64_t* bar_Ve = bar_Vf + ST;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_Ve + ST);

    const int warp = threadIdx.x >> 5;
    const int h = blockIdx.y, b = blockIdx.z, bh = b * gridDim.y + h, hk = h / G;
    if (bh >= BH) return;
    const int q0 = (causal ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * (NT * 128);
    const int ntiles = min(NT, (Nq - q0 + 127) / 128);
    const int nkv_all = (Nkv + BC - 1) / BC;
    auto nkv_of = [&](int t) { return causal ? min(nkv_all, (q0 + t * 128 + 126 + causal) / BC + 1) : nkv_all; };
    const int nkv = nkv_of(ntiles - 1), nkv_first = nkv_of(0);

    if (threadIdx.x == 0) {
        for (int t = 0; t < NT; ++t) {
            mbar_init(&bar_Q[t], 1); mbar_init(&bar_R[t], 1); mbar_init(&bar_O[t], 1);
            for (int u = 0; u < 2; ++u) { mbar_init(&bar_S[2 * t + u], 1); mbar_init(&bar_Pa[2 * t + u], 4); mbar_init(&bar_Pb[2 * t + u], 4); }
        }
        // "stage free" barriers: one commit from the MMA issuer of every active query tile
        for (int s = 0; s < ST; ++s) { mbar_init(&bar_Kf[s], 1); mbar_init(&bar_Ke[s], ntiles); mbar_init(&bar_Vf[s], 1); mbar_init(&bar_Ve[s], ntiles); }
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
        if constexpr (RING) { tma_prefetch_desc(&tmK2); tma_prefetch_desc(&tmV2); }
    }
    if (warp == 2) { tmem_alloc<512>(tmem_slot); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer ================================
        reg_dec<REGS_CTRL>();
        if (elect_one()) {
            for (int t = 0; t < ntiles; ++t) {
                mbar_expect_tx(&bar_Q[t], Q_TILE);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d(sQ + t * Q_TILE + bx * Q_BOX, &tmQ, &bar_Q[t], bx * 64, q0 + t * 128, h, b, kEvictFirst);
            }
            // order K(0) K(1) V(0) K(2) V(1) ...: the score MMAs run two tiles ahead of the PV MMAs
            int arrived = 0;                                    // RING: blocks 0 .. arrived are known to be in place
            auto load_K = [&](int j) {
                const int s = j % ST;
                mbar_wait(&bar_Ke[s], ((j / ST) & 1) ^ 1);
                mbar_expect_tx(&bar_Kf[s], KV_TILE);
                if constexpr (RING) {
                    const int blk = j / ring.tiles_per_block, row = (j - blk * ring.tiles_per_block) * BC;
                    if (blk > arrived) { ring_wait_block(ring.flags + blk, ring.epoch); arrived = blk; }   // (K runs ahead of V)
                    for (int bx = 0; bx < 2; ++bx) {
                        if (blk == 0) tma_load_4d(sK + s * KV_TILE + bx * KV_BOX, &tmK, &bar_Kf[s], bx * 64, row, hk, b, kEvictLast);
                        else tma_load_4d(sK + s * KV_TILE + bx * KV_BOX, &tmK2, &bar_Kf[s], bx * 64, row, bh, blk - 1, kEvictLast);
                    }
                } else {
                    for (int bx = 0; bx < 2; ++bx)
                        tma_load_4d(sK + s * KV_TILE + bx * KV_BOX, &tmK, &bar_Kf[s], bx * 64, j * BC, hk, b, kEvictLast);
                }
            };
            auto load_V = [&](int j) {
                const int s = j % ST;
                mbar_wait(&bar_Ve[s], ((j / ST) & 1) ^ 1);
                mbar_expect_tx(&bar_Vf[s], KV_TILE);
                if constexpr (RING) {
                    const int blk = j / ring.tiles_per_block, row = (j - blk * ring.tiles_per_block) * BC;
                    for (int bx = 0; bx < 2; ++bx) {
                        if (blk == 0) tma_load_4d(sV + s * KV_TILE + bx * KV_BOX, &tmV, &bar_Vf[s], bx * 64, row, hk, b, kEvictLast);
                        else tma_load_4d(sV + s * KV_TILE + bx * KV_BOX, &tmV2, &bar_Vf[s], bx * 64, row, bh, blk - 1, kEvictLast);
                    }
                } else {
                    for (int bx = 0; bx < 2; ++bx)
                        tma_load_4d(sV + s * KV_TILE + bx * KV_BOX, &tmV, &bar_Vf[s], bx * 64, j * BC, hk, b, kEvictLast);
                }
            };
            load_K(0);
            if (nkv > 1) load_K(1);
            for (int j = 0; j < nkv; ++j) {
                load_V(j);
                if (j + 2 < nkv) load_K(j + 2);
            }
        }
    } else if (warp == 1 || warp == 3) {
        // ================================ MMA issuers: warp 1 -> tile 0, warp 3 -> tile 1 ==================
        // One issuing thread PER QUERY TILE.  The MMA queue is shallow: whatever the issuing thread does between two
        // MMAs (a ready barrier wait costs ~66 clk, tools/mma_rate.cu) is a tensor-pipe bubble, and a single thread
        // serving both tiles fell 1.5 steps behind the softmax warps (r1(d) trace); with one thread per tile the
        // other tile's MMAs fill the gaps.  Commits only track the issuing thread's MMAs, so the K/V "stage free"
        // barriers expect one commit from each issuer (initialised with count = ntiles).
        reg_dec<REGS_CTRL>();
        const int t = warp == 1 ? 0 : 1;
        if (t < ntiles && elect_one()) {
            constexpr uint32_t idesc_S = make_idesc(kFmtBF16, 128, BC, false, false);
            constexpr uint32_t idesc_PV = make_idesc(kFmtBF16, 128, D, false, true);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV);
            const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;   // (NT <= 2)

            auto issue_S = [&](int j) {                         // S_t(j) -> score buffer j & 1
                const int s = j % ST, u = j & 1;
                const uint64_t qd = make_sdesc(q_base + t * Q_TILE, 16, 1024, 2);
                const uint64_t kd = make_sdesc(k_base + s * KV_TILE, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk) {
                    const uint32_t inbox = (kk & 3) * 32;
                    umma_f16_ss(tmem + 128 * t + 64 * u, sdesc_advance(qd, (kk >> 2) * Q_BOX + inbox),
                                sdesc_advance(kd, (kk >> 2) * KV_BOX + inbox), idesc_S, kk > 0 ? 1u : 0u);
                }
                umma_commit(&bar_S[2 * t + u]);
            };
            auto issue_PV = [&](int j, int half) {              // half 0 / 1 = first / second 32 keys of P_t(j)
                const int s = j % ST, u = j & 1;
                const uint64_t vd = make_sdesc(v_base + s * KV_TILE, KV_BOX, 1024, 2);
#pragma unroll
                for (int k2 = 0; k2 < 2; ++k2) {
                    const int kk = half * 2 + k2;
                    umma_f16_ts(tmem + 256 + 128 * t, tmem + 128 * t + 64 * u + 8 * kk, sdesc_advance(vd, kk * 16 * 128), idesc_PV,
                                (j > 0 || kk > 0) ? 1u : 0u);
                }
            };

            // prologue: S_t(0) and S_t(1)
            mbar_wait(&bar_Q[t], 0);
            for (int jj = 0; jj < 2 && jj < nkv; ++jj) {
                mbar_wait(&bar_Kf[jj], 0);
                tc_fence_after();
                if (jj < nkv_t) issue_S(jj);
                umma_commit(&bar_Ke[jj]);
            }
            int j0 = 0;
            if constexpr (PEEL) {
                // Steady state (j + 2 < nkv_t): every decision of the general loop below is known.  tools/mma_rate.cu shows
                // that whatever the issuing thread executes between two tcgen05.mma -- a ready barrier wait (~66 clk), even
                // the loop's own compares and branches -- turns into a tensor-pipe bubble, so these iterations carry
                // nothing but the four barrier waits.
                for (; j0 + 2 < nkv_t; ++j0) {
                    const int j = j0, s = j % ST, u = j & 1;
                    const uint32_t ph = (j >> 1) & 1;
                    mbar_wait(&bar_Vf[s], (j / ST) & 1);
                    mbar_wait(&bar_Kf[(j + 2) % ST], ((j + 2) / ST) & 1);
                    mbar_wait(&bar_Pa[2 * t + u], ph);
                    tc_fence_after();
                    issue_PV(j, 0);
                    mbar_wait(&bar_Pb[2 * t + u], ph);
                    tc_fence_after();
                    issue_PV(j, 1);
                    umma_commit(&bar_R[t]);
                    umma_commit(&bar_Ve[s]);
                    issue_S(j + 2);
                    umma_commit(&bar_Ke[(j + 2) % ST]);
                }
            }
            for (int j = j0; j < nkv; ++j) {
                const int s = j % ST, u = j & 1;
                const uint32_t ph = (j >> 1) & 1;
                if (t == 0) FDB_TR(2, j, 0);
                mbar_wait(&bar_Vf[s], (j / ST) & 1);
                if (t == 0) FDB_TR(2, j, 1);
                if (j + 2 < nkv) mbar_wait(&bar_Kf[(j + 2) % ST], ((j + 2) / ST) & 1);
                if (t == 0) FDB_TR(2, j, 2);
                if (j < nkv_t) {
                    mbar_wait(&bar_Pa[2 * t + u], ph);
                    if (t == 0) FDB_TR(2, j, 3);
                    tc_fence_after();
                    if (t == 0) FDB_TR(2, j, 4);
                    issue_PV(j, 0);
                    mbar_wait(&bar_Pb[2 * t + u], ph);
                    tc_fence_after();
                    if (t == 0) FDB_TR(2, j, 5);
                    issue_PV(j, 1);
                    umma_commit(&bar_R[t]);
                }
                umma_commit(&bar_Ve[s]);                        // (an issuer past its own last tile has nothing pending: arrives at once)
                if (t == 0) FDB_TR(2, j, 6);
                if (j + 2 < nkv_t) issue_S(j + 2);              // into the buffer PV_t(j) has just read
                if (j + 1 == nkv_t) umma_commit(&bar_O[t]);
                if (j + 2 < nkv) umma_commit(&bar_Ke[(j + 2) % ST]);
                if (t == 0) FDB_TR(2, j, 7);
            }
        }
    } else if (warp < 4) {
        reg_dec<REGS_CTRL>();
    } else {
        // ================================ softmax warpgroups ==========================
        reg_inc<REGS_SOFTMAX>();
        const int t = (warp - 4) >> 2;
        if (t < ntiles) {
            const int nkv_t = nkv_of(t);
            // tiles entirely below the diagonal (and not the ragged last one) need no mask
            const int n_plain = causal ? min(nkv_t, min((q0 + t * 128 + causal) / BC, Nkv / BC)) : nkv_t - 1;
            dbuf_softmax_role<POLY>(t, tmem, sQ, &bar_S[2 * t], &bar_Pa[2 * t], &bar_Pb[2 * t], &bar_R[t], &bar_O[t], &tmO, Lout, ldL,
                              Nq, Nkv, nkv_t, n_plain, causal, q0, h, b, bh, scale_log2);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

#ifdef B200_TRACE
int dbuf_trace_read(long long* out, int n)
{
    if (n > 3 * 8 * 8) n = 3 * 8 * 8;
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_trace_dbuf, n * sizeof(long long)));
    return n;
}
#endif

template <int POLY, bool PEEL = false>
static int launch_fwd_dbuf_t(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    using namespace fdb;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fwd_dbuf_kernel<POLY, PEEL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const float scale_log2 = a.scale * 1.4426950408889634f;
    CUtensorMap tmQ, tmK, tmV, tmO;
    int trc = 0;
    if ((trc = make_tmap_4d(&tmQ, a.Q, dt, 2, a.B, a.H, a.Nq, D, a.lq.sb, a.lq.sh, a.lq.sn, 128, 64)) ||
        (trc = make_tmap_4d(&tmK, a.K, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lk.sb, a.lk.sh, a.lk.sn, BC, 64)) ||
        (trc = make_tmap_4d(&tmV, a.V, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lv.sb, a.lv.sh, a.lv.sn, BC, 64)) ||
        (trc = make_tmap_4d(&tmO, a.O, dt, 2, a.B, a.H, a.Nq, D, a.lo.sb, a.lo.sh, a.lo.sn, 128, 64)))
        return trc == -2 ? B200_ERR_UNSUPPORTED : B200_ERR_CUDA;   // -2: a stride is zero or not a multiple of 16 bytes
    dim3 grid((a.Nq + NT * 128 - 1) / (NT * 128), a.H, a.B);
    fwd_dbuf_kernel<POLY, PEEL, false><<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmO, a.L, a.ldL, a.Nq, a.Nkv, scale_log2, a.causal ? 1 + (a.Nkv - a.Nq) : 0,
                                                       a.H / a.Hkv, BH_valid, tmK, tmV, RingWalk{nullptr, 0u, 1, 0});
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

// Persistent ring forward: contiguous [BH, Nl, 128] bf16 shards; K/V blocks 1 .. P-1 in `Kslots` / `Vslots`
// ([P-1][BH][Nl][128], filled by the copy engines while the kernel runs); flags[s] reaches `epoch` when block s is in.
bool fwd_ring_persistent_supported(int dtype, int d, int BH, int Nl, int nranks)
{
    return dtype == B200_DTYPE_BF16 && d == fdb::D && Nl >= fdb::BC && Nl % fdb::BC == 0 && BH >= 1 && BH <= 65535 && nranks >= 2 &&
           nranks <= 256;
}
int launch_fwd_ring_persistent(const void* Q, const void* K, const void* V, void* O, float* L, const void* Kslots, const void* Vslots,
                               const uint32_t* flags, uint32_t epoch, int BH, int Nl, int nranks, float scale, cudaStream_t st)
{
    using namespace fdb;
    auto* kern = fwd_dbuf_kernel<4, false, true>;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const Lay lay = lay_contiguous(BH, Nl, D);
    const uint64_t slot = (uint64_t)BH * Nl * D;                 // elements between two receive slots
    CUtensorMap tmQ, tmK, tmV, tmO, tmK2, tmV2;
    int trc = 0;
    if ((trc = make_tmap_4d(&tmQ, Q, dt, 2, 1, BH, Nl, D, lay.sb, lay.sh, lay.sn, 128, 64)) ||
        (trc = make_tmap_4d(&tmK, K, dt, 2, 1, BH, Nl, D, lay.sb, lay.sh, lay.sn, BC, 64)) ||
        (trc = make_tmap_4d(&tmV, V, dt, 2, 1, BH, Nl, D, lay.sb, lay.sh, lay.sn, BC, 64)) ||
        (trc = make_tmap_4d(&tmO, O, dt, 2, 1, BH, Nl, D, lay.sb, lay.sh, lay.sn, 128, 64)) ||
        (trc = make_tmap_4d(&tmK2, Kslots, dt, 2, nranks - 1, BH, Nl, D, slot, lay.sh, lay.sn, BC, 64)) ||
        (trc = make_tmap_4d(&tmV2, Vslots, dt, 2, nranks - 1, BH, Nl, D, slot, lay.sh, lay.sn, BC, 64)))
        return trc == -2 ? B200_ERR_UNSUPPORTED : B200_ERR_CUDA;
    dim3 grid((Nl + NT * 128 - 1) / (NT * 128), BH, 1);
    kern<<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmO, L, (long long)Nl, Nl, nranks * Nl, scale * 1.4426950408889634f, 0, 1, BH,
                                      tmK2, tmV2, RingWalk{flags, epoch, nranks, Nl / BC});
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

int launch_fwd_dbuf(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    if (a.dtype != B200_DTYPE_BF16 || a.d != fdb::D) return B200_ERR_UNSUPPORTED;
#ifdef B200_EXPERIMENTS
    static int poly = -1;                                // side libraries only: B200_DBUF_POLY = pairs of every 16 on the FMA pipe
    if (poly < 0) { const char* e = getenv("B200_DBUF_POLY"); poly = e ? atoi(e) : 4; }
    static int peel = -1;                                // B200_DBUF_PEEL=1: branch-free issue loops (r2: measured 1.5 % slower)
    if (peel < 0) { const char* e = getenv("B200_DBUF_PEEL"); peel = e ? atoi(e) : 0; }
    if (peel) {
        if (poly == 0) return launch_fwd_dbuf_t<0, true>(a, BH_valid, st);
        if (poly == 4) return launch_fwd_dbuf_t<4, true>(a, BH_valid, st);
        return B200_ERR_UNSUPPORTED;
    }
    switch (poly) {
        case 0: return launch_fwd_dbuf_t<0>(a, BH_valid, st);
        case 3: return launch_fwd_dbuf_t<3>(a, BH_valid, st);
        case 4: return launch_fwd_dbuf_t<4>(a, BH_valid, st);
        case 5: return launch_fwd_dbuf_t<5>(a, BH_valid, st);
        case 6: return launch_fwd_dbuf_t<6>(a, BH_valid, st);
        case 8: return launch_fwd_dbuf_t<8>(a, BH_valid, st);
        default: return B200_ERR_UNSUPPORTED;
    }
#else
    return launch_fwd_dbuf_t<4>(a, BH_valid, st);        // the measured configuration
#endif
}

}  // namespace b200
