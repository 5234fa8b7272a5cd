// fwd_pair_sm100.cu -- EXPERIMENTAL forward variant (bf16, d = 128, non-causal), selected with B200_FWD_VARIANT=3:
// the same algorithm and roles as fwd_sm100.cu, run by a CTA PAIR (cluster of 2, `tcgen05.mma.cta_group::2`).
//
// Why it was tried: in fwd_sm100.cu the S = Q K^T MMA (SS, M = 128, N = 128) fetches 8 KB of operands per 64
// tensor clocks -- exactly the 128 B/clk shared-memory port -- and PV another 4 KB, while TMA writes the next K/V
// tiles through the same port: the forward sits on the operand-fetch limit as much as on the tensor pipe.  With a CTA
// pair every MMA is M = 256 (128 query rows in each CTA's TMEM) and the B operand is SPLIT between the two CTAs:
//   S_t  = Q_t K_j^T : each CTA stages 64 of the 128 key rows of K_j      (K-major,  2 boxes of 64 rows x 64 d)
//   O_t += P_t V_j   : each CTA stages 64 of the 128 d-columns of V_j     (MN-major, 1 box  of 128 rows x 64 d)
// so per CTA the K/V traffic (TMA writes and operand reads) halves: 6 KB instead of 8 KB per S MMA, 2 KB instead of
// 4 KB per PV MMA, 32 KB instead of 64 KB of TMA writes per K/V tile pair.
//
// Pair protocol (rank = %cluster_ctarank; rank 0 is the leader):
//   * both CTAs run a TMA producer (their half of every K/V tile, their own Q tiles) whose loads complete on the
//     LEADER's "full" barriers (`cp.async.bulk.tensor ... .cta_group::2`, barrier address mapped with `mapa`);
//     the leader arms those barriers with the bytes of both halves;
//   * only the leader's warp 1 issues MMAs; every `tcgen05.commit` is multicast to the same barrier in both CTAs
//     (S ready, O ready, K/V stage free);
//   * the softmax warpgroups of both CTAs (thread = row, exactly fwd_sm100.cu's non-masked path: speculative first
//     half, lazy rescale, P published in two halves) arrive on the LEADER's P barriers (8 arrivals: 4 warps x 2 CTAs);
//   * TMEM is allocated / freed with `.cta_group::2` by warp 2 of both CTAs; a cluster barrier separates set-up and
//     tear-down, and each producer drains the last multicast "stage free" arrivals before its CTA may exit.
// Shapes: Nq % 512 == 0 and Nkv % 128 == 0 (no ragged tiles, no masks) -- anything else stays on fwd_sm100.cu.
//
// RESULT (r1(d), profiles/r1d_fwd_pair_variant.txt): ran correctly the first time on a B200 (parity against the generic
// kernels at N = 512 / 1024 / 4096) and is 41 % SLOWER than fwd_sm100.cu at C3 (4.69 against 3.32 ms).  The forward is
// bound by the per-tile hand-off chain  S -> softmax -> P-ready -> PV -> S, not by operand traffic: the pair adds a
// multicast commit and a remote mbarrier arrival to every link of that chain (about +650 clk per tile step), and halving
// the shared-memory traffic buys nothing back.  Kept as a working reference for the cta_group::2 protocol (2-CTA TMA
// completion on the leader's barriers, multicast commits, remote arrivals, pair-wide TMEM allocation, tear-down drain);
// never selected by default.  tools/try_fwd_variant.sh 3 re-runs the measurement.
#include "../common.cuh"
#include "../ptx.cuh"
#include "../tmap.h"
#include <stdlib.h>

namespace b200 {

namespace fpr {
constexpr int D = 128, BC = 128, NT = 2;
constexpr int ST = 3;                                        // K/V stages (half tiles: 32 KB per stage and CTA)
constexpr int Q_BOX = 128 * 128, Q_TILE = 2 * Q_BOX;         // 32 KB: [128 rows x 128 d] bf16 as 2 boxes of 64 d
constexpr int KH_BOX = 64 * 128, KH_TILE = 2 * KH_BOX;       // 16 KB: this CTA's 64 key rows of a K tile, 2 boxes of 64 d
constexpr int VH_TILE = 128 * 128;                           // 16 KB: this CTA's 64 d-columns of a V tile, all 128 rows
constexpr int THREADS = 128 + NT * 128;
constexpr int NBARS = NT * 5 + ST * 4;
constexpr int SMEM = 1024 + NT * Q_TILE + ST * (KH_TILE + VH_TILE) + NBARS * 8 + 16;
constexpr int REGS_CTRL = 64, REGS_SOFTMAX = 216;            // setmaxnreg targets, as fwd_sm100.cu
constexpr float kThreshold = 8.0f;                           // lazy-rescale threshold (log2 units), as fwd_sm100.cu
static_assert(128 * REGS_CTRL + 256 * REGS_SOFTMAX <= 384 * 168, "register pool of the CTA");
static_assert(SMEM <= 227 * 1024, "shared memory budget");
}  // namespace fpr

// ------------------------------------------------------------------ cluster / cta_group::2 primitives
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
// shared::cluster address of the location `smem_addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on an mbarrier anywhere in the cluster (address from map_to_cta)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait on a local mbarrier whose arrivals come from both CTAs
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait_cluster(bar, parity)) return;
#if B200_MBAR_TIMEOUT_CYCLES
    uint32_t polls = 0;
    long long t0 = 0;
#endif
    while (!mbar_try_wait_cluster(bar, parity)) {
#if B200_MBAR_TIMEOUT_CYCLES
        if ((++polls & 0xFFFu) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > B200_MBAR_TIMEOUT_CYCLES) {
                printf("b200attn: pair mbarrier wait timed out (block %d,%d thread %d bar@%u parity %u)\n",
                       blockIdx.x, blockIdx.y, threadIdx.x, smem_u32(bar), parity);
                __trap();
            }
        }
#endif
    }
}
// TMA load into THIS CTA's shared memory that completes on a barrier of either CTA of the pair
__device__ __forceinline__ void tma_load_4d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr,
                                                 int c0, int c1, int c2, int c3, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr),
          "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
        : "memory");
}
template <int NCOLS> __device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int NCOLS> __device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}
// arrive (count 1) on the barrier at this offset in BOTH CTAs once every MMA issued so far is complete
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// D[tmem, 128 rows in each CTA] (+)= A[smem of each CTA, 128 rows] * B[smem, N/2 from each CTA]
__device__ __forceinline__ void umma_pair_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[tmem of each CTA, 128 rows] * B[smem, N/2 from each CTA]
__device__ __forceinline__ void umma_pair_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// ------------------------------------------------------------------ softmax warpgroup (one per query tile of a CTA)
// fwd_sm100.cu's softmax_role without the masked path; P-ready arrivals go to the leader CTA's barriers.
__device__ __forceinline__ void pair_softmax_role(int t, uint32_t tmem, uint8_t* sQ, uint64_t* bar_S, uint32_t bar_Pa_leader,
                                                  uint32_t bar_Pb_leader, uint64_t* bar_O, const CUtensorMap* tmO,
                                                  float* __restrict__ Lout, long long ldL, int nkv, int q0, int h, int b, int bh,
                                                  float scale_log2)
{
    using namespace fpr;
    const int wg_thread = threadIdx.x & 127;                       // == TMEM lane == row within the tile
    const int warp_in_wg = (threadIdx.x >> 5) & 3;
    const uint32_t lane_base = uint32_t(warp_in_wg * 32) << 16;
    const uint32_t tS = tmem + 128 * t + lane_base;                // S_t and (aliased) P_t
    const uint32_t tO = tmem + 256 + 128 * t + lane_base;
    const uint64_t scale2 = f2_pack(scale_log2, scale_log2);

    float m = -INFINITY, l = 0.0f;                                 // m in log2 units (already scaled)

    for (int j = 0; j < nkv; ++j) {
        mbar_wait(bar_S, j & 1);
        tc_fence_after();
        uint32_t s[BC];
#pragma unroll
        for (int c = 0; c < BC / 32; ++c) B200_TMEM_LD_32(tS + 32 * c, s + 32 * c);
        tmem_wait_ld();
        uint64_t acc[4];
        auto exp_chunks = [&](const int c0, const int c1, const float m_use) {
            const uint64_t neg_m2 = f2_pack(-m_use, -m_use);
#pragma unroll
            for (int c = c0; c < c1; ++c) {
                uint32_t r[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint64_t x = f2_fma(f2_pack(__uint_as_float(s[32 * c + 2 * i]), __uint_as_float(s[32 * c + 2 * i + 1])), scale2, neg_m2);
                    float x0, x1;
                    f2_unpack(x, x0, x1);
                    const float p0 = fast_exp2(x0), p1 = fast_exp2(x1);
                    if (c == 0 && i < 4) acc[i] = f2_pack(p0, p1); else acc[i & 3] = f2_add(acc[i & 3], f2_pack(p0, p1));
                    r[i] = pack_bf16x2(p0, p1);
                }
                B200_TMEM_ST_16(tS + 16 * c, r);
            }
        };
        // speculate the first half of P on the running max while the row max of this tile is reduced (fwd_sm100.cu)
        exp_chunks(0, BC / 64, m);
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
        if (__any_sync(0xffffffffu, need)) {
            if (need) { alpha = fast_exp2(m - m_new); m = m_new; }
            if (j > 0) {
                // PV_t(j-1) is complete in both CTAs (S_t(j) was committed after it), PV_t(j) is not issued before this
                // warp has arrived on the P barrier: the O_t rows of this CTA are ours.
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
            exp_chunks(0, BC / 64, m);                             // redo the speculated half with the new max
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(bar_Pa_leader);   // first half of P is in TMEM
        exp_chunks(BC / 64, BC / 32, m);
        float a0, a1, a2, a3;
        f2_unpack(f2_add(acc[0], acc[1]), a0, a1);
        f2_unpack(f2_add(acc[2], acc[3]), a2, a3);
        l = fmaf(l, alpha, (a0 + a1) + (a2 + a3));
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(bar_Pb_leader);
    }

    // ---- epilogue: O_t / l -> smem (reusing this CTA's Q tile, 128B-swizzled) -> TMA store; LSE -> global
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
    if (Lout != nullptr)
        Lout[(size_t)bh * ldL + q0 + t * 128 + r] = (m + log2f(l)) * 0.6931471805599453f;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(fpr::THREADS, 1)
fwd_pair_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                float* __restrict__ Lout, long long ldL, int Nkv, float scale_log2, int G, int BH)
{
    using namespace fpr;
    extern __shared__ uint8_t smem_raw[];
    // (the dynamic shared-memory window starts at the same offset in every CTA of the kernel, so the 1024-byte
    // alignment fix-up -- and with it every operand / barrier offset -- is identical in both CTAs of the pair)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + NT * Q_TILE;
    uint8_t* sV = sK + ST * KH_TILE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ST * VH_TILE);
    uint64_t* bar_Q = bars;                 // [NT]  LEADER: Q_t of both CTAs landed (tx bytes of both)
    uint64_t* bar_S = bar_Q + NT;           // [NT]  both: S_t(j) complete (multicast commit)
    uint64_t* bar_Pa = bar_S + NT;          // [NT]  LEADER: first half of P_t(j) written by both CTAs (8 warp arrivals)
    uint64_t* bar_Pb = bar_Pa + NT;         // [NT]  LEADER: all of P_t(j) written by both CTAs
    uint64_t* bar_O = bar_Pb + NT;          // [NT]  both: last PV of tile t complete (multicast commit)
    uint64_t* bar_Kf = bar_O + NT;          // [ST]  LEADER: both halves of the K tile landed
    uint64_t* bar_Ke = bar_Kf + ST;         // [ST]  both: K stage free (multicast commit)
    uint64_t* bar_Vf = bar_Ke + ST;         // [ST]  LEADER
    uint64_t* bar_Ve = bar_Vf + ST;         // [ST]  both
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_Ve + ST);

    const int warp = threadIdx.x >> 5;
    const uint32_t rank = cluster_ctarank();                        // 0 = leader (issues the MMAs), 1 = peer
    const int h = blockIdx.y, b = blockIdx.z, bh = b * gridDim.y + h, hk = h / G;
    if (bh >= BH) return;                                           // both CTAs of a pair share (b, h): the whole cluster leaves
    const int q0 = blockIdx.x * (NT * 128);                         // this CTA's 256 query rows
    const int nkv = Nkv / BC;

    if (threadIdx.x == 0) {
        for (int t = 0; t < NT; ++t) { mbar_init(&bar_Q[t], 1); mbar_init(&bar_S[t], 1); mbar_init(&bar_Pa[t], 8); mbar_init(&bar_Pb[t], 8); mbar_init(&bar_O[t], 1); }
        for (int s = 0; s < ST; ++s) { mbar_init(&bar_Kf[s], 1); mbar_init(&bar_Ke[s], 1); mbar_init(&bar_Vf[s], 1); mbar_init(&bar_Ve[s], 1); }
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
    }
    __syncwarp();
    // barriers of both CTAs must exist before any remote arrival / remote complete_tx / pair-wide allocation
    cluster_sync_all();
    if (warp == 2) { tmem_alloc_pair<512>(tmem_slot); tmem_relinquish_pair(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer (both CTAs) ================================
        reg_dec<REGS_CTRL>();
        if (elect_one()) {
            for (int t = 0; t < NT; ++t) {
                if (rank == 0) mbar_expect_tx(&bar_Q[t], 2 * Q_TILE);
                const uint32_t lb = map_to_cta(smem_u32(&bar_Q[t]), 0);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d_pair(sQ + t * Q_TILE + bx * Q_BOX, &tmQ, lb, bx * 64, q0 + t * 128, h, b, kEvictFirst);
            }
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                const uint32_t ph = (j / ST) & 1;
                mbar_wait(&bar_Ke[s], ph ^ 1);
                if (rank == 0) mbar_expect_tx(&bar_Kf[s], 2 * KH_TILE);
                const uint32_t lk = map_to_cta(smem_u32(&bar_Kf[s]), 0);
                for (int bx = 0; bx < 2; ++bx)       // this CTA's 64 key rows of tile j
                    tma_load_4d_pair(sK + s * KH_TILE + bx * KH_BOX, &tmK, lk, bx * 64, j * BC + (int)rank * 64, hk, b, kEvictLast);
                mbar_wait(&bar_Ve[s], ph ^ 1);
                if (rank == 0) mbar_expect_tx(&bar_Vf[s], 2 * VH_TILE);
                const uint32_t lv = map_to_cta(smem_u32(&bar_Vf[s]), 0);
                // this CTA's 64 d-columns of tile j (all 128 key rows): one box
                tma_load_4d_pair(sV + s * VH_TILE, &tmV, lv, (int)rank * 64, j * BC, hk, b, kEvictLast);
            }
            // drain: the last "stage free" commits are multicast into this CTA's barriers and nothing above waits
            // for them -- the CTA must not exit while such an arrival can still be in flight
            for (int s = 0; s < ST; ++s) {
                const int cnt = (nkv - s + ST - 1) / ST;            // completions of stage s' "free" barriers
                if (cnt > 0) { mbar_wait(&bar_Ke[s], (cnt - 1) & 1); mbar_wait(&bar_Ve[s], (cnt - 1) & 1); }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA only) ============================
        reg_dec<REGS_CTRL>();
        if (rank == 0 && elect_one()) {
            constexpr uint32_t idesc_S = make_idesc(kFmtBF16, 256, BC, false, false);
            constexpr uint32_t idesc_PV = make_idesc(kFmtBF16, 256, D, false, true);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV);

            auto issue_S = [&](int t, int j) {
                const int s = j % ST;
                const uint64_t qd = make_sdesc(q_base + t * Q_TILE, 16, 1024, 2);
                const uint64_t kd = make_sdesc(k_base + s * KH_TILE, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk) {
                    const uint32_t inbox = (kk & 3) * 32;
                    umma_pair_ss(tmem + 128 * t, sdesc_advance(qd, (kk >> 2) * Q_BOX + inbox),
                                 sdesc_advance(kd, (kk >> 2) * KH_BOX + inbox), idesc_S, kk > 0);
                }
                umma_commit_pair(&bar_S[t]);
            };
            auto issue_PV = [&](int t, int j, int half) {       // half 0 / 1 = first / second 64 columns of P
                const int s = j % ST;
                // MN-major B, 64 d-columns per CTA = one 128-byte chunk along N (LBO is never stepped), 8-row K groups 1024 B apart
                const uint64_t vd = make_sdesc(v_base + s * VH_TILE, VH_TILE, 1024, 2);
#pragma unroll
                for (int k2 = 0; k2 < 4; ++k2) {
                    const int kk = half * 4 + k2;
                    umma_pair_ts(tmem + 256 + 128 * t, tmem + 128 * t + 8 * kk, sdesc_advance(vd, kk * 16 * 128), idesc_PV,
                                 (j > 0 || kk > 0));
                }
            };

            mbar_wait_cluster(&bar_Kf[0], 0);
            for (int t = 0; t < NT; ++t) {
                mbar_wait_cluster(&bar_Q[t], 0);
                tc_fence_after();
                issue_S(t, 0);
            }
            umma_commit_pair(&bar_Ke[0]);
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait_cluster(&bar_Vf[s], (j / ST) & 1);
                if (j + 1 < nkv) mbar_wait_cluster(&bar_Kf[(j + 1) % ST], ((j + 1) / ST) & 1);
                for (int t = 0; t < NT; ++t) {
                    mbar_wait_cluster(&bar_Pa[t], j & 1);
                    tc_fence_after();
                    issue_PV(t, j, 0);
                    mbar_wait_cluster(&bar_Pb[t], j & 1);
                    tc_fence_after();
                    issue_PV(t, j, 1);
                    if (t == NT - 1) umma_commit_pair(&bar_Ve[s]);
                    if (j + 1 < nkv) {
                        issue_S(t, j + 1);
                        if (t == NT - 1) umma_commit_pair(&bar_Ke[(j + 1) % ST]);
                    } else {
                        umma_commit_pair(&bar_O[t]);
                    }
                }
            }
        }
    } else if (warp < 4) {
        reg_dec<REGS_CTRL>();
    } else {
        // ================================ softmax warpgroups (both CTAs) ==========================
        reg_inc<REGS_SOFTMAX>();
        const int t = (warp - 4) >> 2;
        pair_softmax_role(t, tmem, sQ, &bar_S[t], map_to_cta(smem_u32(&bar_Pa[t]), 0), map_to_cta(smem_u32(&bar_Pb[t]), 0),
                          &bar_O[t], &tmO, Lout, ldL, nkv, q0, h, b, bh, scale_log2);
    }

    // Tear-down: nobody may leave (or free TMEM) while the other CTA can still read this CTA's shared memory / TMEM
    // through a pair-wide MMA or signal one of its barriers.
    tc_fence_before();
    __syncwarp();
    cluster_sync_all();
    tc_fence_after();
    if (warp == 2) tmem_dealloc_pair<512>(tmem);
}

// Shapes this variant takes (everything else stays on fwd_sm100.cu)
bool fwd_pair_eligible(const FwdArgs& a)
{
    return a.dtype == B200_DTYPE_BF16 && a.d == 128 && !a.causal && a.Nq >= 512 && a.Nq % 512 == 0 && a.Nkv >= 128 &&
           a.Nkv % 128 == 0;
}

int launch_fwd_pair(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    using namespace fpr;
    if (!fwd_pair_eligible(a)) return B200_ERR_UNSUPPORTED;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fwd_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const float scale_log2 = a.scale * 1.4426950408889634f;
    CUtensorMap tmQ, tmK, tmV, tmO;
    if (make_tmap_4d(&tmQ, a.Q, dt, 2, a.B, a.H, a.Nq, D, a.lq.sb, a.lq.sh, a.lq.sn, 128, 64) ||
        make_tmap_4d(&tmK, a.K, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lk.sb, a.lk.sh, a.lk.sn, 64, 64) ||      // half tiles: 64 key rows
        make_tmap_4d(&tmV, a.V, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lv.sb, a.lv.sh, a.lv.sn, 128, 64) ||     // 128 key rows x 64 d-columns
        make_tmap_4d(&tmO, a.O, dt, 2, a.B, a.H, a.Nq, D, a.lo.sb, a.lo.sh, a.lo.sn, 128, 64))
        return B200_ERR_CUDA;
    dim3 grid(a.Nq / (NT * 128), a.H, a.B);                        // even by construction (Nq % 512 == 0): pairs along x
    fwd_pair_kernel<<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmO, a.L, a.ldL, a.Nkv, scale_log2, a.H / a.Hkv, BH_valid);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

}  // namespace b200
