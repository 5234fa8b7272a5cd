// fwd_shs_sm100.cu -- EXPERIMENTAL forward variant (bf16, d = 128), selected with B200_FWD_VARIANT=6:
// ONE score buffer shared by the two query tiles of the CTA, P in its own TMEM columns.
//
// Why: both fwd_sm100.cu (P overwrites S in place, 128-key tiles) and fwd_dbuf_sm100.cu (64-key tiles, two score
// buffers per tile) are limited by the hand-off chain  softmax_t(j) -> PV_t(j) -> S_t(j+1 or j+2) -> softmax_t(...),
// because the buffer the next scores go into is the one PV is still reading P from; the 64-key form pays for its slack
// with operand-fetch-bound N = 64 score MMAs (49 instead of 32 clk per K = 16 step).  The softmax (thread = row) holds a
// whole score row in registers as soon as it has loaded it, so the score buffer is only needed from the MMA to that
// load.  TMEM map (512 columns):
//     O_t  at 128 t      (2 x 128, fp32 accumulators)
//     S    at 256        (128 columns, fp32, shared: S_0(j), S_1(j), S_0(j+1), ... take turns)
//     P_t  at 384 + 64 t (2 x 64, packed bf16)
// and every MMA keeps its efficient shape (M = 128, N = 128).  S_t(j+1) no longer depends on softmax_t(j) or PV_t(j) at
// all -- only on the OTHER tile's softmax having loaded its scores -- so the scores of the next step are ready long
// before a softmax warpgroup finishes the current one; what is left of the chain is  softmax_t(j) -> PV_t(j) ->
// (P_t may be overwritten / O_t may be rescaled), and the softmax computes all of P(j+1) in registers while PV_t(j)
// runs, storing it once bar_R[t] says PV_t(j) has finished reading P_t.
// Roles: warp 0 TMA producer, warp 1 issues the S MMAs, warp 3 issues the PV MMAs (tcgen05.commit tracks the issuing
// thread's MMAs: S-ready / K-stage-free come from warp 1, PV-done / V-stage-free / O-ready from warp 3), warp 2 owns
// TMEM, warps 4-7 / 8-11 are the softmax warpgroups of tile 0 / 1.
// Shapes: everything fwd_sm100.cu takes for bf16 d = 128 (ragged N, causal, strided layouts, grouped K/V heads).
// RESULT (r1(d), profiles/r1d_fwd_shared_s_variant.txt): ran correctly the first time (parity against the generic kernels,
// ragged and causal) and is level with fwd_sm100.cu at C3: 3.33 ms, 3.25 / 3.29 ms with 2 / 4 of 16 exponential pairs on
// the FMA pipe (B200_SHS_POLY), against 3.28-3.32 ms.  The shared buffer is a new serial resource: every score tile
// makes the round trip  issue -> (queue behind the PV MMAs of the in-order tensor pipe) -> 512 clk -> commit -> tcgen05.ld
// before the OTHER tile's score MMA may even be issued, twice per 128 keys; that round trip replaces the chain it
// removed.  (ptxas also spills the 16 speculated P registers per step at the 216-register budget.)  Kept opt-in.
#include "../common.cuh"
#include "../ptx.cuh"
#include "../tmap.h"
#include <stdlib.h>
#include <type_traits>

namespace b200 {

namespace fsh {
constexpr int D = 128, BC = 128, NT = 2, ST = 2;
constexpr int Q_BOX = 128 * 128, Q_TILE = 2 * Q_BOX;         // 32 KB: [128 rows x 128 d] bf16 as 2 boxes of 64 d
constexpr int KV_BOX = BC * 128, KV_TILE = 2 * KV_BOX;       // 32 KB
constexpr int THREADS = 128 + NT * 128;
constexpr int NBARS = NT * 6 + ST * 4;
constexpr int SMEM = 1024 + NT * Q_TILE + 2 * ST * KV_TILE + NBARS * 8 + 16;
constexpr int REGS_CTRL = 64, REGS_SOFTMAX = 216;            // setmaxnreg targets, as fwd_sm100.cu
constexpr float kThreshold = 8.0f;                           // lazy-rescale threshold (log2 units), as fwd_sm100.cu
constexpr uint32_t TM_S = 256, TM_P = 384;                   // TMEM columns (O_t at 128 t)
static_assert(128 * REGS_CTRL + 256 * REGS_SOFTMAX <= 384 * 168, "register pool of the CTA");
static_assert(SMEM <= 227 * 1024, "shared memory budget");
}  // namespace fsh

// One softmax warpgroup (128 threads) serving query tile t.  Thread r owns query row r of the tile.
template <int POLY>
__device__ __forceinline__ void shs_softmax_role(int t, uint32_t tmem, uint8_t* sQ, uint64_t* bar_Sfull, uint64_t* bar_Sfree,
                                                 uint64_t* bar_P, uint64_t* bar_R, uint64_t* bar_O, const CUtensorMap* tmO,
                                                 float* __restrict__ Lout, long long ldL, int Nq, int Nkv, int nkv, int n_plain,
                                                 int causal, int q0, int h, int b, int bh, float scale_log2)
{
    using namespace fsh;
    const int wg_thread = threadIdx.x & 127;                       // == TMEM lane == row within the tile
    const int warp_in_wg = (threadIdx.x >> 5) & 3;
    const uint32_t lane_base = uint32_t(warp_in_wg * 32) << 16;
    const uint32_t tS = tmem + TM_S + lane_base;                   // the shared score buffer
    const uint32_t tP = tmem + TM_P + 64 * t + lane_base;
    const uint32_t tO = tmem + 128 * t + lane_base;
    const int col_end = causal ? min(Nkv, q0 + t * 128 + wg_thread + 1) : Nkv;
    const uint64_t scale2 = f2_pack(scale_log2, scale_log2);

    float m = -INFINITY, l = 0.0f;                                 // m in log2 units (already scaled)

    auto step = [&](auto masked_tag, const int j) {
        constexpr bool kMasked = decltype(masked_tag)::value;
        mbar_wait(bar_Sfull, j & 1);
        tc_fence_after();
        uint32_t s[BC];
#pragma unroll
        for (int c = 0; c < BC / 32; ++c) B200_TMEM_LD_32(tS + 32 * c, s + 32 * c);
        tmem_wait_ld();
        // the row is in registers: hand the score buffer to the other tile's next S MMA
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(bar_Sfree);
        if constexpr (kMasked) {
            const int lim = col_end - j * BC;
#pragma unroll
            for (int i = 0; i < BC; ++i) if (i >= lim) s[i] = 0xff800000u;       // -inf
        }
        // P for the 32 scores of chunk c with the running max m_use, kept in registers (packed bf16x2)
        uint32_t p[BC / 2];
        uint64_t acc[4];
        auto exp_chunk = [&](const int c, const float m_use) {
            const uint64_t neg_m2 = f2_pack(-m_use, -m_use);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const uint64_t x = f2_fma(f2_pack(__uint_as_float(s[32 * c + 2 * i]), __uint_as_float(s[32 * c + 2 * i + 1])), scale2, neg_m2);
                float p0, p1;
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
                p[16 * c + i] = pack_bf16x2(p0, p1);
            }
        };
        // speculate the first chunk on the running max while the row max of this tile is reduced (fwd_sm100.cu)
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
        bool pv_done = (j == 0);                                   // "PV_t(j-1) has completed" observed (warp-uniform)
        if (__any_sync(0xffffffffu, need)) {
            if (need) { alpha = fast_exp2(m - m_new); m = m_new; }
            if (j > 0) {
                // O_t is ours between the completion of PV_t(j-1) and the issue of PV_t(j) (which waits for bar_P below)
                mbar_wait(bar_R, (j - 1) & 1);
                tc_fence_after();
                pv_done = true;
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
            exp_chunk(0, m);                                       // redo the speculated chunk with the new max
        }
#pragma unroll
        for (int c = 1; c < BC / 32; ++c) exp_chunk(c, m);
        float a0, a1, a2, a3;
        f2_unpack(f2_add(acc[0], acc[1]), a0, a1);
        f2_unpack(f2_add(acc[2], acc[3]), a2, a3);
        l = fmaf(l, alpha, (a0 + a1) + (a2 + a3));
        // P_t is still being read by PV_t(j-1) until bar_R says otherwise; all of P(j) was computed meanwhile
        if (!pv_done) {
            mbar_wait(bar_R, (j - 1) & 1);
            tc_fence_after();
        }
#pragma unroll
        for (int c = 0; c < BC / 32; ++c) B200_TMEM_ST_16(tP + 16 * c, p + 16 * c);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(bar_P);
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

template <int POLY>
__global__ void __launch_bounds__(fsh::THREADS, 1)
fwd_shs_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
               float* __restrict__ Lout, long long ldL, int Nq, int Nkv, float scale_log2, int causal, int G, int BH)
{
    using namespace fsh;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + NT * Q_TILE;
    uint8_t* sV = sK + ST * KV_TILE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ST * KV_TILE);
    uint64_t* bar_Q = bars;                 // [NT]  Q tile landed
    uint64_t* bar_Sfull = bar_Q + NT;       // [NT]  S_t(j) complete in the shared score buffer
    uint64_t* bar_Sfree = bar_Sfull + NT;   // [NT]  softmax_t has loaded S_t(j) (one arrival per warp): the buffer is free
    uint64_t* bar_P = bar_Sfree + NT;       // [NT]  P_t(j) written (one arrival per softmax warp)
    uint64_t* bar_R = bar_P + NT;           // [NT]  PV_t(j) complete: P_t may be overwritten, O_t may be rescaled
    uint64_t* bar_O = bar_R + NT;           // [NT]  last PV of tile t complete
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
    auto nkv_of = [&](int t) { return causal ? min(nkv_all, (q0 + t * 128 + 127) / BC + 1) : nkv_all; };
    const int nkv = nkv_of(ntiles - 1), nkv_first = nkv_of(0);

    if (threadIdx.x == 0) {
        for (int t = 0; t < NT; ++t) {
            mbar_init(&bar_Q[t], 1); mbar_init(&bar_Sfull[t], 1); mbar_init(&bar_Sfree[t], 4); mbar_init(&bar_P[t], 4);
            mbar_init(&bar_R[t], 1); mbar_init(&bar_O[t], 1);
        }
        for (int s = 0; s < ST; ++s) { mbar_init(&bar_Kf[s], 1); mbar_init(&bar_Ke[s], 1); mbar_init(&bar_Vf[s], 1); mbar_init(&bar_Ve[s], 1); }
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
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
            auto load_K = [&](int j) {
                const int s = j % ST;
                mbar_wait(&bar_Ke[s], ((j / ST) & 1) ^ 1);
                mbar_expect_tx(&bar_Kf[s], KV_TILE);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d(sK + s * KV_TILE + bx * KV_BOX, &tmK, &bar_Kf[s], bx * 64, j * BC, hk, b, kEvictLast);
            };
            auto load_V = [&](int j) {
                const int s = j % ST;
                mbar_wait(&bar_Ve[s], ((j / ST) & 1) ^ 1);
                mbar_expect_tx(&bar_Vf[s], KV_TILE);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d(sV + s * KV_TILE + bx * KV_BOX, &tmV, &bar_Vf[s], bx * 64, j * BC, hk, b, kEvictLast);
            };
            // the score MMAs run about one K/V tile ahead of the PV MMAs: K(j+1) before V(j)
            load_K(0);
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) load_K(j + 1);
                load_V(j);
            }
        }
    } else if (warp == 1) {
        // ================================ S issuer: S_t(j) = Q_t K_j^T into the shared buffer ===========
        reg_dec<REGS_CTRL>();
        if (elect_one()) {
            constexpr uint32_t idesc_S = make_idesc(kFmtBF16, 128, BC, false, false);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK);
            int prev_t = -1, prev_j = 0;                        // the S MMA whose scores are in the buffer now
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait(&bar_Kf[s], (j / ST) & 1);
                for (int t = 0; t < ntiles; ++t) {
                    const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;   // (NT <= 2)
                    if (j >= nkv_t) continue;
                    if (j == 0) mbar_wait(&bar_Q[t], 0);
                    if (prev_t >= 0) mbar_wait(&bar_Sfree[prev_t], prev_j & 1);
                    tc_fence_after();
                    const uint64_t qd = make_sdesc(q_base + t * Q_TILE, 16, 1024, 2);
                    const uint64_t kd = make_sdesc(k_base + s * KV_TILE, 16, 1024, 2);
#pragma unroll
                    for (int kk = 0; kk < D / 16; ++kk) {
                        const uint32_t inbox = (kk & 3) * 32;
                        umma_f16_ss(tmem + TM_S, sdesc_advance(qd, (kk >> 2) * Q_BOX + inbox),
                                    sdesc_advance(kd, (kk >> 2) * KV_BOX + inbox), idesc_S, kk > 0 ? 1u : 0u);
                    }
                    umma_commit(&bar_Sfull[t]);
                    prev_t = t; prev_j = j;
                }
                umma_commit(&bar_Ke[s]);
            }
        }
    } else if (warp == 3) {
        // ================================ PV issuer: O_t += P_t V_j =====================================
        reg_dec<REGS_CTRL>();
        if (elect_one()) {
            constexpr uint32_t idesc_PV = make_idesc(kFmtBF16, 128, D, false, true);
            const uint32_t v_base = smem_u32(sV);
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait(&bar_Vf[s], (j / ST) & 1);
                for (int t = 0; t < ntiles; ++t) {
                    const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;
                    if (j >= nkv_t) continue;
                    mbar_wait(&bar_P[t], j & 1);
                    tc_fence_after();
                    const uint64_t vd = make_sdesc(v_base + s * KV_TILE, KV_BOX, 1024, 2);
#pragma unroll
                    for (int kk = 0; kk < BC / 16; ++kk)
                        umma_f16_ts(tmem + 128 * t, tmem + TM_P + 64 * t + 8 * kk, sdesc_advance(vd, kk * 16 * 128), idesc_PV,
                                    (j > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&bar_R[t]);
                    if (j + 1 == nkv_t) umma_commit(&bar_O[t]);
                }
                umma_commit(&bar_Ve[s]);
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
            const int n_plain = causal ? min(nkv_t, min((q0 + t * 128 + 1) / BC, Nkv / BC)) : nkv_t - 1;
            shs_softmax_role<POLY>(t, tmem, sQ, &bar_Sfull[t], &bar_Sfree[t], &bar_P[t], &bar_R[t], &bar_O[t], &tmO, Lout, ldL,
                                   Nq, Nkv, nkv_t, n_plain, causal, q0, h, b, bh, scale_log2);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

template <int POLY>
static int launch_fwd_shs_t(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    using namespace fsh;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fwd_shs_kernel<POLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const float scale_log2 = a.scale * 1.4426950408889634f;
    CUtensorMap tmQ, tmK, tmV, tmO;
    if (make_tmap_4d(&tmQ, a.Q, dt, 2, a.B, a.H, a.Nq, D, a.lq.sb, a.lq.sh, a.lq.sn, 128, 64) ||
        make_tmap_4d(&tmK, a.K, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lk.sb, a.lk.sh, a.lk.sn, BC, 64) ||
        make_tmap_4d(&tmV, a.V, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lv.sb, a.lv.sh, a.lv.sn, BC, 64) ||
        make_tmap_4d(&tmO, a.O, dt, 2, a.B, a.H, a.Nq, D, a.lo.sb, a.lo.sh, a.lo.sn, 128, 64))
        return B200_ERR_CUDA;
    dim3 grid((a.Nq + NT * 128 - 1) / (NT * 128), a.H, a.B);
    fwd_shs_kernel<POLY><<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmO, a.L, a.ldL, a.Nq, a.Nkv, scale_log2, a.causal ? 1 : 0,
                                                      a.H / a.Hkv, BH_valid);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

int launch_fwd_shs(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    if (a.dtype != B200_DTYPE_BF16 || a.d != fsh::D) return B200_ERR_UNSUPPORTED;
    static int poly = -1;                                // tuning hook: B200_SHS_POLY = pairs of every 16 on the FMA pipe
    if (poly < 0) { const char* e = getenv("B200_SHS_POLY"); poly = e ? atoi(e) : 0; }
    switch (poly) {
        case 0: return launch_fwd_shs_t<0>(a, BH_valid, st);
        case 2: return launch_fwd_shs_t<2>(a, BH_valid, st);
        case 4: return launch_fwd_shs_t<4>(a, BH_valid, st);
        case 6: return launch_fwd_shs_t<6>(a, BH_valid, st);
        default: return B200_ERR_UNSUPPORTED;
    }
}

}  // namespace b200
