// fwd_split_sm100.cu -- EXPERIMENTAL forward variant (bf16, d = 128), selected with B200_FWD_VARIANT=2.
//
// Same algorithm, tiles and roles as fwd_sm100.cu (two 128-row query tiles per CTA, 128-row K/V tiles, one
// softmax warpgroup per tile, thread = row).  What changes is the per-tile dependency chain
//     softmax(j) -> PV(j) -> S(j+1) -> softmax(j+1)
// which in fwd_sm100.cu is strictly serial because P overwrites S in place.  Here
//   * the softmax holds all 128 scores of its row in registers, so P (64 TMEM columns of packed bf16) is written
//     into columns 64..127 of the tile's S buffer and columns 0..63 are free as soon as they have been loaded;
//   * S(j+1) is issued as two N = 64 MMAs: the LOW key half (columns 0..63) right after that load -- it runs under
//     the rest of softmax(j) -- and the HIGH half (columns 64..127) after PV(j);
//   * the softmax of step j+1 loads the low half first and computes its exponentials with the running (stale) max
//     while PV(j) and S_hi(j+1) are still on the tensor pipe; the lazy-rescale rule (fwd_sm100.cu) makes that
//     speculation right almost always, otherwise the packed low half is rescaled by alpha.
// Non-causal and causal (square), any N; contiguous or strided layouts; passes the forward parity tests.
// RESULT (r1(c)): correct but 14 % SLOWER than fwd_sm100.cu at C3 (3.77 vs 3.29 ms) -- the N = 64 score
// halves read the Q tile twice and are operand-fetch-bound, and the in-order issuing thread has three hand-offs per
// tile step.  Kept for the record and for further experiments; never selected by default.
#include "../common.cuh"
#include "../ptx.cuh"
#include "../tmap.h"
#include <stdlib.h>
#include <type_traits>

namespace b200 {

namespace fsp {
constexpr int D = 128, BC = 128, NT = 2, ST = 2;
constexpr int Q_BOX = 128 * 128, Q_TILE = 2 * Q_BOX;          // 32 KB: [128 rows x 128 d] bf16 as 2 boxes of 64 d
constexpr int KV_BOX = BC * 128, KV_TILE = 2 * KV_BOX;        // 32 KB
constexpr int THREADS = 128 + NT * 128;
constexpr int NBARS = NT * 7 + ST * 4;
constexpr int SMEM = 1024 + NT * Q_TILE + 2 * ST * KV_TILE + NBARS * 8 + 16;
constexpr int REGS_CTRL = 64, REGS_SOFTMAX = 216;
constexpr float kThreshold = 8.0f;                            // lazy-rescale threshold (log2 units), as fwd_sm100.cu
static_assert(SMEM <= 227 * 1024, "shared memory budget");
}  // namespace fsp

// P for 64 scores (registers s, already masked) with running max m_use: packed bf16x2 into p[32], row-sum into acc[4]
__device__ __forceinline__ void fsp_exp64(const uint32_t* s, float scale_log2, float m_use, uint32_t* p, uint64_t* acc, bool first)
{
    const uint64_t scale2 = f2_pack(scale_log2, scale_log2), neg_m2 = f2_pack(-m_use, -m_use);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const uint64_t x = f2_fma(f2_pack(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, neg_m2);
        float x0, x1;
        f2_unpack(x, x0, x1);
        const float p0 = fast_exp2(x0), p1 = fast_exp2(x1);
        if (first && i < 4) acc[i] = f2_pack(p0, p1); else acc[i & 3] = f2_add(acc[i & 3], f2_pack(p0, p1));
        p[i] = pack_bf16x2(p0, p1);
    }
}
__device__ __forceinline__ float fsp_max64(const uint32_t* s)
{
    float mq[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        mq[q] = fmaxf(__uint_as_float(s[16 * q]), __uint_as_float(s[16 * q + 1]));
#pragma unroll
        for (int i = 2; i < 16; i += 2) mq[q] = fmaxf(mq[q], fmaxf(__uint_as_float(s[16 * q + i]), __uint_as_float(s[16 * q + i + 1])));
    }
    return fmaxf(fmaxf(mq[0], mq[1]), fmaxf(mq[2], mq[3]));
}

// One softmax warpgroup (128 threads) serving query tile t; thread r owns query row r of the tile.
__device__ __forceinline__ void fsp_softmax(int t, uint32_t tmem, uint8_t* sQ, uint64_t* bar_Slo, uint64_t* bar_Shi, uint64_t* bar_L,
                                            uint64_t* bar_Pa, uint64_t* bar_Pb, uint64_t* bar_O, const CUtensorMap* tmO,
                                            float* __restrict__ Lout, long long ldL, int Nq, int Nkv, int nkv, int n_plain,
                                            int causal, int q0, int h, int b, int bh, float scale_log2)
{
    using namespace fsp;
    const int r = threadIdx.x & 127;
    const uint32_t lane_base = uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16;
    const uint32_t tS = tmem + 128 * t + lane_base;                // S_t: low key half at +0, high half at +64
    const uint32_t tP = tS + 64;                                   // P_t (packed bf16): aliases the high half of S_t
    const uint32_t tO = tmem + 256 + 128 * t + lane_base;
    const int col_end = causal ? min(Nkv, q0 + t * 128 + r + 1) : Nkv;   // first key column this row must not see
    const bool lane0 = (threadIdx.x & 31) == 0;
    float m = -INFINITY, l = 0.0f;                                 // m in log2 units (already scaled)

    // kSpec: the low half is exponentiated with the running max before the high half (and so the row max) is known
    auto step = [&](auto spec_tag, auto masked_tag, const int j) {
        constexpr bool kSpec = decltype(spec_tag)::value, kMasked = decltype(masked_tag)::value;
        const int lim = col_end - j * BC;
        uint32_t p_lo[32];
        uint64_t acc[4];
        float mx;
        uint32_t lo[64];
        mbar_wait(bar_Slo, j & 1);
        tc_fence_after();
        B200_TMEM_LD_32(tS, lo);
        B200_TMEM_LD_32(tS + 32, lo + 32);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(bar_L);                             // columns 0..63 are free: S_lo(j+1) may be issued
        if constexpr (kMasked) {
#pragma unroll
            for (int i = 0; i < 64; ++i) if (i >= lim) lo[i] = 0xff800000u;      // -inf
        }
        mx = fsp_max64(lo);
        if constexpr (kSpec) fsp_exp64(lo, scale_log2, m, p_lo, acc, true);

        uint32_t hi[64];
        mbar_wait(bar_Shi, j & 1);
        tc_fence_after();
        B200_TMEM_LD_32(tS + 64, hi);
        B200_TMEM_LD_32(tS + 96, hi + 32);
        tmem_wait_ld();
        if constexpr (kMasked) {
#pragma unroll
            for (int i = 0; i < 64; ++i) if (64 + i >= lim) hi[i] = 0xff800000u;
        }
        mx = fmaxf(mx, fsp_max64(hi));
        const float m_new = fmaxf(m, mx * scale_log2);
        const bool need = (m_new - m) > kThreshold;
        float alpha = 1.0f;
        if (__any_sync(0xffffffffu, need)) {
            if (need) { alpha = fast_exp2(m - m_new); m = m_new; }
            if constexpr (kSpec) {
                // PV_t(j-1) is complete (S_hi(j) was issued after it) and PV_t(j) not yet issued: O_t is ours
#pragma unroll
                for (int c = 0; c < D / 32; ++c) {
                    uint32_t o[32];
                    B200_TMEM_LD_32(tO + 32 * c, o);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    B200_TMEM_ST_32(tO + 32 * c, o);
                }
                // the speculated half carries the old max: bring it to the new one (rare path; bf16 double rounding)
                if (need) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        p_lo[i] = pack_bf16x2(__uint_as_float(p_lo[i] << 16) * alpha, __uint_as_float(p_lo[i] & 0xffff0000u) * alpha);
                    const uint64_t a2 = f2_pack(alpha, alpha);
#pragma unroll
                    for (int k = 0; k < 4; ++k) acc[k] = f2_mul(acc[k], a2);
                }
            }
        }
        if constexpr (!kSpec) fsp_exp64(lo, scale_log2, m, p_lo, acc, true);     // first step: no finite max to speculate with
        B200_TMEM_ST_16(tP, p_lo);
        B200_TMEM_ST_16(tP + 16, p_lo + 16);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(bar_Pa);                            // first half of P is in TMEM: PV can start
        fsp_exp64(hi, scale_log2, m, hi, acc, false);              // (in place: hi[i] is read before hi[i/2] is written)
        B200_TMEM_ST_16(tP + 32, hi);
        B200_TMEM_ST_16(tP + 48, hi + 16);
        float a0, a1, a2, a3;
        f2_unpack(f2_add(acc[0], acc[1]), a0, a1);
        f2_unpack(f2_add(acc[2], acc[3]), a2, a3);
        l = fmaf(l, alpha, (a0 + a1) + (a2 + a3));
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane0) mbar_arrive(bar_Pb);
    };

    if (n_plain > 0) step(std::false_type{}, std::false_type{}, 0); else step(std::false_type{}, std::true_type{}, 0);
    for (int j = 1; j < n_plain; ++j) step(std::true_type{}, std::false_type{}, j);
    for (int j = (n_plain > 1 ? n_plain : 1); j < nkv; ++j) step(std::true_type{}, std::true_type{}, j);

    // ---- epilogue: O_t / l -> smem (reusing the Q tile, 128B-swizzled) -> TMA store; LSE -> global
    mbar_wait(bar_O, 0);
    tc_fence_after();
    const float inv_l = 1.0f / l;
    uint8_t* sO = sQ + t * Q_TILE;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        B200_TMEM_LD_32(tO + 32 * c, o);
        tmem_wait_ld();
        uint8_t* rowp = sO + (c >> 1) * Q_BOX + r * 128;           // 32 bf16 = 64 B = 4 chunks: box c/2, chunks (c&1)*4 .. +3
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o[8 * q]) * inv_l, __uint_as_float(o[8 * q + 1]) * inv_l);
            v.y = pack_bf16x2(__uint_as_float(o[8 * q + 2]) * inv_l, __uint_as_float(o[8 * q + 3]) * inv_l);
            v.z = pack_bf16x2(__uint_as_float(o[8 * q + 4]) * inv_l, __uint_as_float(o[8 * q + 5]) * inv_l);
            v.w = pack_bf16x2(__uint_as_float(o[8 * q + 6]) * inv_l, __uint_as_float(o[8 * q + 7]) * inv_l);
            *reinterpret_cast<uint4*>(rowp + ((((c & 1) * 4 + q) ^ (r & 7)) * 16)) = v;
        }
    }
    fence_proxy_async_smem();
    named_bar_sync(1 + t, 128);
    if (r == 0) {
        tma_store_4d(tmO, sO, 0, q0 + t * 128, h, b);
        tma_store_4d(tmO, sO + Q_BOX, 64, q0 + t * 128, h, b);
        tma_store_commit();
        tma_store_wait<0>();
    }
    const int row = q0 + t * 128 + r;
    if (Lout != nullptr && row < Nq) Lout[(size_t)bh * ldL + row] = (m + log2f(l)) * 0.6931471805599453f;
}

__global__ void __launch_bounds__(fsp::THREADS, 1)
fwd_split_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                 float* __restrict__ Lout, long long ldL, int Nq, int Nkv, float scale_log2, int causal, int G, int BH)
{
    using namespace fsp;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + NT * Q_TILE;
    uint8_t* sV = sK + ST * KV_TILE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ST * KV_TILE);
    uint64_t* bar_Q = bars;                 // [NT]  Q tile landed
    uint64_t* bar_Slo = bar_Q + NT;         // [NT]  low key half of S_t(j) complete in TMEM
    uint64_t* bar_Shi = bar_Slo + NT;       // [NT]  high half complete (implies PV_t(j-1) complete)
    uint64_t* bar_L = bar_Shi + NT;         // [NT]  softmax has loaded the low half (one arrival per softmax warp)
    uint64_t* bar_Pa = bar_L + NT;          // [NT]  first half of P_t(j) written
    uint64_t* bar_Pb = bar_Pa + NT;         // [NT]  all of P_t(j) written
    uint64_t* bar_O = bar_Pb + NT;          // [NT]  last PV of tile t complete
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
            mbar_init(&bar_Q[t], 1); mbar_init(&bar_Slo[t], 1); mbar_init(&bar_Shi[t], 1); mbar_init(&bar_L[t], 4);
            mbar_init(&bar_Pa[t], 4); mbar_init(&bar_Pb[t], 4); mbar_init(&bar_O[t], 1);
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
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                const uint32_t ph = (j / ST) & 1;
                mbar_wait(&bar_Ke[s], ph ^ 1);
                mbar_expect_tx(&bar_Kf[s], KV_TILE);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d(sK + s * KV_TILE + bx * KV_BOX, &tmK, &bar_Kf[s], bx * 64, j * BC, hk, b, kEvictLast);
                mbar_wait(&bar_Ve[s], ph ^ 1);
                mbar_expect_tx(&bar_Vf[s], KV_TILE);
                for (int bx = 0; bx < 2; ++bx)
                    tma_load_4d(sV + s * KV_TILE + bx * KV_BOX, &tmV, &bar_Vf[s], bx * 64, j * BC, hk, b, kEvictLast);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        reg_dec<REGS_CTRL>();
        if (elect_one()) {
            constexpr uint32_t idesc_S = make_idesc(kFmtBF16, 128, 64, false, false);     // one key half: N = 64
            constexpr uint32_t idesc_PV = make_idesc(kFmtBF16, 128, D, false, true);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV);

            // S_t(j), key half `half` (rows 64 half .. +63 of the K tile -> TMEM columns 64 half .. +63 of S_t)
            auto issue_S = [&](int t, int j, int half) {
                const int s = j % ST;
                const uint64_t qd = make_sdesc(q_base + t * Q_TILE, 16, 1024, 2);
                const uint64_t kd = make_sdesc(k_base + s * KV_TILE + half * 64 * 128, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk) {
                    const uint32_t inbox = (kk & 3) * 32;
                    umma_f16_ss(tmem + 128 * t + 64 * half, sdesc_advance(qd, (kk >> 2) * Q_BOX + inbox),
                                sdesc_advance(kd, (kk >> 2) * KV_BOX + inbox), idesc_S, kk > 0 ? 1u : 0u);
                }
                umma_commit(half == 0 ? &bar_Slo[t] : &bar_Shi[t]);
            };
            auto issue_PV = [&](int t, int j, int half) {       // half 0 / 1 = first / second 64 keys of P
                const int s = j % ST;
                const uint64_t vd = make_sdesc(v_base + s * KV_TILE, KV_BOX, 1024, 2);
#pragma unroll
                for (int k2 = 0; k2 < 4; ++k2) {
                    const int kk = half * 4 + k2;
                    umma_f16_ts(tmem + 256 + 128 * t, tmem + 128 * t + 64 + 8 * kk, sdesc_advance(vd, kk * 16 * 128), idesc_PV,
                                (j > 0 || kk > 0) ? 1u : 0u);
                }
            };

            mbar_wait(&bar_Kf[0], 0);
            for (int t = 0; t < ntiles; ++t) {
                mbar_wait(&bar_Q[t], 0);
                tc_fence_after();
                issue_S(t, 0, 0);
                issue_S(t, 0, 1);
            }
            umma_commit(&bar_Ke[0]);                            // K(0) is free once both halves of every S_t(0) are done
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait(&bar_Vf[s], (j / ST) & 1);
                if (j + 1 < nkv) mbar_wait(&bar_Kf[(j + 1) % ST], ((j + 1) / ST) & 1);
                for (int t = 0; t < ntiles; ++t) {
                    const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;
                    if (j >= nkv_t) continue;
                    const bool more = j + 1 < nkv_t;
                    if (more) {                                 // low half of S_t(j+1) under the softmax of step j
                        mbar_wait(&bar_L[t], j & 1);
                        tc_fence_after();
                        issue_S(t, j + 1, 0);
                    }
                    mbar_wait(&bar_Pa[t], j & 1);
                    tc_fence_after();
                    issue_PV(t, j, 0);
                    mbar_wait(&bar_Pb[t], j & 1);
                    tc_fence_after();
                    issue_PV(t, j, 1);
                    if (t == ntiles - 1) umma_commit(&bar_Ve[s]);
                    if (more) {
                        issue_S(t, j + 1, 1);                   // high half: its columns held P_t(j) until PV_t(j) above
                        if (t == ntiles - 1) umma_commit(&bar_Ke[(j + 1) % ST]);
                    } else {
                        umma_commit(&bar_O[t]);
                    }
                }
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
            fsp_softmax(t, tmem, sQ, &bar_Slo[t], &bar_Shi[t], &bar_L[t], &bar_Pa[t], &bar_Pb[t], &bar_O[t], &tmO, Lout, ldL,
                        Nq, Nkv, nkv_t, n_plain, causal, q0, h, b, bh, scale_log2);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

int launch_fwd_split(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    using namespace fsp;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fwd_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUtensorMap tmQ, tmK, tmV, tmO;
    if (make_tmap_4d(&tmQ, a.Q, dt, 2, a.B, a.H, a.Nq, D, a.lq.sb, a.lq.sh, a.lq.sn, 128, 64) ||
        make_tmap_4d(&tmK, a.K, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lk.sb, a.lk.sh, a.lk.sn, BC, 64) ||
        make_tmap_4d(&tmV, a.V, dt, 2, a.B, a.Hkv, a.Nkv, D, a.lv.sb, a.lv.sh, a.lv.sn, BC, 64) ||
        make_tmap_4d(&tmO, a.O, dt, 2, a.B, a.H, a.Nq, D, a.lo.sb, a.lo.sh, a.lo.sn, 128, 64))
        return B200_ERR_CUDA;
    dim3 grid((a.Nq + NT * 128 - 1) / (NT * 128), a.H, a.B);
    fwd_split_kernel<<<grid, THREADS, SMEM, st>>>(tmQ, tmK, tmV, tmO, a.L, a.ldL, a.Nq, a.Nkv, a.scale * 1.4426950408889634f,
                                                  a.causal ? 1 : 0, a.H / a.Hkv, BH_valid);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

}  // namespace b200
