// bwd_tf32_sm100.cu -- FlashAttention-2 backward for sm_100a with fp32 I/O (the reference's own type), d = 64:
// kind::tf32 tcgen05 MMAs, every accumulator in TMEM, dQ reduced straight into the caller's fp32 dQ.
//
// This is what sits behind the reference's signature flash_attention_2_backward(const float* ...)
// (src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:249-299; its kernel :47-246 is dQ-stationary
// scalar code with shared + global atomics and only takes head_dim <= 64 -- d = 128 needs 64 KB of dynamic shared
// memory it never opts into).  Same algorithm and roles as bwd_sm100.cu (dK/dV-stationary, everything transposed so a
// thread owns a kv row, five MMAs per iteration, bulk add-reductions for dQ); what differs is forced by tf32:
//   * a [rows x 64] fp32 tile has the bytes of a [rows x 128] bf16 tile (two 128-byte boxes), so the K/V tile stays
//     128 rows and the q tile shrinks to 32 rows (BM = 32, two stages);
//   * MN-major tf32 operands only exist in the 32-byte-atom flavour of the 128-byte swizzle
//     (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, descriptor layout 1, SBO 512) and K-major operands only in the plain one
//     (tools/umma_probe.cu cases 16-27: the K-major x ATOM_32B combinations fault) -- so every tile that is read both
//     ways (Q: M1 K-major / M4 MN-major; dO: M2 / M3; K: M1 / M5) is brought in TWICE by TMA, once per swizzle;
//   * P^T and dS^T are written back to TMEM as tf32 (cvt.rna) over their fp32 sources -- same width, in place;
//   * M5 dQ^T = K^T dS wants M = 128 but d is 64: the A descriptor simply runs on into the V tile, TMEM lanes 64..127 of
//     dQ^T hold garbage that nothing reads;
//   * outputs are fp32: dK / dV go out through fp32 TMA stores, dQ needs no accumulator + cast (the bulk reductions
//     land in the caller's zeroed dQ).
//   M1  S^T  = K Q_i^T      SS  [128 kv x 32 q]  K = 64 (8 steps)     -> TMEM S[b]
//   M2  dP^T = V dO_i^T     SS  [128 kv x 32 q]                       -> TMEM dP
//   M3  dV  += P^T dO_i     TS  [128 kv x 64 d]  K = 32 q (4 steps)   B = dO' (MN-major copy)
//   M4  dK  += dS^T Q_i     TS                                        B = Q'
//   M5  dQ_i^T = K^T dS     SS  [128 (64 real) d x 32 q]  K = 128 kv (16 steps)   A = K' (MN-major copy), B = dS^T (smem)
// TMEM columns (256 allocated): S[0] 0..31, S[1] 32..63, dP 64..95, dQ^T 96..127, dV 128..191, dK 192..255.
// Warps as in bwd_sm100.cu: 0 TMA, 1 MMA, 2 TMEM alloc, 4-7 / 8-11 compute (q columns 0-15 / 16-31), 12-15 dQ drain.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"
#include <stdlib.h>
#include <algorithm>

// bring-up hooks (side libraries only, B200_BWD_DBG): 1 no dQ reduction, 4 no M5, 16 no M1/M2, 64 no M3/M4, 256 no dK/dV store,
// 512 no statistics loads, 1024 no MN-major copies (Q', dO', K')
#ifdef B200_EXPERIMENTS
#define BWT_DBG(bit) ((dbg & (bit)) != 0)
#else
#define BWT_DBG(bit) false
#endif

namespace b200 {

namespace bwt {
constexpr int D = 64, BN = 128, BM = 32, NS = 2;
constexpr int KV_BOX = BN * 128, KV_TILE = 2 * KV_BOX;        // 32 KB: [128 rows x 64 fp32] as 2 boxes of 32 fp32
constexpr int Q_BOX = BM * 128, Q_TILE = 2 * Q_BOX;           // 8 KB
constexpr int DS_TILE = BN * 128;                             // 16 KB: dS^T [128 kv rows x 32 q] fp32
constexpr int STAT_BYTES = BM * 4;
constexpr int STG_BYTES = BM * D * 4;                         // 8 KB: dQ staging [32 q][64 d] fp32
constexpr int THREADS = 512;
constexpr int OFF_K = 0, OFF_KM = OFF_K + KV_TILE, OFF_V = OFF_KM + KV_TILE;      // (V right behind K': M5 reads on into it)
constexpr int OFF_Q = OFF_V + KV_TILE, OFF_QM = OFF_Q + NS * Q_TILE, OFF_DO = OFF_QM + NS * Q_TILE, OFF_DOM = OFF_DO + NS * Q_TILE;
constexpr int OFF_DS = OFF_DOM + NS * Q_TILE, OFF_STAT = OFF_DS + 2 * DS_TILE, OFF_STG = OFF_STAT + NS * 2 * STAT_BYTES;
constexpr int OFF_BAR = OFF_STG + 2 * STG_BYTES;
constexpr int NBARS = 1 + 2 * NS + 2 + 1 + 2 + 1 + 2 + 1;
constexpr int SMEM = 1024 + OFF_BAR + NBARS * 8 + 16;
static_assert(SMEM <= 227 * 1024, "tf32 backward shared memory budget");
constexpr uint32_t TM_S = 0, TM_DP = 64, TM_DQ = 96, TM_DV = 128, TM_DK = 192;
}  // namespace bwt

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void bwt_bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bwt_lds_f32x4(float* dst, uint32_t saddr) {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(dst[0]), "=f"(dst[1]), "=f"(dst[2]), "=f"(dst[3]) : "r"(saddr));
}
__device__ __forceinline__ void bwt_bulk_reduce_add_f32(float* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                 ::"l"(reinterpret_cast<uint64_t>(gdst)), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(bwt::THREADS, 1)
bwd_tf32_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmQm,
                const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmKm,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                const __grid_constant__ CUtensorMap tmdOm, const __grid_constant__ CUtensorMap tmdK,
                const __grid_constant__ CUtensorMap tmdV, float* __restrict__ dQ, const float* __restrict__ L2p,
                const float* __restrict__ Dp, int Nq, int stat_stride, float scale_log2, float scale, int BH, int dbg, int causal)
{
    using namespace bwt;
    (void)dbg;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* bar_KV = bars;
    uint64_t* bar_in_full = bar_KV + 1;          // [NS]
    uint64_t* bar_in_empty = bar_in_full + NS;   // [NS]
    uint64_t* bar_S = bar_in_empty + NS;         // [2]
    uint64_t* bar_dP = bar_S + 2;
    uint64_t* bar_P = bar_dP + 1;                // [2]
    uint64_t* bar_dS = bar_P + 2;
    uint64_t* bar_dQ_full = bar_dS + 1;
    uint64_t* bar_dQ_empty = bar_dQ_full + 1;
    uint64_t* bar_done = bar_dQ_empty + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_done + 1);

    const int warp = threadIdx.x >> 5;
    const int h = blockIdx.y, b = blockIdx.z, kv0 = blockIdx.x * BN;
    const size_t bh = (size_t)b * gridDim.y + h;
    if (bh >= (size_t)BH) return;
    const int n_q = (Nq + BM - 1) / BM;
    // causal = 1 + (Nkv - Nq) (bottom-right aligned mask, as in bwd_sm100.cu) or 0: the q tiles before the diagonal see none
    // of this CTA's keys and are skipped; the tiles that cross it are masked in the compute warps
    const int qt_min = causal ? min(max(kv0 - (causal - 1), 0) / BM, n_q - 1) : 0;
    const int n_it = n_q - qt_min;
    const int i_off = (int)(((long long)blockIdx.x * n_q) / gridDim.x) % n_it;    // staggered start: see bwd_sm100.cu

    if (threadIdx.x == 0) {
        mbar_init(bar_KV, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&bar_in_full[s], 1); mbar_init(&bar_in_empty[s], 1); }
        mbar_init(&bar_S[0], 1); mbar_init(&bar_S[1], 1); mbar_init(bar_dP, 1);
        mbar_init(&bar_P[0], 8); mbar_init(&bar_P[1], 8); mbar_init(bar_dS, 8);
        mbar_init(bar_dQ_full, 1); mbar_init(bar_dQ_empty, 4); mbar_init(bar_done, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmQm); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmKm);
        tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmdOm);
    }
    if (warp == 2) { tmem_alloc<256>(tmem_slot); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (elect_one()) {
            mbar_expect_tx(bar_KV, (BWT_DBG(1024) ? 2 : 3) * KV_TILE);
            for (int bx = 0; bx < 2; ++bx) {
                tma_load_4d(smem + OFF_K + bx * KV_BOX, &tmK, bar_KV, bx * 32, kv0, h, b, kEvictFirst);
                if (!BWT_DBG(1024)) tma_load_4d(smem + OFF_KM + bx * KV_BOX, &tmKm, bar_KV, bx * 32, kv0, h, b, kEvictFirst);
                tma_load_4d(smem + OFF_V + bx * KV_BOX, &tmV, bar_KV, bx * 32, kv0, h, b, kEvictFirst);
            }
            const float* Lrow = L2p + bh * stat_stride;
            const float* Drow = Dp + bh * stat_stride;
            for (int i = 0; i < n_it; ++i) {
                const int s = i % NS;
                int qt = i + i_off;
                if (qt >= n_it) qt -= n_it;
                qt += qt_min;
                mbar_wait(&bar_in_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&bar_in_full[s], (BWT_DBG(1024) ? 2 : 4) * Q_TILE + (BWT_DBG(512) ? 0 : 2 * STAT_BYTES));
                for (int bx = 0; bx < 2; ++bx) {
                    tma_load_4d(smem + OFF_Q + s * Q_TILE + bx * Q_BOX, &tmQ, &bar_in_full[s], bx * 32, qt * BM, h, b, kEvictLast);
                    if (!BWT_DBG(1024)) tma_load_4d(smem + OFF_QM + s * Q_TILE + bx * Q_BOX, &tmQm, &bar_in_full[s], bx * 32, qt * BM, h, b, kEvictLast);
                    tma_load_4d(smem + OFF_DO + s * Q_TILE + bx * Q_BOX, &tmdO, &bar_in_full[s], bx * 32, qt * BM, h, b, kEvictLast);
                    if (!BWT_DBG(1024)) tma_load_4d(smem + OFF_DOM + s * Q_TILE + bx * Q_BOX, &tmdOm, &bar_in_full[s], bx * 32, qt * BM, h, b, kEvictLast);
                }
                if (!BWT_DBG(512)) {
                    bwt_bulk_load_1d(smem + OFF_STAT + s * 2 * STAT_BYTES, Lrow + qt * BM, STAT_BYTES, &bar_in_full[s]);
                    bwt_bulk_load_1d(smem + OFF_STAT + s * 2 * STAT_BYTES + STAT_BYTES, Drow + qt * BM, STAT_BYTES, &bar_in_full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        if (elect_one()) {
            constexpr uint32_t idesc_ss = make_idesc(kFmtTF32, 128, BM, false, false);   // M1, M2
            constexpr uint32_t idesc_ts = make_idesc(kFmtTF32, 128, D, false, true);     // M3, M4
            constexpr uint32_t idesc_dq = make_idesc(kFmtTF32, 128, BM, true, true);     // M5
            const uint32_t sb = smem_u32(smem);
            auto mma_ss = [&](uint32_t a_tile, uint32_t b_tile, uint32_t d_col) {        // D = A[128 x 64] B[32 x 64]^T, both K-major
                const uint64_t ad = make_sdesc(a_tile, 16, 1024, 2), bd = make_sdesc(b_tile, 16, 1024, 2);
                if (BWT_DBG(16)) return;
#pragma unroll
                for (int kk = 0; kk < D / 8; ++kk)
                    umma_tf32_ss(tmem + d_col, sdesc_advance(ad, (kk >> 2) * KV_BOX + (kk & 3) * 32),
                                 sdesc_advance(bd, (kk >> 2) * Q_BOX + (kk & 3) * 32), idesc_ss, kk > 0);
            };
            auto mma_ts = [&](uint32_t a_col, uint32_t b_tile, uint32_t d_col, bool acc) {   // D += A^T-in-TMEM[128 x 32] B[32 q x 64 d]
                const uint64_t bd = make_sdesc(b_tile, Q_BOX, 512, 1);                   // MN-major, 32-byte-atom swizzle
                if (BWT_DBG(64)) return;
#pragma unroll
                for (int kk = 0; kk < BM / 8; ++kk)
                    umma_tf32_ts(tmem + d_col, tmem + a_col + 8 * kk, sdesc_advance(bd, kk * 1024), idesc_ts, (acc || kk > 0) ? 1u : 0u);
            };
            auto M1 = [&](int i) { mma_ss(sb + OFF_K, sb + OFF_Q + (i % NS) * Q_TILE, TM_S + 32 * (i & 1)); umma_commit(&bar_S[i & 1]); };
            auto M2 = [&](int i) { mma_ss(sb + OFF_V, sb + OFF_DO + (i % NS) * Q_TILE, TM_DP); umma_commit(bar_dP); };

            mbar_wait(bar_KV, 0);
            mbar_wait(&bar_in_full[0], 0);
            tc_fence_after();
            M1(0); M2(0);
            for (int i = 0; i < n_it; ++i) {
                const int bb = i & 1, s = i % NS;
                if (i + 1 < n_it) {
                    mbar_wait(&bar_in_full[(i + 1) % NS], ((i + 1) / NS) & 1);
                    tc_fence_after();
                    M1(i + 1);
                }
                mbar_wait(&bar_P[bb], (i >> 1) & 1);
                tc_fence_after();
                mma_ts(TM_S + 32 * bb, sb + OFF_DOM + s * Q_TILE, TM_DV, i > 0);         // M3: dV += P^T dO_i
                mbar_wait(bar_dS, i & 1);
                tc_fence_after();
                mma_ts(TM_DP, sb + OFF_QM + s * Q_TILE, TM_DK, i > 0);                  // M4: dK += dS^T Q_i
                umma_commit(&bar_in_empty[s]);
                if (i + 1 < n_it) M2(i + 1);
                mbar_wait(bar_dQ_empty, (i & 1) ^ 1);
                tc_fence_after();
                if (!BWT_DBG(4)) {                                                      // M5: dQ_i^T = K^T dS
                    const uint64_t ad = make_sdesc(sb + OFF_KM, KV_BOX, 512, 1);
                    const uint64_t bd = make_sdesc(sb + OFF_DS + bb * DS_TILE, DS_TILE, 512, 1);   // (one box along N: LBO unused, kept valid)
#pragma unroll
                    for (int kk = 0; kk < BN / 8; ++kk)
                        umma_tf32_ss(tmem + TM_DQ, sdesc_advance(ad, kk * 1024), sdesc_advance(bd, kk * 1024), idesc_dq, kk > 0);
                }
                umma_commit(bar_dQ_full);
            }
            umma_commit(bar_done);
        }
    } else if (warp >= 4 && warp < 12) {
        // ================================ compute warpgroups ==========================
        const int w = (warp - 4) >> 2;                                     // q columns 16w .. 16w+15
        const int r = threadIdx.x & 127;                                   // kv row == TMEM lane
        const uint32_t lane_base = uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16;
        const bool lane0 = (threadIdx.x & 31) == 0;
        const uint32_t stat_base = smem_u32(smem + OFF_STAT) + 64 * w;
        for (int i = 0; i < n_it; ++i) {
            const int bb = i & 1, s = i % NS;
            const uint32_t tS = tmem + TM_S + 32 * bb + 16 * w + lane_base;
            const uint32_t tP = tmem + TM_DP + 16 * w + lane_base;
            mbar_wait(&bar_in_full[s], (i / NS) & 1);
            float st[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) bwt_lds_f32x4(st + 4 * c, stat_base + s * 2 * STAT_BYTES + 16 * c);
            mbar_wait(&bar_S[bb], (i >> 1) & 1);
            tc_fence_after();
            uint32_t sv[16];
            B200_TMEM_LD_16(tS, sv);
            tmem_wait_ld();
            float p[16];
            uint32_t pk[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                p[c] = fast_exp2(fmaf(__uint_as_float(sv[c]), scale_log2, -st[c]));
            }
            if (causal) {
                // key kv0 + r is visible to the queries q >= kv0 + r - (Nkv - Nq): zero the columns before it (dS inherits the zeros)
                int qt = i + i_off;
                if (qt >= n_it) qt -= n_it;
                const int first_visible = kv0 + r - (causal - 1) - (qt_min + qt) * BM - 16 * w;
                if (first_visible > 0) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) if (c < first_visible) p[c] = 0.0f;
                }
            }
#pragma unroll
            for (int c = 0; c < 16; ++c) pk[c] = to_tf32(p[c]);
            B200_TMEM_ST_16(tS, pk);
#pragma unroll
            for (int c = 0; c < 4; ++c) bwt_lds_f32x4(st + 4 * c, stat_base + s * 2 * STAT_BYTES + STAT_BYTES + 16 * c);   // D_i
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane0) mbar_arrive(&bar_P[bb]);
            // ---- dS^T = scale * P^T o (dP^T - D)
            mbar_wait(bar_dP, i & 1);
            tc_fence_after();
            B200_TMEM_LD_16(tP, sv);
            tmem_wait_ld();
            uint32_t dk[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) dk[c] = to_tf32(p[c] * (__uint_as_float(sv[c]) - st[c]) * scale);
            B200_TMEM_ST_16(tP, dk);
            // shared-memory copy (B operand of M5), MN-major with the 32-byte-atom swizzle: row r = 128 bytes = 4 atoms of
            // 8 q values, atom a stored at a ^ (r & 3); this warpgroup owns atoms 2w, 2w + 1
            uint8_t* rowp = smem + OFF_DS + bb * DS_TILE + r * 128;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                uint4* dst = reinterpret_cast<uint4*>(rowp + (((2 * w + a) ^ (r & 3)) * 32));
                dst[0] = make_uint4(dk[8 * a], dk[8 * a + 1], dk[8 * a + 2], dk[8 * a + 3]);
                dst[1] = make_uint4(dk[8 * a + 4], dk[8 * a + 5], dk[8 * a + 6], dk[8 * a + 7]);
            }
            fence_proxy_async_smem();
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane0) mbar_arrive(bar_dS);
        }
        // ---- epilogue: dV (w = 0) / dK (w = 1): TMEM fp32 [128 kv x 64 d] -> swizzled smem (the dead K / V tile) -> TMA store
        mbar_wait(bar_done, 0);
        tc_fence_after();
        const uint32_t tA = tmem + (w == 0 ? TM_DV : TM_DK) + lane_base;
        uint8_t* out = smem + (w == 0 ? OFF_V : OFF_K);
#pragma unroll
        for (int c = 0; c < 2; ++c) {                                      // box c = d columns 32c .. 32c+31
            uint32_t o[32];
            B200_TMEM_LD_32(tA + 32 * c, o);
            tmem_wait_ld();
            uint8_t* rowp = out + c * KV_BOX + r * 128;
#pragma unroll
            for (int q = 0; q < 8; ++q)
                *reinterpret_cast<uint4*>(rowp + ((q ^ (r & 7)) * 16)) = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        }
        fence_proxy_async_smem();
        named_bar_sync(4 + w, 128);
        if (r == 0 && !BWT_DBG(256)) {
            const CUtensorMap* tm = w == 0 ? &tmdV : &tmdK;
            tma_store_4d(tm, out, 0, kv0, h, b);
            tma_store_4d(tm, out + KV_BOX, 32, kv0, h, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
    } else if (warp >= 12) {
        // ================================ dQ drain ====================================
        const int r = threadIdx.x & 127;                                   // d index == TMEM lane (64 real)
        const uint32_t tQ = tmem + TM_DQ + (uint32_t(((threadIdx.x >> 5) & 3) * 32) << 16);
        const bool lane0 = (threadIdx.x & 31) == 0, leader = r == 0;
        for (int i = 0; i < n_it; ++i) {
            int qt = i + i_off;
            if (qt >= n_it) qt -= n_it;
            qt += qt_min;
            mbar_wait(bar_dQ_full, i & 1);
            tc_fence_after();
            uint32_t v[32];
            B200_TMEM_LD_32(tQ, v);
            tmem_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane0) mbar_arrive(bar_dQ_empty);
            float* stg = reinterpret_cast<float*>(smem + OFF_STG + (i & 1) * STG_BYTES);
            if (leader) tma_store_wait_read<1>();                          // the reduction that last used this buffer has read it
            named_bar_sync(6, 128);
            if (r < D) {
#pragma unroll
                for (int q = 0; q < BM; ++q) stg[q * D + r] = __uint_as_float(v[q]);
            }
            fence_proxy_async_smem();
            named_bar_sync(6, 128);
            if (leader) {
                const int rows = min(BM, Nq - qt * BM);
                if (rows > 0 && !BWT_DBG(1)) bwt_bulk_reduce_add_f32(dQ + (bh * Nq + (size_t)qt * BM) * D, stg, (uint32_t)rows * D * 4);
                tma_store_commit();
            }
        }
        if (leader) tma_store_wait<0>();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<256>(tmem);
}

// ------------------------------------------------------------------------------------ host side
bool tc_bwd_tf32_supported(int dtype, int d, int Nq, int Nkv)
{
    return dtype == B200_DTYPE_F32 && d == bwt::D && Nq >= 1 && Nkv >= 1;
}

// out = round-to-nearest tf32 of in.  The tensor cores TRUNCATE fp32 operands to tf32 (the low 13 mantissa bits are
// ignored), which doubles the operand error of a rounded conversion and biases it; rounding Q, K, V, dO once up front
// (HBM-bound, microseconds) took the worst dQ error on unit-normal data from 2.2e-3 to within north_star's 2e-3.
__global__ void __launch_bounds__(256) round_tf32_kernel(float* __restrict__ out, const float* __restrict__ in, size_t n4)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(in)[i];
        uint4 o;
        o.x = to_tf32(a.x); o.y = to_tf32(a.y); o.z = to_tf32(a.z); o.w = to_tf32(a.w);
        reinterpret_cast<uint4*>(out)[i] = o;
    }
}

static inline size_t bwt_al256(size_t x) { return (x + 255) & ~(size_t)255; }
// workspace: padded statistics 2 x [BH, pad64(Nq)] floats + tf32-rounded copies of Q, dO ([BH, Nq, 64]) and K, V ([BH, Nkv, 64])
size_t tc_bwd_tf32_workspace_bytes(int BH, int Nq, int Nkv)
{
    const size_t Nq_pad = (size_t)(Nq + 63) / 64 * 64;
    return 2 * bwt_al256((size_t)BH * Nq_pad * sizeof(float)) + 2 * bwt_al256((size_t)BH * Nq * bwt::D * sizeof(float)) +
           2 * bwt_al256((size_t)BH * Nkv * bwt::D * sizeof(float));
}

// L2p / Dp: padded statistics as in bwd_sm100.cu (pad = (+inf, 0): padded query rows contribute exactly 0); dQ must be
// zeroed by the caller of this function on the same stream (the kernel add-reduces into it).
int tc_bwd_tf32_launch(const float* Q, const float* K, const float* V, const float* dO, const float* L2p, const float* Dp,
                       float* dQ, float* dK, float* dV, int BH, int Nq, int Nkv, float scale, cudaStream_t st, bool causal)
{
    using namespace bwt;
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(bwd_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int Hm = BH <= 65535 ? BH : 32768, Bm = (BH + Hm - 1) / Hm;
    const Lay lq = lay_contiguous(Hm, Nq, D), lk = lay_contiguous(Hm, Nkv, D);
    const CUtensorMapDataType f32 = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const CUtensorMapSwizzle a32 = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
    CUtensorMap tmQ, tmQm, tmK, tmKm, tmV, tmdO, tmdOm, tmdK, tmdV;
    int trc = 0;
    if ((trc = make_tmap_4d(&tmQ, Q, f32, 4, Bm, Hm, Nq, D, lq.sb, lq.sh, lq.sn, BM, 32)) ||
        (trc = make_tmap_4d(&tmQm, Q, f32, 4, Bm, Hm, Nq, D, lq.sb, lq.sh, lq.sn, BM, 32, a32)) ||
        (trc = make_tmap_4d(&tmdO, dO, f32, 4, Bm, Hm, Nq, D, lq.sb, lq.sh, lq.sn, BM, 32)) ||
        (trc = make_tmap_4d(&tmdOm, dO, f32, 4, Bm, Hm, Nq, D, lq.sb, lq.sh, lq.sn, BM, 32, a32)) ||
        (trc = make_tmap_4d(&tmK, K, f32, 4, Bm, Hm, Nkv, D, lk.sb, lk.sh, lk.sn, BN, 32)) ||
        (trc = make_tmap_4d(&tmKm, K, f32, 4, Bm, Hm, Nkv, D, lk.sb, lk.sh, lk.sn, BN, 32, a32)) ||
        (trc = make_tmap_4d(&tmV, V, f32, 4, Bm, Hm, Nkv, D, lk.sb, lk.sh, lk.sn, BN, 32)) ||
        (trc = make_tmap_4d(&tmdK, dK, f32, 4, Bm, Hm, Nkv, D, lk.sb, lk.sh, lk.sn, BN, 32)) ||
        (trc = make_tmap_4d(&tmdV, dV, f32, 4, Bm, Hm, Nkv, D, lk.sb, lk.sh, lk.sn, BN, 32)))
        return trc == -2 ? B200_ERR_UNSUPPORTED : B200_ERR_CUDA;
    const int Nq_pad = (Nq + 63) / 64 * 64;
    dim3 grid((Nkv + BN - 1) / BN, Hm, Bm);
#ifdef B200_EXPERIMENTS
    static int dbg = -1;
    if (dbg < 0) { const char* e = getenv("B200_BWD_DBG"); dbg = e ? atoi(e) : 0; }
#else
    const int dbg = 0;
#endif
    bwd_tf32_kernel<<<grid, THREADS, SMEM, st>>>(tmQ, tmQm, tmK, tmKm, tmV, tmdO, tmdOm, tmdK, tmdV, dQ, L2p, Dp, Nq, Nq_pad,
                                                 scale * 1.4426950408889634f, scale, BH, dbg, causal ? 1 + (Nkv - Nq) : 0);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}


// D = rowsum(dO o O) is in `Dvec`; `ws` as tc_bwd_tf32_workspace_bytes
int tc_bwd_tf32(const void* Q, const void* K, const void* V, const void* dO, const float* L, const float* Dvec,
                void* dQ, void* dK, void* dV, int BH, int Nq, int Nkv, float scale, void* ws, size_t ws_bytes, cudaStream_t st,
                bool causal)
{
    if (causal && Nq > Nkv) return B200_ERR_INVALID_ARGUMENT;
    if (ws_bytes < tc_bwd_tf32_workspace_bytes(BH, Nq, Nkv)) return B200_ERR_WORKSPACE;
    const size_t Nq_pad = (size_t)(Nq + 63) / 64 * 64;
    const size_t nq = (size_t)BH * Nq * bwt::D, nk = (size_t)BH * Nkv * bwt::D;
    uint8_t* p = reinterpret_cast<uint8_t*>(ws);
    float* L2p = reinterpret_cast<float*>(p); p += bwt_al256((size_t)BH * Nq_pad * sizeof(float));
    float* Dp = reinterpret_cast<float*>(p);  p += bwt_al256((size_t)BH * Nq_pad * sizeof(float));
    float* Qr = reinterpret_cast<float*>(p);  p += bwt_al256(nq * sizeof(float));
    float* dOr = reinterpret_cast<float*>(p); p += bwt_al256(nq * sizeof(float));
    float* Kr = reinterpret_cast<float*>(p);  p += bwt_al256(nk * sizeof(float));
    float* Vr = reinterpret_cast<float*>(p);
    B200_TRY(pad_stats(L, Dvec, L2p, Dp, BH, Nq, st));
    const float* src[4] = {reinterpret_cast<const float*>(Q), reinterpret_cast<const float*>(dO), reinterpret_cast<const float*>(K),
                           reinterpret_cast<const float*>(V)};
    float* dst[4] = {Qr, dOr, Kr, Vr};
    const size_t cnt[4] = {nq, nq, nk, nk};
    for (int i = 0; i < 4; ++i) {
        unsigned grid = (unsigned)std::min<size_t>((cnt[i] / 4 + 255) / 256, (size_t)148 * 8);
        round_tf32_kernel<<<grid, 256, 0, st>>>(dst[i], src[i], cnt[i] / 4);          // (d = 64: counts are multiples of 4)
        B200_CUDA_TRY(cudaGetLastError());
        count_launch();
    }
    B200_CUDA_TRY(cudaMemsetAsync(dQ, 0, nq * sizeof(float), st));
    return tc_bwd_tf32_launch(Qr, Kr, Vr, dOr, L2p, Dp, reinterpret_cast<float*>(dQ), reinterpret_cast<float*>(dK),
                              reinterpret_cast<float*>(dV), BH, Nq, Nkv, scale, st, causal);
}

}  // namespace b200
