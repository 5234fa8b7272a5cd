// fwd_sm100.cu -- FlashAttention-2 forward for sm_100a: TMA -> swizzled smem -> tcgen05.mma with
// accumulators in TMEM, warp-specialised producer / MMA-issuer / softmax roles.
//
// Replaces flash_attention_2_forward_kernel (reference
// src/02_flash_attention_v2_forward/flash_attention_kernel.cu:37-297 and the 32-lane build in
// src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:13-130) and its inner step
// process_kv_block (src/util/attention_helper.h:40-132): S = QK^T, m' = max(m, rowmax),
// P = e^(S-m'), l = e^(m-m') l + rowsum(P), O = e^(m-m') O + PV, end: O /= l, L = m + log l.
//
// One CTA owns NT (1 or 2) query tiles of 128 rows of one (b,h) slab and streams all K/V tiles:
//   warp 0      TMA producer: Q tiles once, then a STAGES-deep ring of K and V tiles
//   warp 1      MMA issuer (one thread): S_t = Q_t K_j^T (SS), O_t += P_t V_j (A = P from TMEM)
//   warp 2      TMEM allocator
//   warps 4..   one softmax warpgroup per query tile: thread = row; tcgen05.ld the S row,
//               rowmax / exp2 / rowsum in registers (no shuffles: a thread owns a whole row),
//               lazy O rescale in TMEM, P back to TMEM (aliasing S), final O/l -> smem -> TMA store.
// With NT = 2 the two tiles ping-pong: while warpgroup 0 does the softmax of S_0(j) the tensor
// pipe runs PV_1(j-1) and S_1(j), and vice versa.
// TMEM map (512 columns): S_t / P_t at 128 t, O_t at 256 + 128 t.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"
#include <type_traits>
#include <stdlib.h>

namespace b200 {

#ifdef B200_TRACE
__device__ long long g_trace[3 * 8 * 8];
#endif
#define B200_TR(role, j, ev) B200_TRACE_EVENT(g_trace, role, j, ev, (role) == 2 || ((threadIdx.x >> 5) & 3) == 0)

// POLY = how many of every 16 (score, score) pairs get their exp2 from the FMA-pipe polynomial
// instead of the MUFU (bf16 only); SPLITP = P is published to the MMA warp in two halves.
// ISSUERS = 2 (experimental, NT = 2 only): one MMA-issuing thread per query tile (warps 1 and 3) instead of one thread
// serving both tiles in turn.
template <typename Elem_, int D_, int NT_, int BC_, int POLY_ = 0, bool SPLITP_ = true, int ISSUERS_ = 1>
struct FwdCfg {
    using Elem = Elem_;
    static constexpr int EB = sizeof(Elem);
    static constexpr bool kF32 = (EB == 4);
    // EB == 1: fp8 e4m3 operands (kind::f8f6f4, K = 32 per MMA), P written back to TMEM as e4m3 (4 per column), O stored
    // as bf16 through its own staging tiles (a bf16 O tile is twice the bytes of an fp8 Q tile)
    static constexpr bool kF8 = (EB == 1);
    static constexpr int OEB = kF8 ? 2 : EB;               // bytes per output element
    static constexpr int OCPB = 128 / OEB, O_BOXES = D_ / OCPB;
    static constexpr int O_TILE = O_BOXES * 128 * 128;
    static constexpr int O_STAGE = kF8 ? NT_ * O_TILE : 0;  // dedicated O staging (fp8 only), after the K/V stages
    static constexpr int D = D_, NT = NT_, BC = BC_, BR = 128, STAGES = 2;
    static constexpr int CPB = 128 / EB;                 // elements per 128-byte swizzle row ("box" width)
    static constexpr int KPER = 32 / EB;                 // K elements per tcgen05.mma (bf16: 16, tf32: 8)
    static constexpr int BOXES = D / CPB;                // boxes along d in a Q / K / V / O tile
    static constexpr int Q_BOX = BR * 128, KV_BOX = BC * 128;
    static constexpr int Q_TILE = BOXES * Q_BOX, KV_TILE = BOXES * KV_BOX;
    static constexpr int THREADS = 128 + NT * 128;
    static constexpr int POLY = POLY_;
    static constexpr bool SPLITP = SPLITP_;
    static constexpr int ISSUERS = ISSUERS_;
    static_assert(ISSUERS == 1 || (ISSUERS == 2 && NT == 2), "one issuer, or one per tile");
    static constexpr int NBARS = NT * 5 + STAGES * 4;
    // setmaxnreg targets (warps 0-3 / softmax).  The pool is what the CTA got at launch -- 384 threads x 168
    // registers = 64512, not the whole 64K file -- and a setmaxnreg.inc that cannot be served waits forever.
    static constexpr int REGS_CTRL = 64, REGS_SOFTMAX = 216;
    static constexpr int SMEM = 1024 + NT * Q_TILE + 2 * STAGES * KV_TILE + O_STAGE + NBARS * 8 + 16;
    static constexpr int PCOLS_PER_CHUNK = kF32 ? 32 : (kF8 ? 8 : 16);   // TMEM columns written per 32 P values
    // V is the MN-major B operand of O += P V.  bf16: plain 128B swizzle.  tf32: the hardware only
    // accepts the 32-byte-atom 128B swizzle for MN-major operands (4-row pattern, SBO = 512).
    static constexpr uint32_t V_LAYOUT = kF32 ? 1u : 2u;
    static constexpr uint32_t V_SBO = kF32 ? 512u : 1024u;
    static_assert(D % CPB == 0 && BC % 64 == 0 && BC <= 128 && D <= 128, "tile shape");
    static_assert(POLY == 0 || (!kF32 && !kF8), "the polynomial exp2 is only used for bf16 P");
    static_assert(NT != 2 || 128 * REGS_CTRL + 256 * REGS_SOFTMAX <= 384 * 168, "register pool of the CTA");
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ uint32_t cvt_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

// element size -> MMA kind: 4 = tf32, 2 = f16 (bf16), 1 = f8f6f4 (e4m3)
template <int EB>
__device__ __forceinline__ void umma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
    if constexpr (EB == 4) umma_tf32_ss(d, a, b, idesc, acc ? 1u : 0u);
    else if constexpr (EB == 1) umma_f8_ss(d, a, b, idesc, acc ? 1u : 0u);
    else umma_f16_ss(d, a, b, idesc, acc ? 1u : 0u);
}
template <int EB>
__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, bool acc) {
    if constexpr (EB == 4) umma_tf32_ts(d, a_tmem, b, idesc, acc ? 1u : 0u);
    else if constexpr (EB == 1) umma_f8_ts(d, a_tmem, b, idesc, acc ? 1u : 0u);
    else umma_f16_ts(d, a_tmem, b, idesc, acc ? 1u : 0u);
}
// two fp32 -> two e4m3 in the low 16 bits (lo in bits 0..7)
__device__ __forceinline__ uint32_t pack_e4m3x2(float lo, float hi) {
    uint16_t r;
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
    return r;
}

// Lazy-rescale threshold in log2 units: the running max is only refreshed (and O rescaled) when
// it would grow by more than this, so stale-max probabilities stay <= 2^8 (exact in the end:
// O and l carry the same stale max, which cancels in O / l and is added back in the LSE).
constexpr float kRescaleThreshold = 8.0f;

// One softmax warpgroup (128 threads) serving query tile t.  Thread r owns query row r of the tile.
template <class C>
__device__ __forceinline__ void softmax_role(int t, uint32_t tmem, uint8_t* sQ, uint64_t* bar_S, uint64_t* bar_Pa,
                                             uint64_t* bar_Pb, uint64_t* bar_O, const CUtensorMap* tmO,
                                             float* __restrict__ Lout, long long ldL, int Nq, int Nkv, int nkv, int n_plain,
                                             int causal, int q0, int h, int b, int bh, float scale_log2)
{
    constexpr int BC = C::BC, D = C::D;
    const int wg_thread = threadIdx.x & 127;                       // == TMEM lane == row within the tile
    const int warp_in_wg = (threadIdx.x >> 5) & 3;
    const uint32_t lane_base = uint32_t(warp_in_wg * 32) << 16;
    const uint32_t tS = tmem + 128 * t + lane_base;                // S_t and (aliased) P_t
    const uint32_t tO = tmem + 256 + 128 * t + lane_base;
    // first key column this row must NOT see: Nkv (ragged last tile), and row + 1 + (Nkv - Nq) under the causal mask
    // (`causal` carries the bottom-right offset: 0 = no mask, 1 + (Nkv - Nq) otherwise)
    const int col_end = causal ? min(Nkv, q0 + t * 128 + wg_thread + causal) : Nkv;
    const uint64_t scale2 = f2_pack(scale_log2, scale_log2);

    float m = -INFINITY, l = 0.0f;                                 // m in log2 units (already scaled)

    // one K/V tile: S_t(j) -> (m, l) update -> P_t(j).  kMasked = the ragged last tile / the causal diagonal tiles.
    auto step = [&](auto masked_tag, const int j) {
        constexpr bool kMasked = decltype(masked_tag)::value;
        mbar_wait(bar_S, j & 1);
        tc_fence_after();
        B200_TR(t, j, 0);
        uint32_t s[BC];
#pragma unroll
        for (int c = 0; c < BC / 32; ++c) B200_TMEM_LD_32(tS + 32 * c, s + 32 * c);
        tmem_wait_ld();
        B200_TR(t, j, 1);
        if constexpr (kMasked) {
            const int lim = col_end - j * BC;
#pragma unroll
            for (int i = 0; i < BC; ++i) if (i >= lim) s[i] = 0xff800000u;       // -inf
        }
        // P for chunks [32 c0, 32 c1) of the row with the running max m; partial row sums into acc
        uint64_t acc[4];
        auto exp_chunks = [&](const int c0, const int c1, const float m_use) {
            const uint64_t neg_m2 = f2_pack(-m_use, -m_use);
#pragma unroll
            for (int c = c0; c < c1; ++c) {
                uint32_t r[32];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint64_t x = f2_fma(f2_pack(__uint_as_float(s[32 * c + 2 * i]), __uint_as_float(s[32 * c + 2 * i + 1])), scale2, neg_m2);
                    float p0, p1;
                    // spread the polynomial pairs evenly through the chunk so MUFU and FMA work interleave
                    const bool poly = !kMasked && C::POLY > 0 && ((i * C::POLY) % 16 < C::POLY);
                    if (poly) {
                        exp2_poly2(x, p0, p1);
                    } else {
                        float x0, x1;
                        f2_unpack(x, x0, x1);
                        p0 = fast_exp2(x0);
                        p1 = fast_exp2(x1);
                    }
                    if (c == 0 && i < 4) acc[i] = f2_pack(p0, p1); else acc[i & 3] = f2_add(acc[i & 3], f2_pack(p0, p1));
                    if constexpr (C::kF32) { r[2 * i] = cvt_tf32(p0); r[2 * i + 1] = cvt_tf32(p1); }
                    else if constexpr (C::kF8) {
                        const uint32_t h2 = pack_e4m3x2(p0, p1);               // four P values per column
                        if (i & 1) r[i >> 1] |= h2 << 16; else r[i >> 1] = h2;
                    }
                    else r[i] = pack_bf16x2(p0, p1);
                }
                if constexpr (C::kF32) B200_TMEM_ST_32(tS + 32 * c, r);
                else if constexpr (C::kF8) B200_TMEM_ST_8(tS + 8 * c, r);
                else B200_TMEM_ST_16(tS + 16 * c, r);
            }
        };
        // Speculate: the first half of P is computed with the running max as it stands, while the row
        // max of this tile (independent ALU work on the same registers) is still being reduced.  The
        // lazy-rescale rule makes the speculation right almost always; when the max does grow by more
        // than the threshold (always at j = 0) the half is simply recomputed with the new max.
        exp_chunks(0, BC / 64, m);
        // row max: four independent chains over contiguous quarters; each FMNMX3 takes an (even, odd)
        // register pair so its three sources never collide on a register bank
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
        const bool need = (m_new - m) > kRescaleThreshold;
        float alpha = 1.0f;
        B200_TR(t, j, 2);
        if (__any_sync(0xffffffffu, need)) {
            if (need) { alpha = fast_exp2(m - m_new); m = m_new; }
            if (j > 0) {
                // PV_t(j-1) is complete (bar_S(j) was committed after it), PV_t(j) not yet issued: O_t is ours.
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
        if ((threadIdx.x & 31) == 0) mbar_arrive(bar_Pa);         // first half of P is in TMEM: PV can start
        B200_TR(t, j, 4);
        exp_chunks(BC / 64, BC / 32, m);
        float a0, a1, a2, a3;
        f2_unpack(f2_add(acc[0], acc[1]), a0, a1);
        f2_unpack(f2_add(acc[2], acc[3]), a2, a3);
        l = fmaf(l, alpha, (a0 + a1) + (a2 + a3));
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(bar_Pb);
        B200_TR(t, j, 7);
    };

    for (int j = 0; j < n_plain; ++j) step(std::false_type{}, j);
    for (int j = n_plain; j < nkv; ++j) step(std::true_type{}, j); // (a full last tile masks nothing)

    // ---- epilogue: O_t / l -> smem (reusing the Q tile, 128B-swizzled) -> TMA store; LSE -> global
    mbar_wait(bar_O, 0);
    tc_fence_after();
    const float inv_l = 1.0f / l;
    // (fp8: the bf16 O tile does not fit the fp8 Q tile; it has its own staging behind the K/V stages)
    uint8_t* sO = C::kF8 ? sQ + C::NT * C::Q_TILE + 2 * C::STAGES * C::KV_TILE + t * C::O_TILE : sQ + t * C::Q_TILE;
    const int r = wg_thread;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        B200_TMEM_LD_32(tO + 32 * c, o);
        tmem_wait_ld();
        if constexpr (C::kF32) {
            // 32 floats = one 128-byte box row: box c, 8 chunks of 16 B
            uint8_t* rowp = sO + c * C::Q_BOX + r * 128;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                float4 v = make_float4(__uint_as_float(o[4 * q]) * inv_l, __uint_as_float(o[4 * q + 1]) * inv_l,
                                       __uint_as_float(o[4 * q + 2]) * inv_l, __uint_as_float(o[4 * q + 3]) * inv_l);
                *reinterpret_cast<float4*>(rowp + ((q ^ (r & 7)) * 16)) = v;
            }
        } else {
            // 32 bf16 = 64 B = 4 chunks: box c/2, chunks (c&1)*4 .. +3
            uint8_t* rowp = sO + (c >> 1) * C::Q_BOX + r * 128;
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
    }
    fence_proxy_async_smem();
    named_bar_sync(1 + t, 128);
    if (wg_thread == 0) {
#pragma unroll
        for (int bx = 0; bx < C::O_BOXES; ++bx) tma_store_4d(tmO, sO + bx * C::Q_BOX, bx * C::OCPB, q0 + t * 128, h, b);
        tma_store_commit();
        tma_store_wait<0>();
    }
    const int row = q0 + t * 128 + r;
    if (Lout != nullptr && row < Nq)
        Lout[(size_t)bh * ldL + row] = (m + log2f(l)) * 0.6931471805599453f;
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
           const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
           float* __restrict__ Lout, long long ldL, int Nq, int Nkv, float scale_log2, int causal, int G, int BH)
{
    constexpr int NT = C::NT, BC = C::BC, D = C::D, ST = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + NT * C::Q_TILE;
    uint8_t* sV = sK + ST * C::KV_TILE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ST * C::KV_TILE + C::O_STAGE);
    uint64_t* bar_Q = bars;                 // [NT]  Q tile landed
    uint64_t* bar_S = bar_Q + NT;           // [NT]  S_t(j) complete in TMEM
    uint64_t* bar_Pa = bar_S + NT;          // [NT]  first half of P_t(j) written (one arrival per softmax warp)
    uint64_t* bar_Pb = bar_Pa + NT;         // [NT]  all of P_t(j) written
    uint64_t* bar_O = bar_Pb + NT;          // [NT]  last PV of tile t complete
    uint64_t* bar_Kf = bar_O + NT;          // [ST]
    uint64_t* bar_Ke = bar_Kf + ST;
    uint64_t* bar_Vf = bar_Ke + ST;
    uint64_t* bar_Ve = bar_Vf + ST;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_Ve + ST);

    const int warp = threadIdx.x >> 5;
    // grid (q blocks, heads of the map, batches of the map); K/V head = h / G (grouped-query attention, G = 1: MHA)
    const int h = blockIdx.y, b = blockIdx.z, bh = b * gridDim.y + h, hk = h / G;
    if (bh >= BH) return;                                          // (a re-split contiguous BH may leave the last batch ragged)
    // causal (bottom-right aligned, Nq <= Nkv): the q blocks near the end of the sequence have the most K/V tiles -- schedule them first
    const int q0 = (causal ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * (NT * 128);
    const int ntiles = min(NT, (Nq - q0 + 127) / 128);
    const int nkv_all = (Nkv + BC - 1) / BC;
    // K/V tiles that query tile t needs (rows q0 + 128 t .. + 127 see keys 0 .. row), and for the whole CTA
    auto nkv_of = [&](int t) { return causal ? min(nkv_all, (q0 + t * 128 + 126 + causal) / BC + 1) : nkv_all; };
    const int nkv = nkv_of(ntiles - 1), nkv_first = nkv_of(0);

    if (threadIdx.x == 0) {
        for (int t = 0; t < NT; ++t) { mbar_init(&bar_Q[t], 1); mbar_init(&bar_S[t], 1); mbar_init(&bar_Pa[t], 4); mbar_init(&bar_Pb[t], 4); mbar_init(&bar_O[t], 1); }
        // "stage free" barriers: one commit per MMA issuer that is active in this CTA
        const int n_free = C::ISSUERS == 2 ? ntiles : 1;
        for (int s = 0; s < ST; ++s) { mbar_init(&bar_Kf[s], 1); mbar_init(&bar_Ke[s], n_free); mbar_init(&bar_Vf[s], 1); mbar_init(&bar_Ve[s], n_free); }
        fence_mbar_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
    }
    if (warp == 2) { tmem_alloc<512>(tmem_slot); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // NT == 2 runs 384 threads: the control warps hand registers to the softmax warps (setmaxnreg is
    // issued inside each role's branch -- ptxas only honours it for the code it dominates).
    if (warp == 0) {
        // ================================ TMA producer ================================
        if constexpr (NT == 2) reg_dec<C::REGS_CTRL>();
        if (elect_one()) {
            for (int t = 0; t < ntiles; ++t) {
                mbar_expect_tx(&bar_Q[t], C::Q_TILE);
                for (int bx = 0; bx < C::BOXES; ++bx)
                    tma_load_4d(sQ + t * C::Q_TILE + bx * C::Q_BOX, &tmQ, &bar_Q[t], bx * C::CPB, q0 + t * 128, h, b, kEvictFirst);
            }
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                const uint32_t ph = (j / ST) & 1;
                mbar_wait(&bar_Ke[s], ph ^ 1);
                mbar_expect_tx(&bar_Kf[s], C::KV_TILE);
                for (int bx = 0; bx < C::BOXES; ++bx)
                    tma_load_4d(sK + s * C::KV_TILE + bx * C::KV_BOX, &tmK, &bar_Kf[s], bx * C::CPB, j * BC, hk, b, kEvictLast);
                mbar_wait(&bar_Ve[s], ph ^ 1);
                mbar_expect_tx(&bar_Vf[s], C::KV_TILE);
                for (int bx = 0; bx < C::BOXES; ++bx)
                    tma_load_4d(sV + s * C::KV_TILE + bx * C::KV_BOX, &tmV, &bar_Vf[s], bx * C::CPB, j * BC, hk, b, kEvictLast);
            }
        }
    } else if (C::ISSUERS == 2 && (warp == 1 || warp == 3)) {
        // ================================ MMA issuers, one per tile (experimental) =====
        // Same MMAs and hand-offs as the single issuer below, but tile t's PV / S MMAs are issued by its own thread
        // (warp 1: tile 0, warp 3: tile 1): a P-ready wait and the fence after it stall only that tile's issue, and
        // the other thread keeps feeding the tensor pipe.  tcgen05.commit tracks the issuing thread's MMAs only, so
        // the K/V "stage free" barriers take one commit from each active issuer.
        reg_dec<C::REGS_CTRL>();
        const int t = warp == 1 ? 0 : 1;
        if (t < ntiles && elect_one()) {
            constexpr uint32_t fmt = C::kF32 ? kFmtTF32 : (C::kF8 ? kFmtE4M3 : kFmtBF16);
            constexpr uint32_t idesc_S = make_idesc(fmt, 128, BC, false, false);
            constexpr uint32_t idesc_PV = make_idesc(fmt, 128, D, false, true);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV);
            const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;
            auto issue_S = [&](int j) {
                const int s = j % ST;
                const uint64_t qd = make_sdesc(q_base + t * C::Q_TILE, 16, 1024, 2);
                const uint64_t kd = make_sdesc(k_base + s * C::KV_TILE, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < D / C::KPER; ++kk) {
                    const uint32_t inbox = (kk & 3) * 32;
                    umma_ss<C::EB>(tmem + 128 * t, sdesc_advance(qd, (kk >> 2) * C::Q_BOX + inbox),
                                     sdesc_advance(kd, (kk >> 2) * C::KV_BOX + inbox), idesc_S, kk > 0);
                }
                umma_commit(&bar_S[t]);
            };
            auto issue_PV = [&](int j, int half) {
                const int s = j % ST;
                const uint64_t vd = make_sdesc(v_base + s * C::KV_TILE, C::KV_BOX, C::V_SBO, C::V_LAYOUT);
                constexpr int KH = BC / C::KPER / 2;
#pragma unroll
                for (int k2 = 0; k2 < KH; ++k2) {
                    const int kk = half * KH + k2;
                    umma_ts<C::EB>(tmem + 256 + 128 * t, tmem + 128 * t + 8 * kk,
                                     sdesc_advance(vd, kk * C::KPER * 128), idesc_PV, (j > 0 || kk > 0));
                }
            };
            mbar_wait(&bar_Kf[0], 0);
            mbar_wait(&bar_Q[t], 0);
            tc_fence_after();
            issue_S(0);
            umma_commit(&bar_Ke[0]);
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait(&bar_Vf[s], (j / ST) & 1);
                if (j + 1 < nkv) mbar_wait(&bar_Kf[(j + 1) % ST], ((j + 1) / ST) & 1);
                if (j < nkv_t) {
                    if (C::SPLITP) {
                        mbar_wait(&bar_Pa[t], j & 1);
                        tc_fence_after();
                        issue_PV(j, 0);
                    }
                    mbar_wait(&bar_Pb[t], j & 1);
                    tc_fence_after();
                    if (!C::SPLITP) issue_PV(j, 0);
                    issue_PV(j, 1);
                }
                umma_commit(&bar_Ve[s]);                 // (an issuer past its own last tile has nothing pending)
                if (j + 1 < nkv_t) issue_S(j + 1);
                else if (j + 1 == nkv_t) umma_commit(&bar_O[t]);
                if (j + 1 < nkv) umma_commit(&bar_Ke[(j + 1) % ST]);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        if constexpr (NT == 2) reg_dec<C::REGS_CTRL>();
        if (elect_one()) {
            constexpr uint32_t fmt = C::kF32 ? kFmtTF32 : (C::kF8 ? kFmtE4M3 : kFmtBF16);
            constexpr uint32_t idesc_S = make_idesc(fmt, 128, BC, false, false);
            constexpr uint32_t idesc_PV = make_idesc(fmt, 128, D, false, true);
            const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV);

            auto issue_S = [&](int t, int j) {
                const int s = j % ST;
                const uint64_t qd = make_sdesc(q_base + t * C::Q_TILE, 16, 1024, 2);
                const uint64_t kd = make_sdesc(k_base + s * C::KV_TILE, 16, 1024, 2);
#pragma unroll
                for (int kk = 0; kk < D / C::KPER; ++kk) {
                    const uint32_t inbox = (kk & 3) * 32;
                    umma_ss<C::EB>(tmem + 128 * t, sdesc_advance(qd, (kk >> 2) * C::Q_BOX + inbox),
                                     sdesc_advance(kd, (kk >> 2) * C::KV_BOX + inbox), idesc_S, kk > 0);
                }
                umma_commit(&bar_S[t]);
            };
            auto issue_PV = [&](int t, int j, int half) {       // half 0 / 1 = first / second BC/2 columns of P
                const int s = j % ST;
                const uint64_t vd = make_sdesc(v_base + s * C::KV_TILE, C::KV_BOX, C::V_SBO, C::V_LAYOUT);
                constexpr int KH = BC / C::KPER / 2;
#pragma unroll
                for (int k2 = 0; k2 < KH; ++k2) {
                    const int kk = half * KH + k2;
                    umma_ts<C::EB>(tmem + 256 + 128 * t, tmem + 128 * t + 8 * kk,
                                     sdesc_advance(vd, kk * C::KPER * 128), idesc_PV, (j > 0 || kk > 0));
                }
            };

            mbar_wait(&bar_Kf[0], 0);
            for (int t = 0; t < ntiles; ++t) {
                mbar_wait(&bar_Q[t], 0);
                tc_fence_after();
                issue_S(t, 0);
            }
            umma_commit(&bar_Ke[0]);
            for (int j = 0; j < nkv; ++j) {
                const int s = j % ST;
                mbar_wait(&bar_Vf[s], (j / ST) & 1);
                if (j + 1 < nkv) mbar_wait(&bar_Kf[(j + 1) % ST], ((j + 1) / ST) & 1);
                for (int t = 0; t < ntiles; ++t) {
                    const int nkv_t = (t == ntiles - 1) ? nkv : nkv_first;   // (NT <= 2)
                    if (j >= nkv_t) continue;
                    if (C::SPLITP) {
                        mbar_wait(&bar_Pa[t], j & 1);
                        tc_fence_after();
                        issue_PV(t, j, 0);
                    }
                    mbar_wait(&bar_Pb[t], j & 1);
                    tc_fence_after();
                    B200_TR(2, j, 2 * t);
                    if (!C::SPLITP) issue_PV(t, j, 0);
                    issue_PV(t, j, 1);
                    if (t == ntiles - 1) umma_commit(&bar_Ve[s]);
                    if (j + 1 < nkv_t) {
                        issue_S(t, j + 1);
                        if (t == ntiles - 1) umma_commit(&bar_Ke[(j + 1) % ST]);
                    } else {
                        umma_commit(&bar_O[t]);
                    }
                    B200_TR(2, j, 2 * t + 1);
                }
            }
        }
    } else if (warp < 4) {
        if constexpr (NT == 2) reg_dec<C::REGS_CTRL>();
    } else {
        // ================================ softmax warpgroups ==========================
        if constexpr (NT == 2) reg_inc<C::REGS_SOFTMAX>();
        const int t = (warp - 4) >> 2;
        if (t < ntiles) {
            const int nkv_t = nkv_of(t);
            // tiles entirely below the diagonal (and not the ragged last one) need no mask
            const int n_plain = causal ? min(nkv_t, min((q0 + t * 128 + causal) / BC, Nkv / BC)) : nkv_t - 1;
            softmax_role<C>(t, tmem, sQ, &bar_S[t], &bar_Pa[t], &bar_Pb[t], &bar_O[t], &tmO, Lout, ldL, Nq, Nkv, nkv_t, n_plain,
                            causal, q0, h, b, bh, scale_log2);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------ host side
template <class C>
static int launch_fwd(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    // the opt-in shared-memory size is a per-device attribute of the function: set it once on every device used
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fwd_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const CUtensorMapDataType dt = C::kF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (C::kF8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
    const CUtensorMapDataType odt = C::kF8 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : dt;     // fp8 in, bf16 out
    const CUtensorMapSwizzle v_swz = C::kF32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    const float scale_log2 = a.scale * 1.4426950408889634f;
    CUtensorMap tmQ, tmK, tmV, tmO;
    int trc = 0;
    if ((trc = make_tmap_4d(&tmQ, a.Q, dt, C::EB, a.B, a.H, a.Nq, C::D, a.lq.sb, a.lq.sh, a.lq.sn, 128, C::CPB)) ||
        (trc = make_tmap_4d(&tmK, a.K, dt, C::EB, a.B, a.Hkv, a.Nkv, C::D, a.lk.sb, a.lk.sh, a.lk.sn, C::BC, C::CPB)) ||
        (trc = make_tmap_4d(&tmV, a.V, dt, C::EB, a.B, a.Hkv, a.Nkv, C::D, a.lv.sb, a.lv.sh, a.lv.sn, C::BC, C::CPB, v_swz)) ||
        (trc = make_tmap_4d(&tmO, a.O, odt, C::OEB, a.B, a.H, a.Nq, C::D, a.lo.sb, a.lo.sh, a.lo.sn, 128, C::OCPB)))
        return trc == -2 ? B200_ERR_UNSUPPORTED : B200_ERR_CUDA;   // -2: a stride is zero or not a multiple of 16 bytes
    dim3 grid((a.Nq + C::NT * 128 - 1) / (C::NT * 128), a.H, a.B);
    fwd_kernel<C><<<grid, C::THREADS, C::SMEM, st>>>(tmQ, tmK, tmV, tmO, a.L, a.ldL, a.Nq, a.Nkv, scale_log2, a.causal ? 1 + (a.Nkv - a.Nq) : 0,
                                                     a.H / a.Hkv, BH_valid);
    B200_CUDA_TRY(cudaGetLastError());
    count_launch();
    return B200_OK;
}

#ifdef B200_TRACE
int fwd_trace_read(long long* out, int n)
{
    if (n > 3 * 8 * 8) n = 3 * 8 * 8;
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_trace, n * sizeof(long long)));
    return n;
}
#endif

bool tc_fwd_supported(int dtype, int d, int Nq, int Nkv)
{
    if (Nq < 1 || Nkv < 1) return false;
    if (dtype == B200_DTYPE_FP8_E4M3) return d == 128;         // forward only: fp8 Q/K/V in, bf16 O out
    if (dtype != B200_DTYPE_BF16 && dtype != B200_DTYPE_F32) return false;
    return d == 64 || d == 128;
}

// Shipped configuration, selected on shape and dtype only:
//   bf16 d = 128 : fwd_dbuf_kernel<4> (fwd_dbuf_sm100.cu: 64-key tiles, double-buffered scores, one issuing thread per
//                  tile, 4 of 16 exp2 pairs on the FMA pipe) -- promoted in r2 after the full -m gpu suite, +5 % burst /
//                  +11 % in the bench loop over this file's 128-key kernel (profiles/r2a_*);
//   bf16 d = 64, fp32 (tf32 MMAs) d = 64 / 128 : fwd_kernel<FwdCfg<...>> of this file.
// The measured-negative variants (split-S, CTA pairs, shared score buffer, two issuers on this kernel, POLY / SPLITP
// sweeps) live in csrc/experiments/ and are only compiled into `make EXPERIMENTS=1` side libraries, where
// B200_FWD_VARIANT selects them for A/B timing; the product library has no environment hooks.
struct fp8_e4m3_t { uint8_t x; };       // (an 8-bit element tag for FwdCfg; the kernel never does arithmetic on it)

static int dispatch_fwd(const FwdArgs& a, int BH_valid, cudaStream_t st)
{
    if (a.dtype == B200_DTYPE_FP8_E4M3) return launch_fwd<FwdCfg<fp8_e4m3_t, 128, 2, 128>>(a, BH_valid, st);
    if (a.dtype == B200_DTYPE_BF16) {
        if (a.d == 128) {
#ifdef B200_EXPERIMENTS
            static int variant = -1;
            if (variant < 0) { const char* e = getenv("B200_FWD_VARIANT"); variant = e ? atoi(e) : 4; }
            if (variant == 2) return launch_fwd_split(a, BH_valid, st);       // split-S pipeline
            if (variant == 3) {                                               // CTA pairs (cta_group::2)
                if (fwd_pair_eligible(a)) return launch_fwd_pair(a, BH_valid, st);
                return launch_fwd<FwdCfg<__nv_bfloat16, 128, 2, 128, 0, true>>(a, BH_valid, st);
            }
            if (variant == 6) return launch_fwd_shs(a, BH_valid, st);         // one shared score buffer, P apart
            if (variant == 5)                                                 // one MMA-issuing thread per tile
                return launch_fwd<FwdCfg<__nv_bfloat16, 128, 2, 128, 0, true, 2>>(a, BH_valid, st);
            if (variant != 4) {
                switch (variant) {                                            // 10 * POLY + SPLITP; 1 = the r1 default
#define B200_V(P, S) case 10 * P + S: return launch_fwd<FwdCfg<__nv_bfloat16, 128, 2, 128, P, S != 0>>(a, BH_valid, st);
                    B200_V(0, 0) B200_V(0, 1) B200_V(4, 1) B200_V(6, 0) B200_V(6, 1) B200_V(8, 1) B200_V(10, 1)
#undef B200_V
                    default: return B200_ERR_UNSUPPORTED;
                }
            }
#endif
            return launch_fwd_dbuf(a, BH_valid, st);
        }
        return launch_fwd<FwdCfg<__nv_bfloat16, 64, 2, 128, 0, true>>(a, BH_valid, st);
    }
    if (a.d == 128) return launch_fwd<FwdCfg<float, 128, 1, 64>>(a, BH_valid, st);
    return launch_fwd<FwdCfg<float, 64, 2, 128>>(a, BH_valid, st);
}

// contiguous [BH, N, d] slabs (the reference layout)
int tc_fwd(const void* Q, const void* K, const void* V, void* O, float* L, int BH, int Nq, int Nkv, int d,
           float scale, int dtype, cudaStream_t st, bool causal)
{
    if (causal && Nq > Nkv) return B200_ERR_INVALID_ARGUMENT;
    const uintptr_t align = (uintptr_t)Q | (uintptr_t)K | (uintptr_t)V | (uintptr_t)O;
    if ((align & 15) != 0 || !(scale > 0.0f))      // TMA needs 16-byte bases; max(S)*scale needs scale > 0
        return simt_fwd(Q, K, V, O, L, nullptr, BH, Nq, Nkv, d, scale, dtype, 0, st, causal ? Nkv - Nq : kNoCausal);
    // the slabs are independent: cut BH into (batches x heads) that fit the grid limits; a ragged last batch is
    // described by the map but its CTAs exit (bh >= BH)
    const int Hm = BH <= 65535 ? BH : 32768, Bm = (BH + Hm - 1) / Hm;
    FwdArgs a{Q, K, V, O, L, lay_contiguous(Hm, Nq, d), lay_contiguous(Hm, Nkv, d), lay_contiguous(Hm, Nkv, d),
              lay_contiguous(Hm, Nq, d), (long long)Nq, Bm, Hm, Hm, Nq, Nkv, d, scale, dtype, causal};
    return dispatch_fwd(a, BH, st);
}

int tc_fwd_strided(const FwdArgs& a, cudaStream_t st)
{
    if (!tc_fwd_supported(a.dtype, a.d, a.Nq, a.Nkv) || !(a.scale > 0.0f)) return B200_ERR_UNSUPPORTED;
    if (a.causal && a.Nq > a.Nkv) return B200_ERR_INVALID_ARGUMENT;
    if (a.B < 1 || a.H < 1 || a.Hkv < 1 || a.H % a.Hkv != 0 || a.B > 65535 || a.H > 65535) return B200_ERR_INVALID_ARGUMENT;
    const uintptr_t align = (uintptr_t)a.Q | (uintptr_t)a.K | (uintptr_t)a.V | (uintptr_t)a.O;
    if ((align & 15) != 0) return B200_ERR_UNSUPPORTED;
    return dispatch_fwd(a, a.B * a.H, st);
}

}  // namespace b200
