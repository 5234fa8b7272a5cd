// ptx.cuh -- thin inline-PTX wrappers for the sm_100a features the attention kernels use:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld / st / fences),
// setmaxnreg, named barriers, and the UMMA shared-memory / instruction descriptors.
//
// Descriptor bit layouts follow the PTX ISA "tcgen05 matrix descriptor" / "instruction
// descriptor" tables (the same fields CuTe names SmemDescriptor / InstrDescriptor).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>

namespace b200 {

#ifndef B200_MBAR_TIMEOUT_CYCLES
// A wait that spins longer than this many SM cycles traps instead of hanging the GPU
// (≈4 s at 1.9 GHz).  0 disables the watchdog.
#define B200_MBAR_TIMEOUT_CYCLES 8000000000ll
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "elect.sync _|P1, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(pred));
    return pred != 0;
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    // generic-proxy smem writes -> visible to the async proxy (TMA store, UMMA operand reads)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
#ifndef B200_MBAR_SUSPEND_NS
// suspend-time hint of mbarrier.try_wait: the hardware may park the polling thread for up to this
// long before it reports "not yet", so a waiting warp costs (almost) no issue slots or power.
#define B200_MBAR_SUSPEND_NS 0
#endif
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
#if B200_MBAR_SUSPEND_NS > 0
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"((uint32_t)B200_MBAR_SUSPEND_NS) : "memory");
#else
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
#endif
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
#if B200_MBAR_TIMEOUT_CYCLES
    // The watchdog reads the clock only once every 4096 polls: the polling loop shares issue slots
    // with the working warps of its SM sub-partition, so it has to stay two instructions long.
    uint32_t polls = 0;
    long long t0 = 0;
#endif
    while (!mbar_try_wait(bar, parity)) {
#if B200_MBAR_TIMEOUT_CYCLES
        if ((++polls & 0xFFFu) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > B200_MBAR_TIMEOUT_CYCLES) {
                printf("b200attn: mbarrier wait timed out (block %d,%d thread %d bar@%u parity %u)\n",
                       blockIdx.x, blockIdx.y, threadIdx.x, smem_u32(bar), parity);
                __trap();
            }
        }
#endif
    }
}

// ------------------------------------------------------------------ named barriers
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ setmaxnreg
template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// ------------------------------------------------------------------ TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// L2 cache-policy constants (same encodings CUTLASS uses for CacheHintSm90)
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictFirst  = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast   = 0x14F0000000000000ull;

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, int c2, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)),
          "r"(c0), "r"(c1), "r"(c2), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, int c2, int c3, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)),
          "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
        ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
        ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// fp32 element-wise add-reduction of an smem tile into global memory (dQ accumulation)
__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
    asm volatile(
        "cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
        ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ------------------------------------------------------------------ tcgen05: TMEM management
template <int NCOLS> __device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS> __device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// arrive (count 1) on an mbarrier once every tcgen05.mma issued so far by this thread is complete
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------ tcgen05: MMA
// D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// kind::f8f6f4 (fp8 e4m3 / e5m2 operands, K = 32 per instruction, fp32 accumulation)
__device__ __forceinline__ void umma_f8_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_f8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// Instruction descriptor (32 bit).  Fields: [4,6) D format (1 = f32); [7,10) A format, [10,13) B format
// (kind::f16: 0 = f16, 1 = bf16; kind::tf32: 2 = tf32); [13] negate A; [14] negate B;
// [15] A major (0 = K-major, 1 = MN-major); [16] B major; [17,23) N>>3; [24,29) M>>4.
enum : uint32_t { kFmtF16 = 0, kFmtBF16 = 1, kFmtTF32 = 2, kFmtE4M3 = 0, kFmtE5M2 = 1 };   // (E4M3 / E5M2: kind::f8f6f4)
__host__ __device__ constexpr uint32_t make_idesc(uint32_t fmt_ab, int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | (fmt_ab << 7) | (fmt_ab << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

// Shared-memory matrix descriptor (64 bit), 128-byte swizzle.
//   [0,14)  start address >> 4          [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset >> 4 [46,48) version = 1 (sm_100)   [61,64) layout: 2 = SWIZZLE_128B
// K-major operand  (rows of 128 B = one swizzle span along K; 8-row groups 1024 B apart): SBO = 1024, LBO unused.
// MN-major operand (rows of 128 B = 64/32 elements along MN; 8 K-rows per 1024 B group):
//   SBO = byte distance between consecutive 8-row K groups, LBO = byte distance between
//   consecutive 128-byte MN chunks.
// layout: 2 = SWIZZLE_128B (16-byte atoms); 1 = SWIZZLE_128B_BASE32B (32-byte atoms, 4-row pattern: the only
// layout the hardware accepts for MN-major tf32 operands; TMA counterpart CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B).
__host__ __device__ constexpr uint64_t make_sdesc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
    return static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4) |
           (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16) |
           (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32) |
           (1ull << 46) | (static_cast<uint64_t>(layout) << 61);
}
__host__ __device__ constexpr uint64_t make_sdesc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return make_sdesc(smem_addr, lbo_bytes, sbo_bytes, 2);
}
// Advancing the start address of an existing descriptor by `bytes` (bytes % 16 == 0, no carry out of 14 bits)
__host__ __device__ constexpr uint64_t sdesc_advance(uint64_t desc, uint32_t bytes) { return desc + (bytes >> 4); }

// ------------------------------------------------------------------ tcgen05: TMEM <-> registers
// 32 lanes x 32 bit, N consecutive columns: thread t of the warp gets lane (warp%4)*32 + t.
#define B200_TMEM_LD_32(taddr, r)                                                                         \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"                                \
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"               \
                 : "=r"((r)[0]), "=r"((r)[1]), "=r"((r)[2]), "=r"((r)[3]), "=r"((r)[4]), "=r"((r)[5]),    \
                   "=r"((r)[6]), "=r"((r)[7]), "=r"((r)[8]), "=r"((r)[9]), "=r"((r)[10]), "=r"((r)[11]),  \
                   "=r"((r)[12]), "=r"((r)[13]), "=r"((r)[14]), "=r"((r)[15]), "=r"((r)[16]),             \
                   "=r"((r)[17]), "=r"((r)[18]), "=r"((r)[19]), "=r"((r)[20]), "=r"((r)[21]),             \
                   "=r"((r)[22]), "=r"((r)[23]), "=r"((r)[24]), "=r"((r)[25]), "=r"((r)[26]),             \
                   "=r"((r)[27]), "=r"((r)[28]), "=r"((r)[29]), "=r"((r)[30]), "=r"((r)[31])              \
                 : "r"(taddr))

#define B200_TMEM_LD_16(taddr, r)                                                                         \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "                                                \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"                        \
                 : "=r"((r)[0]), "=r"((r)[1]), "=r"((r)[2]), "=r"((r)[3]), "=r"((r)[4]), "=r"((r)[5]),    \
                   "=r"((r)[6]), "=r"((r)[7]), "=r"((r)[8]), "=r"((r)[9]), "=r"((r)[10]), "=r"((r)[11]),  \
                   "=r"((r)[12]), "=r"((r)[13]), "=r"((r)[14]), "=r"((r)[15])                             \
                 : "r"(taddr))

#define B200_TMEM_ST_32(taddr, r)                                                                         \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "                                          \
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"                               \
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"                      \
                 ::"r"(taddr), "r"((r)[0]), "r"((r)[1]), "r"((r)[2]), "r"((r)[3]), "r"((r)[4]),           \
                   "r"((r)[5]), "r"((r)[6]), "r"((r)[7]), "r"((r)[8]), "r"((r)[9]), "r"((r)[10]),         \
                   "r"((r)[11]), "r"((r)[12]), "r"((r)[13]), "r"((r)[14]), "r"((r)[15]), "r"((r)[16]),    \
                   "r"((r)[17]), "r"((r)[18]), "r"((r)[19]), "r"((r)[20]), "r"((r)[21]), "r"((r)[22]),    \
                   "r"((r)[23]), "r"((r)[24]), "r"((r)[25]), "r"((r)[26]), "r"((r)[27]), "r"((r)[28]),    \
                   "r"((r)[29]), "r"((r)[30]), "r"((r)[31])                                               \
                 : "memory")

#define B200_TMEM_ST_16(taddr, r)                                                                         \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "                                          \
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"                              \
                 ::"r"(taddr), "r"((r)[0]), "r"((r)[1]), "r"((r)[2]), "r"((r)[3]), "r"((r)[4]),           \
                   "r"((r)[5]), "r"((r)[6]), "r"((r)[7]), "r"((r)[8]), "r"((r)[9]), "r"((r)[10]),         \
                   "r"((r)[11]), "r"((r)[12]), "r"((r)[13]), "r"((r)[14]), "r"((r)[15])                   \
                 : "memory")

#define B200_TMEM_ST_8(taddr, r)                                                                          \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"                 \
                 ::"r"(taddr), "r"((r)[0]), "r"((r)[1]), "r"((r)[2]), "r"((r)[3]), "r"((r)[4]),           \
                   "r"((r)[5]), "r"((r)[6]), "r"((r)[7]) : "memory")

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));   // d.hi = cvt(a), d.lo = cvt(b)
    return r;
}
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ------------------------------------------------------------------ packed fp32x2 arithmetic (sm_100: FFMA2 / FADD2 / FMUL2)
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// 2^x for a pair of fp32 values on the FMA pipe instead of the MUFU: x = n + f with n = round(x),
// f in [-0.5, 0.5]; 2^f by a degree-3 minimax polynomial (max relative error 7.5e-5, far below the
// bf16 rounding the result goes through), 2^n by adding n to the exponent field.  x is clamped at
// -125 so the exponent arithmetic cannot wrap (2^-125 is zero for every purpose here).
__device__ __forceinline__ void exp2_poly2(uint64_t x, float& p_lo, float& p_hi) {
    float lo, hi;
    f2_unpack(x, lo, hi);
    lo = fmaxf(lo, -125.0f);
    hi = fmaxf(hi, -125.0f);
    x = f2_pack(lo, hi);
    const uint64_t t = f2_add(x, f2_pack(12582912.0f, 12582912.0f));            // 1.5 * 2^23: low mantissa bits = n
    const uint64_t n = f2_add(t, f2_pack(-12582912.0f, -12582912.0f));
    const uint64_t f = f2_fma(n, f2_pack(-1.0f, -1.0f), x);
    uint64_t p = f2_fma(f2_pack(0.05517164617776871f, 0.05517164617776871f), f, f2_pack(0.2426111251115799f, 0.2426111251115799f));
    p = f2_fma(p, f, f2_pack(0.6932609677314758f, 0.6932609677314758f));
    p = f2_fma(p, f, f2_pack(0.9999280571937561f, 0.9999280571937561f));
    float tl, th, pl, ph;
    f2_unpack(t, tl, th);
    f2_unpack(p, pl, ph);
    p_lo = __uint_as_float((__float_as_uint(tl) << 23) + __float_as_uint(pl));
    p_hi = __uint_as_float((__float_as_uint(th) << 23) + __float_as_uint(ph));
}

// ------------------------------------------------------------------ optional in-kernel event trace
// Build with EXTRA=-DB200_TRACE (make TRACE=1): CTA (0,0) records clock64() into arr[role][j-16][ev]
// for loop iterations 16..23 at the hand-off points of a kernel; `who` selects the recording warp
// (lane 0 of it writes).  Read back with b200_debug_trace_read; tools/kbench --trace prints it.
// Compiled out otherwise.
#ifdef B200_TRACE
#define B200_TRACE_EVENT(arr, role, j, ev, who)                                                         \
    do { if (blockIdx.x == 0 && blockIdx.y == 0 && (j) >= 16 && (j) < 24 && (threadIdx.x & 31) == 0 && (who)) \
            arr[((role) * 8 + ((j) - 16)) * 8 + (ev)] = clock64(); } while (0)
#else
#define B200_TRACE_EVENT(arr, role, j, ev, who) do {} while (0)
#endif

}  // namespace b200
