// umma_probe.cu -- single-CTA unit probes of the sm_100a building blocks the attention kernels
// rely on: TMA SWIZZLE_128B loads, UMMA shared-memory descriptors (K-major and MN-major),
// A-operand-from-TMEM, tcgen05.ld/st lane mapping, kind::f16 (bf16) and kind::tf32.
//
// Each test multiplies small exactly-representable integers, so the expected error is 0.
// Usage: umma_probe <test-id>   (one test per process: a bad descriptor can poison the context)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>
#include <cuda_fp8.h>
#include "../cuda_flashattention_b200/csrc/ptx.cuh"
#include "../cuda_flashattention_b200/csrc/tmap.h"

using namespace b200;

struct ProbeParams {
    int elem_bytes;        // 2 = bf16, 4 = fp32 (tf32 MMA)
    int a_mode;            // 0 = smem via TMA, 1 = TMEM (tcgen05.st by the threads), 2 = smem written by threads (K-major SW128)
    int a_boxes, a_box_rows, b_boxes, b_box_rows;
    uint32_t a_lbo, a_sbo, a_kadv, a_steps_per_box, a_box_stride;
    uint32_t b_lbo, b_sbo, b_kadv, b_steps_per_box, b_box_stride;
    uint32_t idesc, a_layout, b_layout;
    int ksteps, N, Ktot;
    int a_tmem_cols_per_step;
    const void* A_gmem;    // row-major [128][Ktot] (modes 1, 2)
    float* D;              // [128][N]
};

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, ProbeParams p)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                       // up to 64 KB
    uint8_t* sB = smem + 65536;               // up to 64 KB
    __shared__ uint64_t bar_load, bar_mma;
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_mma, 1);
        fence_mbar_init();
    }
    if (warp == 0) { tmem_alloc<512>(&tmem_base_s); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t tmem_D = tmem;             // columns [0, N)
    const uint32_t tmem_A = tmem + 256;       // A operand staging for the TS mode

    if (tid == 0) {
        uint32_t bytes = 0;
        if (p.a_mode == 0) bytes += p.a_boxes * p.a_box_rows * 128;
        bytes += p.b_boxes * p.b_box_rows * 128;
        mbar_expect_tx(&bar_load, bytes);
        int cols_per_box = 128 / p.elem_bytes;
        if (p.a_mode == 0)
            for (int b = 0; b < p.a_boxes; ++b)
                tma_load_3d(sA + b * p.a_box_stride, &tmA, &bar_load, b * cols_per_box, 0, 0, kEvictNormal);
        for (int b = 0; b < p.b_boxes; ++b)
            tma_load_3d(sB + b * p.b_box_stride, &tmB, &bar_load, b * cols_per_box, 0, 0, kEvictNormal);
    }

    if (p.a_mode == 1) {
        // thread = row; write the row into TMEM columns (bf16: packed pairs; fp32: one per column)
        const uint32_t lane_addr = tmem_A + (uint32_t(warp * 32) << 16);
        if (p.elem_bytes == 2) {
            const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(p.A_gmem) + (size_t)tid * p.Ktot;
            for (int c0 = 0; c0 < p.Ktot / 2; c0 += 16) {
                uint32_t r[16];
                for (int i = 0; i < 16; ++i)
                    r[i] = pack_bf16x2(__bfloat162float(a[2 * (c0 + i)]), __bfloat162float(a[2 * (c0 + i) + 1]));
                B200_TMEM_ST_16(lane_addr + c0, r);
            }
        } else if (p.elem_bytes == 1) {
            // fp8: four K elements per 32-bit column, lowest K index in the lowest byte
            const uint32_t* a = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(p.A_gmem) + (size_t)tid * p.Ktot);
            for (int c0 = 0; c0 < p.Ktot / 4; c0 += 16) {
                uint32_t r[16];
                for (int i = 0; i < 16; ++i) r[i] = a[c0 + i];
                B200_TMEM_ST_16(lane_addr + c0, r);
            }
        } else {
            const float* a = reinterpret_cast<const float*>(p.A_gmem) + (size_t)tid * p.Ktot;
            for (int c0 = 0; c0 < p.Ktot; c0 += 16) {
                uint32_t r[16];
                for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(a[c0 + i]);
                B200_TMEM_ST_16(lane_addr + c0, r);
            }
        }
        tmem_wait_st();
    } else if (p.a_mode == 2) {
        // thread = row r; K-major SWIZZLE_128B: 16-byte chunk c of row r lives at chunk (c ^ (r & 7))
        const uint8_t* a = reinterpret_cast<const uint8_t*>(p.A_gmem) + (size_t)tid * p.Ktot * p.elem_bytes;
        const int row_bytes = p.Ktot * p.elem_bytes;
        for (int off = 0; off < row_bytes; off += 16) {
            int box = off / 128, chunk = (off % 128) / 16;
            uint4 v = *reinterpret_cast<const uint4*>(a + off);
            *reinterpret_cast<uint4*>(sA + box * p.a_box_stride + tid * 128 + ((chunk ^ (tid & 7)) * 16)) = v;
        }
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (tid == 0) {
        mbar_wait(&bar_load, 0);
        tc_fence_after();
        const uint64_t adesc0 = make_sdesc(smem_u32(sA), p.a_lbo, p.a_sbo, p.a_layout);
        const uint64_t bdesc0 = make_sdesc(smem_u32(sB), p.b_lbo, p.b_sbo, p.b_layout);
        for (int s = 0; s < p.ksteps; ++s) {
            uint32_t aoff = (s / p.a_steps_per_box) * p.a_box_stride + (s % p.a_steps_per_box) * p.a_kadv;
            uint32_t boff = (s / p.b_steps_per_box) * p.b_box_stride + (s % p.b_steps_per_box) * p.b_kadv;
            uint64_t bd = sdesc_advance(bdesc0, boff);
            if (p.a_mode == 1) {
                uint32_t at = tmem_A + s * p.a_tmem_cols_per_step;
                if (p.elem_bytes == 2) umma_f16_ts(tmem_D, at, bd, p.idesc, s > 0);
                else if (p.elem_bytes == 1) umma_f8_ts(tmem_D, at, bd, p.idesc, s > 0);
                else                   umma_tf32_ts(tmem_D, at, bd, p.idesc, s > 0);
            } else {
                uint64_t ad = sdesc_advance(adesc0, aoff);
                if (p.elem_bytes == 2) umma_f16_ss(tmem_D, ad, bd, p.idesc, s > 0);
                else if (p.elem_bytes == 1) umma_f8_ss(tmem_D, ad, bd, p.idesc, s > 0);
                else                   umma_tf32_ss(tmem_D, ad, bd, p.idesc, s > 0);
            }
        }
        umma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    {
        const uint32_t lane_addr = tmem_D + (uint32_t(warp * 32) << 16);
        for (int c0 = 0; c0 < p.N; c0 += 16) {
            uint32_t r[16];
            B200_TMEM_LD_16(lane_addr + c0, r);
            tmem_wait_ld();
            for (int i = 0; i < 16; ++i) p.D[(size_t)tid * p.N + c0 + i] = __uint_as_float(r[i]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

static float rnd_small(unsigned& s) { s = s * 1664525u + 1013904223u; return (float)((int)((s >> 16) % 9) - 4) * 0.25f; }

struct Case {
    const char* name;
    int elem_bytes, a_mode;
    bool a_mn, b_mn;          // operand storage: false = [rows=MN][cols=K] (K-major), true = [rows=K][cols=MN] (MN-major)
    int N, K;
    bool swap_a, swap_b;      // alternative hypothesis: swap LBO/SBO roles
    int atom32;               // MN-major operands use the 32-byte-atom 128B swizzle; value = SBO in bytes (0 = off)
    int atomk_a, atomk_b;     // r2: a K-MAJOR operand stored with the 32-byte-atom swizzle (layout 1); value = SBO in bytes
};

int main(int argc, char** argv)
{
    const Case cases[] = {
        {"bf16 NT  SS  (S=QK^T)            N=128 K=128", 2, 0, false, false, 128, 128, false, false},
        {"bf16 NN  SS  (O=PV, V MN-major)  N=128 K=128", 2, 0, false, true, 128, 128, false, false},
        {"bf16 NN  SS  V MN-major, LBO/SBO swapped    ", 2, 0, false, true, 128, 128, false, true},
        {"bf16 NN  TS  (P in TMEM)         N=128 K=128", 2, 1, false, true, 128, 128, false, false},
        {"bf16 NN  SS  A written by threads (swizzle) ", 2, 2, false, true, 128, 128, false, false},
        {"bf16 TN  SS  A MN-major          N=128 K=128", 2, 0, true, false, 128, 128, false, false},
        {"bf16 TN  SS  A MN-major, LBO/SBO swapped    ", 2, 0, true, false, 128, 128, true, false},
        {"tf32 NT  SS                      N=128 K=64 ", 4, 0, false, false, 128, 64, false, false},
        {"tf32 NN  SS  (V MN-major)        N=128 K=64 ", 4, 0, false, true, 128, 64, false, false},
        {"tf32 NN  TS  (P in TMEM)         N=128 K=64 ", 4, 1, false, true, 128, 64, false, false},
        {"tf32 TN  SS  A MN-major          N=128 K=64 ", 4, 0, true, false, 128, 64, false, false},
        {"bf16 NN  SS  N=64 (d=64)               K=128", 2, 0, false, true, 64, 128, false, false},
        {"bf16 TT  SS  A,B MN-major        N=64  K=128", 2, 0, true, true, 64, 128, false, false},
        {"bf16 NT  SS  N=256                     K=64 ", 2, 0, false, false, 256, 64, false, false},
        {"bf16 NT  SS  N=64 K=64 (d=64 S tile)        ", 2, 0, false, false, 64, 64, false, false},
        {"tf32 NN  SS  N=64 (V MN-major)         K=64 ", 4, 0, false, true, 64, 64, false, false},
        {"tf32 NN  SS  V MN-major ATOM_32B SBO=512    ", 4, 0, false, true, 128, 64, false, false, 512},
        {"tf32 NN  SS  V MN-major ATOM_32B SBO=1024   ", 4, 0, false, true, 128, 64, false, false, 1024},
        {"tf32 NN  TS  P in TMEM, V ATOM_32B SBO=512  ", 4, 1, false, true, 128, 64, false, false, 512},
        {"tf32 TN  SS  A MN-major ATOM_32B SBO=512    ", 4, 0, true, false, 128, 64, false, false, 512},
        {"tf32 NN  SS  N=64 V ATOM_32B SBO=512   K=128", 4, 0, false, true, 64, 128, false, false, 512},
        {"tf32 NT  TS  P in TMEM, B K-major           ", 4, 1, false, false, 128, 64, false, false, 0},
        // r2 (tf32 backward): can ONE shared-memory tile serve as a K-major operand of one MMA and as an MN-major operand
        // of another?  MN-major tf32 needs the ATOM_32B swizzle, so the K-major view has to read that layout too.
        {"tf32 NT  SS  B K-major ATOM_32B SBO=1024    ", 4, 0, false, false, 128, 64, false, false, 0, 0, 1024},
        {"tf32 NT  SS  A K-major ATOM_32B SBO=1024    ", 4, 0, false, false, 128, 64, false, false, 0, 1024, 0},
        {"tf32 NT  SS  A,B K-major ATOM_32B SBO=1024  ", 4, 0, false, false, 128, 64, false, false, 0, 1024, 1024},
        {"tf32 NT  SS  B K-major ATOM_32B SBO=512     ", 4, 0, false, false, 128, 64, false, false, 0, 0, 512},
        {"tf32 NT  SS  N=64 A,B K-major ATOM_32B      ", 4, 0, false, false, 64, 64, false, false, 0, 1024, 1024},
        {"tf32 TT  SS  N=64 A,B MN-major ATOM_32B 512 ", 4, 0, true, true, 64, 128, false, false, 512},
        // r2 (fp8 forward, kind::f8f6f4 e4m3): K-major, MN-major B (the V operand of O += P V), A from TMEM (P)
        {"fp8  NT  SS  (S=QK^T)            N=128 K=128", 1, 0, false, false, 128, 128, false, false},
        {"fp8  NN  SS  (V MN-major)        N=128 K=128", 1, 0, false, true, 128, 128, false, false},
        {"fp8  NN  TS  (P in TMEM, V MN)   N=128 K=128", 1, 1, false, true, 128, 128, false, false},
        {"fp8  NN  SS  V MN-major, LBO/SBO swapped    ", 1, 0, false, true, 128, 128, false, true},
        {"fp8  NT  TS  (P in TMEM, B K-major)   K=128 ", 1, 1, false, false, 128, 128, false, false},
    };
    const int ncases = sizeof(cases) / sizeof(cases[0]);
    if (argc < 2) { printf("%d\n", ncases); return 0; }
    const int id = atoi(argv[1]);
    if (id < 0 || id >= ncases) return 1;
    const Case& c = cases[id];
    const int M = 128, N = c.N, K = c.K, eb = c.elem_bytes;
    const int cpb = 128 / eb;          // elements per 128-byte row
    const int kper = 32 / eb;          // K elements per MMA (16 for bf16, 8 for tf32)

    // logical matrices A[M][K], B[N][K] as float; storage per major-ness
    std::vector<float> A((size_t)M * K), B((size_t)N * K);
    unsigned seed = 1234u + id;
    for (auto& v : A) v = rnd_small(seed);
    for (auto& v : B) v = rnd_small(seed);
    auto store = [&](const std::vector<float>& X, int MN, bool mn_major, std::vector<uint8_t>& out, int& rows, int& cols) {
        rows = mn_major ? K : MN; cols = mn_major ? MN : K;
        out.resize((size_t)rows * cols * eb);
        for (int r = 0; r < rows; ++r)
            for (int cc = 0; cc < cols; ++cc) {
                float v = mn_major ? X[(size_t)cc * K + r] : X[(size_t)r * K + cc];
                if (eb == 1) { out[(size_t)r * cols + cc] = (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3); }
                else if (eb == 2) { __nv_bfloat16 h = __float2bfloat16(v); memcpy(&out[((size_t)r * cols + cc) * 2], &h, 2); }
                else memcpy(&out[((size_t)r * cols + cc) * 4], &v, 4);
            }
    };
    std::vector<uint8_t> hA, hB; int ar, ac, br, bc;
    store(A, M, c.a_mn, hA, ar, ac);
    store(B, N, c.b_mn, hB, br, bc);

    void *dA, *dB; float* dD;
    CK(cudaMalloc(&dA, hA.size())); CK(cudaMalloc(&dB, hB.size())); CK(cudaMalloc(&dD, (size_t)M * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xff, (size_t)M * N * 4));

    CUtensorMap tmA, tmB;
    CUtensorMapDataType dt = eb == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUtensorMapSwizzle swA = ((c.atom32 && c.a_mn) || c.atomk_a) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    CUtensorMapSwizzle swB = ((c.atom32 && c.b_mn) || c.atomk_b) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    if (make_tmap_3d(&tmA, dA, dt, eb, 1, ar, ac, ar, cpb, swA) || make_tmap_3d(&tmB, dB, dt, eb, 1, br, bc, br, cpb, swB)) {
        printf("tensor map encode failed\n"); return 2;
    }

    ProbeParams p{};
    p.elem_bytes = eb; p.a_mode = c.a_mode; p.N = N; p.Ktot = K; p.ksteps = K / kper;
    p.A_gmem = dA; p.D = dD;
    auto fill = [&](bool mn_major, int rows, int cols, bool swap, int& boxes, int& box_rows, uint32_t& lbo, uint32_t& sbo,
                    uint32_t& kadv, uint32_t& spb, uint32_t& bstride) {
        boxes = cols / cpb; box_rows = rows; bstride = rows * 128;
        if (!mn_major) {            // K-major: rows = MN, each box = one 128-byte K span
            lbo = 16; sbo = 1024; kadv = 32; spb = cpb / kper;
        } else {                    // MN-major: rows = K, boxes tile the MN extent
            lbo = bstride; sbo = 1024; kadv = kper * 128; spb = 1u << 20;
        }
        if (swap) { uint32_t t = lbo; lbo = sbo; sbo = t; }
    };
    fill(c.a_mn, ar, ac, c.swap_a, p.a_boxes, p.a_box_rows, p.a_lbo, p.a_sbo, p.a_kadv, p.a_steps_per_box, p.a_box_stride);
    fill(c.b_mn, br, bc, c.swap_b, p.b_boxes, p.b_box_rows, p.b_lbo, p.b_sbo, p.b_kadv, p.b_steps_per_box, p.b_box_stride);
    p.a_layout = 2; p.b_layout = 2;
    if (c.atom32 && c.a_mn) { p.a_layout = 1; p.a_sbo = c.atom32; }
    if (c.atom32 && c.b_mn) { p.b_layout = 1; p.b_sbo = c.atom32; }
    if (c.atomk_a && !c.a_mn) { p.a_layout = 1; p.a_sbo = c.atomk_a; }
    if (c.atomk_b && !c.b_mn) { p.b_layout = 1; p.b_sbo = c.atomk_b; }
    p.a_tmem_cols_per_step = 8;    // bf16: 16 K-elements packed 2/column; tf32: 8 K-elements 1/column
    p.idesc = make_idesc(eb == 1 ? kFmtE4M3 : eb == 2 ? kFmtBF16 : kFmtTF32, M, N, c.a_mn, c.b_mn);

    const int smem_bytes = 131072 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    probe_kernel<<<1, 128, smem_bytes>>>(tmA, tmB, p);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());

    std::vector<float> D((size_t)M * N);
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0; int bad = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            float ref = 0;
            for (int k = 0; k < K; ++k) ref += A[(size_t)m * K + k] * B[(size_t)n * K + k];
            double e = fabs((double)ref - (double)D[(size_t)m * N + n]);
            if (!(e <= 1e-3)) ++bad;
            if (e > maxerr || e != e) maxerr = e;
        }
    printf("[probe %2d] %s : max_err=%.4g bad=%d/%d -> %s\n", id, c.name, maxerr, bad, M * N, bad == 0 ? "PASS" : "FAIL");
    return bad == 0 ? 0 : 3;
}
