// kbench.cu -- native micro-harness for the C-ABI (no Python/torch start-up): device-resident
// N(0,1) inputs, CUDA-event timing of b200_fa2_fwd / b200_fa2_bwd, optional in-kernel trace dump
// (library built with EXTRA=-DB200_TRACE).  Not the contract bench (that is bench.py).
//   kbench <fwd|bwd|both> B H N d <bf16|f32> [iters] [--trace] [--causal] [--check] [--nkv M]
//   --nkv M: M key/value rows against N query rows (forward only; cross lengths as in the ring steps)
//   --check: before timing, compare the forward (O, L) of the selected kernel (e.g. B200_FWD_VARIANT) -- and with
//            bwd / both the gradients (e.g. B200_BWD_VARIANT) -- with the library's generic CUDA-core kernels on the
//            same inputs (b200_debug_force_generic); exit 5 on mismatch
//   env B200ATTN_LIB = path of the shared object (default: cuda_flashattention_b200/libb200attn.so)
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <dlfcn.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(2); } } while (0)

__device__ __forceinline__ uint32_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return (uint32_t)x;
}
template <typename T>
__global__ void fill_normal(T* p, size_t n, uint64_t seed, float sigma) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const float u1 = (mix(seed * 0x9E3779B97F4A7C15ull + 2 * i) + 1.0f) * (1.0f / 4294967808.0f);
        const float u2 = mix(seed * 0x9E3779B97F4A7C15ull + 2 * i + 1) * (1.0f / 4294967296.0f);
        const float v = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2) * sigma;
        if constexpr (sizeof(T) == 2) p[i] = __float2bfloat16(v); else p[i] = v;
    }
}

typedef int (*fwd_t)(const void*, const void*, const void*, void*, float*, int, int, int, int, int, float, int, cudaStream_t);
typedef int (*bwd_t)(const void*, const void*, const void*, const void*, const float*, const void*, void*, void*, void*,
                     int, int, int, int, int, float, int, void*, size_t, cudaStream_t);
typedef int (*fwdx_t)(const void*, const void*, const void*, void*, float*, int, int, int, int, int, float, int, int, cudaStream_t);
typedef int (*bwdx_t)(const void*, const void*, const void*, const void*, const float*, const void*, void*, void*, void*,
                      int, int, int, int, int, float, int, int, void*, size_t, cudaStream_t);
typedef size_t (*ws_t)(int, int, int, int, int, int);
typedef int (*trace_t)(long long*, int);
typedef int (*force_t)(int);

int main(int argc, char** argv)
{
    if (argc < 7) { fprintf(stderr, "usage: kbench <fwd|bwd|both> B H N d <bf16|f32> [iters] [--trace]\n"); return 1; }
    const char* mode = argv[1];
    const int B = atoi(argv[2]), H = atoi(argv[3]), N = atoi(argv[4]), d = atoi(argv[5]);
    const bool f32 = strcmp(argv[6], "f32") == 0;
    int iters = 10, M = N; bool trace = false, causal = false, check = false;
    for (int i = 7; i < argc; ++i) {
        if (!strcmp(argv[i], "--trace")) trace = true; else if (!strcmp(argv[i], "--causal")) causal = true;
        else if (!strcmp(argv[i], "--check")) check = true;
        else if (!strcmp(argv[i], "--nkv") && i + 1 < argc) M = atoi(argv[++i]);
        else iters = atoi(argv[i]);
    }
    if (M != N && (strcmp(mode, "fwd") != 0 || causal)) { fprintf(stderr, "--nkv: forward, non-causal only\n"); return 1; }
    const char* path = getenv("B200ATTN_LIB");
    if (!path) path = "cuda_flashattention_b200/libb200attn.so";
    void* h = dlopen(path, RTLD_NOW);
    if (!h) { fprintf(stderr, "dlopen %s: %s\n", path, dlerror()); return 2; }
    fwd_t fwd = (fwd_t)dlsym(h, "b200_fa2_fwd");
    bwd_t bwd = (bwd_t)dlsym(h, "b200_fa2_bwd");
    fwdx_t fwdx = (fwdx_t)dlsym(h, "b200_fa2_fwd_ex");
    bwdx_t bwdx = (bwdx_t)dlsym(h, "b200_fa2_bwd_ex");
    ws_t wsb = (ws_t)dlsym(h, "b200_fa2_bwd_workspace_bytes");
    trace_t tr = (trace_t)dlsym(h, strcmp(mode, "bwd") == 0 ? "b200_debug_trace_read_bwd" : "b200_debug_trace_read");
    if (!fwd || !bwd || !wsb || !fwdx || !bwdx) { fprintf(stderr, "missing symbols\n"); return 2; }

    const size_t n = (size_t)B * H * N * d, nk = (size_t)B * H * M * d, eb = f32 ? 4 : 2;
    void *Q, *K, *V, *O, *dO, *dQ, *dK, *dV, *ws; float* L;
    CK(cudaMalloc(&Q, n * eb)); CK(cudaMalloc(&K, nk * eb)); CK(cudaMalloc(&V, nk * eb)); CK(cudaMalloc(&O, n * eb));
    CK(cudaMalloc(&dO, n * eb)); CK(cudaMalloc(&dQ, n * eb)); CK(cudaMalloc(&dK, n * eb)); CK(cudaMalloc(&dV, n * eb));
    CK(cudaMalloc(&L, (size_t)B * H * N * 4));
    const int dt = f32 ? 1 : 0;
    const size_t wsn = wsb(B, H, N, N, d, dt);
    CK(cudaMalloc(&ws, wsn ? wsn : 16));
    void* bufs[4] = {Q, K, V, dO};
    for (int i = 0; i < 4; ++i) {
        const size_t ni = (i == 1 || i == 2) ? nk : n;
        if (f32) fill_normal<float><<<148 * 8, 256>>>((float*)bufs[i], ni, 1234 + i, 1.0f);
        else fill_normal<__nv_bfloat16><<<148 * 8, 256>>>((__nv_bfloat16*)bufs[i], ni, 1234 + i, 1.0f);
    }
    CK(cudaDeviceSynchronize());
    const bool do_f = strcmp(mode, "bwd") != 0, do_b = strcmp(mode, "fwd") != 0;
    auto run_f = [&]() { int s = causal ? fwdx(Q, K, V, O, L, B, H, N, N, d, 0.0f, dt, 1, 0) : fwd(Q, K, V, O, L, B, H, N, M, d, 0.0f, dt, 0); if (s) { fprintf(stderr, "fwd status %d\n", s); exit(3); } };
    auto run_b = [&]() { int s = causal ? bwdx(Q, K, V, O, L, dO, dQ, dK, dV, B, H, N, N, d, 0.0f, dt, 1, ws, wsn, 0)
                                          : bwd(Q, K, V, O, L, dO, dQ, dK, dV, B, H, N, N, d, 0.0f, dt, ws, wsn, 0); if (s) { fprintf(stderr, "bwd status %d\n", s); exit(3); } };
    run_f();
    if (check) {
        // forward parity of the selected tensor-core kernel against the generic CUDA-core kernels (fp32 math)
        force_t force = (force_t)dlsym(h, "b200_debug_force_generic");
        if (!force) { fprintf(stderr, "missing b200_debug_force_generic\n"); return 2; }
        CK(cudaDeviceSynchronize());
        std::vector<uint8_t> o1(n * eb), o2(n * eb);
        std::vector<float> l1((size_t)B * H * N), l2((size_t)B * H * N);
        CK(cudaMemcpy(o1.data(), O, n * eb, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(l1.data(), L, l1.size() * 4, cudaMemcpyDeviceToHost));
        force(1); run_f(); force(0);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(o2.data(), O, n * eb, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(l2.data(), L, l2.size() * 4, cudaMemcpyDeviceToHost));
        auto val = [&](const std::vector<uint8_t>& v, size_t i) -> double {
            if (f32) return ((const float*)v.data())[i];
            uint32_t u = (uint32_t)((const uint16_t*)v.data())[i] << 16; float f; memcpy(&f, &u, 4); return f;
        };
        double max_o = 0, max_l = 0, num = 0, den = 0; size_t bad = 0;
        for (size_t i = 0; i < n; ++i) {
            const double a = val(o1, i), b = val(o2, i), e = fabs(a - b);
            if (!(e <= 1e30)) ++bad;
            if (e > max_o) max_o = e;
            num += e * e; den += b * b;
        }
        for (size_t i = 0; i < l1.size(); ++i) { const double e = fabs((double)l1[i] - l2[i]); if (!(e <= 1e30)) ++bad; if (e > max_l) max_l = e; }
        const double rel = sqrt(num / (den > 0 ? den : 1));
        const double tol = f32 ? 2e-3 : 2e-2;
        const bool ok = bad == 0 && max_o <= tol && max_l <= tol && rel <= 1e-2;
        printf("check vs generic kernels: max|dO| %.3e  rel-L2 %.3e  max|dL| %.3e  non-finite %zu  -> %s\n", max_o, rel, max_l, bad,
               ok ? "PASS" : "FAIL");
        if (!ok) return 5;
        run_f();
        if (do_b) {
            // backward parity on the same (O, L): selected kernel (e.g. B200_BWD_VARIANT) against the generic kernels
            run_b();
            CK(cudaDeviceSynchronize());
            void* g[3] = {dQ, dK, dV};
            std::vector<uint8_t> g1[3], g2[3];
            for (int t = 0; t < 3; ++t) { g1[t].resize(n * eb); g2[t].resize(n * eb); CK(cudaMemcpy(g1[t].data(), g[t], n * eb, cudaMemcpyDeviceToHost)); }
            force(1); run_b(); force(0);
            CK(cudaDeviceSynchronize());
            const char* names[3] = {"dQ", "dK", "dV"};
            bool okb = true;
            for (int t = 0; t < 3; ++t) {
                CK(cudaMemcpy(g2[t].data(), g[t], n * eb, cudaMemcpyDeviceToHost));
                double mx = 0, num2 = 0, den2 = 0, ref_max = 1.0; size_t bad2 = 0;   // (N(0,1) dO: gradients are not unit-scale)
                for (size_t i = 0; i < n; ++i) {
                    const double a = val(g1[t], i), b = val(g2[t], i), e = fabs(a - b);
                    if (!(e <= 1e30)) ++bad2;
                    if (e > mx) mx = e;
                    if (fabs(b) > ref_max) ref_max = fabs(b);
                    num2 += e * e; den2 += b * b;
                }
                const double rel2 = sqrt(num2 / (den2 > 0 ? den2 : 1));
                const bool o = bad2 == 0 && mx <= tol * ref_max && rel2 <= 1e-2;
                printf("check %s vs generic kernels: max|d| %.3e  rel-L2 %.3e  non-finite %zu  -> %s\n", names[t], mx, rel2, bad2, o ? "PASS" : "FAIL");
                okb = okb && o;
            }
            if (!okb) return 5;
        }
    }
    for (int i = 0; i < 3; ++i) { if (do_f) run_f(); if (do_b) run_b(); }
    CK(cudaDeviceSynchronize());
    std::vector<float> tf, tb;
    cudaEvent_t e0, e1, e2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    for (int i = 0; i < iters; ++i) {
        CK(cudaEventRecord(e0));
        if (do_f) run_f();
        CK(cudaEventRecord(e1));
        if (do_b) run_b();
        CK(cudaEventRecord(e2));
        CK(cudaEventSynchronize(e2));
        float a, b;
        CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b, e1, e2));
        tf.push_back(a); tb.push_back(b);
    }
    auto med = [](std::vector<float> v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
    auto mn = [](std::vector<float> v) { return *std::min_element(v.begin(), v.end()); };
    // causal: the visible half of the score matrix only (N (N + 1) / 2 pairs)
    const double pairs = causal ? 0.5 * N * ((double)N + 1) : (double)N * M;
    const double ff = 4.0 * B * H * pairs * d, fb = 10.0 * B * H * pairs * d;
    printf("B=%d H=%d N=%d", B, H, N);
    if (M != N) printf(" Nkv=%d", M);
    printf(" d=%d %s%s:", d, f32 ? "f32" : "bf16", causal ? " causal" : "");
    if (do_f) printf("  fwd med %.3f ms (%.1f TF/s) best %.3f ms (%.1f)", med(tf), ff / med(tf) / 1e9, mn(tf), ff / mn(tf) / 1e9);
    if (do_b) printf("  bwd med %.3f ms (%.1f TF/s) best %.3f ms (%.1f)", med(tb), fb / med(tb) / 1e9, mn(tb), fb / mn(tb) / 1e9);
    printf("\n");
    if (trace) {
        if (!tr) { fprintf(stderr, "library was not built with -DB200_TRACE\n"); return 4; }
        long long t[3 * 8 * 8];
        tr(t, 3 * 8 * 8);
        const long long t0 = t[0];
        const char* names_f[3] = {"softmax0", "softmax1", "mma"}, *names_b[3] = {"compute0", "drain", "mma"};
        const char** names = strcmp(mode, "bwd") == 0 ? names_b : names_f;
        for (int r = 0; r < 3; ++r)
            for (int j = 0; j < 8; ++j) {
                printf("%-8s j=%2d:", names[r], 16 + j);
                for (int e = 0; e < 8; ++e) printf(" %7lld", t[(r * 8 + j) * 8 + e] ? t[(r * 8 + j) * 8 + e] - t0 : -1);
                printf("\n");
            }
    }
    return 0;
}
