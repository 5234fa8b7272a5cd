// pcie_probe.cu -- pinned host <-> device copy rates on this box: H2D alone, D2H alone, both at once,
// for the chunk sizes the host pipeline (csrc/api.cu) uses.  The e2e bench leg is bounded by these.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(2); } } while (0)

int main()
{
    const size_t total = (size_t)1 << 30;
    void *h_in, *h_out, *d_in, *d_out;
    CK(cudaMallocHost(&h_in, total)); CK(cudaMallocHost(&h_out, total));
    CK(cudaMalloc(&d_in, total)); CK(cudaMalloc(&d_out, total));
    cudaStream_t s0, s1;
    CK(cudaStreamCreate(&s0)); CK(cudaStreamCreate(&s1));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const size_t chunks[] = {(size_t)8 << 20, (size_t)32 << 20, (size_t)128 << 20, total};
    for (size_t chunk : chunks) {
        for (int mode = 0; mode < 3; ++mode) {          // 0 = H2D, 1 = D2H, 2 = both
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0, s0));
                CK(cudaStreamWaitEvent(s1, e0, 0));
                for (size_t off = 0; off < total; off += chunk) {
                    if (mode != 1) CK(cudaMemcpyAsync((char*)d_in + off, (char*)h_in + off, chunk, cudaMemcpyHostToDevice, s0));
                    if (mode != 0) CK(cudaMemcpyAsync((char*)h_out + off, (char*)d_out + off, chunk, cudaMemcpyDeviceToHost, s1));
                }
                CK(cudaEventRecord(e1, s1));
                CK(cudaStreamWaitEvent(s0, e1, 0));
                CK(cudaEventRecord(e1, s0));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (ms < best) best = ms;
            }
            printf("chunk %4zu MB  %-5s  %.2f ms  %.1f GB/s per direction\n", chunk >> 20, mode == 0 ? "H2D" : mode == 1 ? "D2H" : "both", best,
                   total / best / 1e6);
        }
    }
    return 0;
}
