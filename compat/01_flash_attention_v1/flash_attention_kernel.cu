// Source-compat shim for src/01_flash_attention_v1/flash_attention_kernel.cu.
// The 01 driver defines the host wrapper flash_attention() ITSELF (main.cu:7-70) and launches
// `flash_attention_kernel<<<Tr, Br, smem>>>(Q,K,V,O,l,m,N,d,Br,Bc,scale)` from it, so the drop-in
// unit here is the __global__ symbol.  FA1's algorithm (HBM round trip of O/l/m per KV tile,
// flash_attention_kernel.cu:6-164) is out of scope: this stand-in keeps the launch
// contract -- one thread per query row, outputs normalised O plus l = sum exp(s - m) and m = rowmax --
// with a single-pass online softmax.  Library users call b200_fa1_fwd_f32 instead.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <cfloat>
__global__ void flash_attention_kernel(const float* Q, const float* K, const float* V, float* O, float* l, float* m,
                                       int N, int d, int Br, int Bc, float scale)
{
    (void)Bc;
    const int row = blockIdx.x * Br + threadIdx.x;
    if (row >= N || threadIdx.x >= Br) return;
    float mx = -FLT_MAX, sum = 0.0f;
    for (int c = 0; c < d; ++c) O[(size_t)row * d + c] = 0.0f;
    for (int j = 0; j < N; ++j) {
        float s = 0.0f;
        for (int c = 0; c < d; ++c) s += Q[(size_t)row * d + c] * K[(size_t)j * d + c];
        s *= scale;
        const float m_new = fmaxf(mx, s);
        const float alpha = expf(mx - m_new), p = expf(s - m_new);
        sum = sum * alpha + p;
        for (int c = 0; c < d; ++c) O[(size_t)row * d + c] = O[(size_t)row * d + c] * alpha + p * V[(size_t)j * d + c];
        mx = m_new;
    }
    for (int c = 0; c < d; ++c) O[(size_t)row * d + c] /= sum;
    l[row] = sum; m[row] = mx;
}
