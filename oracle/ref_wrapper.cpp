// ref_wrapper.cpp -- C-linkage shim around the UNMODIFIED reference CPU code.
//
// TEST INFRASTRUCTURE ONLY (see naive_oracle.c header).  This file contains no attention math:
// it #includes the reference sources where they lie (REF_ROOT, default /root/reference) and
// re-exports their functions with C linkage so that python (ctypes) can call them.  The built
// library goes to oracle/_ref/ (git-ignored, travels to the GPU box).  No reference source is
// copied into this repository.
//
// Also hosts the reference's *test fixtures* (glibc rand() streams) so that golden vectors can
// be regenerated here: each fixture cites the reference test it mirrors.
#include <cmath>
#include <cfloat>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "src/util/naive_attention.h"        // naive_forward_pass, naive_attention_forward, naive_attention_backward
#define main ref00_main_unused
#include "src/00_naive_attention/main.cpp"   // naive_attention (row-wise variant)
#undef main

extern "C" {

void ref_naive_forward_pass(const float* Q, const float* K, const float* V, float* O, float* L, int N, int d, float scale)
{ naive_forward_pass(Q, K, V, O, L, N, d, scale); }

void ref_naive_attention_forward(const float* Q, const float* K, const float* V, float* O, int N, int d, float scale)
{ naive_attention_forward(Q, K, V, O, N, d, scale); }

void ref_naive_attention_backward(const float* Q, const float* K, const float* V, const float* O, const float* L,
                                  const float* dO, float* dQ, float* dK, float* dV, int N, int d, float scale)
{ naive_attention_backward(Q, K, V, O, L, dO, dQ, dK, dV, N, d, scale); }

void ref_naive_attention(const float* Q, const float* K, const float* V, float* O, int N, int d)
{ naive_attention(Q, K, V, O, N, d); }

// Fixture G3: src/02_flash_attention_v2_forward/main.cu:26-33  (srand(42); Q,K,V interleaved per index)
void ref_fixture_fwd_random(float* Q, float* K, float* V, int n_elems)
{
    srand(42);
    for (int i = 0; i < n_elems; i++) {
        Q[i] = (rand() % 1000) / 1000.0f - 0.5f;
        K[i] = (rand() % 1000) / 1000.0f - 0.5f;
        V[i] = (rand() % 1000) / 1000.0f - 0.5f;
    }
}

// Fixture G4: src/02_flash_attention_v2_backward/main.cu:221-227
void ref_fixture_bwd_random(float* Q, float* K, float* V, float* dO, int n_elems)
{
    srand(42);
    for (int i = 0; i < n_elems; i++) {
        Q[i] = ((rand() % 2000) / 1000.0f - 1.0f) * 0.5f;
        K[i] = ((rand() % 2000) / 1000.0f - 1.0f) * 0.5f;
        V[i] = ((rand() % 2000) / 1000.0f - 1.0f) * 0.5f;
        dO[i] = ((rand() % 2000) / 1000.0f - 1.0f) * 0.2f;
    }
}

// Fixture FA1 test 7: src/01_flash_attention_v1/main.cu:302-307
void ref_fixture_fa1_random(float* Q, float* K, float* V, int n_elems)
{
    srand(42);
    for (int i = 0; i < n_elems; i++) {
        Q[i] = (float)rand() / RAND_MAX;
        K[i] = (float)rand() / RAND_MAX;
        V[i] = (float)rand() / RAND_MAX * 100;
    }
}

}  // extern "C"
