// Host-side subset of the reference's src/util/attention_helper.h: the fixture and comparison
// helpers its ring / 1-GPU drivers call (create_simple_test_data :151-173, compare_outputs :174-208,
// print_matrix :136-149).  Re-stated, not copied; the device functions there are replaced kernels.
#pragma once
#include <vector>
#include <cmath>
#include <cstdio>
#include <algorithm>
#include "cuda_helper.h"

static inline void print_matrix(const char* name, const float* data, int rows, int cols, int max_rows = 4, int max_cols = 8) {
    printf("%s (%dx%d):\n", name, rows, cols);
    for (int i = 0; i < std::min(rows, max_rows); ++i) {
        for (int j = 0; j < std::min(cols, max_cols); ++j) printf("%8.4f ", data[i * cols + j]);
        printf(cols > max_cols ? "...\n" : "\n");
    }
    if (rows > max_rows) printf("...\n");
}

static inline void create_simple_test_data(std::vector<float>& Q, std::vector<float>& K, std::vector<float>& V, int seq_len, int head_dim) {
    Q.assign((size_t)seq_len * head_dim, 0.0f);
    V.resize((size_t)seq_len * head_dim);
    for (int i = 0; i < seq_len && i < head_dim; ++i) Q[(size_t)i * head_dim + i] = 1.0f;
    K = Q;
    for (int i = 0; i < seq_len; ++i)
        for (int j = 0; j < head_dim; ++j) V[(size_t)i * head_dim + j] = i * 4.0f + j + 1.0f;
}

static inline bool compare_outputs(const float* ref, const float* test, int size, float rtol = 1e-3, float atol = 1.0) {
    int bad = 0; float worst = 0.0f; int worst_i = 0;
    for (int i = 0; i < size; ++i) {
        float diff = std::fabs(ref[i] - test[i]);
        float rel = diff / (std::fabs(ref[i]) + 1e-8f);
        if (rel > rtol && diff > atol) {
            if (bad < 10) printf("Diff at index %d: ref=%.4f, test=%.4f, diff=%.4f (rel=%.6f)\n", i, ref[i], test[i], diff, rel);
            ++bad;
            if (diff > worst) { worst = diff; worst_i = i; }
        }
    }
    if (bad) printf("Total significant differences: %d out of %d (max %.4f at %d)\n", bad, size, worst, worst_i);
    else printf("All outputs match within tolerance (rtol=%.1e, atol=%.1f)\n", rtol, atol);
    return bad == 0;
}
