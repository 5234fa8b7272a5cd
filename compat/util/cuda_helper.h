// Host-side subset of the reference's src/util/cuda_helper.h (macros and helpers its drivers use).
// The device helpers there (warp reductions, init_array) belong to the replaced kernels and are not needed.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>
#include <math.h>
#include <cfloat>
#include <assert.h>
#define FULL_MASK 0xffffffff
#ifndef WARP_SIZE
#define WARP_SIZE 32
#endif
static inline int div_up(int a, int b) { return (a + b - 1) / b; }
static inline void swap_float_ptrs(float** a, float** b) { float* t = *a; *a = *b; *b = t; }
#define CHECK_CUDA(call)                                                                        \
    do {                                                                                        \
        cudaError_t err__ = (call);                                                             \
        if (err__ != cudaSuccess) {                                                             \
            fprintf(stderr, "CUDA error at %s:%d - %s\n", __FILE__, __LINE__, cudaGetErrorString(err__)); \
            assert(0);                                                                          \
        }                                                                                       \
    } while (0)
