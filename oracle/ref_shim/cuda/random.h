// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// OptiX SDK 8.0.0 SDK/cuda/random.h is not under /root/reference; this
// restates its published definitions: tea<N> (identical to the reference's in-tree copy,
// lib/random.cuh:7-19), lcg (Numerical Recipes constants, low 24 bits), rnd.
#pragma once
#include "../vector_types.h"

template <unsigned int N>
static inline unsigned int tea(unsigned int val0, unsigned int val1) {
    unsigned int v0 = val0, v1 = val1, s0 = 0;
    for (unsigned int n = 0; n < N; n++) {
        s0 += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v0;
}
static inline unsigned int lcg(unsigned int& prev) {
    const unsigned int LCG_A = 1664525u;
    const unsigned int LCG_C = 1013904223u;
    prev = (LCG_A * prev + LCG_C);
    return prev & 0x00FFFFFF;
}
static inline float rnd(unsigned int& prev) { return ((float)lcg(prev) / (float)0x01000000); }
