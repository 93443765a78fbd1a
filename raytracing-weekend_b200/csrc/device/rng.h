// rng.h — the reference's random number generators.
//   tea<N>, xorshift32, randf : reference lib/random.cuh:7-38
//   lcg, rnd                  : OptiX SDK 8.0.0 cuda/random.h (not under
//                               /root/reference; published recurrence)
#pragma once
#include "vec.h"

namespace rtw {

template <unsigned int N>
RTW_HD uint32_t tea(uint32_t s0, uint32_t s1) {
    uint32_t t = 0;
#pragma unroll
    for (unsigned int n = 0; n < N; n++) {
        t += 0x9E3779B9;
        s0 += ((s1 << 4) + 0xa341316c) ^ (s1 + t) ^ ((s1 >> 5) + 0xc8013ea4);
        s1 += ((s0 << 4) + 0xad90777d) ^ (s0 + t) ^ ((s0 >> 5) + 0x7e95761e);
    }
    return s0;
}

RTW_HD uint32_t xorshift32(uint32_t& s) {
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    return s;
}

// lib/random.cuh:31-38: the `1.0f` branch returns an *integer* constant
// converted to float (1065353215.f); kept literally.
RTW_HD float randf(uint32_t& s) {
    float r = ((float)xorshift32(s)) / 4294967296.0f;
    if (r != 1.0f) return r;
    return (float)0x3F7FFFFF;
}

RTW_HD uint32_t lcg(uint32_t& prev) {
    prev = 1664525u * prev + 1013904223u;
    return prev & 0x00FFFFFF;
}

RTW_HD float rnd(uint32_t& prev) { return ((float)lcg(prev) / (float)0x01000000); }

// Q23 (normalised): free-flight state of a constant medium hit by a ray whose
// path seed at ray start is `ray_seed`; the path seed itself is never mutated
// inside traversal.
RTW_HD uint32_t medium_state(uint32_t ray_seed, uint32_t prim) { return tea<4>(ray_seed, prim); }

}  // namespace rtw
