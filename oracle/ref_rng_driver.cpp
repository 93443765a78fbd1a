// ORACLE — test infrastructure only.
// Driver that compiles the reference's OWN lib/random.cuh (the single reference
// source that builds without the OptiX SDK) and prints known-answer vectors for
// tea<N>, xorshift32 and randf.  The include below resolves through -I to
// /root/reference/RestOfLife/lib/random.cuh; no reference source is copied.
// Output (JSON on stdout) is committed as tests/golden/rng_kat.json by
// oracle/make_rng_kat.sh.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include "lib/random.cuh"

static uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main() {
    printf("{\n \"source\": \"RestOfLife/lib/random.cuh compiled with g++ (oracle/ref_rng_driver.cpp)\",\n");
    // tea<64>, tea<16>, tea<4> on a grid of inputs
    const uint32_t ins[][2] = {{0, 0}, {1, 0}, {0, 1}, {359999, 0}, {359999, 1023}, {123456789u, 987654321u},
                               {0xffffffffu, 0xffffffffu}, {600 * 300 + 300, 7}};
    printf(" \"tea\": [\n");
    for (size_t i = 0; i < sizeof(ins) / sizeof(ins[0]); i++) {
        printf("  {\"s0\": %u, \"s1\": %u, \"tea64\": %u, \"tea16\": %u, \"tea4\": %u}%s\n", ins[i][0], ins[i][1],
               tea<64>(ins[i][0], ins[i][1]), tea<16>(ins[i][0], ins[i][1]), tea<4>(ins[i][0], ins[i][1]),
               i + 1 < sizeof(ins) / sizeof(ins[0]) ? "," : "");
    }
    printf(" ],\n \"randf_streams\": [\n");
    const uint32_t seeds[] = {0x6314759u, 0x314759u, 1u, 0xC0FFEEu, 0xfffffffeu};
    for (size_t i = 0; i < sizeof(seeds) / sizeof(seeds[0]); i++) {
        uint32_t s = seeds[i], s2 = seeds[i];
        printf("  {\"seed\": %u, \"xorshift32\": [", seeds[i]);
        for (int k = 0; k < 16; k++) printf("%u%s", xorshift32(s), k < 15 ? ", " : "");
        printf("], \"randf_bits\": [");
        for (int k = 0; k < 16; k++) printf("%u%s", fbits(randf(s2)), k < 15 ? ", " : "");
        printf("]}%s\n", i + 1 < sizeof(seeds) / sizeof(seeds[0]) ? "," : "");
    }
    // the state whose quotient rounds to 1.0f: randf returns (float)0x3F7FFFFF (random.cuh:31-38)
    {
        // find s such that xorshift32(s) == 0xffffffff is not needed: test the branch via the output function
        // by scanning a few states whose xorshift output is >= 0xffffff80
        printf(" ],\n \"randf_one_branch\": [");
        int found = 0;
        for (uint32_t s0 = 1; s0 < 400000000u && found < 2; s0++) {
            uint32_t s = s0;
            uint32_t x = xorshift32(s);
            if (x >= 0xffffff80u) {
                uint32_t s2 = s0;
                printf("%s{\"seed\": %u, \"xorshift32\": %u, \"randf_bits\": %u}", found ? ", " : "", s0, x, fbits(randf(s2)));
                found++;
            }
        }
        printf("]\n}\n");
    }
    return 0;
}
