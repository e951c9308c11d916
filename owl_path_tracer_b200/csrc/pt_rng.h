// pt_rng.h — per-pixel random stream of the reference (path_tracer/src/random.hpp:34-85).
// Seed: 4 rounds of TEA over (x, y) -> state.  Draw: 32-bit LCG, value = state * 2^-32
// rounded to binary32 (so 1.0f is reachable).
#pragma once
#include "pt_defs.h"

namespace pt {

PT_HD uint32_t rng_seed(uint32_t u, uint32_t v) {
    uint32_t s = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        s += 0x9e3779b9u;
        u += ((v << 4) + 0xa341316cu) ^ (v + s) ^ ((v >> 5) + 0xc8013ea4u);
        v += ((u << 4) + 0xad90777du) ^ (u + s) ^ ((u >> 5) + 0x7e95761eu);
    }
    return u;
}

PT_HD float rng_next(uint32_t& state) {
    state = 16807u * state + 1013904223u;
    // ldexpf((float)state, -32): the int->float conversion rounds to nearest even, the
    // scaling by 2^-32 is exact.
    return (float)state * 2.3283064365386963e-10f;
}

}  // namespace pt
