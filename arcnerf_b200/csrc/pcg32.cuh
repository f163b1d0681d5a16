// PCG32 (O'Neill's permuted congruential generator, XSH-RR 64/32) with logarithmic jump-ahead.
// The occupancy samplers draw their start jitter from it; the stream must match the generator the
// reference uses (arcnerf/ops/include/pcg32.h: multiplier 0x5851f42d4c957f2d, seed(initstate, initseq=1),
// next_float = mantissa trick on the top 23 bits, advance = Brown's arbitrary-stride algorithm).
#pragma once
#include <stdint.h>

struct Pcg32 {
    uint64_t state;
    uint64_t inc;

    static constexpr uint64_t kMult = 0x5851f42d4c957f2dULL;

    __host__ __device__ Pcg32() : state(0x853c49e6748fea9bULL), inc(0xda3e39cb94b95bdbULL) {}
    __host__ __device__ explicit Pcg32(uint64_t initstate, uint64_t initseq = 1u) {
        state = 0u;
        inc = (initseq << 1u) | 1u;
        next_u32();
        state += initstate;
        next_u32();
    }

    __host__ __device__ uint32_t next_u32() {
        const uint64_t old = state;
        state = old * kMult + inc;
        const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31u));
    }

    // Uniform in [0,1): 23 random mantissa bits under exponent 0, minus one.
    __host__ __device__ float next_f32() {
        union { uint32_t u; float f; } x;
        x.u = (next_u32() >> 9) | 0x3f800000u;
        return x.f - 1.0f;
    }

    // state <- state advanced by `delta` draws in O(log delta).
    __host__ __device__ void advance(uint64_t delta) {
        uint64_t cur_mult = kMult, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
        while (delta > 0) {
            if (delta & 1u) {
                acc_mult *= cur_mult;
                acc_plus = acc_plus * cur_mult + cur_plus;
            }
            cur_plus = (cur_mult + 1u) * cur_plus;
            cur_mult *= cur_mult;
            delta >>= 1;
        }
        state = acc_mult * state + acc_plus;
    }
};
